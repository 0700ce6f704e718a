// Exact-fp32 SIMT GEMM with generic operand strides, used by nn.Linear forward / input gradient / weight gradient
// and by the LSTM projections (fc1/fc2/heads, avse_model.py:550,553,618-627).
//   C[i, j] = act( sum_l A(i, l) * B(j, l) + bias[j] )        A(i,l) = A[i*a_si + l*a_sl], B(j,l) = B[j*b_sj + l*b_sl]
// Tiles BM x BN x 16 with BM*BN = 4096, 256 threads, 4x4 outputs per thread.  gridDim.z > 1 splits the reduction
// (l) range and accumulates with atomics into a pre-zeroed C.
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

struct GemmArgs {
    const float* A;
    long long a_si, a_sl;
    const float* B;
    long long b_sj, b_sl;
    float* C;
    long long c_si, c_sj;
    int M, N, K;
    const float* bias;   // per j (may be null)
    float* colsum;       // optional: colsum[i] (+)= sum_l A(i,l), computed by the blockIdx.x == 0 column (bias gradient)
    int act;
    float slope;
    int accumulate;      // C += result
    int k_per_split;
};

constexpr int BK = 16;

__device__ __forceinline__ float act_apply_g(float v, int act, float slope) {
    switch (act) {
        case MAAVSS_ACT_TANH: return tanhf(v);
        case MAAVSS_ACT_LEAKY: return v > 0.f ? v : v * slope;
        case MAAVSS_ACT_SIGMOID: return 1.f / (1.f + expf(-v));
        default: return v;
    }
}

// A (ROWS x BK) operand tile travels global -> registers (fetch_tile, issued one tile ahead so the loads overlap the
// FMAs of the current tile) -> smem as [BK][ROWS + 4] (stash_tile); the thread mapping walks the unit-stride global
// dimension with consecutive threads.
template <int ROWS>
__device__ __forceinline__ void fetch_tile(float (&reg)[ROWS * BK / 256], const float* __restrict__ P, long long s_r, long long s_l,
                                           int r0, int l0, int R, int l_end) {
    constexpr int PER = ROWS * BK / 256;
    if (s_l == 1) {
#pragma unroll
        for (int q = 0; q < PER; ++q) {
            const int e = threadIdx.x + q * 256;
            const int l = e % BK, r = e / BK;
            const int gr = r0 + r, gl = l0 + l;
            reg[q] = (gr < R && gl < l_end) ? __ldg(P + (long long)gr * s_r + gl) : 0.f;
        }
    } else {
#pragma unroll
        for (int q = 0; q < PER; ++q) {
            const int e = threadIdx.x + q * 256;
            const int r = e % ROWS, l = e / ROWS;
            const int gr = r0 + r, gl = l0 + l;
            reg[q] = (gr < R && gl < l_end) ? __ldg(P + (long long)gr * s_r + (long long)gl * s_l) : 0.f;
        }
    }
}

template <int ROWS>
__device__ __forceinline__ void stash_tile(float (*dst)[ROWS + 4], const float (&reg)[ROWS * BK / 256], long long s_l) {
    constexpr int PER = ROWS * BK / 256;
#pragma unroll
    for (int q = 0; q < PER; ++q) {
        const int e = threadIdx.x + q * 256;
        if (s_l == 1) dst[e % BK][e / BK] = reg[q];
        else dst[e / ROWS][e % ROWS] = reg[q];
    }
}

template <int BM>
__global__ void __launch_bounds__(256) sgemm_kernel(GemmArgs a) {
    constexpr int BN = 4096 / BM;
    constexpr int TI = BM / 4;           // threads along i
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN + 4];
    const int i0 = blockIdx.y * BM, j0 = blockIdx.x * BN;
    const int ti = threadIdx.x % TI, tj = threadIdx.x / TI;
    const int l_begin = blockIdx.z * a.k_per_split;
    const int l_end = min(a.K, l_begin + a.k_per_split);
    float acc[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) acc[u][v] = 0.f;
    float rs[4] = {0.f, 0.f, 0.f, 0.f};
    float ra[BM * BK / 256], rb[BN * BK / 256];
    fetch_tile<BM>(ra, a.A, a.a_si, a.a_sl, i0, l_begin, a.M, l_end);
    fetch_tile<BN>(rb, a.B, a.b_sj, a.b_sl, j0, l_begin, a.N, l_end);
    for (int l0 = l_begin; l0 < l_end; l0 += BK) {
        stash_tile<BM>(As, ra, a.a_sl);
        stash_tile<BN>(Bs, rb, a.b_sl);
        __syncthreads();
        if (l0 + BK < l_end) {
            fetch_tile<BM>(ra, a.A, a.a_si, a.a_sl, i0, l0 + BK, a.M, l_end);
            fetch_tile<BN>(rb, a.B, a.b_sj, a.b_sl, j0, l0 + BK, a.N, l_end);
        }
#pragma unroll
        for (int l = 0; l < BK; ++l) {
            const float4 av = *reinterpret_cast<const float4*>(&As[l][ti * 4]);
            const float4 bv = *reinterpret_cast<const float4*>(&Bs[l][tj * 4]);
            const float ar[4] = {av.x, av.y, av.z, av.w}, br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                rs[u] += ar[u];
#pragma unroll
                for (int v = 0; v < 4; ++v) acc[u][v] = fmaf(ar[u], br[v], acc[u][v]);
            }
        }
        __syncthreads();
    }
    const bool split = gridDim.z > 1;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int i = i0 + ti * 4 + u;
        if (i >= a.M) continue;
        if (a.colsum && blockIdx.x == 0 && tj == 0) {
            if (split) atomicAdd(a.colsum + i, rs[u]);
            else a.colsum[i] = a.accumulate ? a.colsum[i] + rs[u] : rs[u];
        }
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            const int j = j0 + tj * 4 + v;
            if (j >= a.N) continue;
            float* cp = a.C + (long long)i * a.c_si + (long long)j * a.c_sj;
            if (split) {
                atomicAdd(cp, acc[u][v]);
            } else {
                float r = acc[u][v] + (a.bias ? a.bias[j] : 0.f);
                r = act_apply_g(r, a.act, a.slope);
                *cp = a.accumulate ? *cp + r : r;
            }
        }
    }
}

int launch_gemm(GemmArgs a, int splits, cudaStream_t st) {
    if (a.M == 0 || a.N == 0) return MAAVSS_OK;
    if (splits < 1) splits = 1;
    a.k_per_split = ((a.K + splits - 1) / splits + BK - 1) / BK * BK;
    if (a.k_per_split < BK) a.k_per_split = BK;
    splits = (a.K + a.k_per_split - 1) / a.k_per_split;
    if (splits < 1) splits = 1;
    if (splits > 1 && (a.bias || a.act != MAAVSS_ACT_NONE)) {
        maavss_set_error("gemm: split reduction cannot fuse bias/activation");
        return MAAVSS_ERR_INVALID;
    }
    if (a.M <= 16) {
        dim3 grid(ceil_div(a.N, 256), ceil_div(a.M, 16), splits);
        sgemm_kernel<16><<<grid, 256, 0, st>>>(a);
    } else {
        dim3 grid(ceil_div(a.N, 64), ceil_div(a.M, 64), splits);
        sgemm_kernel<64><<<grid, 256, 0, st>>>(a);
    }
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

}  // namespace

int maavss_skinny_fwd(const float*, const float*, const float*, float*, int, int, int, int, float, cudaStream_t);
int maavss_skinny_dgrad(const float*, const float*, float*, int, int, int, cudaStream_t);
int maavss_skinny_wgrad(const float*, const float*, float*, float*, int, int, int, int, cudaStream_t);

// Mid-size products (LSTM input projections: 512 x 1024 x 512) have ~128 output tiles, under one CTA per SM: split the
// reduction so that >= 2 waves of CTAs are resident and accumulate with atomics into a zeroed output.
static int auto_splits(int M, int N, int K) {
    const int tiles = (M <= 16) ? ceil_div(N, 256) * ceil_div(M, 16) : ceil_div(N, 64) * ceil_div(M, 64);
    if (tiles >= 296) return 1;
    int splits = (296 + tiles - 1) / tiles;
    const int max_splits = (K + 63) / 64;
    if (splits > max_splits) splits = max_splits;
    return splits < 1 ? 1 : splits;
}

extern "C" int maavss_linear_fwd(const float* x, const float* w, const float* bias, float* y, int M, int N, int K, int act,
                                 float slope, void* stream) {
    if (M == 0 || N == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && w && y && K > 0, "linear_fwd: bad argument");
    {
        const int rc = maavss_skinny_fwd(x, w, bias, y, M, N, K, act, slope, (cudaStream_t)stream);
        if (rc != 0) return rc < 0 ? rc : MAAVSS_OK;
    }
    GemmArgs a{};
    a.A = x; a.a_si = K; a.a_sl = 1;
    a.B = w; a.b_sj = K; a.b_sl = 1;
    a.C = y; a.c_si = N; a.c_sj = 1;
    a.M = M; a.N = N; a.K = K;
    a.bias = bias; a.act = act; a.slope = slope;
    int splits = 1;
    if (!bias && act == MAAVSS_ACT_NONE) {
        splits = auto_splits(M, N, K);
        if (splits > 1) MAAVSS_CUDA(cudaMemsetAsync(y, 0, sizeof(float) * (size_t)M * N, (cudaStream_t)stream));
    }
    return launch_gemm(a, splits, (cudaStream_t)stream);
}

extern "C" int maavss_linear_dgrad(const float* dz, const float* w, float* dx, int M, int N, int K, void* stream) {
    if (M == 0 || K == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dz && w && dx && N > 0, "linear_dgrad: bad argument");
    {
        const int rc = maavss_skinny_dgrad(dz, w, dx, M, N, K, (cudaStream_t)stream);
        if (rc != 0) return rc < 0 ? rc : MAAVSS_OK;
    }
    // dx[m, k] = sum_n dz[m, n] * w[n, k]:  i = m, j = k, l = n
    GemmArgs a{};
    a.A = dz; a.a_si = N; a.a_sl = 1;
    a.B = w; a.b_sj = 1; a.b_sl = K;
    a.C = dx; a.c_si = K; a.c_sj = 1;
    a.M = M; a.N = K; a.K = N;
    cudaStream_t st = (cudaStream_t)stream;
    // few output tiles when M is skinny: split the reduction so that >= ~2 waves of CTAs stream w
    const int splits = auto_splits(M, K, N);
    if (splits > 1) MAAVSS_CUDA(cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)M * K, st));
    return launch_gemm(a, splits, st);
}

extern "C" int maavss_linear_wgrad(const float* dz, const float* x, float* dw, float* dbias, int M, int N, int K, int accumulate,
                                   void* stream) {
    if (N == 0 || K == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dz && x && dw, "linear_wgrad: bad argument");
    {
        const int rc = maavss_skinny_wgrad(dz, x, dw, dbias, M, N, K, accumulate, (cudaStream_t)stream);
        if (rc != 0) return rc < 0 ? rc : MAAVSS_OK;
    }
    // dw[n, k] = sum_m dz[m, n] * x[m, k]:  i = n, j = k, l = m ; dbias[n] = sum_m dz[m, n] = row sums of A
    GemmArgs a{};
    a.A = dz; a.a_si = 1; a.a_sl = N;
    a.B = x; a.b_sj = 1; a.b_sl = K;
    a.C = dw; a.c_si = K; a.c_sj = 1;
    a.M = N; a.N = K; a.K = M;
    a.colsum = dbias;
    a.accumulate = accumulate;
    const int splits = auto_splits(N, K, M);
    if (splits > 1 && !accumulate) {
        MAAVSS_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)N * K, (cudaStream_t)stream));
        if (dbias) MAAVSS_CUDA(cudaMemsetAsync(dbias, 0, sizeof(float) * (size_t)N, (cudaStream_t)stream));
    }
    return launch_gemm(a, splits, (cudaStream_t)stream);
}
