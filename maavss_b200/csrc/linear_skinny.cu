// Weight-streaming kernels for the skinny nn.Linear layers of the fusion models (M = clip batch per GPU <= 16 rows,
// avse_model.py:550,553,618-627; avse_model_final.py:141,144,203-213).  With M this small the layer is a pure
// HBM stream over the (N x K) weight matrix: each weight is read (fwd, dgrad) or read+written (wgrad accumulate)
// exactly once, vectorised 16 B per lane and coalesced along K, the M activations ride in shared memory /
// registers.  Split reductions are combined deterministically through distributed shared memory of a thread-block
// cluster (fwd) so bias + activation stay fused; dgrad uses red.global.add on a pre-zeroed dx.
#include <cooperative_groups.h>

#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace cg = cooperative_groups;

namespace skinny {

constexpr int MT = 16;          // activation rows per CTA
constexpr int FWD_ROWS = 32;    // weight rows per CTA (8 warps x 4)
constexpr int FWD_KC = 512;     // K chunk staged in shared memory

__device__ __forceinline__ float act_apply(float v, int act, float slope) {
    switch (act) {
        case MAAVSS_ACT_TANH: return tanhf(v);
        case MAAVSS_ACT_LEAKY: return v > 0.f ? v : v * slope;
        case MAAVSS_ACT_SIGMOID: return 1.f / (1.f + expf(-v));
        default: return v;
    }
}

__device__ __forceinline__ float4 ldg_stream(const float* p) {
    // weights are touched once per pass: keep them out of L1
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// Butterfly reduction of 64 per-lane partials over the 32 lanes of a warp in 62 shuffles: after the 5 rounds lane l
// owns the totals of flat indices fwd_owner_base(l) and fwd_owner_base(l) + 1.
__device__ __forceinline__ void warp_reduce64(float (&v)[64], int lane) {
#pragma unroll
    for (int round = 0; round < 5; ++round) {
        const int half = 32 >> round;              // values kept after this round
        const int mask = 16 >> round;
        const bool up = (lane & mask) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const float keep = up ? v[i + half] : v[i];
            const float send = up ? v[i] : v[i + half];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, mask);
        }
    }
}
__device__ __forceinline__ int owner_base(int lane) {
    return ((lane >> 4) & 1) * 32 + ((lane >> 3) & 1) * 16 + ((lane >> 2) & 1) * 8 + ((lane >> 1) & 1) * 4 + (lane & 1) * 2;
}

// y[m, n] = act(sum_k x[m, k] w[n, k] + bias[n]).  grid (ceil(N/32), ceil(M/16), KS), cluster (1, 1, KS).
__global__ void __launch_bounds__(256, 2) lin_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                         const float* __restrict__ bias, float* __restrict__ y, int M, int N,
                                                         int K, int kslice, int act, float slope) {
    __shared__ __align__(16) float xs[MT][FWD_KC];
    __shared__ __align__(16) float red[FWD_ROWS * MT];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * FWD_ROWS, m0 = blockIdx.y * MT;
    const int kbeg = blockIdx.z * kslice, kend = min(K, kbeg + kslice);
    const int nrow = n0 + warp * 4;
    float acc[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) acc[i] = 0.f;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);

    for (int kc = kbeg; kc < kend; kc += FWD_KC) {
        __syncthreads();
        // stage x[m0 .. m0+16) x [kc, kc+512)
#pragma unroll
        for (int q = 0; q < MT * FWD_KC / 4 / 256; ++q) {
            const int e = threadIdx.x + q * 256;
            const int m = e / (FWD_KC / 4), k4 = (e % (FWD_KC / 4)) * 4;
            const int gk = kc + k4;
            float4 v = zero4;
            if (m0 + m < M && gk < kend) v = *reinterpret_cast<const float4*>(x + (size_t)(m0 + m) * K + gk);
            *reinterpret_cast<float4*>(&xs[m][k4]) = v;
        }
        __syncthreads();
#pragma unroll
        for (int it = 0; it < FWD_KC / 128; it += 2) {
            float4 wv[2][4];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int gk = kc + ((it + u) * 32 + lane) * 4;
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    wv[u][r] = (gk < kend && nrow + r < N) ? ldg_stream(w + (size_t)(nrow + r) * K + gk) : zero4;
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int k4 = ((it + u) * 32 + lane) * 4;
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                    const float4 xv = *reinterpret_cast<const float4*>(&xs[m][k4]);
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        float a = acc[r * MT + m];
                        a = fmaf(wv[u][r].x, xv.x, a);
                        a = fmaf(wv[u][r].y, xv.y, a);
                        a = fmaf(wv[u][r].z, xv.z, a);
                        a = fmaf(wv[u][r].w, xv.w, a);
                        acc[r * MT + m] = a;
                    }
                }
            }
        }
    }
    warp_reduce64(acc, lane);
    const int ob = owner_base(lane);
    red[warp * 64 + ob] = acc[0];
    red[warp * 64 + ob + 1] = acc[1];
    // red[(warp*4 + r) * 16 + m] == partial of row n0 + warp*4 + r, activation row m
    const int KS = gridDim.z;
    if (KS == 1) {
        __syncthreads();
        for (int o = threadIdx.x; o < FWD_ROWS * MT; o += 256) {
            const int m = o / FWD_ROWS, r = o % FWD_ROWS;
            if (m0 + m < M && n0 + r < N) {
                const float v = red[r * MT + m] + (bias ? bias[n0 + r] : 0.f);
                y[(size_t)(m0 + m) * N + n0 + r] = act_apply(v, act, slope);
            }
        }
    } else {
        cg::cluster_group cluster = cg::this_cluster();
        cluster.sync();
        const int per = FWD_ROWS * MT / KS;
        const int rank = blockIdx.z;
        for (int q = threadIdx.x; q < per; q += 256) {
            const int o = rank * per + q;
            const int m = o / FWD_ROWS, r = o % FWD_ROWS;
            float v = 0.f;
            for (int j = 0; j < KS; ++j) v += cluster.map_shared_rank(red, j)[r * MT + m];
            if (m0 + m < M && n0 + r < N) {
                v += bias ? bias[n0 + r] : 0.f;
                y[(size_t)(m0 + m) * N + n0 + r] = act_apply(v, act, slope);
            }
        }
        cluster.sync();
    }
}


// dx[m, k] = sum_n dz[m, n] w[n, k].  grid (ceil(K/128), ceil(M/16), NS); the 8 warps of a CTA split the CTA's n slice,
// lane = one float4 of k.  NS > 1: red.global.add into a pre-zeroed dx.
constexpr int DG_NCH = 256;   // rows of dz staged per chunk
__global__ void __launch_bounds__(256, 2) lin_dgrad_kernel(const float* __restrict__ dz, const float* __restrict__ w,
                                                           float* __restrict__ dx, int M, int N, int K, int nslice) {
    __shared__ __align__(16) float sm[4 * MT * 128];   // 32 KB: dz chunk [n][16] (16 KB) then the cross-warp reduction
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k4 = (blockIdx.x * 32 + lane) * 4;
    const bool kok = k4 < K;
    const int m0 = blockIdx.y * MT;
    const int nbeg = blockIdx.z * nslice, nend = min(N, nbeg + nslice);
    float acc[64];   // [m][4 k]
#pragma unroll
    for (int i = 0; i < 64; ++i) acc[i] = 0.f;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int nc = nbeg; nc < nend; nc += DG_NCH) {
        __syncthreads();
        // stage dz[m0 + m, nc + j] -> sm[j * 16 + m]   (coalesced along n, transposed into smem)
#pragma unroll
        for (int q = 0; q < MT * DG_NCH / 256; ++q) {
            const int e = threadIdx.x + q * 256;
            const int j = e % DG_NCH, m = e / DG_NCH;
            float v = 0.f;
            if (m0 + m < M && nc + j < nend) v = dz[(size_t)(m0 + m) * N + nc + j];
            sm[j * MT + m] = v;
        }
        __syncthreads();
        const int cnt = min(DG_NCH, nend - nc);
        for (int j0 = warp * 8; j0 < cnt; j0 += 64) {     // each warp: 8 consecutive rows per trip
            float4 wv[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) wv[u] = (kok && j0 + u < cnt) ? ldg_stream(w + (size_t)(nc + j0 + u) * K + k4) : zero4;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float4* dzr = reinterpret_cast<const float4*>(&sm[(j0 + u) * MT]);
#pragma unroll
                for (int mq = 0; mq < 4; ++mq) {
                    const float4 d = dzr[mq];
                    const float dm[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                    for (int mm = 0; mm < 4; ++mm) {
                        const int m = mq * 4 + mm;
                        acc[m * 4 + 0] = fmaf(dm[mm], wv[u].x, acc[m * 4 + 0]);
                        acc[m * 4 + 1] = fmaf(dm[mm], wv[u].y, acc[m * 4 + 1]);
                        acc[m * 4 + 2] = fmaf(dm[mm], wv[u].z, acc[m * 4 + 2]);
                        acc[m * 4 + 3] = fmaf(dm[mm], wv[u].w, acc[m * 4 + 3]);
                    }
                }
            }
        }
    }
    // cross-warp tree: 8 -> 4 -> 2 -> 1 through smem; slot layout [warp][m][lane] float4 (conflict-free)
    float4* red = reinterpret_cast<float4*>(sm);
#pragma unroll
    for (int half = 4; half >= 1; half >>= 1) {
        __syncthreads();
        if (warp >= half && warp < 2 * half) {
#pragma unroll
            for (int m = 0; m < MT; ++m)
                red[((warp - half) * MT + m) * 32 + lane] = make_float4(acc[m * 4], acc[m * 4 + 1], acc[m * 4 + 2], acc[m * 4 + 3]);
        }
        __syncthreads();
        if (warp < half) {
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const float4 o = red[(warp * MT + m) * 32 + lane];
                acc[m * 4] += o.x; acc[m * 4 + 1] += o.y; acc[m * 4 + 2] += o.z; acc[m * 4 + 3] += o.w;
            }
        }
    }
    if (warp == 0 && kok) {
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            if (m0 + m >= M) break;
            float* dst = dx + (size_t)(m0 + m) * K + k4;
            if (gridDim.z == 1) {
                *reinterpret_cast<float4*>(dst) = make_float4(acc[m * 4], acc[m * 4 + 1], acc[m * 4 + 2], acc[m * 4 + 3]);
            } else {
                atomicAdd(dst, acc[m * 4]); atomicAdd(dst + 1, acc[m * 4 + 1]);
                atomicAdd(dst + 2, acc[m * 4 + 2]); atomicAdd(dst + 3, acc[m * 4 + 3]);
            }
        }
    }
}

// dw[n, k] (+)= sum_m dz[m, n] x[m, k];  dbias[n] (+)= sum_m dz[m, n]   (M <= 16).
// block (KX, 256/KX): thread = one float4 of k (x column kept in registers for all m) x a strided set of rows n.
constexpr int WG_ROWS = 64;    // rows of dw per CTA
__global__ void __launch_bounds__(256, 2) lin_wgrad_kernel(const float* __restrict__ dz, const float* __restrict__ x,
                                                           float* __restrict__ dw, float* __restrict__ dbias, int M, int N, int K,
                                                           int accumulate, int kx) {
    __shared__ __align__(16) float dzs[WG_ROWS * MT];   // [row][m]
    const int tx = threadIdx.x % kx, ty = threadIdx.x / kx, ny = 256 / kx;
    const int k4 = (blockIdx.x * kx + tx) * 4;
    const bool kok = k4 < K;
    const int n0 = blockIdx.y * WG_ROWS;
    for (int e = threadIdx.x; e < WG_ROWS * MT; e += 256) {
        const int j = e % WG_ROWS, m = e / WG_ROWS;
        dzs[j * MT + m] = (m < M && n0 + j < N) ? dz[(size_t)m * N + n0 + j] : 0.f;
    }
    float4 xr[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) xr[m] = (kok && m < M) ? *reinterpret_cast<const float4*>(x + (size_t)m * K + k4) : make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    if (dbias && blockIdx.x == 0 && threadIdx.x < WG_ROWS && n0 + threadIdx.x < N) {
        float s = 0.f;
#pragma unroll
        for (int m = 0; m < MT; ++m) s += dzs[threadIdx.x * MT + m];
        dbias[n0 + threadIdx.x] = accumulate ? dbias[n0 + threadIdx.x] + s : s;
    }
    if (!kok) return;
    const int rows = min(WG_ROWS, N - n0);
    for (int j0 = ty; j0 < rows; j0 += 4 * ny) {
        float4 old[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int j = j0 + u * ny;
            old[u] = (accumulate && j < rows) ? ldg_stream(dw + (size_t)(n0 + j) * K + k4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int j = j0 + u * ny;
            if (j >= rows) break;
            const float4* dzr = reinterpret_cast<const float4*>(&dzs[j * MT]);
            float4 a = old[u];
#pragma unroll
            for (int mq = 0; mq < 4; ++mq) {
                const float4 d = dzr[mq];
                const float dm[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                for (int mm = 0; mm < 4; ++mm) {
                    const float4 xv = xr[mq * 4 + mm];
                    a.x = fmaf(dm[mm], xv.x, a.x); a.y = fmaf(dm[mm], xv.y, a.y);
                    a.z = fmaf(dm[mm], xv.z, a.z); a.w = fmaf(dm[mm], xv.w, a.w);
                }
            }
            *reinterpret_cast<float4*>(dw + (size_t)(n0 + j) * K + k4) = a;
        }
    }
}

// Same product for 16 < M <= 64 rows (the four accumulation windows of a 16-clip batch stacked along M,
// train_step.FusionTrainStep): the activations no longer fit in registers for all m, so a thread keeps 8 rows x
// one float4 of dw as accumulators and walks m in chunks of 16 (x chunk in registers, L1/L2 resident; dz tile in
// shared memory).  dw is still read (when accumulating) and written exactly once.
constexpr int WG_MMAX = 64;
__global__ void __launch_bounds__(256, 2) lin_wgrad_wide_kernel(const float* __restrict__ dz, const float* __restrict__ x,
                                                                float* __restrict__ dw, float* __restrict__ dbias, int M, int N,
                                                                int K, int accumulate, int kx) {
    __shared__ __align__(16) float dzs[WG_ROWS * WG_MMAX];   // [row][m]
    const int tx = threadIdx.x % kx, ty = threadIdx.x / kx, ny = 256 / kx;
    const int k4 = (blockIdx.x * kx + tx) * 4;
    const bool kok = k4 < K;
    const int n0 = blockIdx.y * WG_ROWS;
    for (int e = threadIdx.x; e < WG_ROWS * WG_MMAX; e += 256) {
        const int j = e % WG_ROWS, m = e / WG_ROWS;
        dzs[j * WG_MMAX + m] = (m < M && n0 + j < N) ? dz[(size_t)m * N + n0 + j] : 0.f;
    }
    __syncthreads();
    if (dbias && blockIdx.x == 0 && threadIdx.x < WG_ROWS && n0 + threadIdx.x < N) {
        float s = 0.f;
        for (int m = 0; m < M; ++m) s += dzs[threadIdx.x * WG_MMAX + m];
        dbias[n0 + threadIdx.x] = accumulate ? dbias[n0 + threadIdx.x] + s : s;
    }
    if (!kok) return;
    const int rows = min(WG_ROWS, N - n0);
    const int mchunks = (M + MT - 1) / MT;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j0 = ty; j0 < rows; j0 += 8 * ny) {
        float4 a[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int j = j0 + u * ny;
            a[u] = (accumulate && j < rows) ? ldg_stream(dw + (size_t)(n0 + j) * K + k4) : zero4;
        }
        for (int mc = 0; mc < mchunks; ++mc) {
            float4 xr[MT];
#pragma unroll
            for (int mm = 0; mm < MT; ++mm) {
                const int m = mc * MT + mm;
                xr[mm] = (m < M) ? *reinterpret_cast<const float4*>(x + (size_t)m * K + k4) : zero4;
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int j = min(j0 + u * ny, rows - 1);      // rows past the end recompute the last one and are not stored
                const float4* dzr = reinterpret_cast<const float4*>(&dzs[j * WG_MMAX + mc * MT]);
#pragma unroll
                for (int mq = 0; mq < 4; ++mq) {
                    const float4 d = dzr[mq];
                    const float dm[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
                    for (int mm = 0; mm < 4; ++mm) {
                        const float4 xv = xr[mq * 4 + mm];
                        a[u].x = fmaf(dm[mm], xv.x, a[u].x); a[u].y = fmaf(dm[mm], xv.y, a[u].y);
                        a[u].z = fmaf(dm[mm], xv.z, a[u].z); a[u].w = fmaf(dm[mm], xv.w, a[u].w);
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int j = j0 + u * ny;
            if (j < rows) *reinterpret_cast<float4*>(dw + (size_t)(n0 + j) * K + k4) = a[u];
        }
    }
}

}  // namespace skinny

// ------------------------------------------------------------------------------------------------ host launchers
// Each returns 1 when it handled the call, 0 when the shape is outside the skinny envelope (the caller falls back to
// the tiled SIMT GEMM), < 0 on error.
static int pow2_at_most(int v, int cap) {
    int p = 1;
    while (p * 2 <= v && p * 2 <= cap) p *= 2;
    return p;
}

int maavss_skinny_fwd(const float* x, const float* w, const float* bias, float* y, int M, int N, int K, int act, float slope,
                      cudaStream_t st) {
    if (M > 128 || K % 4 != 0 || K < 128) return 0;
    if (((uintptr_t)x | (uintptr_t)w) & 15) return 0;
    const int nb = ceil_div(N, skinny::FWD_ROWS), mb = ceil_div(M, skinny::MT);
    int ks = 1;
    while (ks < 8 && nb * mb * ks < 2 * maavss_num_sms() && K / (ks * 2) >= 256) ks *= 2;
    int kslice = ((K + ks - 1) / ks + 127) / 128 * 128;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nb, mb, ks);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = ks;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    MAAVSS_CUDA(cudaLaunchKernelEx(&cfg, skinny::lin_fwd_kernel, x, w, bias, y, M, N, K, kslice, act, slope));
    maavss_count_launch();
    return 1;
}

int maavss_skinny_dgrad(const float* dz, const float* w, float* dx, int M, int N, int K, cudaStream_t st) {
    if (M > 128 || K % 4 != 0) return 0;
    if (((uintptr_t)w | (uintptr_t)dx) & 15) return 0;
    const int kb = ceil_div(K, 128), mb = ceil_div(M, skinny::MT);
    int ns = 1;
    while (kb * mb * ns < 2 * maavss_num_sms() && N / (ns * 2) >= 64) ns *= 2;
    const int nslice = (N + ns - 1) / ns;
    if (ns > 1) MAAVSS_CUDA(cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)M * K, st));
    skinny::lin_dgrad_kernel<<<dim3(kb, mb, ns), 256, 0, st>>>(dz, w, dx, M, N, K, nslice);
    MAAVSS_LAUNCH_CHECK();
    return 1;
}

int maavss_skinny_wgrad(const float* dz, const float* x, float* dw, float* dbias, int M, int N, int K, int accumulate,
                        cudaStream_t st) {
    if (M > skinny::WG_MMAX || K % 4 != 0) return 0;
    if (((uintptr_t)x | (uintptr_t)dw) & 15) return 0;
    const int kx = pow2_at_most(K / 4 > 0 ? K / 4 : 1, 256);
    dim3 grid(ceil_div(K / 4, kx), ceil_div(N, skinny::WG_ROWS));
    if (M <= skinny::MT) skinny::lin_wgrad_kernel<<<grid, 256, 0, st>>>(dz, x, dw, dbias, M, N, K, accumulate, kx);
    else skinny::lin_wgrad_wide_kernel<<<grid, 256, 0, st>>>(dz, x, dw, dbias, M, N, K, accumulate, kx);
    MAAVSS_LAUNCH_CHECK();
    return 1;
}
