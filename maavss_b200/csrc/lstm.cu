// Bidirectional single-layer LSTM without biases (nn.LSTM(bias=False, batch_first=True, bidirectional=True),
// avse_model.py:542-543, avse_model_final.py:132-133): the recurrent part, one launch per time step covering both
// directions.  The input projections x w_ih^T and all weight gradients are GEMMs (gemm_simt.cu).
// Layouts: gin/gates/dgates (2, B, T, 4H) with gate order i,f,g,o; cells (2, B, T, H); y/dy (B, T, 2H).
#include <stdlib.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

__device__ __forceinline__ float sigmoidf_(float v) { return 1.f / (1.f + expf(-v)); }

// grid (H / FU, 2), 256 threads.  A CTA owns FU = 4 hidden units j: its 4 x FU rows of w_hh and the previous hidden state
// of a 64-row batch chunk sit in shared memory (h rows H + 4 floats apart: conflict-free float4 reads across b, the w
// rows are warp-uniform broadcasts); thread (b, u) accumulates the four gate pre-activations of (b, j0 + u) over k with
// 5 LDS.128 per 16 FMA, then finishes the cell update.
constexpr int FU = 4, FB = 64;
__global__ void __launch_bounds__(256) lstm_step_fwd_kernel(const float* __restrict__ gin, const float* __restrict__ w0,
                                                            const float* __restrict__ w1, float* __restrict__ gates,
                                                            float* __restrict__ cells, float* y, int B, int T, int H, int step) {
    extern __shared__ __align__(16) float sm[];
    const int HS = H + 4, H4 = H >> 2;
    float* ws = sm;                  // [4 * FU][H], row g * FU + u
    float* hs = sm + 4 * FU * H;     // [FB][HS]
    const int dir = blockIdx.y, j0 = blockIdx.x * FU;
    const int t = dir ? T - 1 - step : step;
    const int tp = dir ? t + 1 : t - 1;
    const int bl = threadIdx.x & (FB - 1), u = threadIdx.x >> 6, j = j0 + u;
    if (step > 0) {
        const float* w = dir ? w1 : w0;
        for (int e = threadIdx.x; e < 4 * FU * H4; e += 256) {
            const int row = e / H4, c4 = e - row * H4;
            const int g = row / FU, uu = row - g * FU;
            reinterpret_cast<float4*>(ws)[e] = __ldg(reinterpret_cast<const float4*>(w + (size_t)(g * H + j0 + uu) * H) + c4);
        }
    }
    for (int b0 = 0; b0 < B; b0 += FB) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        if (step > 0) {
            __syncthreads();         // the previous chunk's reads of hs
            const int nb = min(FB, B - b0);
            for (int e = threadIdx.x; e < FB * H4; e += 256) {
                const int r = e / H4, c4 = e - r * H4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < nb) v = *(reinterpret_cast<const float4*>(y + ((size_t)(b0 + r) * T + tp) * 2 * H + dir * H) + c4);
                *reinterpret_cast<float4*>(hs + r * HS + 4 * c4) = v;
            }
            __syncthreads();
            const float4* hp = reinterpret_cast<const float4*>(hs + bl * HS);
            const float4* wp = reinterpret_cast<const float4*>(ws + u * H);
#pragma unroll 2
            for (int k = 0; k < H4; ++k) {
                const float4 h = hp[k];
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const float4 wv = wp[g * FU * H4 + k];
                    acc[g] = fmaf(h.x, wv.x, acc[g]);
                    acc[g] = fmaf(h.y, wv.y, acc[g]);
                    acc[g] = fmaf(h.z, wv.z, acc[g]);
                    acc[g] = fmaf(h.w, wv.w, acc[g]);
                }
            }
        }
        const int b = b0 + bl;
        if (b < B) {
            const size_t gb = (((size_t)dir * B + b) * T + t) * 4 * H;
            const float ig = sigmoidf_(acc[0] + gin[gb + 0 * H + j]);
            const float fg = sigmoidf_(acc[1] + gin[gb + 1 * H + j]);
            const float gg = tanhf(acc[2] + gin[gb + 2 * H + j]);
            const float og = sigmoidf_(acc[3] + gin[gb + 3 * H + j]);
            const float cp = step > 0 ? cells[(((size_t)dir * B + b) * T + tp) * H + j] : 0.f;
            const float c = fmaf(fg, cp, ig * gg);
            gates[gb + 0 * H + j] = ig;
            gates[gb + 1 * H + j] = fg;
            gates[gb + 2 * H + j] = gg;
            gates[gb + 3 * H + j] = og;
            cells[(((size_t)dir * B + b) * T + t) * H + j] = c;
            y[((size_t)b * T + t) * 2 * H + dir * H + j] = og * tanhf(c);
        }
    }
}

// Backward step.  grid (H/32, 2, ceil(B/8)), 512 threads = 32 hidden units j x 16 splits of the 4H gate rows.
// dh[b, j] = dy[b, t, j] + sum_k dgates[b, t_next, k] w_hh[k, j]: every w_hh element is read once per CTA column
// (coalesced over j), the 8-batch chunk of dgates(t_next) sits transposed in shared memory ([k][b], float4
// broadcast reads), the 16 partial sums meet in shared memory, then 2 (b, j) cells per thread finish the gate gradients.
constexpr int LB = 8;    // batch rows per chunk
constexpr int KS = 16;   // splits of the 4H reduction
constexpr int BT = 32 * KS;
__global__ void __launch_bounds__(BT) lstm_step_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ gates,
                                                            const float* __restrict__ cells, const float* __restrict__ w0,
                                                            const float* __restrict__ w1, float* __restrict__ dgates,
                                                            float* __restrict__ dcell, int B, int T, int H, int step) {
    extern __shared__ __align__(16) float sm[];
    float* dgn = sm;                        // [4H][LB]
    float* red = sm + 4 * H * LB;           // [KS][LB][32]
    const int dir = blockIdx.y;
    const int jl = threadIdx.x & 31, ks = threadIdx.x >> 5;
    const int j = blockIdx.x * 32 + jl;
    const int t = dir ? T - 1 - step : step;
    const int tn = dir ? t - 1 : t + 1;   // time index processed at step + 1
    const int tp = dir ? t + 1 : t - 1;   // time index processed at step - 1
    const bool has_next = step + 1 < T;
    const float* w = (dir ? w1 : w0);
    const int H4 = 4 * H, kper = H4 / KS;
    {
        const int b0 = blockIdx.z * LB;
        const int nb = min(LB, B - b0);
        if (has_next) {
            __syncthreads();
            for (int e = threadIdx.x; e < LB * H4; e += BT) {
                const int k = e % H4, b = e / H4;
                dgn[k * LB + b] = b < nb ? dgates[(((size_t)dir * B + b0 + b) * T + tn) * H4 + k] : 0.f;
            }
            __syncthreads();
            float acc[LB];
#pragma unroll
            for (int b = 0; b < LB; ++b) acc[b] = 0.f;
            const int k0 = ks * kper;
#pragma unroll 8
            for (int k = k0; k < k0 + kper; ++k) {
                const float wv = __ldg(w + (size_t)k * H + j);
                const float4* d = reinterpret_cast<const float4*>(dgn + k * LB);
#pragma unroll
                for (int q = 0; q < LB / 4; ++q) {
                    const float4 dv = d[q];
                    acc[q * 4 + 0] = fmaf(dv.x, wv, acc[q * 4 + 0]);
                    acc[q * 4 + 1] = fmaf(dv.y, wv, acc[q * 4 + 1]);
                    acc[q * 4 + 2] = fmaf(dv.z, wv, acc[q * 4 + 2]);
                    acc[q * 4 + 3] = fmaf(dv.w, wv, acc[q * 4 + 3]);
                }
            }
#pragma unroll
            for (int b = 0; b < LB; ++b) red[(ks * LB + b) * 32 + jl] = acc[b];
            __syncthreads();
        }
        for (int e = threadIdx.x; e < nb * 32; e += BT) {
            const int jj = e & 31, bl = e >> 5;
            const int b = b0 + bl, jg = blockIdx.x * 32 + jj;
            float dh = dy[((size_t)b * T + t) * 2 * H + dir * H + jg];
            if (has_next) {
                float s = 0.f;
#pragma unroll
                for (int q = 0; q < KS; ++q) s += red[(q * LB + bl) * 32 + jj];
                dh += s;
            }
            const size_t gb = (((size_t)dir * B + b) * T + t) * H4;
            const float ig = gates[gb + jg], fg = gates[gb + H + jg], gg = gates[gb + 2 * H + jg], og = gates[gb + 3 * H + jg];
            const float c = cells[(((size_t)dir * B + b) * T + t) * H + jg];
            const float cp = step > 0 ? cells[(((size_t)dir * B + b) * T + tp) * H + jg] : 0.f;
            const float tc = tanhf(c);
            float* dcp = dcell + ((size_t)dir * B + b) * H + jg;
            const float dc = fmaf(dh * og, 1.f - tc * tc, *dcp);
            dgates[gb + jg] = dc * gg * ig * (1.f - ig);
            dgates[gb + H + jg] = dc * cp * fg * (1.f - fg);
            dgates[gb + 2 * H + jg] = dc * ig * (1.f - gg * gg);
            dgates[gb + 3 * H + jg] = dh * tc * og * (1.f - og);
            *dcp = dc * fg;
        }
    }
}

}  // namespace

extern "C" int maavss_lstm_step_fwd(const float* gin, const float* w_hh_fwd, const float* w_hh_rev, float* gates, float* cells,
                                    float* y, int B, int T, int H, int step, void* stream) {
    if (B == 0 || T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(gin && w_hh_fwd && w_hh_rev && gates && cells && y, "lstm_step_fwd: null argument");
    MAAVSS_CHECK_ARG(H % 32 == 0 && H >= 32 && H <= 512, "lstm_step_fwd: H=%d must be a multiple of 32 in [32, 512]", H);
    MAAVSS_CHECK_ARG(step >= 0 && step < T && B <= 4096, "lstm_step_fwd: bad step/batch");
    dim3 grid(H / FU, 2);
    const size_t smem = sizeof(float) * ((size_t)4 * FU * H + (size_t)FB * (H + 4));
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        MAAVSS_CUDA(cudaFuncSetAttribute(lstm_step_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    lstm_step_fwd_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(gin, w_hh_fwd, w_hh_rev, gates, cells, y, B, T, H, step);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_lstm_step_bwd(const float* dy, const float* gates, const float* cells, const float* w_hh_fwd,
                                    const float* w_hh_rev, float* dgates, float* dcell, int B, int T, int H, int step,
                                    void* stream) {
    if (B == 0 || T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dy && gates && cells && w_hh_fwd && w_hh_rev && dgates && dcell, "lstm_step_bwd: null argument");
    MAAVSS_CHECK_ARG(step >= 0 && step < T, "lstm_step_bwd: bad step");
    MAAVSS_CHECK_ARG(H % 32 == 0 && H >= 32 && H <= 512, "lstm_step_bwd: H=%d must be a multiple of 32 in [32, 512]", H);
    dim3 grid(H / 32, 2, ceil_div(B, LB));
    const size_t smem = sizeof(float) * (4 * H * LB + KS * LB * 32);
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        MAAVSS_CUDA(cudaFuncSetAttribute(lstm_step_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    lstm_step_bwd_kernel<<<grid, BT, smem, (cudaStream_t)stream>>>(dy, gates, cells, w_hh_fwd, w_hh_rev, dgates, dcell, B, T, H,
                                                                   step);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// Whole-layer entry points (bilstm_fwd / bilstm_bwd): input projections, the T recurrent steps, and in the
// backward the h_{t-1} gather, the four weight gradients and the input gradient, all enqueued by one C call.
namespace {

// hprev[d][b][t][:] = h of the step BEFORE t in direction d (zeros at the direction's first step), from y (B, T, 2H)
__global__ void __launch_bounds__(256) lstm_hprev_kernel(const float* __restrict__ y, float* __restrict__ hprev, int B, int T, int H4) {
    const long long n = 2LL * B * T * H4;            // float4 units
    for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < n; e += (long long)gridDim.x * 256) {
        const int c4 = (int)(e % H4);
        long long r = e / H4;
        const int t = (int)(r % T);
        r /= T;
        const int b = (int)(r % B), d = (int)(r / B);
        const int tp = d ? t + 1 : t - 1;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tp >= 0 && tp < T) v = __ldg(reinterpret_cast<const float4*>(y + ((size_t)b * T + tp) * 8 * H4 + (size_t)d * 4 * H4) + c4);
        reinterpret_cast<float4*>(hprev)[e] = v;
    }
}

inline size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

// one launch for all T steps (lstm_persist.cu) unless MAAVSS_LSTM_STEPS=1 asks for the per-step kernels (benchmarks)
inline bool use_recurrent_kernel(int B, int T, int H, int backward) {
    // bit 0: per-step kernels in the forward, bit 1: in the backward.  Default 2: measured on the headline step (same box,
    // 150 steps each) 2.147 ms with the one-launch forward + per-step backward, 2.170 the other way round, 2.215 both in one
    // launch, 2.240 both per step -- the one-launch backward (a full register file per CTA) cannot share SMs with the
    // weight-gradient GEMMs and the optimizer tail that run beside it.
    static const int steps = [] {
        const char* e = getenv("MAAVSS_LSTM_STEPS");
        return e ? atoi(e) : 2;
    }();
    return !(steps & (backward ? 2 : 1)) && maavss_bilstm_recurrent_supported(B, T, H);
}

// y[M, N] = x w^T on the 3xTF32 tensor-core GEMM when the shape allows, else the SIMT kernels
int proj_fwd(const float* x, const float* w, float* y, int M, int N, int K, void* ws, long long ws_bytes, void* st) {
    if (M >= 32 && maavss_gemm_tf32x3_supported(N, M, K, 0, K, 0, K, N, 1) && maavss_gemm_tf32x3_workspace_bytes(N, M, K) <= ws_bytes)
        return maavss_gemm_tf32x3(w, 0, K, x, 0, K, N, M, K, y, N, 1, nullptr, 0, MAAVSS_ACT_NONE, 0.f, 0, ws, ws_bytes, st);
    return maavss_linear_fwd(x, w, nullptr, y, M, N, K, MAAVSS_ACT_NONE, 0.f, st);
}
int proj_wgrad(const float* dz, const float* x, float* dw, int M, int N, int K, int accumulate, void* ws, long long ws_bytes, void* st) {
    if (M >= 32 && maavss_gemm_tf32x3_supported(N, K, M, 1, N, 1, K, K, 0) && maavss_gemm_tf32x3_workspace_bytes(N, K, M) <= ws_bytes)
        return maavss_gemm_tf32x3(dz, 1, N, x, 1, K, N, K, M, dw, K, 0, nullptr, 0, MAAVSS_ACT_NONE, 0.f, accumulate, ws, ws_bytes, st);
    return maavss_linear_wgrad(dz, x, dw, nullptr, M, N, K, accumulate, st);
}

}  // namespace

extern "C" long long maavss_bilstm_workspace_bytes(int B, int T, int I, int H, int backward) {
    const int M = B * T;
    long long g = maavss_gemm_tf32x3_workspace_bytes(4 * H, M, I);
    if (backward) {
        const long long c[] = {maavss_gemm_tf32x3_workspace_bytes(4 * H, I, M), maavss_gemm_tf32x3_workspace_bytes(4 * H, H, M),
                               maavss_gemm_tf32x3_workspace_bytes(I, M, 4 * H)};
        for (long long v : c) g = v > g ? v : g;
        // + dcell (2, B, H), hprev (2, B, T, H), dx partial (B*T, I)
        return (long long)(al256((size_t)g) + al256(sizeof(float) * 2 * (size_t)B * H) + al256(sizeof(float) * 2 * (size_t)M * H) +
                           al256(sizeof(float) * (size_t)M * I)) + 256;
    }
    return (long long)al256((size_t)g) + 256;
}

extern "C" int maavss_bilstm_fwd(const float* x, const float* w_ih_fwd, const float* w_hh_fwd, const float* w_ih_rev, const float* w_hh_rev,
                                 float* gin, float* gates, float* cells, float* y, int B, int T, int I, int H, void* workspace,
                                 long long workspace_bytes, void* stream) {
    if (B == 0 || T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && w_ih_fwd && w_hh_fwd && w_ih_rev && w_hh_rev && gin && gates && cells && y, "bilstm_fwd: null argument");
    MAAVSS_CHECK_ARG(workspace_bytes >= maavss_bilstm_workspace_bytes(B, T, I, H, 0), "bilstm_fwd: workspace too small");
    void* ws = (void*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    const long long wsb = workspace_bytes - 256;
    const int M = B * T;
    int rc = proj_fwd(x, w_ih_fwd, gin, M, 4 * H, I, ws, wsb, stream);
    if (rc) return rc;
    rc = proj_fwd(x, w_ih_rev, gin + (size_t)M * 4 * H, M, 4 * H, I, ws, wsb, stream);
    if (rc) return rc;
    if (use_recurrent_kernel(B, T, H, 0)) return maavss_bilstm_recurrent_fwd(gin, w_hh_fwd, w_hh_rev, gates, cells, y, B, T, H, stream);
    for (int s = 0; s < T; ++s) {
        rc = maavss_lstm_step_fwd(gin, w_hh_fwd, w_hh_rev, gates, cells, y, B, T, H, s, stream);
        if (rc) return rc;
    }
    return MAAVSS_OK;
}

extern "C" int maavss_bilstm_wgrad(const float* x, const float* y, const float* dgates, float* dw_ih_fwd, float* dw_hh_fwd, float* dw_ih_rev,
                                   float* dw_hh_rev, int accumulate, int B, int T, int I, int H, void* workspace, long long workspace_bytes,
                                   void* stream) {
    if (B == 0 || T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && y && dgates, "bilstm_wgrad: null argument");
    MAAVSS_CHECK_ARG(workspace_bytes >= maavss_bilstm_workspace_bytes(B, T, I, H, 1), "bilstm_wgrad: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int M = B * T;
    uint8_t* base = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    long long g = maavss_bilstm_workspace_bytes(B, T, I, H, 1) - 256 - (long long)(al256(sizeof(float) * 2 * (size_t)B * H) +
                  al256(sizeof(float) * 2 * (size_t)M * H) + al256(sizeof(float) * (size_t)M * I));
    void* gws = base;
    float* hprev = (float*)(base + g + al256(sizeof(float) * 2 * (size_t)B * H));
    {
        const long long n4 = 2LL * M * (H / 4);
        int blocks = (int)((n4 + 255) / 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        lstm_hprev_kernel<<<blocks, 256, 0, st>>>(y, hprev, B, T, H / 4);
        MAAVSS_LAUNCH_CHECK();
    }
    float* dw_ih[2] = {dw_ih_fwd, dw_ih_rev};
    float* dw_hh[2] = {dw_hh_fwd, dw_hh_rev};
    for (int d = 0; d < 2; ++d) {
        const float* dg = dgates + (size_t)d * M * 4 * H;
        int rc;
        if (dw_ih[d]) {
            rc = proj_wgrad(dg, x, dw_ih[d], M, 4 * H, I, accumulate, gws, g, stream);
            if (rc) return rc;
        }
        if (dw_hh[d]) {
            rc = proj_wgrad(dg, hprev + (size_t)d * M * H, dw_hh[d], M, 4 * H, H, accumulate, gws, g, stream);
            if (rc) return rc;
        }
    }
    return MAAVSS_OK;
}

extern "C" int maavss_bilstm_bwd(const float* dy, const float* x, const float* w_ih_fwd, const float* w_hh_fwd, const float* w_ih_rev,
                                 const float* w_hh_rev, const float* gates, const float* cells, const float* y, float* dgates, float* dx,
                                 float* dw_ih_fwd, float* dw_hh_fwd, float* dw_ih_rev, float* dw_hh_rev, int accumulate, int B, int T,
                                 int I, int H, void* workspace, long long workspace_bytes, void* stream) {
    if (B == 0 || T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dy && x && w_ih_fwd && w_hh_fwd && w_ih_rev && w_hh_rev && gates && cells && y && dgates, "bilstm_bwd: null argument");
    MAAVSS_CHECK_ARG(workspace_bytes >= maavss_bilstm_workspace_bytes(B, T, I, H, 1), "bilstm_bwd: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int M = B * T;
    uint8_t* base = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    long long g = maavss_bilstm_workspace_bytes(B, T, I, H, 1) - 256 - (long long)(al256(sizeof(float) * 2 * (size_t)B * H) +
                  al256(sizeof(float) * 2 * (size_t)M * H) + al256(sizeof(float) * (size_t)M * I));
    void* gws = base;
    float* dcell = (float*)(base + g);
    float* hprev = (float*)((uint8_t*)dcell + al256(sizeof(float) * 2 * (size_t)B * H));
    int rc;
    if (use_recurrent_kernel(B, T, H, 1)) {
        rc = maavss_bilstm_recurrent_bwd(dy, gates, cells, w_hh_fwd, w_hh_rev, dgates, B, T, H, stream);
        if (rc) return rc;
    } else {
        MAAVSS_CUDA(cudaMemsetAsync(dcell, 0, sizeof(float) * 2 * (size_t)B * H, st));
        for (int s = T - 1; s >= 0; --s) {
            rc = maavss_lstm_step_bwd(dy, gates, cells, w_hh_fwd, w_hh_rev, dgates, dcell, B, T, H, s, stream);
            if (rc) return rc;
        }
    }
    // the input gradient first: it is what the rest of the backward pass waits for
    const float* w_ih[2] = {w_ih_fwd, w_ih_rev};
    for (int d = 0; d < 2 && dx; ++d) {
        const float* dg = dgates + (size_t)d * M * 4 * H;
        // dx[M, I] (+)= dg[M, 4H] w_ih[4H, I]: the second direction accumulates into the first one's result
        if (M >= 32 && maavss_gemm_tf32x3_supported(I, M, 4 * H, 1, I, 0, 4 * H, I, 1) && maavss_gemm_tf32x3_workspace_bytes(I, M, 4 * H) <= g) {
            rc = maavss_gemm_tf32x3(w_ih[d], 1, I, dg, 0, 4 * H, I, M, 4 * H, dx, I, 1, nullptr, 0, MAAVSS_ACT_NONE, 0.f, d, gws, g, stream);
        } else if (d == 0) {
            rc = maavss_linear_dgrad(dg, w_ih[d], dx, M, 4 * H, I, stream);
        } else {
            float* part = (float*)((uint8_t*)hprev + al256(sizeof(float) * 2 * (size_t)M * H));
            rc = maavss_linear_dgrad(dg, w_ih[d], part, M, 4 * H, I, stream);
            if (!rc) rc = maavss_add_noise(dx, part, 1.0f, dx, (long long)M * I, stream);      // dx += part (unfused a + b * 1)
        }
        if (rc) return rc;
    }
    if (dw_ih_fwd || dw_hh_fwd || dw_ih_rev || dw_hh_rev)
        return maavss_bilstm_wgrad(x, y, dgates, dw_ih_fwd, dw_hh_fwd, dw_ih_rev, dw_hh_rev, accumulate, B, T, I, H, workspace, workspace_bytes,
                                   stream);
    return MAAVSS_OK;
}
