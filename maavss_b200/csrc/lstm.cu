// Bidirectional single-layer LSTM without biases (nn.LSTM(bias=False, batch_first=True, bidirectional=True),
// avse_model.py:542-543, avse_model_final.py:132-133): the recurrent part, one launch per time step covering both
// directions.  The input projections x w_ih^T and all weight gradients are GEMMs (gemm_simt.cu).
// Layouts: gin/gates/dgates (2, B, T, 4H) with gate order i,f,g,o; cells (2, B, T, H); y/dy (B, T, 2H).
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
