// Bidirectional single-layer LSTM without biases (nn.LSTM(bias=False, batch_first=True, bidirectional=True),
// avse_model.py:542-543, avse_model_final.py:132-133): the recurrent part, one launch per time step covering both
// directions.  The input projections x w_ih^T and all weight gradients are GEMMs (gemm_simt.cu).
// Layouts: gin/gates/dgates (2, B, T, 4H) with gate order i,f,g,o; cells (2, B, T, H); y/dy (B, T, 2H).
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

__device__ __forceinline__ float sigmoidf_(float v) { return 1.f / (1.f + expf(-v)); }

// grid (H, 2), 128 threads: warp g owns row g*H + j of w_hh (kept in registers), dots it with h_prev[b, :] for
// every b, then threads b < B finish the cell update of (b, j).
template <int HPL>   // H / 32 values per lane
__global__ void __launch_bounds__(128) lstm_step_fwd_kernel(const float* __restrict__ gin, const float* __restrict__ w0,
                                                            const float* __restrict__ w1, float* __restrict__ gates,
                                                            float* __restrict__ cells, float* __restrict__ y, int B, int T,
                                                            int H, int step) {
    extern __shared__ float pre[];   // [4][B]
    const int j = blockIdx.x, dir = blockIdx.y;
    const int t = dir ? T - 1 - step : step;
    const int tp = dir ? t + 1 : t - 1;
    const int g = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* w = (dir ? w1 : w0) + (size_t)(g * H + j) * H;
    float wr[HPL];
#pragma unroll
    for (int q = 0; q < HPL; ++q) wr[q] = w[lane + 32 * q];
    for (int b = 0; b < B; ++b) {
        float s = 0.f;
        if (step > 0) {
            const float* hp = y + ((size_t)b * T + tp) * 2 * H + dir * H;
#pragma unroll
            for (int q = 0; q < HPL; ++q) s = fmaf(wr[q], hp[lane + 32 * q], s);
            s = warp_sum(s);
        }
        if (lane == 0) pre[g * B + b] = s;
    }
    __syncthreads();
    for (int b = threadIdx.x; b < B; b += 128) {
        const size_t gb = (((size_t)dir * B + b) * T + t) * 4 * H;
        const float ig = sigmoidf_(pre[0 * B + b] + gin[gb + 0 * H + j]);
        const float fg = sigmoidf_(pre[1 * B + b] + gin[gb + 1 * H + j]);
        const float gg = tanhf(pre[2 * B + b] + gin[gb + 2 * H + j]);
        const float og = sigmoidf_(pre[3 * B + b] + gin[gb + 3 * H + j]);
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

// Backward step.  grid (H/32, 2, ceil(B/16)), 256 threads = 32 hidden units j x 8 splits of the 4H gate rows.
// dh[b, j] = dy[b, t, j] + sum_k dgates[b, t_next, k] w_hh[k, j]: every w_hh element is read once per CTA column
// (coalesced over j), the 16-batch chunk of dgates(t_next) sits transposed in shared memory ([k][b], float4
// broadcast reads), the 8 partial sums meet in shared memory, then 2 (b, j) cells per thread finish the gate gradients.
constexpr int LB = 16;   // batch rows per chunk
__global__ void __launch_bounds__(256) lstm_step_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ gates,
                                                            const float* __restrict__ cells, const float* __restrict__ w0,
                                                            const float* __restrict__ w1, float* __restrict__ dgates,
                                                            float* __restrict__ dcell, int B, int T, int H, int step) {
    extern __shared__ __align__(16) float sm[];
    float* dgn = sm;                        // [4H][LB]
    float* red = sm + 4 * H * LB;           // [8][LB][32]
    const int dir = blockIdx.y;
    const int jl = threadIdx.x & 31, ks = threadIdx.x >> 5;
    const int j = blockIdx.x * 32 + jl;
    const int t = dir ? T - 1 - step : step;
    const int tn = dir ? t - 1 : t + 1;   // time index processed at step + 1
    const int tp = dir ? t + 1 : t - 1;   // time index processed at step - 1
    const bool has_next = step + 1 < T;
    const float* w = (dir ? w1 : w0);
    const int H4 = 4 * H, kper = H4 / 8;
    {
        const int b0 = blockIdx.z * LB;
        const int nb = min(LB, B - b0);
        if (has_next) {
            __syncthreads();
            for (int e = threadIdx.x; e < LB * H4; e += 256) {
                const int k = e % H4, b = e / H4;
                dgn[k * LB + b] = b < nb ? dgates[(((size_t)dir * B + b0 + b) * T + tn) * H4 + k] : 0.f;
            }
            __syncthreads();
            float acc[LB];
#pragma unroll
            for (int b = 0; b < LB; ++b) acc[b] = 0.f;
            const int k0 = ks * kper;
#pragma unroll 4
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
        for (int e = threadIdx.x; e < nb * 32; e += 256) {
            const int jj = e & 31, bl = e >> 5;
            const int b = b0 + bl, jg = blockIdx.x * 32 + jj;
            float dh = dy[((size_t)b * T + t) * 2 * H + dir * H + jg];
            if (has_next) {
                float s = 0.f;
#pragma unroll
                for (int q = 0; q < 8; ++q) s += red[(q * LB + bl) * 32 + jj];
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
    dim3 grid(H, 2);
    const size_t smem = sizeof(float) * 4 * B;
    cudaStream_t st = (cudaStream_t)stream;
#define LSTM_CASE(HPL_) case HPL_: lstm_step_fwd_kernel<HPL_><<<grid, 128, smem, st>>>(gin, w_hh_fwd, w_hh_rev, gates, cells, y, B, T, H, step); break;
    switch (H / 32) {
        LSTM_CASE(1) LSTM_CASE(2) LSTM_CASE(3) LSTM_CASE(4) LSTM_CASE(5) LSTM_CASE(6) LSTM_CASE(7) LSTM_CASE(8)
        LSTM_CASE(9) LSTM_CASE(10) LSTM_CASE(11) LSTM_CASE(12) LSTM_CASE(13) LSTM_CASE(14) LSTM_CASE(15) LSTM_CASE(16)
    }
#undef LSTM_CASE
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
    const size_t smem = sizeof(float) * (4 * H * LB + 8 * LB * 32);
    static size_t smem_set = 0;
    if (smem > 48 * 1024 && smem > smem_set) {
        MAAVSS_CUDA(cudaFuncSetAttribute(lstm_step_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    lstm_step_bwd_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(dy, gates, cells, w_hh_fwd, w_hh_rev, dgates, dcell, B, T, H,
                                                                   step);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
