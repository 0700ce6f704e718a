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

// grid (ceil(H/256), B, 2): thread (b, j).  dh = dy + dgates[t_next] w_hh (coalesced over j), then the gate gradients.
__global__ void __launch_bounds__(256) lstm_step_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ gates,
                                                            const float* __restrict__ cells, const float* __restrict__ w0,
                                                            const float* __restrict__ w1, float* __restrict__ dgates,
                                                            float* __restrict__ dcell, int B, int T, int H, int step) {
    extern __shared__ float dgn[];   // [4H] gate gradients of the recurrence's next step for this b
    const int b = blockIdx.y, dir = blockIdx.z;
    const int j = blockIdx.x * 256 + threadIdx.x;
    const int t = dir ? T - 1 - step : step;
    const int tn = dir ? t - 1 : t + 1;   // time index processed at step + 1
    const int tp = dir ? t + 1 : t - 1;   // time index processed at step - 1
    const bool has_next = step + 1 < T;
    if (has_next) {
        const float* src = dgates + (((size_t)dir * B + b) * T + tn) * 4 * H;
        for (int k = threadIdx.x; k < 4 * H; k += 256) dgn[k] = src[k];
    }
    __syncthreads();
    if (j >= H) return;
    float dh = dy[((size_t)b * T + t) * 2 * H + dir * H + j];
    if (has_next) {
        const float* w = (dir ? w1 : w0) + j;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        for (int k = 0; k < 4 * H; k += 4) {
            s0 = fmaf(dgn[k], w[(size_t)k * H], s0);
            s1 = fmaf(dgn[k + 1], w[(size_t)(k + 1) * H], s1);
            s2 = fmaf(dgn[k + 2], w[(size_t)(k + 2) * H], s2);
            s3 = fmaf(dgn[k + 3], w[(size_t)(k + 3) * H], s3);
        }
        dh += (s0 + s1) + (s2 + s3);
    }
    const size_t gb = (((size_t)dir * B + b) * T + t) * 4 * H;
    const float ig = gates[gb + j], fg = gates[gb + H + j], gg = gates[gb + 2 * H + j], og = gates[gb + 3 * H + j];
    const float c = cells[(((size_t)dir * B + b) * T + t) * H + j];
    const float cp = step > 0 ? cells[(((size_t)dir * B + b) * T + tp) * H + j] : 0.f;
    const float tc = tanhf(c);
    float* dcp = dcell + ((size_t)dir * B + b) * H + j;
    const float dc = fmaf(dh * og, 1.f - tc * tc, *dcp);
    dgates[gb + j] = dc * gg * ig * (1.f - ig);
    dgates[gb + H + j] = dc * cp * fg * (1.f - fg);
    dgates[gb + 2 * H + j] = dc * ig * (1.f - gg * gg);
    dgates[gb + 3 * H + j] = dh * tc * og * (1.f - og);
    *dcp = dc * fg;
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
    MAAVSS_CHECK_ARG(step >= 0 && step < T && B <= 65535 && H <= 2048, "lstm_step_bwd: bad step/batch");
    dim3 grid(ceil_div(H, 256), B, 2);
    lstm_step_bwd_kernel<<<grid, 256, sizeof(float) * 4 * H, (cudaStream_t)stream>>>(dy, gates, cells, w_hh_fwd, w_hh_rev, dgates,
                                                                                    dcell, B, T, H, step);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
