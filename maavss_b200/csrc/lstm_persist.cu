// The recurrent part of the bidirectional LSTM (avse_model.py:542-543) as ONE launch per direction pair instead of one
// launch per time step: lstm.cu's step kernels cost ~11 us per step of which ~8 us is launch / drain / re-reading w_hh.
//
// One thread-block cluster of 8 CTAs per (direction, chunk of 8 batch rows); rows of the batch are independent, so clusters
// never talk to each other (no grid-wide barrier).  CTA r of a cluster owns hidden units [32r, 32r + 32): its 128 rows of
// w_hh (forward) or its 32 columns (backward) -- 128 KB -- live in REGISTERS, 64 floats per thread, for the whole sequence;
// per step the CTA multiplies them with the previous hidden state / next gate gradients of the 8 batch rows (read from shared
// memory as warp-wide broadcasts), finishes its 32 units and writes its slice of the new state into the shared memory of all
// 8 CTAs through DSMEM; one cluster.sync() per step.  H = 256 only (what the reference builds); other sizes keep lstm.cu.
#include <cooperative_groups.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace cg = cooperative_groups;

namespace lp {

constexpr int H = 256, BC = 8, CL = 8, UN = H / CL, ROWS = 4 * UN, THREADS = 512;

__device__ __forceinline__ float sigmoidf_(float v) { return 1.f / (1.f + expf(-v)); }

// gin / gates (2, B, T, 4H), cells (2, B, T, H), y (B, T, 2H); grid = 8 CTAs x 2 directions x ceil(B / 8) chunks.
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(THREADS, 1)
    bilstm_fwd_kernel(const float* __restrict__ gin, const float* __restrict__ w0, const float* __restrict__ w1, float* __restrict__ gates,
                      float* __restrict__ cells, float* __restrict__ y, int B, int T) {
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int cid = blockIdx.x / CL;
    const int dir = cid & 1, b0 = (cid >> 1) * BC;
    const int nb = min(BC, B - b0);
    __shared__ __align__(16) float hs[2][BC][H];          // previous hidden state of the chunk, double buffered
    __shared__ float red[4][ROWS][BC + 1];                // partial dot products of the 4 k-quarters
    __shared__ float gact[ROWS][BC + 1];                  // gate pre-activations of this CTA's units
    const int tid = threadIdx.x, row = tid & (ROWS - 1), kq = tid >> 7;
    const int g = row >> 5, u = row & 31;
    float wr[64];                                         // w_hh[g*H + 32*rank + u][64*kq .. +63]
    {
        const float4* wp = reinterpret_cast<const float4*>((dir ? w1 : w0) + (size_t)(g * H + rank * UN + u) * H + kq * 64);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float4 v = __ldg(wp + i);
            wr[4 * i] = v.x; wr[4 * i + 1] = v.y; wr[4 * i + 2] = v.z; wr[4 * i + 3] = v.w;
        }
    }
    for (int e = tid; e < 2 * BC * H; e += THREADS) (&hs[0][0][0])[e] = 0.f;
    cluster.sync();                                       // every CTA of the cluster is running and its hs is initialised
    float c_prev = 0.f;                                   // cell state of (unit tid & 31, batch row tid >> 5), threads < 256
    for (int s = 0; s < T; ++s) {
        const int t = dir ? T - 1 - s : s;
        const int cur = s & 1;
        // this step's input projections: in flight during the recurrent product (thread handles (row, b) pairs tid, tid + 512)
        float gpre[2];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int e = tid + q * THREADS, r2 = e & (ROWS - 1), b = e >> 7;
            gpre[q] = b < nb ? __ldg(gin + (((size_t)dir * B + b0 + b) * T + t) * 4 * H + (r2 >> 5) * H + rank * UN + (r2 & 31)) : 0.f;
        }
        if (s > 0) {
            float acc[BC];
#pragma unroll
            for (int b = 0; b < BC; ++b) {
                const float4* hp = reinterpret_cast<const float4*>(&hs[cur][b][kq * 64]);
                float a = 0.f;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float4 h = hp[i];               // same address in every lane of the warp: one broadcast
                    a = fmaf(h.x, wr[4 * i], a);
                    a = fmaf(h.y, wr[4 * i + 1], a);
                    a = fmaf(h.z, wr[4 * i + 2], a);
                    a = fmaf(h.w, wr[4 * i + 3], a);
                }
                acc[b] = a;
            }
#pragma unroll
            for (int b = 0; b < BC; ++b) red[kq][row][b] = acc[b];
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int e = tid + q * THREADS, r2 = e & (ROWS - 1), b = e >> 7;
            float v = gpre[q];
            if (s > 0) v += (red[0][r2][b] + red[1][r2][b]) + (red[2][r2][b] + red[3][r2][b]);
            gact[r2][b] = v;
        }
        __syncthreads();
        float o_ig = 0.f, o_fg = 0.f, o_gg = 0.f, o_og = 0.f, o_c = 0.f, o_h = 0.f;
        bool owner = false;
        if (tid < UN * BC) {
            const int u2 = tid & 31, b = tid >> 5;
            if (b < nb) {
                const float ig = sigmoidf_(gact[u2][b]), fg = sigmoidf_(gact[UN + u2][b]);
                const float gg = tanhf(gact[2 * UN + u2][b]), og = sigmoidf_(gact[3 * UN + u2][b]);
                const float c = fmaf(fg, c_prev, ig * gg);
                const float h = og * tanhf(c);
                c_prev = c;
                float* mine = &hs[cur ^ 1][b][rank * UN + u2];
#pragma unroll
                for (int r = 0; r < CL; ++r) *cluster.map_shared_rank(mine, r) = h;
                owner = true;
                o_ig = ig; o_fg = fg; o_gg = gg; o_og = og; o_c = c; o_h = h;
            }
        }
        // The barrier's release orders only what was written BEFORE the arrive: the hidden-state exchange.  The global stores
        // (what the backward pass needs) follow the arrive, so the barrier does not wait for them to reach L2.
        cluster.barrier_arrive();
        if (owner) {
            const int u2 = tid & 31, b = tid >> 5, j = rank * UN + u2;
            const size_t gb = (((size_t)dir * B + b0 + b) * T + t) * 4 * H;
            gates[gb + j] = o_ig;
            gates[gb + H + j] = o_fg;
            gates[gb + 2 * H + j] = o_gg;
            gates[gb + 3 * H + j] = o_og;
            cells[(((size_t)dir * B + b0 + b) * T + t) * H + j] = o_c;
            y[((size_t)(b0 + b) * T + t) * 2 * H + dir * H + j] = o_h;
        }
        cluster.barrier_wait();                           // step s is complete in every CTA's hs[cur ^ 1]; red / gact are free again
    }
}

// Backward: dgates (2, B, T, 4H) out.  Thread (u = tid & 31, ks = tid >> 5) keeps w_hh[64*ks .. +63][32*rank + u].
constexpr int KS = 16;
constexpr size_t BWD_SMEM = sizeof(float) * (2 * BC * 4 * H + KS * UN * (BC + 1));
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(THREADS, 1)
    bilstm_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ gates, const float* __restrict__ cells,
                      const float* __restrict__ w0, const float* __restrict__ w1, float* __restrict__ dgates, int B, int T) {
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) float sm[];
    float* dgn = sm;                                      // [2][BC][4H]: gate gradients of the step processed before this one
    float* red = sm + 2 * BC * 4 * H;                     // [KS][UN][BC + 1]
    const int rank = (int)cluster.block_rank();
    const int cid = blockIdx.x / CL;
    const int dir = cid & 1, b0 = (cid >> 1) * BC;
    const int nb = min(BC, B - b0);
    const int tid = threadIdx.x, u = tid & 31, ks = tid >> 5;
    float wr[64];
    {
        const float* wp = (dir ? w1 : w0) + (size_t)(ks * 64) * H + rank * UN + u;
#pragma unroll
        for (int i = 0; i < 64; ++i) wr[i] = __ldg(wp + (size_t)i * H);
    }
    for (int e = tid; e < 2 * BC * 4 * H; e += THREADS) dgn[e] = 0.f;
    cluster.sync();
    float dc = 0.f;                                       // running cell gradient of (unit tid & 31, batch row tid >> 5), threads < 256
    for (int s = T - 1; s >= 0; --s) {
        const int t = dir ? T - 1 - s : s;
        const int tp = dir ? t + 1 : t - 1;               // time index of step s - 1
        const bool has_next = s + 1 < T;
        const int cur = s & 1;
        // this step's saved activations: in flight during the product
        const bool owner = tid < UN * BC && (tid >> 5) < nb;
        float p_dy = 0.f, ig = 0.f, fg = 0.f, gg = 0.f, og = 0.f, c = 0.f, cp = 0.f;
        size_t gb = 0;
        const int j = rank * UN + (tid & 31);
        if (owner) {
            const int b = tid >> 5;
            gb = (((size_t)dir * B + b0 + b) * T + t) * 4 * H;
            p_dy = __ldg(dy + ((size_t)(b0 + b) * T + t) * 2 * H + dir * H + j);
            ig = __ldg(gates + gb + j); fg = __ldg(gates + gb + H + j); gg = __ldg(gates + gb + 2 * H + j); og = __ldg(gates + gb + 3 * H + j);
            c = __ldg(cells + (((size_t)dir * B + b0 + b) * T + t) * H + j);
            cp = s > 0 ? __ldg(cells + (((size_t)dir * B + b0 + b) * T + tp) * H + j) : 0.f;
        }
        if (has_next) {
#pragma unroll
            for (int b = 0; b < BC; ++b) {
                const float4* dp = reinterpret_cast<const float4*>(dgn + (size_t)(cur * BC + b) * 4 * H + ks * 64);
                float a = 0.f;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float4 d = dp[i];
                    a = fmaf(d.x, wr[4 * i], a);
                    a = fmaf(d.y, wr[4 * i + 1], a);
                    a = fmaf(d.z, wr[4 * i + 2], a);
                    a = fmaf(d.w, wr[4 * i + 3], a);
                }
                red[(ks * UN + u) * (BC + 1) + b] = a;
            }
        }
        __syncthreads();
        float d4[4] = {0.f, 0.f, 0.f, 0.f};
        if (owner) {
            const int u2 = tid & 31, b = tid >> 5;
            float dh = p_dy;
            if (has_next) {
                float sum = 0.f;
#pragma unroll
                for (int q = 0; q < KS; ++q) sum += red[(q * UN + u2) * (BC + 1) + b];
                dh += sum;
            }
            const float tc = tanhf(c);
            const float dcv = fmaf(dh * og, 1.f - tc * tc, dc);
            d4[0] = dcv * gg * ig * (1.f - ig);
            d4[1] = dcv * cp * fg * (1.f - fg);
            d4[2] = dcv * ig * (1.f - gg * gg);
            d4[3] = dh * tc * og * (1.f - og);
            dc = dcv * fg;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float* mine = dgn + (size_t)((cur ^ 1) * BC + b) * 4 * H + q * H + j;
#pragma unroll
                for (int r = 0; r < CL; ++r) *cluster.map_shared_rank(mine, r) = d4[q];
            }
        }
        cluster.barrier_arrive();                         // releases the exchange; the global stores below are not waited for
        if (owner) {
#pragma unroll
            for (int q = 0; q < 4; ++q) dgates[gb + q * H + j] = d4[q];
        }
        cluster.barrier_wait();
    }
}

}  // namespace lp

extern "C" int maavss_bilstm_recurrent_supported(int B, int T, int H) { return H == lp::H && B > 0 && T > 0 && (long long)((B + lp::BC - 1) / lp::BC) * 2 * lp::CL < (1LL << 31); }

extern "C" int maavss_bilstm_recurrent_fwd(const float* gin, const float* w_hh_fwd, const float* w_hh_rev, float* gates, float* cells, float* y,
                                           int B, int T, int H, void* stream) {
    if (B == 0 || T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(gin && w_hh_fwd && w_hh_rev && gates && cells && y, "bilstm_recurrent_fwd: null argument");
    MAAVSS_CHECK_ARG(maavss_bilstm_recurrent_supported(B, T, H), "bilstm_recurrent_fwd: H=%d is not 256 (use maavss_lstm_step_fwd)", H);
    MAAVSS_CHECK_ARG((((uintptr_t)w_hh_fwd | (uintptr_t)w_hh_rev) & 15) == 0, "bilstm_recurrent_fwd: w_hh must be 16-byte aligned");
    const unsigned grid = (unsigned)(((B + lp::BC - 1) / lp::BC) * 2 * lp::CL);
    lp::bilstm_fwd_kernel<<<grid, lp::THREADS, 0, (cudaStream_t)stream>>>(gin, w_hh_fwd, w_hh_rev, gates, cells, y, B, T);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bilstm_recurrent_bwd(const float* dy, const float* gates, const float* cells, const float* w_hh_fwd, const float* w_hh_rev,
                                           float* dgates, int B, int T, int H, void* stream) {
    if (B == 0 || T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dy && gates && cells && w_hh_fwd && w_hh_rev && dgates, "bilstm_recurrent_bwd: null argument");
    MAAVSS_CHECK_ARG(maavss_bilstm_recurrent_supported(B, T, H), "bilstm_recurrent_bwd: H=%d is not 256 (use maavss_lstm_step_bwd)", H);
    static bool attr_set = false;
    if (!attr_set) {
        MAAVSS_CUDA(cudaFuncSetAttribute(lp::bilstm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lp::BWD_SMEM));
        attr_set = true;
    }
    const unsigned grid = (unsigned)(((B + lp::BC - 1) / lp::BC) * 2 * lp::CL);
    lp::bilstm_bwd_kernel<<<grid, lp::THREADS, lp::BWD_SMEM, (cudaStream_t)stream>>>(dy, gates, cells, w_hh_fwd, w_hh_rev, dgates, B, T);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
