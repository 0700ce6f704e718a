// Register-accumulator ("direct") convolutions -- forward, data gradient, weight gradient -- for the small layers of the fusion encoders
// (avse_model.py:427-446,474-502; avse_model_final.py:75-107): few channels, few taps, many positions.
//
// The implicit-GEMM kernels (conv_igemm.cu) pay a gather + two barriers per 16 reduction indices and, for the
// data gradient of a strided layer, multiply 3 zeros out of 4.  For these layers the products are cheap and the
// index work is what costs, so:
//   WGRAD  a thread keeps a CO x J block of dW (CO = 4 output channels, J = 18 (ci, tap) columns) in registers and
//          walks positions; per position it loads 4 dy values and 18 gathered x values for 72 FMAs.  Blocks of dW
//          are spread over blockIdx.y/z, positions over blockIdx.x; one shuffle + shared-memory reduction and one
//          atomicAdd per dW element per CTA at the end.
//   DGRAD  a thread keeps CI input channels of one input position in registers and walks only the taps whose
//          stride phase matches its position (1/(SD*SH*SW) of them), weights read from a shared-memory copy.
//   FWD    (end of file) a thread keeps CO output channels of one output position, weights staged as for DGRAD.
// Each returns 1 when it handled the call and 0 when the layer is outside its envelope (caller falls back to
// the implicit-GEMM path).
#include <stdlib.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace direct {

struct Geo {
    maavss_conv_desc d;
    int KV, KHW, O, OHW, I, HW;
};

static Geo make_geo(const maavss_conv_desc* d) {
    Geo g;
    g.d = *d;
    g.KV = d->KD * d->KH * d->KW;
    g.KHW = d->KH * d->KW;
    g.OHW = d->OH * d->OW;
    g.O = d->OD * g.OHW;
    g.HW = d->H * d->W;
    g.I = d->D * g.HW;
    return g;
}

// ---------------------------------------------------------------------------------------------- WGRAD
constexpr int WG_J = 18, WG_T = 256;

// CO = output channels per thread block of dW: 4, or 2 for the C1->2 layers (half the accumulators: one more CTA per SM).
// NQ positions per loop trip (2 for CO = 2; 1 for CO = 4, whose 72 accumulators leave no registers), all their loads issued before the FMAs: the loop is one memory latency per trip otherwise
// (measured ~1 us per trip at 16 warps per SM), and the grid is sized to ONE wave of resident CTAs by the launcher -- 512 CTAs
// at 2 per SM ran as two waves, which was most of the 33 us of the C1->2 layer.
template <int CO, int NQ>
__global__ void __launch_bounds__(WG_T, 2) wgrad_kernel(Geo g, const float* __restrict__ x, const float* __restrict__ dy,
                                                     float* __restrict__ dw, int P, int pos_per_cta, int J) {
    __shared__ int4 jtab[WG_J];                         // {x offset of (ci, tap) or -1, kd, kh, kw}
    __shared__ float red[WG_T / 32][CO * WG_J];
    const maavss_conv_desc& d = g.d;
    const int j0 = blockIdx.y * WG_J, co0 = blockIdx.z * CO;
    if (threadIdx.x < WG_J) {
        const int j = j0 + threadIdx.x;
        int4 e = make_int4(-1, 0, 0, 0);
        if (j < J) {
            const int ci = j / g.KV, tap = j - ci * g.KV;
            const int kd = tap / g.KHW, r = tap - kd * g.KHW;
            const int kh = r / d.KW, kw = r - kh * d.KW;
            e = make_int4(ci * g.I + kd * g.HW + kh * d.W + kw, kd, kh, kw);
        }
        jtab[threadIdx.x] = e;
    }
    __syncthreads();
    float acc[CO][WG_J];
#pragma unroll
    for (int u = 0; u < CO; ++u)
#pragma unroll
        for (int jj = 0; jj < WG_J; ++jj) acc[u][jj] = 0.f;
    const int p_begin = blockIdx.x * pos_per_cta, p_end = min(P, p_begin + pos_per_cta);
    for (int p0 = p_begin + (int)threadIdx.x; p0 < p_end; p0 += NQ * WG_T) {
        float dyv[NQ][CO], xv[NQ][WG_J];
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const int p = p0 + q * WG_T;
            const bool live = p < p_end;
            const int pp = live ? p : p0;
            const int n = pp / g.O;
            const int o = pp - n * g.O;
            const int a = o / g.OHW;
            const int r = o - a * g.OHW;
            const int b = r / d.OW, c = r - b * d.OW;
            const int id0 = a * d.SD - d.PD, ih0 = b * d.SH - d.PH, iw0 = c * d.SW - d.PW;
            const float* xp = x + ((size_t)n * d.Cin * g.I + id0 * g.HW + ih0 * d.W + iw0);
#pragma unroll
            for (int u = 0; u < CO; ++u)
                dyv[q][u] = (live && co0 + u < d.Cout) ? __ldg(dy + ((size_t)n * d.Cout + co0 + u) * g.O + o) : 0.f;
#pragma unroll
            for (int jj = 0; jj < WG_J; ++jj) {
                const int4 e = jtab[jj];
                float v = 0.f;
                if (live && e.x >= 0 && (unsigned)(id0 + e.y) < (unsigned)d.D && (unsigned)(ih0 + e.z) < (unsigned)d.H &&
                    (unsigned)(iw0 + e.w) < (unsigned)d.W)
                    v = __ldg(xp + e.x);
                xv[q][jj] = v;
            }
        }
#pragma unroll
        for (int q = 0; q < NQ; ++q)
#pragma unroll
            for (int jj = 0; jj < WG_J; ++jj)
#pragma unroll
                for (int u = 0; u < CO; ++u) acc[u][jj] = fmaf(dyv[q][u], xv[q][jj], acc[u][jj]);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int u = 0; u < CO; ++u)
#pragma unroll
        for (int jj = 0; jj < WG_J; ++jj) {
            const float s = warp_sum(acc[u][jj]);
            if (lane == 0) red[warp][u * WG_J + jj] = s;
        }
    __syncthreads();
    if (threadIdx.x < CO * WG_J) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < WG_T / 32; ++w) s += red[w][threadIdx.x];
        const int u = threadIdx.x / WG_J, jj = threadIdx.x - u * WG_J;
        if (co0 + u < d.Cout && j0 + jj < J) atomicAdd(dw + (size_t)(co0 + u) * J + j0 + jj, s);
    }
}

template <int CO, int NQ>
static int resident_ctas_per_sm() {
    static int n = 0;
    if (n == 0) {
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, wgrad_kernel<CO, NQ>, WG_T, 0) != cudaSuccess || n < 1) n = 2;
    }
    return n;
}

}  // namespace direct

int maavss_conv_wgrad_direct(const maavss_conv_desc* d, const float* x, const float* dy, float* dw, int accumulate,
                             cudaStream_t st) {
    direct::Geo g = direct::make_geo(d);
    const long long P = (long long)d->N * g.O;
    const int J = d->Cin * g.KV;
    const int CO = d->Cout <= 2 ? 2 : 4;
    const int gy = ceil_div(J, direct::WG_J), gz = ceil_div(d->Cout, CO);
    // envelope (measured on B200, bench.py --breakdown): every (column block, channel block) pair walks all positions
    // and re-reads dy / x, so the kernel wins for the thin layers (<= 24 block pairs: C1->2 ... C8->16 with 1x9 taps,
    // C2->8 with 5x5) by 4-30x and loses 20-30 % to the tiled implicit GEMM beyond that
    if (P >= (1LL << 31) || (long long)gy * gz > 24 || (long long)gy * gz * P > (16LL << 20)) return 0;
    const int sms = maavss_num_sms();
    const int occ = CO == 2 ? direct::resident_ctas_per_sm<2, 2>() : direct::resident_ctas_per_sm<4, 1>();
    static const int waves_pct = [] {                    // tuning knob (benchmarks only): MAAVSS_WGRAD_WAVE_PCT, % of one resident wave
        const char* e = getenv("MAAVSS_WGRAD_WAVE_PCT");
        return e ? atoi(e) : 100;
    }();
    long long want_x = ((long long)occ * sms * waves_pct / 100) / ((long long)gy * gz);   // one wave of resident CTAs overall
    if (want_x < 1) want_x = 1;
    long long ppc = (P + want_x - 1) / want_x;
    if (ppc < 8 * direct::WG_T) ppc = 8 * direct::WG_T;                                   // >= 8 positions per thread
    ppc = (ppc + direct::WG_T - 1) / direct::WG_T * direct::WG_T;
    if (!accumulate) {
        cudaError_t e = cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)d->Cout * J, st);
        if (e != cudaSuccess) {
            maavss_set_error("conv_wgrad: %s", cudaGetErrorString(e));
            return MAAVSS_ERR_CUDA;
        }
    }
    dim3 grid(ceil_div(P, ppc), gy, gz);
    if (CO == 2) direct::wgrad_kernel<2, 2><<<grid, direct::WG_T, 0, st>>>(g, x, dy, dw, (int)P, (int)ppc, J);
    else direct::wgrad_kernel<4, 1><<<grid, direct::WG_T, 0, st>>>(g, x, dy, dw, (int)P, (int)ppc, J);
    MAAVSS_LAUNCH_CHECK();
    return 1;
}

// ---------------------------------------------------------------------------------------------- DGRAD
namespace direct {

constexpr int DG_T = 256, DG_WS = 8192;     // threads (= input positions) per CTA, floats of staged weights

__device__ __forceinline__ int div_small(int t, int S) { return S == 1 ? t : (S == 2 ? t >> 1 : t / S); }

// dx[n, ci, q] = sum_{co} sum_{o : q + pad - o*stride is a tap} dy[n, co, o] * w[co, ci, tap]  (+ bias[ci]).
// A thread owns one input position q and CI_T input channels.  The output positions that reach q form a small
// box [od_min..od_max] x [oh_min..oh_max] x [ow_min..ow_max] computed once per thread, so the inner loops run over
// contributing taps only (1/(SD*SH*SW) of the kernel volume for a strided layer) with no divisibility tests.
// Weights are staged per group of output channels as ws[co][tap][CI_T] so a tap is one LDS.128 per 4 channels.
template <int CI_T, int SW_T>   // SW_T: compile-time stride along W (0 = runtime)
__global__ void __launch_bounds__(DG_T, 4) dgrad_kernel(Geo g, const float* __restrict__ dy, const float* __restrict__ w,
                                                     const float* __restrict__ bias, float* __restrict__ dx, int Q,
                                                     int co_chunk) {
    __shared__ __align__(16) float ws[DG_WS];
    const maavss_conv_desc& d = g.d;
    const int ci0 = blockIdx.y * CI_T;
    // Threads are dealt to the CTA's 256 consecutive positions so that a warp holds positions of ONE residue mod SW: its
    // lanes then walk the same kernel taps in lock step (the staged-weight LDS.128 is a broadcast instead of SW..KW
    // distinct addresses) and read consecutive dy elements.
    int tl = threadIdx.x;
    if (d.SW > 1 && DG_T % d.SW == 0) tl = (tl % (DG_T / d.SW)) * d.SW + tl / (DG_T / d.SW);
    const int q = blockIdx.x * DG_T + tl;
    const bool valid = q < Q;
    const int qq = valid ? q : 0;
    const int n = qq / g.I;
    const int pos = qq - n * g.I;
    const int id = pos / g.HW;
    const int r = pos - id * g.HW;
    const int ih = r / d.W, iw = r - ih * d.W;
    const int td = id + d.PD, th = ih + d.PH, tw = iw + d.PW;
    const int od_lo = (td - d.KD + 1 <= 0) ? 0 : div_small(td - d.KD + d.SD, d.SD), od_hi = min(d.OD - 1, div_small(td, d.SD));
    const int oh_lo = (th - d.KH + 1 <= 0) ? 0 : div_small(th - d.KH + d.SH, d.SH), oh_hi = min(d.OH - 1, div_small(th, d.SH));
    const int ow_lo = (tw - d.KW + 1 <= 0) ? 0 : div_small(tw - d.KW + d.SW, d.SW), ow_hi = min(d.OW - 1, div_small(tw, d.SW));
    float acc[CI_T];
#pragma unroll
    for (int c = 0; c < CI_T; ++c) acc[c] = 0.f;
    for (int co0 = 0; co0 < d.Cout; co0 += co_chunk) {
        const int cc = min(co_chunk, d.Cout - co0);
        __syncthreads();
        for (int e = threadIdx.x; e < cc * g.KV * CI_T; e += DG_T) {
            const int c = e % CI_T, rr = e / CI_T;
            const int tap = rr % g.KV, col = rr / g.KV;
            const int ci = ci0 + c;
            ws[e] = (ci < d.Cin) ? __ldg(w + ((size_t)(co0 + col) * d.Cin + ci) * g.KV + tap) : 0.f;
        }
        __syncthreads();
        if (!valid) continue;
        // (od, oh) outside, output channels inside: the row pointers advance by constants per channel and the tap loop
        // (<= 5 contributing taps along W for the 1x9 / 5x5 stride-2 layers) is unrolled with loop-invariant predicates
        const int wstep = (SW_T ? SW_T : d.SW) * CI_T, nw = ow_hi - ow_lo + 1;
        for (int od = od_lo; od <= od_hi; ++od) {
            const int kd = td - od * d.SD;
            for (int oh = oh_lo; oh <= oh_hi; ++oh) {
                const int kh = th - oh * d.SH;
                const float* dyr = dy + ((size_t)n * d.Cout + co0) * g.O + od * g.OHW + oh * d.OW + ow_lo;
                const float* wr = ws + ((kd * d.KH + kh) * d.KW + tw) * CI_T - ow_lo * wstep;
                for (int col = 0; col < cc; ++col, dyr += g.O, wr += g.KV * CI_T) {
#pragma unroll
                    for (int k = 0; k < 5; ++k) {
                        if (k < nw) {
                            const float dyv = __ldg(dyr + k);
                            const float4* w4 = reinterpret_cast<const float4*>(wr - k * wstep);
#pragma unroll
                            for (int c4 = 0; c4 < CI_T / 4; ++c4) {
                                const float4 wv = w4[c4];
                                acc[4 * c4 + 0] = fmaf(dyv, wv.x, acc[4 * c4 + 0]);
                                acc[4 * c4 + 1] = fmaf(dyv, wv.y, acc[4 * c4 + 1]);
                                acc[4 * c4 + 2] = fmaf(dyv, wv.z, acc[4 * c4 + 2]);
                                acc[4 * c4 + 3] = fmaf(dyv, wv.w, acc[4 * c4 + 3]);
                            }
                        }
                    }
                    for (int k = 5; k < nw; ++k) {
                        const float dyv = __ldg(dyr + k);
                        const float4* w4 = reinterpret_cast<const float4*>(wr - k * wstep);
#pragma unroll
                        for (int c4 = 0; c4 < CI_T / 4; ++c4) {
                            const float4 wv = w4[c4];
                            acc[4 * c4 + 0] = fmaf(dyv, wv.x, acc[4 * c4 + 0]);
                            acc[4 * c4 + 1] = fmaf(dyv, wv.y, acc[4 * c4 + 1]);
                            acc[4 * c4 + 2] = fmaf(dyv, wv.z, acc[4 * c4 + 2]);
                            acc[4 * c4 + 3] = fmaf(dyv, wv.w, acc[4 * c4 + 3]);
                        }
                    }
                }
            }
        }
    }
    if (!valid) return;
#pragma unroll
    for (int c = 0; c < CI_T; ++c) {
        const int ci = ci0 + c;
        if (ci < d.Cin) dx[((size_t)n * d.Cin + ci) * g.I + pos] = acc[c] + (bias ? __ldg(bias + ci) : 0.f);
    }
}

}  // namespace direct

int maavss_conv_dgrad_direct(const maavss_conv_desc* d, const float* dy, const float* w, const float* bias, float* dx,
                             cudaStream_t st) {
    direct::Geo g = direct::make_geo(d);
    const long long Q = (long long)d->N * g.I;
    // channels per thread: 16 when that still gives every SM a CTA, fewer for the deep 8x8 / 8x4 layers
    int CI_T = d->Cin >= 16 ? 16 : (d->Cin > 4 ? 8 : 4);
    const int sms = maavss_num_sms();
    static const int cta_target = [] {                   // tuning knob (benchmarks only): MAAVSS_DGRAD_CTA_TARGET, in SMs
        const char* e = getenv("MAAVSS_DGRAD_CTA_TARGET");
        return e ? atoi(e) : 1;
    }();
    while (CI_T > 4 && (long long)ceil_div(Q, direct::DG_T) * ceil_div(d->Cin, CI_T) < (long long)cta_target * sms) CI_T /= 2;
    const int co_chunk = min(d->Cout, direct::DG_WS / (g.KV * CI_T));
    // envelope: the staged weight block must fit, and the layer must be small enough that re-reading dy per group of
    // CI_T input channels (no shared-memory reuse of dy) is cheaper than the implicit GEMM's index work
    const double fma = (double)Q * d->Cin * d->Cout * g.KV / ((double)d->SD * d->SH * d->SW);
    static const double fma_max = [] {                   // tuning knob (benchmarks only): MAAVSS_DGRAD_DIRECT_MAX_FMA
        const char* e = getenv("MAAVSS_DGRAD_DIRECT_MAX_FMA");
        return e ? atof(e) : 8e9;
    }();
    if (co_chunk < 1 || Q >= (1LL << 31) || fma > fma_max || ceil_div(d->Cin, CI_T) > 65535) return 0;
    dim3 grid(ceil_div(Q, direct::DG_T), ceil_div(d->Cin, CI_T));
#define DG_LAUNCH(C_, S_) direct::dgrad_kernel<C_, S_><<<grid, direct::DG_T, 0, st>>>(g, dy, w, bias, dx, (int)Q, co_chunk)
#define DG_STRIDE(C_)                        \
    do {                                     \
        if (d->SW == 1) DG_LAUNCH(C_, 1);    \
        else if (d->SW == 2) DG_LAUNCH(C_, 2); \
        else DG_LAUNCH(C_, 0);               \
    } while (0)
    if (CI_T == 16) DG_STRIDE(16);
    else if (CI_T == 8) DG_STRIDE(8);
    else DG_STRIDE(4);
#undef DG_STRIDE
#undef DG_LAUNCH
    MAAVSS_LAUNCH_CHECK();
    return 1;
}

// ---------------------------------------------------------------------------------------------- FWD
namespace direct {

// y[n, co, p] = bias[co] + sum_{ci, tap} x[n, ci, p*stride - pad + tap] * w[co, ci, tap].
// A thread owns one output position p and CO_T output channels.  Weights are staged per group of input channels as
// ws[ci][tap][CO_T] (one LDS.128 broadcast per 4 channels and tap); x is read straight from global memory (consecutive
// lanes read addresses SW apart, neighbouring taps hit the same lines in L1).  (kd, kh) rows run outside the channel
// loop so that the row pointers advance by constants; the taps along W are unrolled with loop-invariant predicates.
template <int CO_T, int KW_T>   // KW_T: compile-time kernel width (0 = runtime)
__global__ void __launch_bounds__(DG_T, 4) fwd_kernel(Geo g, const float* __restrict__ x, const float* __restrict__ w,
                                                      const float* __restrict__ bias, float* __restrict__ y, int P, int ci_chunk) {
    __shared__ __align__(16) float ws[DG_WS];
    const maavss_conv_desc& d = g.d;
    const int KW = KW_T ? KW_T : d.KW;
    const int co0 = blockIdx.y * CO_T;
    const int p = blockIdx.x * DG_T + threadIdx.x;
    const bool valid = p < P;
    const int pp = valid ? p : 0;
    const int n = pp / g.O;
    const int pos = pp - n * g.O;
    const int od = pos / g.OHW;
    const int r = pos - od * g.OHW;
    const int oh = r / d.OW, ow = r - oh * d.OW;
    const int id0 = od * d.SD - d.PD, ih0 = oh * d.SH - d.PH, iw0 = ow * d.SW - d.PW;
    const int kd_lo = max(0, -id0), kd_hi = min(d.KD - 1, d.D - 1 - id0);
    const int kh_lo = max(0, -ih0), kh_hi = min(d.KH - 1, d.H - 1 - ih0);
    const int kw_lo = max(0, -iw0), kw_hi = min(KW - 1, d.W - 1 - iw0);
    float acc[CO_T];
#pragma unroll
    for (int c = 0; c < CO_T; ++c) acc[c] = 0.f;
    for (int ci0 = 0; ci0 < d.Cin; ci0 += ci_chunk) {
        const int cc = min(ci_chunk, d.Cin - ci0);
        __syncthreads();
        for (int e = threadIdx.x; e < cc * g.KV * CO_T; e += DG_T) {
            const int c = e % CO_T, rr = e / CO_T;
            const int tap = rr % g.KV, ci = rr / g.KV;
            const int co = co0 + c;
            ws[e] = (co < d.Cout) ? __ldg(w + ((size_t)co * d.Cin + ci0 + ci) * g.KV + tap) : 0.f;
        }
        __syncthreads();
        if (!valid) continue;
        for (int kd = kd_lo; kd <= kd_hi; ++kd) {
            for (int kh = kh_lo; kh <= kh_hi; ++kh) {
                const float* xr = x + ((size_t)n * d.Cin + ci0) * g.I + (size_t)(id0 + kd) * g.HW + (ih0 + kh) * d.W + iw0;
                const float* wr = ws + ((kd * d.KH + kh) * KW) * CO_T;
                for (int ci = 0; ci < cc; ++ci, xr += g.I, wr += g.KV * CO_T) {
#pragma unroll
                    for (int k = 0; k < (KW_T ? KW_T : 9); ++k) {
                        if (k >= kw_lo && k <= kw_hi) {
                            const float xv = __ldg(xr + k);
                            const float4* w4 = reinterpret_cast<const float4*>(wr + k * CO_T);
#pragma unroll
                            for (int c4 = 0; c4 < CO_T / 4; ++c4) {
                                const float4 wv = w4[c4];
                                acc[4 * c4 + 0] = fmaf(xv, wv.x, acc[4 * c4 + 0]);
                                acc[4 * c4 + 1] = fmaf(xv, wv.y, acc[4 * c4 + 1]);
                                acc[4 * c4 + 2] = fmaf(xv, wv.z, acc[4 * c4 + 2]);
                                acc[4 * c4 + 3] = fmaf(xv, wv.w, acc[4 * c4 + 3]);
                            }
                        }
                    }
                    if (!KW_T) {
                        for (int k = max(9, kw_lo); k <= kw_hi; ++k) {
                            const float xv = __ldg(xr + k);
                            const float4* w4 = reinterpret_cast<const float4*>(wr + k * CO_T);
#pragma unroll
                            for (int c4 = 0; c4 < CO_T / 4; ++c4) {
                                const float4 wv = w4[c4];
                                acc[4 * c4 + 0] = fmaf(xv, wv.x, acc[4 * c4 + 0]);
                                acc[4 * c4 + 1] = fmaf(xv, wv.y, acc[4 * c4 + 1]);
                                acc[4 * c4 + 2] = fmaf(xv, wv.z, acc[4 * c4 + 2]);
                                acc[4 * c4 + 3] = fmaf(xv, wv.w, acc[4 * c4 + 3]);
                            }
                        }
                    }
                }
            }
        }
    }
    if (!valid) return;
#pragma unroll
    for (int c = 0; c < CO_T; ++c) {
        const int co = co0 + c;
        if (co < d.Cout) y[((size_t)n * d.Cout + co) * g.O + pos] = acc[c] + (bias ? __ldg(bias + co) : 0.f);
    }
}

}  // namespace direct

int maavss_conv_fwd_direct(const maavss_conv_desc* d, const float* x, const float* w, const float* bias, float* y, cudaStream_t st) {
    direct::Geo g = direct::make_geo(d);
    const long long P = (long long)d->N * g.O;
    int CO_T = d->Cout >= 16 ? 16 : (d->Cout > 4 ? 8 : 4);
    const int sms = maavss_num_sms();
    while (CO_T > 4 && (long long)ceil_div(P, direct::DG_T) * ceil_div(d->Cout, CO_T) < sms) CO_T /= 2;
    const int ci_chunk = min(d->Cin, direct::DG_WS / (g.KV * CO_T));
    // envelope (bench.py --breakdown on B200, profiles/r1_step_breakdown_*.txt): layers up to 5e8 FMAs with >= 4096
    // output positions run 5-40 % faster here (thin layers 1.5-2x); the implicit GEMM's shared-memory reuse of x wins
    // beyond that and for the 8x4 / 8x8 maps whose 2048 positions leave most SMs without a CTA
    const double fma = (double)P * d->Cin * d->Cout * g.KV;
    static const double fma_max = [] {                   // tuning knob (benchmarks only): MAAVSS_FWD_DIRECT_MAX_FMA
        const char* e = getenv("MAAVSS_FWD_DIRECT_MAX_FMA");
        return e ? atof(e) : 5e8;
    }();
    // (a single group of output channels reads x exactly once: any size, e.g. the C1->16 first Conv3d of the Frames model)
    const bool single_group = ceil_div(d->Cout, CO_T) == 1 && d->Cin <= 4;
    if (ci_chunk < 1 || P >= (1LL << 31) || P < 4096 || (fma > fma_max && !single_group) || ceil_div(d->Cout, CO_T) > 65535) return 0;
    dim3 grid(ceil_div(P, direct::DG_T), ceil_div(d->Cout, CO_T));
#define FW_LAUNCH(C_, K_) direct::fwd_kernel<C_, K_><<<grid, direct::DG_T, 0, st>>>(g, x, w, bias, y, (int)P, ci_chunk)
#define FW_WIDTH(C_)                           \
    do {                                       \
        if (d->KW == 9) FW_LAUNCH(C_, 9);      \
        else if (d->KW == 5) FW_LAUNCH(C_, 5); \
        else FW_LAUNCH(C_, 0);                 \
    } while (0)
    if (CO_T == 16) FW_WIDTH(16);
    else if (CO_T == 8) FW_WIDTH(8);
    else FW_WIDTH(4);
#undef FW_WIDTH
#undef FW_LAUNCH
    MAAVSS_LAUNCH_CHECK();
    return 1;
}
