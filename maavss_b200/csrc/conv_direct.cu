// Register-accumulator ("direct") convolution gradients for the small layers of the fusion encoders
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
// Both return 1 when they handled the call and 0 when the layer is outside their envelope (caller falls back to
// the implicit-GEMM path).
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
constexpr int WG_CO = 4, WG_J = 18, WG_T = 256;

__global__ void __launch_bounds__(WG_T) wgrad_kernel(Geo g, const float* __restrict__ x, const float* __restrict__ dy,
                                                     float* __restrict__ dw, int P, int pos_per_cta, int J) {
    __shared__ int4 jtab[WG_J];                         // {x offset of (ci, tap) or -1, kd, kh, kw}
    __shared__ float red[WG_T / 32][WG_CO * WG_J];
    const maavss_conv_desc& d = g.d;
    const int j0 = blockIdx.y * WG_J, co0 = blockIdx.z * WG_CO;
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
    float acc[WG_CO][WG_J];
#pragma unroll
    for (int u = 0; u < WG_CO; ++u)
#pragma unroll
        for (int jj = 0; jj < WG_J; ++jj) acc[u][jj] = 0.f;
    const int p_begin = blockIdx.x * pos_per_cta, p_end = min(P, p_begin + pos_per_cta);
    for (int p = p_begin + (int)threadIdx.x; p < p_end; p += WG_T) {
        const int n = p / g.O;
        const int o = p - n * g.O;
        const int a = o / g.OHW;
        const int r = o - a * g.OHW;
        const int b = r / d.OW, c = r - b * d.OW;
        const int id0 = a * d.SD - d.PD, ih0 = b * d.SH - d.PH, iw0 = c * d.SW - d.PW;
        const float* xp = x + ((size_t)n * d.Cin * g.I + id0 * g.HW + ih0 * d.W + iw0);
        float dyv[WG_CO];
#pragma unroll
        for (int u = 0; u < WG_CO; ++u)
            dyv[u] = (co0 + u < d.Cout) ? __ldg(dy + ((size_t)n * d.Cout + co0 + u) * g.O + o) : 0.f;
#pragma unroll
        for (int jj = 0; jj < WG_J; ++jj) {
            const int4 e = jtab[jj];
            float xv = 0.f;
            if (e.x >= 0 && (unsigned)(id0 + e.y) < (unsigned)d.D && (unsigned)(ih0 + e.z) < (unsigned)d.H &&
                (unsigned)(iw0 + e.w) < (unsigned)d.W)
                xv = __ldg(xp + e.x);
#pragma unroll
            for (int u = 0; u < WG_CO; ++u) acc[u][jj] = fmaf(dyv[u], xv, acc[u][jj]);
        }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int u = 0; u < WG_CO; ++u)
#pragma unroll
        for (int jj = 0; jj < WG_J; ++jj) {
            const float s = warp_sum(acc[u][jj]);
            if (lane == 0) red[warp][u * WG_J + jj] = s;
        }
    __syncthreads();
    if (threadIdx.x < WG_CO * WG_J) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < WG_T / 32; ++w) s += red[w][threadIdx.x];
        const int u = threadIdx.x / WG_J, jj = threadIdx.x - u * WG_J;
        if (co0 + u < d.Cout && j0 + jj < J) atomicAdd(dw + (size_t)(co0 + u) * J + j0 + jj, s);
    }
}

}  // namespace direct

int maavss_conv_wgrad_direct(const maavss_conv_desc* d, const float* x, const float* dy, float* dw, int accumulate,
                             cudaStream_t st) {
    direct::Geo g = direct::make_geo(d);
    const long long P = (long long)d->N * g.O;
    const int J = d->Cin * g.KV;
    const int gy = ceil_div(J, direct::WG_J), gz = ceil_div(d->Cout, direct::WG_CO);
    // cost model: every (column block, channel block) pair walks all positions; beyond ~16 M position-passes the
    // tiled implicit GEMM (operand reuse through shared memory) wins
    if (P >= (1LL << 31) || (long long)gy * gz * P > (16LL << 20) || gy > 65535 || gz > 65535) return 0;
    const int sms = maavss_num_sms();
    long long want_x = (4LL * sms + (long long)gy * gz - 1) / ((long long)gy * gz);      // ~4 CTAs per SM overall
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
    direct::wgrad_kernel<<<grid, direct::WG_T, 0, st>>>(g, x, dy, dw, (int)P, (int)ppc, J);
    MAAVSS_LAUNCH_CHECK();
    return 1;
}
