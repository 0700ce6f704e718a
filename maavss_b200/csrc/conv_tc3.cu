// fp32-accurate tensor-core path for the mid / deep Conv2d layers of the fusion encoders (avse_model.py:427-446,474-502:
// Cin >= 8, 1x9 stride (1,2) and 5x5 stride (2,2)/(1,2) kernels): the three products of a layer are GEMMs on the 3xTF32
// tcgen05 kernel (tc_gemm3.cu) over an explicit im2col matrix.
//
//   forward   col[pos][ci*taps + tap] = x gathered (zero padding)            im2col_kernel        (HBM / L2 streaming)
//             y[n][co][o] = sum_k col[pos][k] w[co][k] + bias[co]            gemm, blocked transposed output = NCHW
//   dgrad     dyr[pos][co] = dy[n][co][o]                                    rows_kernel (32 x 32 smem transpose)
//             dcol[pos][tap*Cin + ci] = sum_co dyr[pos][co] w[co][ci][tap]   gemm against the tap-major copy of w (permute_w_kernel)
//             dx[n][ci][q] = sum over the taps that reach q of dcol          col2im_kernel (coalesced gather, deterministic)
//   wgrad     dw[co][k] (+)= sum_pos dyr[pos][co] col[pos][k]                gemm, both operands MN-major, split over pos
//
// The im2col matrix (<= 38 MB per layer at the headline shapes) is written once in the forward, stays mostly L2-resident for
// the GEMM that follows, and is kept for the weight gradient.  The SIMT kernels (conv_direct.cu / conv_igemm.cu) did these
// layers at ~10 TFLOP/s; here the GEMMs are tensor-pipe work and the layer is bound by the col / dcol streams.
#include <stdint.h>
#include <stdlib.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace tc3 {

struct Geo {
    int N, Cin, H, W, Cout, KH, KW, SH, SW, PH, PW, OH, OW;
    int taps, Kc, OHW, I;
    long long pos;
};

static Geo make_geo(const maavss_conv_desc* d) {
    Geo g;
    g.N = d->N; g.Cin = d->Cin; g.H = d->H; g.W = d->W; g.Cout = d->Cout; g.KH = d->KH; g.KW = d->KW;
    g.SH = d->SH; g.SW = d->SW; g.PH = d->PH; g.PW = d->PW; g.OH = d->OH; g.OW = d->OW;
    g.taps = g.KH * g.KW;
    g.Kc = g.Cin * g.taps;
    g.OHW = g.OH * g.OW;
    g.I = g.H * g.W;
    g.pos = (long long)g.N * g.OHW;
    return g;
}

// One thread per float4 of col: 4 consecutive k of one output position.
__global__ void __launch_bounds__(256) im2col_kernel(Geo g, const float* __restrict__ x, float* __restrict__ col, long long total4) {
    const int K4 = g.Kc >> 2;
    for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < total4; e += (long long)gridDim.x * 256) {
        const long long pos = e / K4;
        const int k = (int)(e - pos * K4) << 2;
        const int n = (int)(pos / g.OHW);
        const int r = (int)(pos - (long long)n * g.OHW);
        const int oh = r / g.OW, ow = r - oh * g.OW;
        int ci = k / g.taps;
        int tap = k - ci * g.taps;
        int kh = tap / g.KW, kw = tap - kh * g.KW;
        const int ih0 = oh * g.SH - g.PH, iw0 = ow * g.SW - g.PW;
        float v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int ih = ih0 + kh, iw = iw0 + kw;
            v[u] = (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W) ? __ldg(x + (((size_t)n * g.Cin + ci) * g.H + ih) * g.W + iw) : 0.f;
            if (++kw == g.KW) {
                kw = 0;
                if (++kh == g.KH) {
                    kh = 0;
                    ++ci;
                }
            }
        }
        reinterpret_cast<float4*>(col)[e] = make_float4(v[0], v[1], v[2], v[3]);
    }
}

// Staged variant (the one used when the input window of a tile fits shared memory): CTA = (n, oh, OWT consecutive ow).  The
// Cin x KH x span input window (span = (OWT-1)*SW + KW, zero padding materialised) is staged with coalesced loads, offs[k] =
// window offset of reduction index k is tabulated once per CTA, and every float4 of col costs one LDS.128 of offsets, four
// LDS of values and one coalesced STG.128 -- the plain kernel spends ~100 instructions per float4 on index divisions
// (measured 27 us for 37.7 MB; instruction bound).
__global__ void __launch_bounds__(256) im2col_staged_kernel(Geo g, const float* __restrict__ x, float* __restrict__ col, int OWT, int span) {
    extern __shared__ __align__(16) float sm[];
    int* offs = reinterpret_cast<int*>(sm);                 // [Kc]
    float* win = sm + g.Kc;                                 // [Cin][KH][span]
    const int tiles_w = (g.OW + OWT - 1) / OWT;
    const int tw = blockIdx.x % tiles_w, oh = (blockIdx.x / tiles_w) % g.OH, n = blockIdx.x / (tiles_w * g.OH);
    const int ow0 = tw * OWT, nw = min(OWT, g.OW - ow0);
    const int ih0 = oh * g.SH - g.PH, iw0 = ow0 * g.SW - g.PW;
    const int rows = g.Cin * g.KH;
    for (int k = threadIdx.x; k < g.Kc; k += 256) {
        const int ci = k / g.taps, tap = k - ci * g.taps;
        const int kh = tap / g.KW, kw = tap - kh * g.KW;
        offs[k] = (ci * g.KH + kh) * span + kw;
    }
    // (flat loop on purpose: a warp-per-row walk without the divisions measured SLOWER -- st4 forward 48 vs 35 us -- short rows
    // leave lanes idle and serialise the load latency per row)
    const float inv_span = 1.0f / (float)span, inv_kh = 1.0f / (float)g.KH;
    for (int e = threadIdx.x; e < rows * span; e += 256) {
        // e / span and r / KH through the fp32 reciprocal + one correction step (exact: e < 2^16): the integer divisions were
        // ~40 of the ~50 issue slots per staged element
        int r = __float2int_rz((float)e * inv_span);
        r += (r + 1) * span <= e;
        r -= r * span > e;
        const int j = e - r * span;
        int ci = __float2int_rz((float)r * inv_kh);
        ci += (ci + 1) * g.KH <= r;
        ci -= ci * g.KH > r;
        const int kh = r - ci * g.KH;
        const int ih = ih0 + kh, iw = iw0 + j;
        win[e] = (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W) ? __ldg(x + (((size_t)n * g.Cin + ci) * g.H + ih) * g.W + iw) : 0.f;
    }
    __syncthreads();
    const int K4 = g.Kc >> 2;
    float4* dst = reinterpret_cast<float4*>(col + ((size_t)(n * g.OH + oh) * g.OW + ow0) * g.Kc);
    const int total = nw * K4;
    // (position, quad) pairs dealt to the threads in order: consecutive threads write consecutive float4 of col
    int pos = threadIdx.x / K4, kq = threadIdx.x - pos * K4;
    const int dpos = 256 / K4, dkq = 256 - dpos * K4;
    for (int e = threadIdx.x; e < total; e += 256) {
        const int4 o = *reinterpret_cast<const int4*>(offs + 4 * kq);
        const float* w0 = win + pos * g.SW;
        dst[e] = make_float4(w0[o.x], w0[o.y], w0[o.z], w0[o.w]);
        kq += dkq;
        pos += dpos;
        if (kq >= K4) {
            kq -= K4;
            ++pos;
        }
    }
}

// w_perm[co][tap * Cin + ci] = w[co][ci * taps + tap]: the data-gradient product writes dcol with the channel index fastest,
// so that col2im reads 128 contiguous bytes per (position, tap).
__global__ void __launch_bounds__(256) permute_w_kernel(const float* __restrict__ w, float* __restrict__ wp, int Cout, int Cin, int taps) {
    const int Kc = Cin * taps, total = Cout * Kc;
    for (int e = blockIdx.x * 256 + threadIdx.x; e < total; e += gridDim.x * 256) {
        const int co = e / Kc, k = e - co * Kc;
        const int tap = k / Cin, ci = k - tap * Cin;
        wp[e] = __ldg(w + (size_t)co * Kc + ci * taps + tap);
    }
}

// out[(n * inner + q) * C + c] = in[(n * C + c) * inner + q]; grid (inner tiles, C tiles, N), 32 x 32 tile through shared memory.
__global__ void __launch_bounds__(256) rows_kernel(const float* __restrict__ in, float* __restrict__ out, int C, int inner) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z, q0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
        const int c = c0 + r, q = q0 + tx;
        tile[r][tx] = (c < C && q < inner) ? __ldg(in + ((size_t)n * C + c) * inner + q) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
        const int q = q0 + r, c = c0 + tx;
        if (q < inner && c < C) out[((size_t)n * inner + q) * C + c] = tile[tx][r];
    }
}

// dx[n][ci][q] = sum over (kh, kw) with (ih + PH - kh) % SH == 0, (iw + PW - kw) % SW == 0 of dcol[(n, oh, ow)][tap*Cin + ci]
// (dcol from the product with the permuted weights: channel index fastest).  CTA = 32 consecutive input positions q of one
// sample x CL input channels; lanes run along ci, so every load of a warp is one contiguous 64 / 128-byte piece of a dcol
// row; the sums meet in a shared-memory tile and leave coalesced along q.  Deterministic (a gather, no atomics).
// SHT / SWT: compile-time strides (1 or 2; 0 = read them from g).  ncu on the first version: 73 % issue-active, ~200
// instructions per output element, most of them runtime divisions by the strides and by W -- here the strides are shifts and
// (ih, iw) of the thread's positions advance by constants.
template <int CL, int SHT, int SWT>
__global__ void __launch_bounds__(256) col2im_kernel(Geo g, const float* __restrict__ dcol, float* __restrict__ dx) {
    __shared__ float tile[CL][33];
    constexpr int JP = 256 / CL, NJ = 32 / JP;
    const int SH = SHT ? SHT : g.SH, SW = SWT ? SWT : g.SW;
    const int c = threadIdx.x % CL, j0 = threadIdx.x / CL;
    const int n = blockIdx.z, q0 = blockIdx.x * 32, cb = blockIdx.y * CL;
    const int ci = cb + c;
    const float* base = dcol + (size_t)n * g.OHW * g.Kc + ci;
    const int step = g.Kc - SW * g.Cin;                                     // ow + 1: next dcol row, kw - SW
    const long long rstep = (long long)g.OW * g.Kc - (long long)SH * g.KW * g.Cin;      // oh + 1: next output row, kh - SH
    int q = q0 + j0;
    int ih = q / g.W, iw = q - ih * g.W;                                    // one division per thread; then += JP with carry
    const int dq_h = JP / g.W, dq_w = JP - dq_h * g.W;
    float acc[NJ];
    // the thread's NJ positions side by side: their loads (up to 5 per row of taps) are all in flight together
#pragma unroll
    for (int u = 0; u < NJ; ++u) {
        acc[u] = 0.f;
        if (q < g.I && ci < g.Cin) {
            // output rows / columns that reach (ih, iw): oh in [oh_lo, oh_hi], kh = ih + PH - oh*SH, likewise along W
            const int th = ih + g.PH, tw = iw + g.PW;
            const int oh_hi = min(g.OH - 1, th / SH), ow_hi = min(g.OW - 1, tw / SW);
            const int oh_lo = th - g.KH + 1 <= 0 ? 0 : (th - g.KH + SH) / SH;
            const int ow_lo = tw - g.KW + 1 <= 0 ? 0 : (tw - g.KW + SW) / SW;
            const int cnt = ow_hi - ow_lo + 1;
            const float* p0 = base + (size_t)(oh_lo * g.OW + ow_lo) * g.Kc + ((th - oh_lo * SH) * g.KW + tw - ow_lo * SW) * g.Cin;
            for (int oh = oh_lo; oh <= oh_hi; ++oh, p0 += rstep) {
                float v[5];
#pragma unroll
                for (int i = 0; i < 5; ++i) v[i] = i < cnt ? __ldg(p0 + i * step) : 0.f;
#pragma unroll
                for (int i = 0; i < 5; ++i) acc[u] += v[i];
                for (int i = 5; i < cnt; ++i) acc[u] += __ldg(p0 + i * step);
            }
        }
        q += JP;
        ih += dq_h;
        iw += dq_w;
        if (iw >= g.W) {
            iw -= g.W;
            ++ih;
        }
    }
#pragma unroll
    for (int u = 0; u < NJ; ++u) tile[c][j0 + u * JP] = acc[u];
    __syncthreads();
    for (int e = threadIdx.x; e < CL * 32; e += 256) {
        const int cc = e >> 5, jj = e & 31;
        if (cb + cc < g.Cin && q0 + jj < g.I) dx[((size_t)n * g.Cin + cb + cc) * g.I + q0 + jj] = tile[cc][jj];
    }
}

template <int CL>
static void launch_col2im(const Geo& g, const float* dcol, float* dx, cudaStream_t st) {
    dim3 grid((g.I + 31) / 32, (g.Cin + CL - 1) / CL, g.N);
    if (g.SH == 1 && g.SW == 2) col2im_kernel<CL, 1, 2><<<grid, 256, 0, st>>>(g, dcol, dx);
    else if (g.SH == 2 && g.SW == 2) col2im_kernel<CL, 2, 2><<<grid, 256, 0, st>>>(g, dcol, dx);
    else if (g.SH == 1 && g.SW == 1) col2im_kernel<CL, 1, 1><<<grid, 256, 0, st>>>(g, dcol, dx);
    else col2im_kernel<CL, 0, 0><<<grid, 256, 0, st>>>(g, dcol, dx);
}

static inline size_t al256(size_t v) { return (v + 255) & ~(size_t)255; }

static inline int grid_for(long long n_threads) {
    long long b = (n_threads + 255) / 256;
    if (b > 148 * 16) b = 148 * 16;
    return (int)(b < 1 ? 1 : b);
}

}  // namespace tc3

extern "C" int maavss_conv_tc3_supported(const maavss_conv_desc* d) {
    if (!d) return 0;
    if (d->D != 1 || d->KD != 1 || d->OD != 1 || d->SD != 1 || d->PD != 0) return 0;
    if (d->N <= 0 || d->Cin < 8 || d->Cout < 16 || d->Cout % 4) return 0;
    if (d->Cin < 16 && d->Cout < 32) return 0;        // C8 -> C16 (measured at the headline shapes): the direct kernels win
    const tc3::Geo g = tc3::make_geo(d);
    if (g.Kc % 4 || g.OHW % 32 || g.N > 65535) return 0;
    if (g.pos < 128 || g.pos * (long long)g.Kc > (1LL << 27)) return 0;       // im2col matrix <= 512 MB
    return 1;
}

extern "C" long long maavss_conv_tc3_col_elems(const maavss_conv_desc* d) {
    const tc3::Geo g = tc3::make_geo(d);
    return g.pos * g.Kc;
}

extern "C" long long maavss_conv_tc3_workspace_bytes(const maavss_conv_desc* d, int mode) {
    const tc3::Geo g = tc3::make_geo(d);
    const int P = (int)g.pos;
    if (mode == 0) return maavss_gemm_tf32x3_workspace_bytes(P, g.Cout, g.Kc) + 256;
    if (mode == 1) return (long long)tc3::al256((size_t)g.pos * g.Kc * 4) + maavss_gemm_tf32x3_workspace_bytes(P, g.Kc, g.Cout) + 256;
    return maavss_gemm_tf32x3_workspace_bytes(g.Cout, g.Kc, P) + 256;
}

extern "C" int maavss_nchw_to_rows(const float* in, int N, int C, long long inner, float* out, void* stream) {
    if ((long long)N * C * inner == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(in && out && N <= 65535 && inner < (1LL << 31), "nchw_to_rows: bad argument");
    dim3 grid((unsigned)((inner + 31) / 32), (unsigned)((C + 31) / 32), (unsigned)N);
    tc3::rows_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in, out, C, (int)inner);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_conv_tc3_fwd(const maavss_conv_desc* d, const float* x, const float* w, const float* bias, float* y, float* col,
                                   void* workspace, long long workspace_bytes, void* stream) {
    MAAVSS_CHECK_ARG(maavss_conv_tc3_supported(d), "conv_tc3_fwd: layer outside the envelope (maavss_conv_tc3_supported)");
    MAAVSS_CHECK_ARG(x && w && y && col, "conv_tc3_fwd: null argument");
    MAAVSS_CHECK_ARG(workspace_bytes >= maavss_conv_tc3_workspace_bytes(d, 0), "conv_tc3_fwd: workspace too small");
    const tc3::Geo g = tc3::make_geo(d);
    cudaStream_t st = (cudaStream_t)stream;
    const long long total4 = g.pos * (g.Kc / 4);
    int OWT = g.OW < 64 ? g.OW : 64;
    while (OWT > 8 && (long long)g.N * g.OH * ((g.OW + OWT - 1) / OWT) < 1500) OWT /= 2;      // >= ~10 CTAs per SM: the staging phase is latency bound
    int span = (OWT - 1) * g.SW + g.KW;
    size_t smem = sizeof(float) * ((size_t)g.Kc + (size_t)g.Cin * g.KH * span);
    // (ncu: this kernel is instruction-issue bound, 72 % issue-active, ~50 instructions per float4 of col.)
    if (smem <= 48 * 1024 && (long long)g.N * g.OH * ((g.OW + OWT - 1) / OWT) < (1LL << 31)) {
        const unsigned grid = (unsigned)(g.N * g.OH * ((g.OW + OWT - 1) / OWT));
        tc3::im2col_staged_kernel<<<grid, 256, smem, st>>>(g, x, col, OWT, span);
    } else {
        tc3::im2col_kernel<<<(unsigned)((total4 + 255) / 256), 256, 0, st>>>(g, x, col, total4);      // one float4 per thread
    }
    MAAVSS_LAUNCH_CHECK();
    void* ws = (void*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    return maavss_gemm_tf32x3_blocked(col, 0, g.Kc, w, 0, g.Kc, (int)g.pos, g.Cout, g.Kc, y, g.OHW, 1, g.OHW, (long long)g.Cout * g.OHW, bias, 1,
                                      MAAVSS_ACT_NONE, 0.f, 0, ws, workspace_bytes - 256, stream);
}

extern "C" int maavss_conv_tc3_permute_weights(const maavss_conv_desc* d, const float* w, float* w_perm, void* stream) {
    MAAVSS_CHECK_ARG(d && w && w_perm, "conv_tc3_permute_weights: null argument");
    const tc3::Geo g = tc3::make_geo(d);
    tc3::permute_w_kernel<<<tc3::grid_for((long long)g.Cout * g.Kc), 256, 0, (cudaStream_t)stream>>>(w, w_perm, g.Cout, g.Cin, g.taps);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_conv_tc3_dgrad(const maavss_conv_desc* d, const float* dy_rows, const float* w_perm, float* dx, void* workspace,
                                     long long workspace_bytes, void* stream) {
    MAAVSS_CHECK_ARG(maavss_conv_tc3_supported(d), "conv_tc3_dgrad: layer outside the envelope (maavss_conv_tc3_supported)");
    MAAVSS_CHECK_ARG(dy_rows && w_perm && dx, "conv_tc3_dgrad: null argument");
    MAAVSS_CHECK_ARG(workspace_bytes >= maavss_conv_tc3_workspace_bytes(d, 1), "conv_tc3_dgrad: workspace too small");
    const tc3::Geo g = tc3::make_geo(d);
    cudaStream_t st = (cudaStream_t)stream;
    float* dcol = (float*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    const size_t dcol_bytes = tc3::al256((size_t)g.pos * g.Kc * 4);
    void* gws = (uint8_t*)dcol + dcol_bytes;
    int rc = maavss_gemm_tf32x3(dy_rows, 0, g.Cout, w_perm, 1, g.Kc, (int)g.pos, g.Kc, g.Cout, dcol, g.Kc, 0, nullptr, 0, MAAVSS_ACT_NONE, 0.f, 0, gws,
                                workspace_bytes - 256 - (long long)dcol_bytes, stream);
    if (rc) return rc;
    if (g.Cin >= 32) tc3::launch_col2im<32>(g, dcol, dx, st);
    else if (g.Cin <= 8) tc3::launch_col2im<8>(g, dcol, dx, st);
    else tc3::launch_col2im<16>(g, dcol, dx, st);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_conv_tc3_wgrad(const maavss_conv_desc* d, const float* col, const float* dy_rows, float* dw, int accumulate,
                                     void* workspace, long long workspace_bytes, void* stream) {
    MAAVSS_CHECK_ARG(maavss_conv_tc3_supported(d), "conv_tc3_wgrad: layer outside the envelope (maavss_conv_tc3_supported)");
    MAAVSS_CHECK_ARG(col && dy_rows && dw, "conv_tc3_wgrad: null argument");
    MAAVSS_CHECK_ARG(workspace_bytes >= maavss_conv_tc3_workspace_bytes(d, 2), "conv_tc3_wgrad: workspace too small");
    const tc3::Geo g = tc3::make_geo(d);
    void* ws = (void*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    return maavss_gemm_tf32x3(dy_rows, 1, g.Cout, col, 1, g.Kc, g.Cout, g.Kc, (int)g.pos, dw, g.Kc, 0, nullptr, 0, MAAVSS_ACT_NONE, 0.f,
                              accumulate, ws, workspace_bytes - 256, stream);
}
