// utilities.video_phasegram (/root/reference/utilities.py:206-228), the op between the attention maps and the
// model on every train.py micro-step (train.py:139):
//   antialiased bilinear resize (H, W) -> (R, R), fft2, fftshift over ALL dims (batch and time axes roll too --
//   reproduced), atan2, flatten, cumsum / (2*pi*R*R), temporal diff with a zero first frame, and
//   division by the global max |.|.
// Kernel 0 (resize, HBM bound: it reads the raw frames once): one CTA per (frame, strip of output rows) stages the input
// rows the strip needs in shared memory with coalesced float4 loads, filters horizontally, then vertically; ~1000 small
// CTAs, 4 per SM, hide the load latency.  Kernel 1: one CTA per resized frame keeps everything in shared memory (row FFTs
// -> column FFTs -> angle -> block scan) and writes R*R floats; kernel 2: temporal diff + global abs-max; kernel 3: scale.
#include <math.h>
#include <stdint.h>
#include "common.cuh"
#include "fft_core.cuh"
#include "../../include/maavss_b200.h"

namespace {

constexpr int kPgThreads = 256;
constexpr int kMaxTaps = 34;

// torch _upsample_bilinear2d_aa weights for output index i (align_corners = False), all in fp32 like ATen.
__device__ __forceinline__ void aa_weights(int i, int in_size, float scale, int& xmin, int& xsize, float* w) {
    const float support = scale >= 1.f ? scale : 1.f;
    const float invscale = scale >= 1.f ? 1.f / scale : 1.f;
    const float center = scale * ((float)i + 0.5f);
    xmin = max((int)(center - support + 0.5f), 0);
    xsize = min((int)(center + support + 0.5f), in_size) - xmin;
    if (xsize > kMaxTaps) xsize = kMaxTaps;
    float total = 0.f;
    for (int j = 0; j < xsize; ++j) {
        float x = fabsf(((float)(j + xmin) - center + 0.5f) * invscale);
        const float v = x < 1.f ? 1.f - x : 0.f;
        w[j] = v;
        total += v;
    }
    if (total != 0.f)
        for (int j = 0; j < xsize; ++j) w[j] /= total;
}

// sum_j p[j * stride] * w[j], j ascending (ATen's order).  n is warp-uniform; the common down-scales (<= 5.5x) have at
// most 12 taps and take the unrolled, predicated form.
__device__ __forceinline__ float aa_filter(const float* p, int stride, const float* w, int n) {
    float acc = p[0] * w[0];
    if (n <= 12) {
#pragma unroll
        for (int j = 1; j < 12; ++j)
            if (j < n) acc += p[j * stride] * w[j];
    } else {
        for (int j = 1; j < n; ++j) acc += p[j * stride] * w[j];
    }
    return acc;
}

struct PgArgs {
    const float* frames;   // (B, T, H, W)
    float* pflat;          // (B, T, R*R)
    int B, T, H, W, cumulative;
    long long batch_stride;   // elements between consecutive clips of `frames` (>= T*H*W: a time-slice view needs no copy)
};

struct RsArgs {
    const float* frames;   // (B, T, H, W), clips batch_stride apart
    float* resized;        // (B*T, R, R)
    int T, H, W, orc, rows_cap;   // orc output rows per CTA; rows_cap input rows of staging
    long long batch_stride;
};

constexpr int kRsMaxOrc = 8;

template <int R>
__global__ void __launch_bounds__(kPgThreads) phasegram_resize_kernel(RsArgs a) {
    extern __shared__ float smem[];
    __shared__ float wx[R][kMaxTaps], wy[kRsMaxOrc][kMaxTaps];
    __shared__ int xmin_s[R], xsz_s[R], ymin_s[kRsMaxOrc], ysz_s[kRsMaxOrc];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int frame = blockIdx.y, b = frame / a.T, t = frame % a.T;
    const int o0 = blockIdx.x * a.orc, no = min(a.orc, R - o0);
    const float* src = a.frames + (size_t)b * a.batch_stride + (size_t)t * a.H * a.W;
    const int SW = a.W + 1, TS = R + 1;
    float* stage = smem;                       // rows_cap x (W+1): bank-conflict free for a warp reading 32 rows of one column
    float* tmp = smem + a.rows_cap * SW;       // rows_cap x (R+1)
    if (tid < R) aa_weights(tid, a.W, (float)a.W / (float)R, xmin_s[tid], xsz_s[tid], wx[tid]);
    else if (tid < R + no) aa_weights(o0 + tid - R, a.H, (float)a.H / (float)R, ymin_s[tid - R], ysz_s[tid - R], wy[tid - R]);
    __syncthreads();
    const int ylo = ymin_s[0];
    const int rows = ymin_s[no - 1] + ysz_s[no - 1] - ylo;
    if (rows > a.rows_cap) __trap();           // cannot happen: rows_cap is the analytic bound + 2 (launch_resize)
    const float* p = src + (size_t)ylo * a.W;
    if ((a.W & 3) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        const int q4 = a.W >> 2, n4 = rows * q4;
        const int sh = (q4 & (q4 - 1)) == 0 ? 31 - __clz(q4) : -1;
        const float4* p4 = reinterpret_cast<const float4*>(p);
        for (int f0 = tid; f0 < n4; f0 += 4 * kPgThreads) {       // four independent 16-byte loads in flight per thread
            float4 v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (f0 + k * kPgThreads < n4) v[k] = __ldg(p4 + f0 + k * kPgThreads);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int f = f0 + k * kPgThreads;
                if (f < n4) {
                    const int r = sh >= 0 ? (f >> sh) : f / q4;
                    float* d = stage + r * SW + 4 * (f - r * q4);
                    d[0] = v[k].x; d[1] = v[k].y; d[2] = v[k].z; d[3] = v[k].w;
                }
            }
        }
    } else {
        for (int r = warp; r < rows; r += kPgThreads / 32)
            for (int c = lane; c < a.W; c += 32) stage[r * SW + c] = __ldg(p + (size_t)r * a.W + c);
    }
    __syncthreads();
    for (int oc = warp; oc < R; oc += kPgThreads / 32) {          // horizontal pass (last dim first, as ATen)
        const int n = xsz_s[oc], x0 = xmin_s[oc];
        for (int r = lane; r < rows; r += 32) tmp[r * TS + oc] = aa_filter(stage + r * SW + x0, 1, wx[oc], n);
    }
    __syncthreads();
    float* dst = a.resized + ((size_t)frame * R + o0) * R;
    for (int e = tid; e < no * R; e += kPgThreads) {              // vertical pass
        const int ol = e / R, oc = e % R;
        dst[e] = aa_filter(tmp + (ymin_s[ol] - ylo) * TS + oc, TS, wy[ol], ysz_s[ol]);
    }
}

template <int R>
__global__ void __launch_bounds__(kPgThreads) phasegram_frame_kernel(PgArgs a) {
    static_assert(R == 64, "one pass of 256 threads covers R FFTs of R/16 threads each only for R = 64");
    constexpr int TPF = R / 16;                 // threads per 1-D FFT
    constexpr int FPB = kPgThreads / TPF;       // FFTs per pass of the block
    constexpr int BL = fft_buf_len(R);
    extern __shared__ float smem[];
    float* img = smem;                                          // R*R resized image
    float2* tw = reinterpret_cast<float2*>(img + R * R);        // R entries exp(-2*pi*i*q/(2R))
    float2* rowbuf = tw + R;                                    // R rows  x BL
    float2* colbuf = rowbuf + R * BL;                           // R cols  x BL
    __shared__ float warp_tot[kPgThreads / 32];

    const int tid = threadIdx.x;
    const int b = blockIdx.x / a.T, t = blockIdx.x % a.T;
    const float* src = a.frames + (size_t)b * a.batch_stride + (size_t)t * a.H * a.W;
    for (int q = tid; q < R; q += kPgThreads) {
        float s, c;
        sincospif(-(float)q / (float)R, &s, &c);
        tw[q] = make_float2(c, s);
    }
    for (int e = tid; e < R * R; e += kPgThreads) img[e] = src[e];
    __syncthreads();
    const int fi = tid / TPF, i = tid % TPF;
    // row FFTs: rowbuf[r][pad(v)] = sum_x img[r][x] e^{-2 pi i v x / R}
    for (int r0 = 0; r0 < R; r0 += FPB) {
        const int r = r0 + fi;
        float2 v[16];
#pragma unroll
        for (int m = 0; m < 16; ++m) v[m] = make_float2(img[r * R + i + m * TPF], 0.f);
        fft_forward_block<R>(v, i, tw, rowbuf + r * BL);
    }
    // column FFTs: colbuf[v][pad(u)] = sum_r rowbuf[r][v] e^{-2 pi i u r / R}
    for (int c0 = 0; c0 < R; c0 += FPB) {
        const int c = c0 + fi;
        float2 v[16];
#pragma unroll
        for (int m = 0; m < 16; ++m) v[m] = rowbuf[(i + m * TPF) * BL + pad_idx(c)];
        fft_forward_block<R>(v, i, tw, colbuf + c * BL);
    }
    // angle of the spatially fft-shifted spectrum, flattened row-major, then the inclusive scan
    constexpr int PER = R * R / kPgThreads;
    float ang[PER];
    float run = 0.f;
#pragma unroll
    for (int q = 0; q < PER; ++q) {
        const int n = tid * PER + q;
        const int y = n / R, x = n % R;
        const int u = (y + R / 2) & (R - 1), vv = (x + R / 2) & (R - 1);
        float2 z = colbuf[vv * BL + pad_idx(u)];
        if ((u == 0 || u == R / 2) && (vv == 0 || vv == R / 2)) z.y = 0.f;   // self-conjugate bins of a real image: exactly real
        const float an = atan2f(z.y, z.x);
        if (a.cumulative) {
            run += an;
            ang[q] = run;
        } else {
            ang[q] = an;
        }
    }
    float offset = 0.f;
    if (a.cumulative) {
        const int lane = tid & 31, wid = tid >> 5;
        float inc = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float nb = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += nb;
        }
        if (lane == 31) warp_tot[wid] = inc;
        __syncthreads();
        float woff = 0.f;
        for (int k = 0; k < wid; ++k) woff += warp_tot[k];
        offset = woff + (inc - run);
    }
    // stored at the frame's own (b, t): the batch / time rolls of the all-dims fftshift are index arithmetic of the
    // window kernels below (a clip's frames are shared by its overlapping accumulation windows)
    float* dst = a.pflat + ((size_t)b * a.T + t) * R * R + (size_t)tid * PER;
    const float two_pi_n = (float)(2.0 * M_PI * (double)(R * R));
    const float pi_f = (float)M_PI, two_pi = (float)(M_PI * 2.0);
#pragma unroll
    for (int q = 0; q < PER; ++q)
        dst[q] = a.cumulative ? __fdiv_rn(ang[q] + offset, two_pi_n) : __fdiv_rn(ang[q] + pi_f, two_pi);
}

// ---- window half.  A "window" j is the reference's attn[:, :, j:j+Tw] (train.py:138-139): out_j[bd, td] is built from
// the per-frame flat phases pf[(bd - roll) mod B][j + (td - Tw/2) mod Tw] (the batch and time rolls of the all-dims
// fftshift, utilities.py:210), temporal diff with a zero first frame (:219-222), division by the window's max |.| (:226-227).
struct WinArgs {
    const float* pf;        // (B, Tc, n)
    float* out;             // (G, B, Tw, n)
    unsigned int* mx;       // (G) float bits of max |.| per window
    int B, Tc, Tw, G, roll, diff, normalize;
    int n4;                 // n / 4
};

__device__ __forceinline__ float4 win_value(const WinArgs& a, long long q4, int& j) {
    const int pix = (int)(q4 % a.n4);
    long long f = q4 / a.n4;
    const int td = (int)(f % a.Tw);
    f /= a.Tw;
    const int bd = (int)(f % a.B);
    j = (int)(f / a.B);
    const int sb = (bd - a.roll + a.B) % a.B;
    const int st = (td + a.Tw - a.Tw / 2) % a.Tw;
    const float4* base = reinterpret_cast<const float4*>(a.pf) + (size_t)sb * a.Tc * a.n4 + pix;
    float4 v = __ldg(base + (size_t)(j + st) * a.n4);
    if (a.diff) {
        if (td == 0) return make_float4(0.f, 0.f, 0.f, 0.f);
        const int sp = (td - 1 + a.Tw - a.Tw / 2) % a.Tw;
        const float4 u = __ldg(base + (size_t)(j + sp) * a.n4);
        v.x -= u.x; v.y -= u.y; v.z -= u.z; v.w -= u.w;
    }
    return v;
}

// one CTA never straddles two windows: per_win4 is a multiple of the 256-thread tile (n4 = 1024)
__global__ void __launch_bounds__(256) pg_win_max_kernel(WinArgs a, long long per_win4, int ctas_per_win) {
    const int j = blockIdx.x / ctas_per_win, c = blockIdx.x % ctas_per_win;
    float m = 0.f;
    for (long long q = (long long)c * 256 + threadIdx.x; q < per_win4; q += (long long)ctas_per_win * 256) {
        int jj;
        const float4 v = win_value(a, (long long)j * per_win4 + q, jj);
        m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
    }
    m = warp_max(m);
    __shared__ float sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) m = fmaxf(m, sm[k]);
        atomicMax(a.mx + j, __float_as_uint(m));
    }
}

__global__ void __launch_bounds__(256) pg_win_write_kernel(WinArgs a, long long total4) {
    for (long long q = (long long)blockIdx.x * 256 + threadIdx.x; q < total4; q += (long long)gridDim.x * 256) {
        int j;
        float4 v = win_value(a, q, j);
        if (a.normalize) {
            const float inv = 1.0f / __uint_as_float(__ldg(a.mx + j));
            v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
        }
        reinterpret_cast<float4*>(a.out)[q] = v;
    }
}

template <int R>
int launch_resize(const float* frames, float* resized, int B, int T, int H, int W, long long batch_stride, cudaStream_t st) {
    // input rows one strip of `orc` output rows touches: (orc - 1) * scale + 2 * support + 1, +2 of rounding slack
    const float scale = (float)H / (float)R, support = scale >= 1.f ? scale : 1.f;
    int orc = kRsMaxOrc, rows_cap = 0;
    size_t smem = 0;
    for (;; orc /= 2) {
        rows_cap = (int)ceilf((orc - 1) * scale + 2.f * support) + 3;
        if (rows_cap > H) rows_cap = H;
        smem = sizeof(float) * (size_t)rows_cap * ((W + 1) + (R + 1));
        if (smem <= 56 * 1024 || orc == 1) break;
    }
    MAAVSS_CHECK_ARG(smem + 12 * 1024 <= 227 * 1024, "video_phasegram: frame %dx%d too large for the shared-memory resize", H, W);
    RsArgs a{frames, resized, T, H, W, orc, rows_cap, batch_stride};
    MAAVSS_CUDA(cudaFuncSetAttribute(phasegram_resize_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MAAVSS_CHECK_ARG((long long)B * T <= 65535, "video_phasegram: B*T > 65535");
    phasegram_resize_kernel<R><<<dim3((R + orc - 1) / orc, B * T), kPgThreads, smem, st>>>(a);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

template <int R>
int launch_frames(const PgArgs& a, cudaStream_t st) {
    const size_t smem = sizeof(float) * R * R + sizeof(float2) * (R + 2 * R * fft_buf_len(R));
    MAAVSS_CUDA(cudaFuncSetAttribute(phasegram_frame_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    phasegram_frame_kernel<R><<<a.B * a.T, kPgThreads, smem, st>>>(a);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

}  // namespace

static int pg_check(int B, int T, int H, int W, int R, long long batch_stride) {
    MAAVSS_CHECK_ARG(R == 64, "video_phasegram: resize target %d unsupported (p_size 64, run_config.py:35)", R);
    const int do_resize = !(H == R && W == R);
    MAAVSS_CHECK_ARG(!do_resize || (H >= R && W >= R && H <= 16 * R && W <= 16 * R), "video_phasegram: only down-scaling by <= 16 is supported");
    MAAVSS_CHECK_ARG(batch_stride >= (long long)T * H * W, "video_phasegram: batch_stride smaller than one clip");
    return MAAVSS_OK;
}

extern "C" int maavss_phasegram_frames(const float* frames, float* pflat, float* scratch, int B, int T, int H, int W, int R,
                                       long long batch_stride, int cumulative, void* stream) {
    if (B == 0 || T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(frames && pflat && scratch, "phasegram_frames: null argument");
    int rc = pg_check(B, T, H, W, R, batch_stride);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int do_resize = !(H == R && W == R);
    if (do_resize) {
        rc = launch_resize<64>(frames, scratch, B, T, H, W, batch_stride, st);
        if (rc) return rc;
    }
    PgArgs a{do_resize ? scratch : frames, pflat, B, T, R, R, cumulative, do_resize ? (long long)T * R * R : batch_stride};
    return launch_frames<64>(a, st);
}

static int pg_win_args(WinArgs& a, const float* pflat, float* out, void* max_ws, int B, int Tc, int Tw, int G, int R, int batch_roll,
                       int diff, int normalize) {
    MAAVSS_CHECK_ARG(pflat && max_ws, "phasegram_windows: null argument");
    MAAVSS_CHECK_ARG(B > 0 && Tw > 0 && G > 0 && G - 1 + Tw <= Tc, "phasegram_windows: %d windows of %d frames do not fit %d frames", G, Tw, Tc);
    MAAVSS_CHECK_ARG((R * R) % 1024 == 0, "phasegram_windows: R*R must be a multiple of 1024");
    a = WinArgs{pflat, out, (unsigned int*)max_ws, B, Tc, Tw, G, ((batch_roll % B) + B) % B, diff, normalize, R * R / 4};
    return MAAVSS_OK;
}

extern "C" int maavss_phasegram_windows_max(const float* pflat, void* max_ws, int B, int Tc, int Tw, int G, int R, int batch_roll,
                                            int diff, void* stream) {
    WinArgs a;
    int rc = pg_win_args(a, pflat, nullptr, max_ws, B, Tc, Tw, G, R, batch_roll, diff, 1);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    MAAVSS_CUDA(cudaMemsetAsync(max_ws, 0, sizeof(unsigned int) * G, st));
    const long long per_win4 = (long long)B * Tw * a.n4;
    int cpw = (int)((per_win4 + 1023) / 1024);
    const int cap = (148 * 8 + G - 1) / G;
    if (cpw > cap) cpw = cap;
    pg_win_max_kernel<<<G * cpw, 256, 0, st>>>(a, per_win4, cpw);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_phasegram_windows_write(const float* pflat, const void* max_ws, float* out, int B, int Tc, int Tw, int G, int R,
                                              int batch_roll, int diff, int normalize, void* stream) {
    WinArgs a;
    MAAVSS_CHECK_ARG(out, "phasegram_windows_write: null output");
    int rc = pg_win_args(a, pflat, out, const_cast<void*>(max_ws), B, Tc, Tw, G, R, batch_roll, diff, normalize);
    if (rc) return rc;
    const long long total4 = (long long)G * B * Tw * a.n4;
    int blocks = (int)((total4 + 511) / 512);
    if (blocks > 148 * 8) blocks = 148 * 8;
    pg_win_write_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(a, total4);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_video_phasegram(const float* frames, float* out, float* scratch, void* max_ws, int B, int T, int H, int W,
                                      int R, long long batch_stride, int diff, int cumulative, int normalize, void* stream) {
    if (B == 0 || T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(frames && out && scratch && max_ws, "video_phasegram: null argument");
    // `out` is free until the window kernel writes it: it holds the resized frames (B*T, R, R) meanwhile
    int rc = maavss_phasegram_frames(frames, scratch, out, B, T, H, W, R, batch_stride, cumulative, stream);
    if (rc) return rc;
    rc = maavss_phasegram_windows_max(scratch, max_ws, B, T, T, 1, R, B / 2, diff, stream);
    if (rc) return rc;
    return maavss_phasegram_windows_write(scratch, max_ws, out, B, T, T, 1, R, B / 2, diff, normalize, stream);
}
