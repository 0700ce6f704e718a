// utilities.video_phasegram (/root/reference/utilities.py:206-228), the op between the attention maps and the
// model on every train.py micro-step (train.py:139):
//   antialiased bilinear resize (H, W) -> (R, R), fft2, fftshift over ALL dims (batch and time axes roll too --
//   reproduced), atan2, flatten, cumsum / (2*pi*R*R), temporal diff with a zero first frame, and
//   division by the global max |.|.
// Kernel 1: one CTA per frame keeps everything in shared memory (resize -> row FFTs -> column FFTs -> angle ->
// block scan) and writes R*R floats; kernel 2: temporal diff + global abs-max; kernel 3: scale.
#include <math.h>
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

struct PgArgs {
    const float* frames;   // (B, T, H, W)
    float* pflat;          // (B, T, R*R)
    int B, T, H, W, do_resize, cumulative;
    long long batch_stride;   // elements between consecutive clips of `frames` (>= T*H*W: a time-slice view needs no copy)
};

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
    float* tmp = reinterpret_cast<float*>(colbuf + R * BL);     // H x R horizontal pass (resize only)
    __shared__ float wx[R][kMaxTaps], wy[R][kMaxTaps];
    __shared__ int xmin_s[R], xsz_s[R], ymin_s[R], ysz_s[R];
    __shared__ float warp_tot[kPgThreads / 32];

    const int tid = threadIdx.x;
    const int b = blockIdx.x / a.T, t = blockIdx.x % a.T;
    const float* src = a.frames + (size_t)b * a.batch_stride + (size_t)t * a.H * a.W;
    for (int q = tid; q < R; q += kPgThreads) {
        float s, c;
        sincospif(-(float)q / (float)R, &s, &c);
        tw[q] = make_float2(c, s);
    }
    if (a.do_resize) {
        if (tid < R) aa_weights(tid, a.W, (float)a.W / (float)R, xmin_s[tid], xsz_s[tid], wx[tid]);
        else if (tid < 2 * R) aa_weights(tid - R, a.H, (float)a.H / (float)R, ymin_s[tid - R], ysz_s[tid - R], wy[tid - R]);
        __syncthreads();
        for (int e = tid; e < a.H * R; e += kPgThreads) {      // horizontal pass (last dim first, as ATen)
            const int row = e / R, oc = e % R;
            const float* p = src + (size_t)row * a.W + xmin_s[oc];
            float acc = p[0] * wx[oc][0];
            for (int j = 1; j < xsz_s[oc]; ++j) acc += p[j] * wx[oc][j];
            tmp[e] = acc;
        }
        __syncthreads();
        for (int e = tid; e < R * R; e += kPgThreads) {        // vertical pass
            const int orow = e / R, oc = e % R;
            const float* p = tmp + (size_t)ymin_s[orow] * R + oc;
            float acc = p[0] * wy[orow][0];
            for (int j = 1; j < ysz_s[orow]; ++j) acc += p[(size_t)j * R] * wy[orow][j];
            img[e] = acc;
        }
    } else {
        for (int e = tid; e < R * R; e += kPgThreads) img[e] = src[e];
    }
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
    // destination frame after the batch / time fftshift rolls
    const int bd = (b + a.B / 2) % a.B, td = (t + a.T / 2) % a.T;
    float* dst = a.pflat + ((size_t)bd * a.T + td) * R * R + (size_t)tid * PER;
    const float two_pi_n = (float)(2.0 * M_PI * (double)(R * R));
    const float pi_f = (float)M_PI, two_pi = (float)(M_PI * 2.0);
#pragma unroll
    for (int q = 0; q < PER; ++q)
        dst[q] = a.cumulative ? __fdiv_rn(ang[q] + offset, two_pi_n) : __fdiv_rn(ang[q] + pi_f, two_pi);
}

__global__ void __launch_bounds__(256) pg_diff_max_kernel(const float* __restrict__ p, float* __restrict__ out, int T, long long n,
                                                          long long total, int diff, unsigned int* __restrict__ mx) {
    float m = 0.f;
    for (long long q = (long long)blockIdx.x * 256 + threadIdx.x; q < total; q += (long long)gridDim.x * 256) {
        float v = p[q];
        if (diff) {
            const long long fr = q / n;
            v = (fr % T == 0) ? 0.f : v - p[q - n];
        }
        out[q] = v;
        m = fmaxf(m, fabsf(v));
    }
    m = warp_max(m);
    __shared__ float sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) m = fmaxf(m, sm[k]);
        atomicMax(mx, __float_as_uint(m));
    }
}

__global__ void __launch_bounds__(256) pg_scale_kernel(float* __restrict__ out, long long total, const unsigned int* __restrict__ mx) {
    const float inv = 1.0f / __uint_as_float(mx[0]);
    for (long long q = (long long)blockIdx.x * 256 + threadIdx.x; q < total; q += (long long)gridDim.x * 256) out[q] *= inv;
}

template <int R>
int launch_frames(const PgArgs& a, cudaStream_t st) {
    size_t smem = sizeof(float) * R * R + sizeof(float2) * (R + 2 * R * fft_buf_len(R));
    if (a.do_resize) smem += sizeof(float) * (size_t)a.H * R;
    MAAVSS_CHECK_ARG(smem + 20 * 1024 <= 227 * 1024, "video_phasegram: frame %dx%d too large for the shared-memory resize", a.H, a.W);
    MAAVSS_CUDA(cudaFuncSetAttribute(phasegram_frame_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    phasegram_frame_kernel<R><<<a.B * a.T, kPgThreads, smem, st>>>(a);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

}  // namespace

extern "C" int maavss_video_phasegram(const float* frames, float* out, float* scratch, void* max_ws, int B, int T, int H, int W,
                                      int R, long long batch_stride, int diff, int cumulative, int normalize, void* stream) {
    if (B == 0 || T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(frames && out && scratch && max_ws, "video_phasegram: null argument");
    MAAVSS_CHECK_ARG(R == 64, "video_phasegram: resize target %d unsupported (p_size 64, run_config.py:35)", R);
    const int do_resize = !(H == R && W == R);
    MAAVSS_CHECK_ARG(!do_resize || (H >= R && W >= R && H <= 16 * R && W <= 16 * R), "video_phasegram: only down-scaling by <= 16 is supported");
    cudaStream_t st = (cudaStream_t)stream;
    MAAVSS_CHECK_ARG(batch_stride >= (long long)T * H * W, "video_phasegram: batch_stride smaller than one clip");
    PgArgs a{frames, scratch, B, T, H, W, do_resize, cumulative, batch_stride};
    int rc = launch_frames<64>(a, st);
    if (rc) return rc;
    const long long n = (long long)R * R, total = n * B * T;
    MAAVSS_CUDA(cudaMemsetAsync(max_ws, 0, sizeof(unsigned int), st));
    int blocks = (int)((total + 2047) / 2048);
    if (blocks > 148 * 8) blocks = 148 * 8;
    pg_diff_max_kernel<<<blocks, 256, 0, st>>>(scratch, out, T, n, total, diff, (unsigned int*)max_ws);
    MAAVSS_LAUNCH_CHECK();
    if (normalize) {
        pg_scale_kernel<<<blocks, 256, 0, st>>>(out, total, (const unsigned int*)max_ws);
        MAAVSS_LAUNCH_CHECK();
    }
    return MAAVSS_OK;
}
