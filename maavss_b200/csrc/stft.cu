// Fused STFT / iSTFT kernels (/root/reference/av_dataset.py:157-201).
//
// STFT:  reflect-pad + frame + Hamming window + real FFT (packed as an n/2-point complex FFT) +
//        1/sqrt(sum w^2) + last-frame / Nyquist trim + (2,T,F) plane layout (+ optional mag/phase),
//        one pass over HBM: reads 4*N bytes, writes 8*F*T bytes per clip.
// iSTFT: Nyquist zero re-pad + Hermitian unpack + inverse FFT + sqrt(n) + window + gather-style
//        (deterministic, atomics-free) overlap-add in shared memory + 1/sum w^2 envelope + centre trim.
#include <math.h>
#include <stdlib.h>
#include "common.cuh"
#include "stft_phases.cuh"
#include "../../include/maavss_b200.h"

struct maavss_stft_plan {
    int n_fft, hop;
    float2* tw;      // device, n_fft/2 entries: exp(-2*pi*i*q/n_fft)
    float* window;   // device, n_fft
    double sum_w2;
    int device;
};

namespace {

constexpr int kThreads = 256;

struct StftArgs {
    const float* audio;
    float* planes;
    const float2* tw;
    const float* window;
    int n_samples, hop, T_out, F_out, batches_per_cta, polar;
    float scale;
};

template <int M>
__global__ void __launch_bounds__(kThreads) stft_fwd_kernel(StftArgs a) {
    constexpr int TPF = M / 16;
    constexpr int P = kThreads / TPF;
    extern __shared__ float2 smem[];
    float2* tw = smem;
    float2* bufs = smem + M;
    for (int q = threadIdx.x; q < M; q += kThreads) tw[q] = a.tw[q];
    const int fi = threadIdx.x / TPF, i = threadIdx.x % TPF;
    float2* buf = bufs + fi * fft_buf_len(M);
    const float* clip = a.audio + (size_t)blockIdx.y * a.n_samples;
    float* out = a.planes + (size_t)blockIdx.y * 2 * a.T_out * a.F_out;
    __syncthreads();
    for (int bt = 0; bt < a.batches_per_cta; ++bt) {
        const int t0 = (blockIdx.x * a.batches_per_cta + bt) * P;
        if (t0 >= a.T_out) break;   // uniform across the block
        const int t = t0 + fi;
        const bool active = t < a.T_out;
        float2 v[16];
        if (active) {
            stft_load_frame<M>(v, i, clip, a.window, t, a.hop, a.n_samples);
        } else {
#pragma unroll
            for (int m = 0; m < 16; ++m) v[m] = make_float2(0.f, 0.f);
        }
        fft_forward_block<M>(v, i, tw, buf);
        if (active) {
            float* row0 = out + (size_t)t * a.F_out;
            stft_post_frame<M>(i, buf, tw, row0, row0 + (size_t)a.T_out * a.F_out, a.F_out, a.scale, a.polar);
        }
        __syncthreads();
    }
}

struct IstftArgs {
    const float* planes;
    float* audio;
    const float2* tw;
    const float* window;
    int hop, T, F_in, G, polar, out_len;
    float scale;
};

template <int M>
__global__ void __launch_bounds__(kThreads) istft_kernel(IstftArgs a) {
    constexpr int TPF = M / 16;
    constexpr int P = kThreads / TPF;
    constexpr int N = 2 * M;
    extern __shared__ float2 smem[];
    float2* tw = smem;
    float2* bufs = smem + M;
    float* acc = reinterpret_cast<float*>(bufs + P * fft_buf_len(M));
    const int tid = threadIdx.x;
    const int fi = tid / TPF, i = tid % TPF;
    const int hop = a.hop;
    const int s0 = blockIdx.x * a.G * hop;
    const int S_out = min(a.G * hop, a.out_len - s0);
    const int u0 = s0 + M;  // padded coordinate of the first output sample of this strip
    int t_lo = (u0 - N + 1 > 0) ? (u0 - N + 1 + hop - 1) / hop : 0;
    int t_hi = min(a.T - 1, (u0 + S_out - 1) / hop);
    for (int q = tid; q < M; q += kThreads) tw[q] = a.tw[q];
    for (int q = tid; q < S_out; q += kThreads) acc[q] = 0.f;
    const float* in0 = a.planes + (size_t)blockIdx.y * 2 * a.T * a.F_in;
    const float* in1 = in0 + (size_t)a.T * a.F_in;
    __syncthreads();
    for (int tb = t_lo; tb <= t_hi; tb += P) {
        const int t = tb + fi;
        const int t_last = min(tb + P - 1, t_hi);
        float2 v[16];
        if (t <= t_hi) {
            istft_load_frame<M>(v, i, in0 + (size_t)t * a.F_in, in1 + (size_t)t * a.F_in, a.F_in, tw, a.polar);
        } else {
#pragma unroll
            for (int m = 0; m < 16; ++m) v[m] = make_float2(0.f, 0.f);
        }
        fft_forward_block<M>(v, i, tw, bufs + fi * fft_buf_len(M));
        const int lo = max(u0, tb * hop);
        const int hi = min(u0 + S_out, t_last * hop + N);
        for (int u = lo + tid; u < hi; u += kThreads) {
            int tf = (u - N + 1 > 0) ? (u - N + 1 + hop - 1) / hop : 0;
            tf = max(tf, tb);
            const int tl = min(t_last, u / hop);
            float s = 0.f;
            for (int tt = tf; tt <= tl; ++tt) {
                const int j = u - tt * hop;
                s += istft_frame_sample(bufs + (tt - tb) * fft_buf_len(M), j) * __ldg(a.window + j);
            }
            acc[u - u0] += s;
        }
        __syncthreads();
    }
    float* out = a.audio + (size_t)blockIdx.y * a.out_len + s0;
    for (int q = tid; q < S_out; q += kThreads) {
        const int u = u0 + q;
        const int tf = (u - N + 1 > 0) ? (u - N + 1 + hop - 1) / hop : 0;
        const int tl = min(a.T - 1, u / hop);
        float env = 0.f;
        for (int tt = tf; tt <= tl; ++tt) {
            const float w = __ldg(a.window + (u - tt * hop));
            env = fmaf(w, w, env);
        }
        out[q] = acc[q] * a.scale / env;
    }
}

__global__ void absmax_kernel(const float* __restrict__ x, long long per_clip, float eps, unsigned int* __restrict__ ws) {
    const float* p = x + (size_t)blockIdx.y * per_clip;
    float m = 0.f;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < per_clip; q += (long long)gridDim.x * blockDim.x)
        m = fmaxf(m, fabsf(p[q]) + eps);
    m = warp_max(m);
    __shared__ float sm[32];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = (threadIdx.x < (blockDim.x >> 5)) ? sm[threadIdx.x] : 0.f;
        m = warp_max(m);
        if (threadIdx.x == 0) atomicMax(ws + blockIdx.y, __float_as_uint(m));   // m >= 0: uint order == float order
    }
}

__global__ void scale_by_inv_kernel(float* __restrict__ x, long long per_clip, const unsigned int* __restrict__ ws) {
    float* p = x + (size_t)blockIdx.y * per_clip;
    const float inv = 1.0f / __uint_as_float(ws[blockIdx.y]);
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < per_clip; q += (long long)gridDim.x * blockDim.x)
        p[q] *= inv;
}

__global__ void axpy_kernel(const float* __restrict__ y, const float* __restrict__ noise, float s, float* __restrict__ x, long long n) {
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x)
        x[q] = __fadd_rn(y[q], __fmul_rn(noise[q], s));   // unfused, bit-exact with tensor + noise * std
}

template <int M>
int launch_stft(const maavss_stft_plan* p, const StftArgs& a0, int batch, cudaStream_t st) {
    StftArgs a = a0;
    constexpr int P = kThreads / (M / 16);
    const int batches = ceil_div(a.T_out, P);
    a.batches_per_cta = ((long long)batches * batch >= 148LL * 32) ? 4 : 1;
    dim3 grid(ceil_div(batches, a.batches_per_cta), batch);
    const size_t smem = sizeof(float2) * (M + (size_t)P * fft_buf_len(M));
    MAAVSS_CUDA(cudaFuncSetAttribute(stft_fwd_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    stft_fwd_kernel<M><<<grid, kThreads, smem, st>>>(a);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

template <int M>
int launch_istft(const IstftArgs& a0, int batch, cudaStream_t st) {
    IstftArgs a = a0;
    constexpr int P = kThreads / (M / 16);
    int G = 16384 / a.hop;
    if (G < 1) G = 1;
    if (G > 64) G = 64;
    a.G = G;
    dim3 grid(ceil_div(a.out_len, (long long)G * a.hop), batch);
    const size_t smem = sizeof(float2) * (M + (size_t)P * fft_buf_len(M)) + sizeof(float) * (size_t)G * a.hop;
    MAAVSS_CUDA(cudaFuncSetAttribute(istft_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    istft_kernel<M><<<grid, kThreads, smem, st>>>(a);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

bool supported_nfft(int n) { return n >= 64 && n <= 4096 && (n & (n - 1)) == 0; }

}  // namespace

extern "C" int maavss_stft_plan_create(maavss_stft_plan** out, int n_fft, int hop, const float* window_host) {
    MAAVSS_CHECK_ARG(out && window_host, "stft_plan_create: null argument");
    MAAVSS_CHECK_ARG(supported_nfft(n_fft), "stft_plan_create: n_fft=%d unsupported (power of two in [64, 4096])", n_fft);
    MAAVSS_CHECK_ARG(hop >= 1 && hop <= n_fft, "stft_plan_create: hop=%d must be in [1, n_fft]", hop);
    maavss_stft_plan* p = (maavss_stft_plan*)calloc(1, sizeof(maavss_stft_plan));
    p->n_fft = n_fft;
    p->hop = hop;
    const int M = n_fft / 2;
    float2* tw_h = (float2*)malloc(sizeof(float2) * M);
    for (int q = 0; q < M; ++q) {
        const double ang = -2.0 * M_PI * (double)q / (double)n_fft;
        tw_h[q] = make_float2((float)cos(ang), (float)sin(ang));
    }
    double s = 0.0;
    for (int j = 0; j < n_fft; ++j) s += (double)window_host[j] * (double)window_host[j];
    p->sum_w2 = s;
    cudaGetDevice(&p->device);
    cudaError_t e = cudaMalloc(&p->tw, sizeof(float2) * M);
    if (e == cudaSuccess) e = cudaMalloc(&p->window, sizeof(float) * n_fft);
    if (e == cudaSuccess) e = cudaMemcpy(p->tw, tw_h, sizeof(float2) * M, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->window, window_host, sizeof(float) * n_fft, cudaMemcpyHostToDevice);
    free(tw_h);
    if (e != cudaSuccess) {
        maavss_set_error("stft_plan_create: %s", cudaGetErrorString(e));
        maavss_stft_plan_destroy(p);
        return MAAVSS_ERR_CUDA;
    }
    *out = p;
    return MAAVSS_OK;
}

extern "C" void maavss_stft_plan_destroy(maavss_stft_plan* p) {
    if (!p) return;
    if (p->tw) cudaFree(p->tw);
    if (p->window) cudaFree(p->window);
    free(p);
}

extern "C" int maavss_stft_out_dims(int n_samples, int n_fft, int hop, int trim_stft_end, int* T_out, int* F_out) {
    MAAVSS_CHECK_ARG(T_out && F_out, "stft_out_dims: null argument");
    MAAVSS_CHECK_ARG(n_samples > n_fft / 2, "stft_out_dims: reflect padding needs n_samples (%d) > n_fft/2 (%d)", n_samples, n_fft / 2);
    *T_out = n_samples / hop;                       // 1 + N//hop frames, last one dropped (av_dataset.py:171-174)
    *F_out = trim_stft_end ? n_fft / 2 : n_fft / 2 + 1;
    return MAAVSS_OK;
}

#define DISPATCH_M(M_, CALL)                        \
    switch (M_) {                                   \
        case 32: { constexpr int MM = 32; CALL; } break;     \
        case 64: { constexpr int MM = 64; CALL; } break;     \
        case 128: { constexpr int MM = 128; CALL; } break;   \
        case 256: { constexpr int MM = 256; CALL; } break;   \
        case 512: { constexpr int MM = 512; CALL; } break;   \
        case 1024: { constexpr int MM = 1024; CALL; } break; \
        case 2048: { constexpr int MM = 2048; CALL; } break; \
        default: maavss_set_error("unsupported n_fft"); return MAAVSS_ERR_UNSUPPORTED; \
    }

extern "C" int maavss_stft_fwd(const maavss_stft_plan* p, const float* audio, float* planes, int batch,
                               int n_samples, int trim_stft_end, int normalized, int polar, void* stream) {
    if (batch == 0) return MAAVSS_OK;   // empty batch: torch hands out null data pointers
    MAAVSS_CHECK_ARG(p && audio && planes, "stft_fwd: null argument");
    MAAVSS_CHECK_ARG(batch > 0 && batch <= 65535, "stft_fwd: batch=%d out of range [0, 65535]", batch);
    StftArgs a;
    int rc = maavss_stft_out_dims(n_samples, p->n_fft, p->hop, trim_stft_end, &a.T_out, &a.F_out);
    if (rc) return rc;
    if (a.T_out == 0) return MAAVSS_OK;
    a.audio = audio;
    a.planes = planes;
    a.tw = p->tw;
    a.window = p->window;
    a.n_samples = n_samples;
    a.hop = p->hop;
    a.polar = polar;
    a.scale = normalized ? (float)(1.0 / sqrt(p->sum_w2)) : 1.0f;
    a.batches_per_cta = 1;
    cudaStream_t st = (cudaStream_t)stream;
    DISPATCH_M(p->n_fft / 2, rc = launch_stft<MM>(p, a, batch, st));
    return rc;
}

extern "C" int maavss_istft_fwd(const maavss_stft_plan* p, const float* planes, float* audio, int batch,
                                int n_frames, int trim_stft_end, int normalized, int polar, void* stream) {
    if (batch == 0 || n_frames <= 1) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(p && planes && audio, "istft_fwd: null argument");
    MAAVSS_CHECK_ARG(batch > 0 && batch <= 65535, "istft_fwd: batch=%d out of range [0, 65535]", batch);
    IstftArgs a;
    a.planes = planes;
    a.audio = audio;
    a.tw = p->tw;
    a.window = p->window;
    a.hop = p->hop;
    a.T = n_frames;
    a.F_in = trim_stft_end ? p->n_fft / 2 : p->n_fft / 2 + 1;
    a.polar = polar;
    a.out_len = p->hop * (n_frames - 1);
    a.scale = (float)((normalized ? sqrt((double)p->n_fft) : 1.0) / (double)(p->n_fft / 2));
    a.G = 1;
    cudaStream_t st = (cudaStream_t)stream;
    int rc = MAAVSS_OK;
    DISPATCH_M(p->n_fft / 2, rc = launch_istft<MM>(a, batch, st));
    return rc;
}

extern "C" int maavss_absmax_normalise(float* x, int batch, long long per_clip, float eps, void* workspace, void* stream) {
    MAAVSS_CHECK_ARG(x && workspace, "absmax_normalise: null argument");
    if (batch == 0 || per_clip == 0) return MAAVSS_OK;
    cudaStream_t st = (cudaStream_t)stream;
    MAAVSS_CUDA(cudaMemsetAsync(workspace, 0, sizeof(unsigned int) * batch, st));
    int bx = (int)min((long long)ceil_div(per_clip, 256 * 8), 296LL);
    dim3 grid(bx, batch);
    absmax_kernel<<<grid, 256, 0, st>>>(x, per_clip, eps, (unsigned int*)workspace);
    MAAVSS_LAUNCH_CHECK();
    scale_by_inv_kernel<<<grid, 256, 0, st>>>(x, per_clip, (const unsigned int*)workspace);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_add_noise(const float* y, const float* noise, float noise_std, float* x, long long n, void* stream) {
    MAAVSS_CHECK_ARG(y && noise && x, "add_noise: null argument");
    if (n == 0) return MAAVSS_OK;
    int blocks = (int)min((long long)ceil_div(n, 256 * 4), 148LL * 16);
    axpy_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(y, noise, noise_std, x, n);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
