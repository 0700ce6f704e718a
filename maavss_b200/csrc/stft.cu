// Fused STFT / iSTFT kernels (/root/reference/av_dataset.py:157-201).
//
// STFT:  reflect-pad + frame + Hamming window + real FFT (packed as an n/2-point complex FFT) +
//        1/sqrt(sum w^2) + last-frame / Nyquist trim + (2,T,F) plane layout (+ optional mag/phase),
//        one pass over HBM: reads 4*N bytes, writes 8*F*T bytes per clip.
// iSTFT: Nyquist zero re-pad + Hermitian unpack + inverse FFT + sqrt(n) + window + gather-style
//        (deterministic, atomics-free) overlap-add in a shared-memory ring + 1/sum w^2 envelope + centre trim.
//
// Both kernels are instruction-issue bound before they are HBM bound (an fp32 FFT costs about as many
// issue slots per frame as the frame's bytes cost HBM cycles), so the design minimises instructions:
// 16 points per thread, radix-16 register butterflies, every twiddle / window value a thread needs is
// thread-constant and read with one LDS.64 from a per-thread table, scaling is folded into the tables,
// and a frame's TPF = n_fft/32 threads synchronise among themselves (warp sync or a named barrier),
// never with the whole CTA in the forward kernel.
#include <math.h>
#include <stdlib.h>
#include "common.cuh"

// ---- bounds-instrumented debug build (python -m maavss_b200.build --variant stftdbg): every shared / global access of the
// forward kernel is checked against the CTA's dynamic shared window and the caller's buffers; the first violation is
// recorded (and the access redirected to a dummy word) instead of faulting.  Used to root-cause the n_fft 4096 fault.
#ifdef MAAVSS_STFT_BOUNDS
struct MaavssStftDbg {
    unsigned long long lo[6], hi[6];     // valid global ranges [lo, hi)
    int n_ranges;
    unsigned int count;                  // violations seen
    int rec[12];                         // first violation: line, blockIdx.x, blockIdx.y, threadIdx.x, is_shared, offset lo/hi, limit, size
    float4 dummy[2];
};
__device__ MaavssStftDbg g_stft_dbg;
extern __shared__ float2 maavss_dbg_smem[];
template <typename T>
__device__ __noinline__ T* maavss_dbg_chk(T* p, int line) {
    bool ok = false;
    long long off = 0, limit = 0;
    int shared = 0;
    if (__isShared(p)) {
        shared = 1;
        unsigned int dyn;
        asm volatile("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
        off = (const char*)p - (const char*)maavss_dbg_smem;
        limit = dyn;
        ok = off >= 0 && off + (long long)sizeof(T) <= limit;
    } else {
        const unsigned long long a = (unsigned long long)p;
        for (int r = 0; r < g_stft_dbg.n_ranges; ++r)
            if (a >= g_stft_dbg.lo[r] && a + sizeof(T) <= g_stft_dbg.hi[r]) ok = true;
        off = (long long)a;
    }
    if (ok) return p;
    if (atomicAdd(&g_stft_dbg.count, 1u) == 0) {
        int* r = g_stft_dbg.rec;
        r[0] = line; r[1] = blockIdx.x; r[2] = blockIdx.y; r[3] = threadIdx.x; r[4] = shared;
        r[5] = (int)(off & 0xffffffffll); r[6] = (int)(off >> 32); r[7] = (int)limit; r[8] = (int)sizeof(T);
    }
    return reinterpret_cast<T*>(const_cast<float4*>(g_stft_dbg.dummy));
}
#define MAAVSS_AT(p, i) (*maavss_dbg_chk(&(p)[(i)], __LINE__))
#define MAAVSS_DEREF(ptr) (*maavss_dbg_chk((ptr), __LINE__))
#endif
#ifndef MAAVSS_STFT_MINBLOCKS
#define MAAVSS_STFT_MINBLOCKS 0      // 0: per-size choice (stft_fwd_kernel); the debug variants pin it
#endif

#include "stft_phases.cuh"
#include "../../include/maavss_b200.h"

struct maavss_stft_plan {
    int n_fft, hop;
    float2* tw;      // device, n_fft/2 entries: exp(-2*pi*i*q/n_fft)
    float2* ttab;    // device, per-thread pass twiddles (fft_build_ttab)
    float* window;   // device, n_fft
    float* ienv;     // device, hop entries: 1 / sum_q w^2[r + q*hop]
    double sum_w2;
    int device;
};

namespace {

static inline int round_up(int a, int b) { return (a + b - 1) / b * b; }

template <int M>
struct Cfg {
    static constexpr int TPF = M / 16;
    static constexpr int NT = (M >= 2048) ? 512 : 256;        // threads per CTA (inverse); forward uses 256
    static constexpr int TTAB = FftShape<M>::NSLOT * TPF;       // float2 entries
};

// Synchronise the TPF threads that share one frame: a warp sync, or one named barrier per frame group (ids 1..4; a CTA holds
// at most 4 groups of more than one warp).  The id is an IMMEDIATE on purpose.  With `bar.sync %r, n` ptxas cannot tell which
// barriers the kernel uses and reserves all 16 of the CTA ("used 16 barriers"); the n_fft 4096 / 2048 forward kernels built
// that way for 128 registers raised an illegal-address error as soon as two CTAs shared an SM and a CTA ran more than one
// batch of frames (round 1's "known issue").  Bisected on a B200 in round 2 (gpurun_out -> profiles/r2_stft4096_rootcause.txt):
// every shared / global index of the kernel is in range (bounds-instrumented build, 0 violations), and the fault disappears
// with immediate ids (3 barriers reserved), with one CTA per SM, or with a CTA-wide barrier -- nothing else was changed.
template <int TPF>
__device__ __forceinline__ void group_sync(int fi) {
    if constexpr (TPF <= 32) {
        __syncwarp();
    } else {
#ifdef MAAVSS_STFT_REG_BARRIER     // debug variants only (build.py): the register-indexed form that faulted
        asm volatile("bar.sync %0, %1;" ::"r"(fi + 1), "n"(TPF) : "memory");
#else
        switch (fi) {
            case 0: asm volatile("bar.sync 1, %0;" ::"n"(TPF) : "memory"); break;
            case 1: asm volatile("bar.sync 2, %0;" ::"n"(TPF) : "memory"); break;
            case 2: asm volatile("bar.sync 3, %0;" ::"n"(TPF) : "memory"); break;
            default: asm volatile("bar.sync 4, %0;" ::"n"(TPF) : "memory"); break;
        }
#endif
    }
}

// M-point forward FFT of v (elements i + m*TPF); result in natural order in buf (padded layout when
// PADFINAL, plain otherwise).  Only the frame's own threads synchronise.
template <int M, bool PADFINAL>
__device__ __forceinline__ void fft_group(float2 (&v)[16], int i, int fi, const float2* __restrict__ ttab,
                                          float2* __restrict__ buf) {
    using S = FftShape<M>;
    constexpr int TPF = S::TPF;
    fft_pass_t<M, 16, 1, 0, true>(v, i, ttab, buf);
    group_sync<TPF>(fi);
    fft_reload<M>(v, i, buf);
    group_sync<TPF>(fi);
    if constexpr (S::R2 > 1) {
        fft_pass_t<M, S::R1, 16, 0, true>(v, i, ttab, buf);
        group_sync<TPF>(fi);
        fft_reload<M>(v, i, buf);
        group_sync<TPF>(fi);
        fft_pass_t<M, S::R2, 16 * S::R1, S::SLOTS1, PADFINAL>(v, i, ttab, buf);
    } else {
        fft_pass_t<M, S::R1, 16, 0, PADFINAL>(v, i, ttab, buf);
    }
    group_sync<TPF>(fi);
}

constexpr int kFwdThreads = 256;

struct StftArgs {
    const float* audio;
    float* planes;
    const float2* tw;
    const float2* ttab;
    const float* window;
    int n_samples, hop, T_out, F_out, batches_per_cta, polar;
    float scale;
};

template <int M>
// CTAs per SM the register budget is set for: shared memory holds 2 CTAs at M = 2048 (97 KB each), so 128 registers cost
// nothing there and remove the spills of the 80-register build
__global__ void __launch_bounds__(kFwdThreads, (M <= 256) ? 4 : (MAAVSS_STFT_MINBLOCKS ? MAAVSS_STFT_MINBLOCKS : (M >= 2048 ? 2 : 3)))
stft_fwd_kernel(StftArgs a) {
    constexpr int TPF = M / 16;
    constexpr int P = kFwdThreads / TPF;
    extern __shared__ float2 smem[];
    float2* tw = smem;                       // M
    float2* wsm = tw + M;                    // M   (window pairs, pre-scaled)
    float2* ttab = wsm + M;                  // Cfg<M>::TTAB
    float2* bufs = ttab + Cfg<M>::TTAB;      // P * fft_buf_len(M)
    const float hs = 0.5f * a.scale;
    for (int q = threadIdx.x; q < M; q += kFwdThreads) {
        MAAVSS_AT(tw, q) = MAAVSS_AT(a.tw, q);
        const float2 w = MAAVSS_DEREF(reinterpret_cast<const float2*>(a.window + 2 * q));
        MAAVSS_AT(wsm, q) = make_float2(w.x * hs, w.y * hs);
    }
    for (int q = threadIdx.x; q < Cfg<M>::TTAB; q += kFwdThreads) MAAVSS_AT(ttab, q) = MAAVSS_AT(a.ttab, q);
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
            stft_load_frame<M>(v, i, clip, wsm, t, a.hop, a.n_samples);
        } else {
#pragma unroll
            for (int m = 0; m < 16; ++m) v[m] = make_float2(0.f, 0.f);
        }
        fft_group<M, true>(v, i, fi, ttab, buf);
        if (active) {
            float* row0 = out + (size_t)t * a.F_out;
            stft_post_frame<M>(i, buf, tw, row0, row0 + (size_t)a.T_out * a.F_out, a.F_out, a.polar);
        }
        group_sync<TPF>(fi);
    }
}

struct IstftArgs {
    const float* planes;
    float* audio;
    const float2* tw;
    const float2* ttab;
    const float* window;
    const float* ienv;
    int hop, T, F_in, G, polar, out_len, log2C;
    float scale;
};

// One CTA owns G hop-slots of one clip: padded positions [t_a*hop, t_b*hop).  It walks the frames
// [t_a - (N-1)/hop, min(t_b, T)) in batches of P, inverse-transforms each batch into shared memory
// and folds it into the overlap-add ring (istft_gather_elem).
// V = vector width of the generic gather (any hop); R > 0 selects the column-owning gather (hop a power of two,
// R = n_fft / hop, V = 4).
template <int M, int V, int R>
__global__ void __launch_bounds__(Cfg<M>::NT, (M <= 256) ? 4 : (M <= 1024 ? 2 : 1)) istft_kernel(IstftArgs a) {
    constexpr int TPF = M / 16;
    constexpr int NT = Cfg<M>::NT;
    constexpr int P = NT / TPF;
    constexpr int N = 2 * M;
    constexpr int FS = 2 * fft_buf_len(M);          // floats between frames
    extern __shared__ float2 smem[];
    float2* tw = smem;                               // M
    float2* ttab = tw + M;                           // TTAB
    float2* bufs = ttab + Cfg<M>::TTAB;              // P * fft_buf_len(M)
    float* wg = reinterpret_cast<float*>(bufs + P * fft_buf_len(M));   // N
    float* acc = wg + N;                             // L
    const int tid = threadIdx.x;
    const int fi = tid / TPF, i = tid % TPF;
    const int hop = a.hop;
    IstftGather c;
    c.hop = hop; c.N = N; c.M = M; c.T = a.T;
    c.PH = P * hop;
    c.L = (P - 1) * hop + N;
    c.frame_stride = FS;
    c.inv_hop = 1.0f / (float)hop;
    const int t_a = blockIdx.x * a.G, t_b = t_a + a.G;
    c.own_lo = max(M, t_a * hop);
    c.own_hi = min(M + a.out_len, t_b * hop);
    const int t_lim = min(t_b, a.T);                 // frames >= t_lim are not needed / do not exist
    const int t_start = max(0, t_a - (N - 1) / hop);
    for (int q = tid; q < M; q += NT) tw[q] = a.tw[q];
    for (int q = tid; q < Cfg<M>::TTAB; q += NT) ttab[q] = a.ttab[q];
    const float sc = 0.5f * a.scale;
    for (int q = tid; q < N; q += NT) wg[q] = a.window[q] * ((q & 1) ? -sc : sc);
    for (int q = tid; q < c.L; q += NT) acc[q] = 0.f;
    const float inv_scale2 = 1.0f / (sc * sc);
    const float* in0 = a.planes + (size_t)blockIdx.y * 2 * a.T * a.F_in;
    const float* in1 = in0 + (size_t)a.T * a.F_in;
    float* out_clip = a.audio + (size_t)blockIdx.y * a.out_len;
    float2* buf = bufs + fi * fft_buf_len(M);
    c.off = 0;
    constexpr int RR = (R > 0) ? R : 1;
    constexpr bool WREG = (R > 0 && R <= 4);
    float4 wr[RR];
    float4 ie = make_float4(0.f, 0.f, 0.f, 0.f);
    if constexpr (R > 0) {
        const int col4 = (tid & ((hop >> 2) - 1)) * 4;
        ie = *reinterpret_cast<const float4*>(a.ienv + col4);
#pragma unroll
        for (int s = 0; s < RR; ++s) {
            const float4 w = *reinterpret_cast<const float4*>(a.window + col4 + s * hop);
            wr[s] = make_float4(w.x * sc, -w.y * sc, w.z * sc, -w.w * sc);   // == wg[col4 + s*hop ..]
        }
    }
    IstftRaw x;
    istft_fetch_raw<M>(x, i, in0 + (size_t)(t_start + fi) * a.F_in, in1 + (size_t)(t_start + fi) * a.F_in, a.F_in,
                       t_start + fi < t_lim);
    __syncthreads();
    for (int tb = t_start; tb * hop < c.own_hi; tb += P) {
        // a warp (or barrier group) whose frames are all beyond t_lim skips the transform altogether
        const int t_group0 = (TPF >= 32) ? tb + fi : tb + (tid & ~31) / TPF;
        if (t_group0 < t_lim) {
            istft_build_z<M>(x, i, tw, a.polar, buf);
            group_sync<TPF>(fi);
            float2 v[16];
            fft_reload<M>(v, i, buf);
            group_sync<TPF>(fi);
            fft_group<M, false>(v, i, fi, ttab, buf);
        }
        __syncthreads();
        {   // raw bins of the next batch travel while this batch is folded into the ring
            const int tn = tb + P + fi;
            istft_fetch_raw<M>(x, i, in0 + (size_t)tn * a.F_in, in1 + (size_t)tn * a.F_in, a.F_in,
                               tn < t_lim && (tb + P) * hop < c.own_hi);
        }
        c.base_u = tb * hop;
        c.nvalid = max(0, min(P, t_lim - tb));
        if constexpr (R > 0) {
            istft_gather_cols<RR, WREG>(tid, NT, a.log2C, P, c, acc, reinterpret_cast<const float*>(bufs), wg, wr, ie,
                                        inv_scale2, out_clip);
        } else {
            for (int e = tid; e < c.L / V; e += NT)
                istft_gather_elem<V>(e, c, acc, reinterpret_cast<const float*>(bufs), wg, a.ienv, inv_scale2, out_clip);
        }
        c.off += c.PH;
        if (c.off >= c.L) c.off -= c.L;
        __syncthreads();
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
    constexpr int P = kFwdThreads / (M / 16);
    const int batches = ceil_div(a.T_out, P);
    // enough CTAs for ~8 waves of 3 resident CTAs per SM; each CTA amortises its table prologue over bpc batches
    long long bpc = ceil_div((long long)batches * batch, 148LL * 3 * 8);
    if (bpc < 1) bpc = 1;
    if (bpc > 16) bpc = 16;
    if (bpc > batches) bpc = batches;
#ifdef MAAVSS_STFT_MINBLOCKS_DEBUG   // bisect switch of the debug variants only
    if (const char* e = getenv("MAAVSS_STFT_BPC")) bpc = atoi(e) > 0 ? atoi(e) : bpc;
#endif
    a.batches_per_cta = (int)bpc;
    dim3 grid(ceil_div(batches, a.batches_per_cta), batch);
    size_t smem = sizeof(float2) * (2 * (size_t)M + Cfg<M>::TTAB + (size_t)P * fft_buf_len(M));
#ifdef MAAVSS_STFT_MINBLOCKS_DEBUG   // bisect switch: pad the allocation so that only one CTA fits an SM
    if (const char* e = getenv("MAAVSS_STFT_SMEM_PAD")) smem += (size_t)atoi(e);
#endif
    MAAVSS_CUDA(cudaFuncSetAttribute(stft_fwd_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    stft_fwd_kernel<M><<<grid, kFwdThreads, smem, st>>>(a);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

template <int M, int V, int R>
int launch_istft_v(const IstftArgs& a, int batch, cudaStream_t st) {
    constexpr int NT = Cfg<M>::NT;
    constexpr int P = NT / (M / 16);
    const int slots = ceil_div(M + a.out_len, a.hop);
    dim3 grid(ceil_div(slots, a.G), batch);
    const size_t L = (size_t)(P - 1) * a.hop + 2 * M;
    const size_t smem = sizeof(float2) * ((size_t)M + Cfg<M>::TTAB + (size_t)P * fft_buf_len(M)) + sizeof(float) * (2 * M + L);
    MAAVSS_CUDA(cudaFuncSetAttribute(istft_kernel<M, V, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    istft_kernel<M, V, R><<<grid, NT, smem, st>>>(a);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

template <int M>
int launch_istft(const IstftArgs& a0, int batch, cudaStream_t st) {
    IstftArgs a = a0;
    constexpr int P = Cfg<M>::NT / (M / 16);
    // G hop-slots per CTA: the (n_fft-1)/hop warm-up frames of a strip are recomputed by its neighbour, so
    // strips are long (<= 6 % recompute) unless the launch would leave SMs idle.
    const int slots = ceil_div(M + a.out_len, a.hop);
    const int warm = (2 * M - 1) / a.hop;
    int G = round_up(max(128, 16 * warm), P);
    while (G > P && (long long)ceil_div(slots, G) * batch < 148LL * 4) G = round_up(G / 2, P);
    a.G = G;
    const uintptr_t al = reinterpret_cast<uintptr_t>(a.audio);
    // the column-owning gather keeps R window vectors in registers: it pays for M >= 512 (measured), where the
    // kernel is not held to 64 registers, and for R <= 4
    if constexpr (M >= 512) {
        const bool pow2 = (a.hop & (a.hop - 1)) == 0;
        if (pow2 && a.hop >= 4 && a.hop / 4 <= Cfg<M>::NT && al % 16 == 0 && (2 * M) % a.hop == 0) {
            a.log2C = 0;
            while ((4 << a.log2C) < a.hop) ++a.log2C;
            switch (2 * M / a.hop) {
                case 1: return launch_istft_v<M, 4, 1>(a, batch, st);
                case 2: return launch_istft_v<M, 4, 2>(a, batch, st);
                case 4: return launch_istft_v<M, 4, 4>(a, batch, st);
                default: break;
            }
        }
    }
    if (a.hop % 4 == 0 && al % 16 == 0) return launch_istft_v<M, 4, 0>(a, batch, st);
    if (a.hop % 2 == 0 && al % 8 == 0) return launch_istft_v<M, 2, 0>(a, batch, st);
    return launch_istft_v<M, 1, 0>(a, batch, st);
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
    const int n_tt = fft_ttab_len(M);
    float2* tw_h = (float2*)malloc(sizeof(float2) * M);
    float2* tt_h = (float2*)malloc(sizeof(float2) * n_tt);
    float* ie_h = (float*)malloc(sizeof(float) * hop);
    for (int q = 0; q < M; ++q) {
        const double ang = -2.0 * M_PI * (double)q / (double)n_fft;
        tw_h[q] = make_float2((float)cos(ang), (float)sin(ang));
    }
    fft_build_ttab(M, tt_h);
    double s = 0.0;
    for (int j = 0; j < n_fft; ++j) s += (double)window_host[j] * (double)window_host[j];
    p->sum_w2 = s;
    for (int r = 0; r < hop; ++r) {
        double e = 0.0;
        for (int j = r; j < n_fft; j += hop) e += (double)window_host[j] * (double)window_host[j];
        ie_h[r] = (float)(1.0 / e);
    }
    cudaGetDevice(&p->device);
    cudaError_t e = cudaMalloc(&p->tw, sizeof(float2) * M);
    if (e == cudaSuccess) e = cudaMalloc(&p->ttab, sizeof(float2) * n_tt);
    if (e == cudaSuccess) e = cudaMalloc(&p->window, sizeof(float) * n_fft);
    if (e == cudaSuccess) e = cudaMalloc(&p->ienv, sizeof(float) * hop);
    if (e == cudaSuccess) e = cudaMemcpy(p->tw, tw_h, sizeof(float2) * M, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->ttab, tt_h, sizeof(float2) * n_tt, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->window, window_host, sizeof(float) * n_fft, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->ienv, ie_h, sizeof(float) * hop, cudaMemcpyHostToDevice);
    free(tw_h);
    free(tt_h);
    free(ie_h);
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
    if (p->ttab) cudaFree(p->ttab);
    if (p->ienv) cudaFree(p->ienv);
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
    a.ttab = p->ttab;
    a.window = p->window;
    a.n_samples = n_samples;
    a.hop = p->hop;
    a.polar = polar;
    a.scale = normalized ? (float)(1.0 / sqrt(p->sum_w2)) : 1.0f;
    a.batches_per_cta = 1;
    cudaStream_t st = (cudaStream_t)stream;
#ifdef MAAVSS_STFT_BOUNDS
    {
        const int M = p->n_fft / 2;
        const void* ptr[5] = {audio, planes, p->tw, p->ttab, p->window};
        const size_t len[5] = {sizeof(float) * (size_t)batch * n_samples, sizeof(float) * (size_t)batch * 2 * a.T_out * a.F_out,
                               sizeof(float2) * M, sizeof(float2) * fft_ttab_len(M), sizeof(float) * p->n_fft};
        MaavssStftDbg cur;
        MAAVSS_CUDA(cudaMemcpyFromSymbol(&cur, g_stft_dbg, sizeof(cur)));
        for (int r = 0; r < 5; ++r) {
            cur.lo[r] = (unsigned long long)ptr[r];
            cur.hi[r] = cur.lo[r] + len[r];
        }
        cur.n_ranges = 5;
        MAAVSS_CUDA(cudaMemcpyToSymbol(g_stft_dbg, &cur, sizeof(cur)));
    }
#endif
    DISPATCH_M(p->n_fft / 2, rc = launch_stft<MM>(p, a, batch, st));
    return rc;
}

#ifdef MAAVSS_STFT_BOUNDS
// debug build only (not in the public header): out[0] = violations so far, out[1..9] = the first one's record
extern "C" int maavss_stft_debug_read(int* out10, int reset) {
    MaavssStftDbg cur;
    MAAVSS_CUDA(cudaDeviceSynchronize());
    MAAVSS_CUDA(cudaMemcpyFromSymbol(&cur, g_stft_dbg, sizeof(cur)));
    out10[0] = (int)cur.count;
    for (int q = 0; q < 9; ++q) out10[1 + q] = cur.rec[q];
    if (reset) {
        cur.count = 0;
        MAAVSS_CUDA(cudaMemcpyToSymbol(g_stft_dbg, &cur, sizeof(cur)));
    }
    return MAAVSS_OK;
}
#endif

extern "C" int maavss_istft_fwd(const maavss_stft_plan* p, const float* planes, float* audio, int batch,
                                int n_frames, int trim_stft_end, int normalized, int polar, void* stream) {
    if (batch == 0 || n_frames <= 1) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(p && planes && audio, "istft_fwd: null argument");
    MAAVSS_CHECK_ARG(batch > 0 && batch <= 65535, "istft_fwd: batch=%d out of range [0, 65535]", batch);
    IstftArgs a;
    a.planes = planes;
    a.audio = audio;
    a.tw = p->tw;
    a.ttab = p->ttab;
    a.window = p->window;
    a.ienv = p->ienv;
    a.hop = p->hop;
    a.T = n_frames;
    a.F_in = trim_stft_end ? p->n_fft / 2 : p->n_fft / 2 + 1;
    a.polar = polar;
    a.out_len = p->hop * (n_frames - 1);
    a.scale = (float)((normalized ? sqrt((double)p->n_fft) : 1.0) / (double)(p->n_fft / 2));
    a.G = 1;
    a.log2C = 0;
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
