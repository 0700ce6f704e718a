// In-register / shared-memory Stockham FFT building blocks (fp32, forward sign exp(-2*pi*i*..)).
//
// One complex FFT of M points is done by TPF = M/16 threads, 16 points per thread, as
// 2 or 3 autosort passes of radix 16 / R1 / R2 with one shared-memory exchange between
// passes.  In every pass thread i owns elements {i + m*TPF, m = 0..15} of the current
// array, whatever the radix, so the load pattern is always the same coalesced stride.
// Twiddles come from a shared-memory table tw[q] = exp(-2*pi*i*q/(2M)), q in [0, M).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

// Every shared / global access of the STFT phase functions goes through these two hooks: plain indexing in the product,
// range-checked accessors in the bounds-instrumented debug build of stft.cu (-DMAAVSS_STFT_BOUNDS, see DESIGN.md
// "STFT n_fft 4096 fault").
#ifndef MAAVSS_AT
#define MAAVSS_AT(p, i) ((p)[(i)])
#define MAAVSS_DEREF(ptr) (*(ptr))
#endif

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ float2 cmul_negi(float2 a) { return make_float2(a.y, -a.x); }  // a * (-i)

// Shared-memory exchange buffers are indexed through pad_idx: one float2 of padding after every 16
// keeps the stride-16 scatter of the first radix-16 pass (and the stride-NS scatters after it) off a
// single bank pair.  A buffer for M points needs fft_buf_len(M) float2.
__device__ __forceinline__ int pad_idx(int x) { return x + (x >> 4); }
constexpr int fft_buf_len(int M) { return M + M / 16; }

template <int R>
struct Dft;

template <>
struct Dft<2> {
    __device__ static __forceinline__ void run(float2* u) {
        float2 a = u[0], b = u[1];
        u[0] = cadd(a, b);
        u[1] = csub(a, b);
    }
};

template <>
struct Dft<4> {
    __device__ static __forceinline__ void run(float2* u) {
        float2 t0 = cadd(u[0], u[2]), t1 = csub(u[0], u[2]);
        float2 t2 = cadd(u[1], u[3]), t3 = cmul_negi(csub(u[1], u[3]));
        u[0] = cadd(t0, t2);
        u[1] = cadd(t1, t3);
        u[2] = csub(t0, t2);
        u[3] = csub(t1, t3);
    }
};

template <>
struct Dft<8> {
    __device__ static __forceinline__ void run(float2* u) {
        float2 e[4] = {u[0], u[2], u[4], u[6]};
        float2 o[4] = {u[1], u[3], u[5], u[7]};
        Dft<4>::run(e);
        Dft<4>::run(o);
        const float h = 0.70710678118654752440f;
        o[1] = make_float2(h * (o[1].x + o[1].y), h * (o[1].y - o[1].x));    // * exp(-i*pi/4)
        o[2] = cmul_negi(o[2]);                                             // * exp(-i*pi/2)
        o[3] = make_float2(h * (o[3].y - o[3].x), -h * (o[3].x + o[3].y));   // * exp(-3i*pi/4)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            u[k] = cadd(e[k], o[k]);
            u[k + 4] = csub(e[k], o[k]);
        }
    }
};

template <>
struct Dft<16> {
    __device__ static __forceinline__ void run(float2* u) {
        float2 e[8], o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            e[k] = u[2 * k];
            o[k] = u[2 * k + 1];
        }
        Dft<8>::run(e);
        Dft<8>::run(o);
        // W16^k = (cos(pi*k/8), -sin(pi*k/8))
        const float c1 = 0.92387953251128675613f, s1 = 0.38268343236508977173f;
        const float h = 0.70710678118654752440f;
        o[1] = cmul(o[1], make_float2(c1, -s1));
        o[2] = make_float2(h * (o[2].x + o[2].y), h * (o[2].y - o[2].x));
        o[3] = cmul(o[3], make_float2(s1, -c1));
        o[4] = cmul_negi(o[4]);
        o[5] = cmul(o[5], make_float2(-s1, -c1));
        o[6] = make_float2(h * (o[6].y - o[6].x), -h * (o[6].x + o[6].y));
        o[7] = cmul(o[7], make_float2(-c1, -s1));
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            u[k] = cadd(e[k], o[k]);
            u[k + 8] = csub(e[k], o[k]);
        }
    }
};

// tw table lookup for exp(-2*pi*i*q/(2M)) with q in [0, 2M): second half is the negated first.
template <int M>
__device__ __forceinline__ float2 tw_lookup(const float2* __restrict__ tw, int q) {
    float2 w = tw[q & (M - 1)];
    if (q & M) w = make_float2(-w.x, -w.y);
    return w;
}

// One Stockham pass: radix R, NS = product of the radices of earlier passes.
// v[16]: this thread's elements {i + m*TPF}.  Results are scattered to out (shared memory).
template <int M, int R, int NS>
__device__ __forceinline__ void fft_pass(const float2 (&v)[16], int i, const float2* __restrict__ tw,
                                         float2* __restrict__ out) {
    constexpr int TPF = M / 16;
    constexpr int NB = 16 / R;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        const int j = i + b * TPF;
        const int k = j & (NS - 1);
        float2 u[R];
#pragma unroll
        for (int r = 0; r < R; ++r) u[r] = v[b + r * NB];
        if constexpr (NS > 1) {
            // W^(r*k), W = exp(-2*pi*i/(NS*R)); table resolution is 2M.
            const int q1 = k * ((2 * M) / (NS * R));
            float2 w[R];
            w[1] = tw_lookup<M>(tw, q1);
            if constexpr (R > 2) {
                w[2] = tw_lookup<M>(tw, 2 * q1);
                w[3] = cmul(w[1], w[2]);
            }
            if constexpr (R > 4) {
                w[4] = tw_lookup<M>(tw, 4 * q1);
                w[5] = cmul(w[1], w[4]);
                w[6] = cmul(w[2], w[4]);
                w[7] = cmul(w[3], w[4]);
            }
            if constexpr (R > 8) {
                w[8] = tw_lookup<M>(tw, 8 * q1);
#pragma unroll
                for (int r = 9; r < 16; ++r) w[r] = cmul(w[r - 8], w[8]);
            }
#pragma unroll
            for (int r = 1; r < R; ++r) u[r] = cmul(u[r], w[r]);
        }
        Dft<R>::run(u);
        const int j0 = (j / NS) * (NS * R) + k;
#pragma unroll
        for (int q = 0; q < R; ++q) out[pad_idx(j0 + q * NS)] = u[q];
    }
}

template <int M>
__device__ __forceinline__ void fft_reload(float2 (&v)[16], int i, const float2* __restrict__ buf) {
#pragma unroll
    for (int m = 0; m < 16; ++m) v[m] = MAAVSS_AT(buf, pad_idx(i + m * (M / 16)));
}

// Full M-point forward FFT of the 16 values per thread in v (elements i + m*M/16); the natural-
// order result is left in buf[pad_idx(0..M)).  All threads of the block must call (contains barriers).
template <int M>
__device__ __forceinline__ void fft_forward_block(float2 (&v)[16], int i, const float2* __restrict__ tw,
                                                  float2* __restrict__ buf) {
    constexpr int R1 = (M / 16 >= 16) ? 16 : (M / 16);
    constexpr int R2 = M / (16 * R1);
    fft_pass<M, 16, 1>(v, i, tw, buf);
    __syncthreads();
    fft_reload<M>(v, i, buf);
    __syncthreads();
    fft_pass<M, R1, 16>(v, i, tw, buf);
    if constexpr (R2 > 1) {
        __syncthreads();
        fft_reload<M>(v, i, buf);
        __syncthreads();
        fft_pass<M, R2, 16 * R1>(v, i, tw, buf);
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------
// Table-twiddle variant.  Every twiddle a thread needs in passes >= 2 depends only on its thread
// index i (element j = i + b*TPF is fixed), so the plan tabulates them per thread:
//     ttab[(slot * TPF) + i],  slot = SLOT0 + b*(R-1) + (r-1)  ->  exp(-2*pi*i * r*k / (NS*R)),  k = j & (NS-1)
// One LDS.64 per twiddle replaces the index arithmetic + cmul chains of fft_pass.
template <int M>
struct FftShape {
    static constexpr int TPF = M / 16;
    static constexpr int R1 = (M / 16 >= 16) ? 16 : (M / 16);
    static constexpr int R2 = M / (16 * R1);
    static constexpr int SLOTS1 = 16 - 16 / R1;
    static constexpr int SLOTS2 = (R2 > 1) ? 16 - 16 / R2 : 0;
    static constexpr int NSLOT = SLOTS1 + SLOTS2;
};

// Host side: fill ttab (NSLOT * TPF float2) for an M-point FFT.  M is a runtime value here so the
// plan and the CPU emulation share it.
static inline int fft_ttab_len(int M) {
    const int R1 = (M / 16 >= 16) ? 16 : (M / 16), R2 = M / (16 * R1);
    return ((16 - 16 / R1) + (R2 > 1 ? 16 - 16 / R2 : 0)) * (M / 16);
}
static inline void fft_build_ttab(int M, float2* ttab) {
    const int TPF = M / 16;
    const int R1 = (M / 16 >= 16) ? 16 : (M / 16), R2 = M / (16 * R1);
    int slot0 = 0;
    const int radix[2] = {R1, R2};
    int NS = 16;
    for (int p = 0; p < 2; ++p) {
        const int R = radix[p];
        if (R <= 1) break;
        const int NB = 16 / R;
        for (int b = 0; b < NB; ++b)
            for (int r = 1; r < R; ++r)
                for (int i = 0; i < TPF; ++i) {
                    const int k = (i + b * TPF) & (NS - 1);
                    const double ang = -2.0 * 3.14159265358979323846 * (double)(r * k) / (double)(NS * R);
                    ttab[(slot0 + b * (R - 1) + (r - 1)) * TPF + i] = make_float2((float)cos(ang), (float)sin(ang));
                }
        slot0 += NB * (R - 1);
        NS *= R;
    }
}

// One Stockham pass with tabulated twiddles.  PAD selects the padded exchange layout (needed for
// every pass but the last, whose scatter is already lane-contiguous) or a plain natural-order array.
template <int M, int R, int NS, int SLOT0, bool PAD>
__device__ __forceinline__ void fft_pass_t(const float2 (&v)[16], int i, const float2* __restrict__ ttab,
                                           float2* __restrict__ out) {
    constexpr int TPF = M / 16;
    constexpr int NB = 16 / R;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        const int j = i + b * TPF;
        const int k = j & (NS - 1);
        float2 u[R];
#pragma unroll
        for (int r = 0; r < R; ++r) u[r] = v[b + r * NB];
        if constexpr (NS > 1) {
#pragma unroll
            for (int r = 1; r < R; ++r) u[r] = cmul(u[r], MAAVSS_AT(ttab, (SLOT0 + b * (R - 1) + (r - 1)) * TPF + i));
        }
        Dft<R>::run(u);
        const int j0 = (j / NS) * (NS * R) + k;
#pragma unroll
        for (int q = 0; q < R; ++q) {
            const int x = j0 + q * NS;
            MAAVSS_AT(out, PAD ? pad_idx(x) : x) = u[q];
        }
    }
}
