// Per-thread phases of the fused STFT / iSTFT kernels (shared by the CUDA kernels in stft.cu and
// by the host emulation in tests/emu/, which runs the same code thread-by-thread on the CPU).
//
// Reference semantics: AV_Dataset.stft / istft (/root/reference/av_dataset.py:157-201), i.e.
// torch.stft(center=True, pad_mode="reflect", onesided) / sqrt(sum w^2)   and
// torch.istft(normalized=True) = irfft(X*sqrt(n)) * w, overlap-add, / sum w^2, centre trim.
#pragma once
#include "fft_core.cuh"

// Source index into the un-padded clip for padded position p (= t*hop + j - n_fft/2).
__device__ __forceinline__ int reflect_idx(int p, int n) {
    if (p < 0) p = -p;
    if (p >= n) p = 2 * (n - 1) - p;
    return p;
}

// Load frame t as the packed complex sequence z[j] = x[2j]*w[2j] + i*x[2j+1]*w[2j+1]; thread i keeps
// elements j = i + m*TPF.
template <int M>
__device__ __forceinline__ void stft_load_frame(float2 (&v)[16], int i, const float* __restrict__ clip,
                                                const float* __restrict__ window, int t, int hop,
                                                int n_samples) {
    constexpr int TPF = M / 16;
    const int base = t * hop - M;  // pad = n_fft/2 = M
    const bool interior = (base >= 0) && (base + 2 * M <= n_samples);
#pragma unroll
    for (int m = 0; m < 16; ++m) {
        const int j = i + m * TPF;
        const float2 w = *reinterpret_cast<const float2*>(window + 2 * j);
        float x0, x1;
        if (interior) {
            x0 = clip[base + 2 * j];
            x1 = clip[base + 2 * j + 1];
        } else {
            x0 = clip[reflect_idx(base + 2 * j, n_samples)];
            x1 = clip[reflect_idx(base + 2 * j + 1, n_samples)];
        }
        v[m] = make_float2(x0 * w.x, x1 * w.y);
    }
}

// Z (M-point FFT of the packed sequence, natural order in buf) -> one-sided spectrum X[0..M] of the
// 2M-point real FFT, scaled, written as one row of the re plane and one row of the im plane
// (or |X| / atan2 when polar).  Bins >= F_out are dropped (trim_stft_end removes the Nyquist bin).
template <int M>
__device__ __forceinline__ void stft_post_frame(int i, const float2* __restrict__ buf,
                                                const float2* __restrict__ tw, float* __restrict__ row0,
                                                float* __restrict__ row1, int F_out, float scale,
                                                int polar) {
    constexpr int TPF = M / 16;
    const float hs = 0.5f * scale;
#pragma unroll
    for (int m = 0; m < 9; ++m) {
        int k = i + m * TPF;            // m < 8: k in [0, M/2); m == 8: the self-paired bin M/2
        if (m == 8) {
            if (i != 0) break;
            k = M / 2;
        }
        const float2 A = buf[pad_idx(k)];
        const float2 B = buf[pad_idx((M - k) & (M - 1))];
        const float2 S = make_float2(A.x + B.x, A.y - B.y);   // A + conj(B)
        const float2 D = make_float2(A.x - B.x, A.y + B.y);   // A - conj(B)
        const float2 WD = cmul(tw[k], D);
        float2 X0 = make_float2(hs * (S.x + WD.y), hs * (S.y - WD.x));
        float2 X1 = make_float2(hs * (S.x - WD.y), hs * (-S.y - WD.x));
        if (polar) {
            X0 = make_float2(sqrtf(X0.x * X0.x + X0.y * X0.y), atan2f(X0.y, X0.x));
            X1 = make_float2(sqrtf(X1.x * X1.x + X1.y * X1.y), atan2f(X1.y, X1.x));
        }
        if (k < F_out) {
            row0[k] = X0.x;
            row1[k] = X0.y;
        }
        const int k1 = M - k;
        if (k1 != k && k1 < F_out) {
            row0[k1] = X1.x;
            row1[k1] = X1.y;
        }
    }
}

// Build conj(Z) for the inverse: Z[k] = 0.5*((X[k] + conj(X[M-k])) + i*exp(+2*pi*i*k/n)*(X[k] - conj(X[M-k])))
// with the imaginary parts of the DC and Nyquist bins ignored (c2r semantics).  The M-point FFT of
// conj(Z) is conj(M * z) where z[j] = x[2j] + i*x[2j+1] is the real frame.
template <int M>
__device__ __forceinline__ float2 istft_fetch_bin(const float* __restrict__ row0, const float* __restrict__ row1,
                                                  int k, int F_in, int polar) {
    if (k >= F_in) return make_float2(0.f, 0.f);   // trimmed Nyquist bin = zero padding (F.pad(stft,(0,1)))
    float2 x = make_float2(row0[k], row1[k]);
    if (polar) {
        float s, c;
        sincosf(x.y, &s, &c);
        x = make_float2(x.x * c, x.x * s);
    }
    if (k == 0 || k == M) x.y = 0.f;
    return x;
}

template <int M>
__device__ __forceinline__ void istft_load_frame(float2 (&v)[16], int i, const float* __restrict__ row0,
                                                 const float* __restrict__ row1, int F_in,
                                                 const float2* __restrict__ tw, int polar) {
    constexpr int TPF = M / 16;
#pragma unroll
    for (int m = 0; m < 16; ++m) {
        const int k = i + m * TPF;
        const float2 A = istft_fetch_bin<M>(row0, row1, k, F_in, polar);
        const float2 B = istft_fetch_bin<M>(row0, row1, M - k, F_in, polar);
        const float2 S = make_float2(A.x + B.x, A.y - B.y);
        const float2 D = make_float2(A.x - B.x, A.y + B.y);
        const float2 w = tw[k];                              // exp(-2*pi*i*k/n); need its conjugate
        const float2 WD = make_float2(w.x * D.x + w.y * D.y, w.x * D.y - w.y * D.x);   // conj(w)*D
        // Z = 0.5*(S + i*WD) ; store conj(Z)
        v[m] = make_float2(0.5f * (S.x - WD.y), -0.5f * (S.y + WD.x));
    }
}

// Sample s (0 <= s < 2M) of the real frame whose FFT(conj Z) sits in buf (natural order), before
// the 1/M factor:  x[2j] = Re, x[2j+1] = -Im.
__device__ __forceinline__ float istft_frame_sample(const float2* __restrict__ buf, int s) {
    const float2 z = buf[pad_idx(s >> 1)];
    return (s & 1) ? -z.y : z.x;
}
