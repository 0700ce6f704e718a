// Per-thread phases of the fused STFT / iSTFT kernels (shared by the CUDA kernels in stft.cu and
// by the host emulation in tests/emu/, which runs the same code thread-by-thread on the CPU).
//
// Reference semantics: AV_Dataset.stft / istft (/root/reference/av_dataset.py:157-201), i.e.
// torch.stft(center=True, pad_mode="reflect", onesided) / sqrt(sum w^2)   and
// torch.istft(normalized=True) = irfft(X*sqrt(n)) * w, overlap-add, / sum w^2, centre trim.
//
// Scaling is folded into the tables the kernels keep in shared memory:
//   forward  wsm[j]  = (w[2j], w[2j+1]) * 0.5 * scale            (0.5 = the real-FFT unpack factor)
//   inverse  wg[s]   = w[s] * 0.5 * scale * (s odd ? -1 : +1)    (the FFT runs on conj(2Z); odd samples come out negated)
//            ienv[r] = 1 / sum_q w^2[r + q*hop]                   (steady-state overlap-add envelope)
#pragma once
#include "fft_core.cuh"

// Source index into the un-padded clip for padded position p (= t*hop + j - n_fft/2).
__device__ __forceinline__ int reflect_idx(int p, int n) {
    if (p < 0) p = -p;
    if (p >= n) p = 2 * (n - 1) - p;
    return p;
}

// Load frame t as the packed complex sequence z[j] = x[2j]*w[2j] + i*x[2j+1]*w[2j+1]; thread i keeps
// elements j = i + m*TPF.  wsm is the (pre-scaled) window as float2 pairs.
template <int M>
__device__ __forceinline__ void stft_load_frame(float2 (&v)[16], int i, const float* __restrict__ clip,
                                                const float2* __restrict__ wsm, int t, int hop, int n_samples) {
    constexpr int TPF = M / 16;
    const int base = t * hop - M;  // pad = n_fft/2 = M
    const bool interior = (base >= 0) && (base + 2 * M <= n_samples);
    if (interior) {
        const float* p = clip + base + 2 * i;
        if ((reinterpret_cast<uintptr_t>(p) & 7) == 0) {
#pragma unroll
            for (int m = 0; m < 16; ++m) {
                const float2 x = MAAVSS_DEREF(reinterpret_cast<const float2*>(p + 2 * m * TPF));
                const float2 w = MAAVSS_AT(wsm, i + m * TPF);
                v[m] = make_float2(x.x * w.x, x.y * w.y);
            }
        } else {
#pragma unroll
            for (int m = 0; m < 16; ++m) {
                const float2 w = MAAVSS_AT(wsm, i + m * TPF);
                v[m] = make_float2(MAAVSS_AT(p, 2 * m * TPF) * w.x, MAAVSS_AT(p, 2 * m * TPF + 1) * w.y);
            }
        }
    } else {
#pragma unroll
        for (int m = 0; m < 16; ++m) {
            const int j = i + m * TPF;
            const float2 w = MAAVSS_AT(wsm, j);
            v[m] = make_float2(MAAVSS_AT(clip, reflect_idx(base + 2 * j, n_samples)) * w.x,
                               MAAVSS_AT(clip, reflect_idx(base + 2 * j + 1, n_samples)) * w.y);
        }
    }
}

// Z (M-point FFT of the packed sequence, natural order in the padded buf) -> one-sided spectrum X[0..M]
// of the 2M-point real FFT, written as one row of the re plane and one row of the im plane (or |X| /
// atan2 when polar).  Bins >= F_out are dropped (trim_stft_end removes the Nyquist bin).  tw[k] =
// exp(-2*pi*i*k/(2M)).
template <int M>
__device__ __forceinline__ void stft_post_frame(int i, const float2* __restrict__ buf,
                                                const float2* __restrict__ tw, float* __restrict__ row0,
                                                float* __restrict__ row1, int F_out, int polar) {
    constexpr int TPF = M / 16;
#pragma unroll
    for (int m = 0; m < 9; ++m) {
        int k = i + m * TPF;            // m < 8: k in [0, M/2); m == 8: the self-paired bin M/2
        if (m == 8) {
            if (i != 0) break;
            k = M / 2;
        }
        const float2 A = MAAVSS_AT(buf, pad_idx(k));
        const float2 B = MAAVSS_AT(buf, pad_idx((M - k) & (M - 1)));
        const float2 S = make_float2(A.x + B.x, A.y - B.y);   // A + conj(B)
        const float2 D = make_float2(A.x - B.x, A.y + B.y);   // A - conj(B)
        const float2 WD = cmul(MAAVSS_AT(tw, k), D);
        float2 X0 = make_float2(S.x + WD.y, S.y - WD.x);
        float2 X1 = make_float2(S.x - WD.y, -S.y - WD.x);
        if (polar) {
            X0 = make_float2(sqrtf(X0.x * X0.x + X0.y * X0.y), atan2f(X0.y, X0.x));
            X1 = make_float2(sqrtf(X1.x * X1.x + X1.y * X1.y), atan2f(X1.y, X1.x));
        }
        if (k < F_out) {
            MAAVSS_AT(row0, k) = X0.x;
            MAAVSS_AT(row1, k) = X0.y;
        }
        const int k1 = M - k;
        if (k1 != k && k1 < F_out) {
            MAAVSS_AT(row0, k1) = X1.x;
            MAAVSS_AT(row1, k1) = X1.y;
        }
    }
}

// Build conj(Z) for the inverse: Z[k] = 0.5*((X[k] + conj(X[M-k])) + i*exp(+2*pi*i*k/n)*(X[k] - conj(X[M-k])))
// with the imaginary parts of the DC and Nyquist bins ignored (c2r semantics).  The M-point FFT of
// conj(Z) is conj(M * z) where z[j] = x[2j] + i*x[2j+1] is the real frame.
// Inverse input stage.  Thread i owns the bin pairs (k, M-k) for k = i + m*TPF, m < 8 (thread 0 also the
// self-paired bin M/2), so every bin of the frame is loaded exactly once and the Hermitian unpack shares
// S, D and W*D between Z[k] and Z[M-k].  The raw values are fetched into registers one batch ahead
// (istft_fetch_raw, pure loads), turned into conj(2Z) and scattered into the padded exchange buffer
// (istft_build_z); after a group barrier fft_reload hands every thread its column i + m*TPF.
// Bin M (Nyquist) only exists when F_in = M + 1; trim_stft_end re-pads it with zero
// (F.pad(stft,(0,1)), av_dataset.py:184).
struct IstftRaw {
    float a[16];   // (re, im) of X[k_m]
    float b[16];   // (re, im) of X[M - k_m]
    float s[2];    // thread 0: X[M/2]
};

template <int M>
__device__ __forceinline__ void istft_fetch_raw(IstftRaw& x, int i, const float* __restrict__ row0,
                                                const float* __restrict__ row1, int F_in, bool active) {
    constexpr int TPF = M / 16;
    if (!active) {
#pragma unroll
        for (int m = 0; m < 16; ++m) x.a[m] = x.b[m] = 0.f;
        x.s[0] = x.s[1] = 0.f;
        return;
    }
    const float* a0 = row0 + i;
    const float* a1 = row1 + i;
    const float* b0 = row0 + (M - i);
    const float* b1 = row1 + (M - i);
#pragma unroll
    for (int m = 0; m < 8; ++m) {
        x.a[2 * m] = a0[m * TPF];
        x.a[2 * m + 1] = a1[m * TPF];
        if (m == 0 && i == 0) {
            x.b[0] = (F_in > M) ? b0[0] : 0.f;
            x.b[1] = (F_in > M) ? b1[0] : 0.f;
        } else {
            x.b[2 * m] = b0[-m * TPF];
            x.b[2 * m + 1] = b1[-m * TPF];
        }
    }
    x.s[0] = x.s[1] = 0.f;
    if (i == 0) {
        x.s[0] = row0[M / 2];
        x.s[1] = row1[M / 2];
    }
}

__device__ __forceinline__ float2 polar_to_cart(float2 x) {
    float s, c;
    sincosf(x.y, &s, &c);
    return make_float2(x.x * c, x.x * s);
}

// conj(2 Z[k]) and conj(2 Z[M-k]) from A = X[k], B = X[M-k], w = exp(-2*pi*i*k/n):
//   S = A + conj(B), D = A - conj(B), WD = conj(w) * D,  2Z[k] = S + i*WD,  2Z[M-k] = conj(S) + i*conj(WD)
__device__ __forceinline__ void istft_pair(float2 A, float2 B, float2 w, float2& zk, float2& zp) {
    const float2 S = make_float2(A.x + B.x, A.y - B.y);
    const float2 D = make_float2(A.x - B.x, A.y + B.y);
    const float2 WD = make_float2(w.x * D.x + w.y * D.y, w.x * D.y - w.y * D.x);
    zk = make_float2(S.x - WD.y, -S.y - WD.x);
    zp = make_float2(S.x + WD.y, S.y - WD.x);
}

template <int M>
__device__ __forceinline__ void istft_build_z(const IstftRaw& x, int i, const float2* __restrict__ tw, int polar,
                                              float2* __restrict__ buf) {
    constexpr int TPF = M / 16;
#pragma unroll
    for (int m = 0; m < 8; ++m) {
        const int k = i + m * TPF;
        float2 A = make_float2(x.a[2 * m], x.a[2 * m + 1]);
        float2 B = make_float2(x.b[2 * m], x.b[2 * m + 1]);
        if (polar) {
            A = polar_to_cart(A);
            B = polar_to_cart(B);
        }
        if (m == 0 && i == 0) A.y = B.y = 0.f;      // c2r ignores the imaginary part of DC and Nyquist
        float2 zk, zp;
        istft_pair(A, B, tw[k], zk, zp);
        buf[pad_idx(k)] = zk;
        if (!(m == 0 && i == 0)) buf[pad_idx(M - k)] = zp;
    }
    if (i == 0) {
        float2 A = make_float2(x.s[0], x.s[1]);
        if (polar) A = polar_to_cart(A);
        float2 zk, zp;
        istft_pair(A, A, tw[M / 2], zk, zp);
        buf[pad_idx(M / 2)] = zk;
    }
}

// ---------------------------------------------------------------------------------- overlap-add gather
// One batch of P consecutive frames [tb, tb+P) sits in shared memory as real sample arrays
// (frame f at bufs + f*frame_stride, 2M floats, odd samples negated).  acc is a ring of
// L = (P-1)*hop + N floats holding the partial overlap-add of padded positions [tb*hop, tb*hop + L);
// position u lives at acc[(u - tb*hop + off) mod L].  After the batch has been added, the first P*hop
// positions are complete: they are divided by the envelope, stored and cleared; the rest stay in the
// ring for the next batch (whose window starts P*hop further on).  Each position is touched by
// exactly one thread per batch, so the overlap-add is deterministic and atomics-free.
struct IstftGather {
    int hop, N, M, T;          // geometry (N = n_fft = 2M, T = number of frames)
    int PH, L;                 // P*hop, ring length
    int off;                   // ring offset of this batch
    int nvalid;                // frames of this batch that exist and are needed (others hold garbage)
    int frame_stride;          // floats between frames in bufs
    int base_u;                // tb*hop
    int own_lo, own_hi;        // padded positions this CTA stores
    float inv_hop;
};

template <int V>
struct VecF;
template <>
struct VecF<1> {
    float a[1];
    __device__ __forceinline__ void load(const float* p) { a[0] = p[0]; }
    __device__ __forceinline__ void store(float* p) const { p[0] = a[0]; }
};
template <>
struct VecF<2> {
    float a[2];
    __device__ __forceinline__ void load(const float* p) {
        const float2 t = *reinterpret_cast<const float2*>(p);
        a[0] = t.x; a[1] = t.y;
    }
    __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float2*>(p) = make_float2(a[0], a[1]); }
};
template <>
struct VecF<4> {
    float a[4];
    __device__ __forceinline__ void load(const float* p) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        a[0] = t.x; a[1] = t.y; a[2] = t.z; a[3] = t.w;
    }
    __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = make_float4(a[0], a[1], a[2], a[3]); }
};

// Envelope for one padded position, the slow way (clip edges only).
__device__ __forceinline__ float istft_env_edge(int u, int hop, int N, int T, const float* __restrict__ wg) {
    const int tf = (u - N + 1 > 0) ? (u - N + 1 + hop - 1) / hop : 0;
    const int tl = min(T - 1, u / hop);
    float env = 0.f;
    for (int tt = tf; tt <= tl; ++tt) {
        const float w = wg[u - tt * hop];
        env = fmaf(w, w, env);
    }
    return env;   // wg carries scale: caller divides by scale^2 (see istft_gather_elem)
}

// Element e of the batch: V consecutive positions starting at u_rel = e*V.
template <int V>
__device__ __forceinline__ void istft_gather_elem(int e, const IstftGather& c, float* __restrict__ acc,
                                                  const float* __restrict__ bufs, const float* __restrict__ wg,
                                                  const float* __restrict__ ienv, float inv_scale2,
                                                  float* __restrict__ out_clip) {
    const int u_rel = e * V;
    const int q = (int)(((float)u_rel + 0.5f) * c.inv_hop);     // floor(u_rel / hop), exact for these ranges
    int tr = min(q, c.nvalid - 1);
    int j = u_rel - tr * c.hop;
    int pos = u_rel + c.off;
    if (pos >= c.L) pos -= c.L;
    VecF<V> val;
    val.load(acc + pos);
    const float* fb = bufs + tr * c.frame_stride + j;
    for (; tr >= 0 && j < c.N; --tr, j += c.hop, fb += c.hop - c.frame_stride) {
        VecF<V> s, w;
        s.load(fb);
        w.load(wg + j);
#pragma unroll
        for (int x = 0; x < V; ++x) val.a[x] = fmaf(s.a[x], w.a[x], val.a[x]);
    }
    if (u_rel < c.PH) {
        const int u = c.base_u + u_rel;
        if (u >= c.own_lo && u < c.own_hi) {
            VecF<V> o;
            if (u >= c.N - c.hop && u + V - 1 < c.T * c.hop) {
                VecF<V> ie;
                ie.load(ienv + (u_rel - q * c.hop));
#pragma unroll
                for (int x = 0; x < V; ++x) o.a[x] = val.a[x] * ie.a[x];
            } else {
#pragma unroll
                for (int x = 0; x < V; ++x)
                    o.a[x] = val.a[x] / (istft_env_edge(u + x, c.hop, c.N, c.T, wg) * inv_scale2);
            }
            o.store(out_clip + (u - c.M));
        }
        VecF<V> z;
#pragma unroll
        for (int x = 0; x < V; ++x) z.a[x] = 0.f;
        z.store(acc + pos);
    } else {
        val.store(acc + pos);
    }
}

// Column-owning variant for power-of-two hops that divide n_fft (R = n_fft / hop frames overlap).  A thread
// keeps one float4 column (4*col .. 4*col+3 within the hop) for the whole kernel, so its R window vectors
// and its envelope vector stay in registers (wr, ie) and the frame loop is fully unrolled: one LDS.128
// per (position, frame) instead of two plus loop control.  rows = P - 1 + R hop-rows per batch.
template <int R, bool WREG>
__device__ __forceinline__ void istft_gather_cols(int tid, int nthreads, int log2C, int P, const IstftGather& c,
                                                  float* __restrict__ acc, const float* __restrict__ bufs,
                                                  const float* __restrict__ wg, const float4 (&wr)[R], float4 ie,
                                                  float inv_scale2, float* __restrict__ out_clip) {
    const int C = 1 << log2C;
    const int col4 = (tid & (C - 1)) * 4;
    const int rows_per_pass = nthreads >> log2C;
    const int nrows = P - 1 + R;
    for (int q = tid >> log2C; q < nrows; q += rows_per_pass) {
        const int u_rel = q * c.hop + col4;
        int pos = u_rel + c.off;
        if (pos >= c.L) pos -= c.L;
        float4 val = *reinterpret_cast<const float4*>(acc + pos);
        const float* fb = bufs + q * c.frame_stride + col4;
#pragma unroll
        for (int s = 0; s < R; ++s) {
            const int tr = q - s;
            if (tr >= 0 && tr < c.nvalid) {
                const float4 f = *reinterpret_cast<const float4*>(fb + s * (c.hop - c.frame_stride));
                float4 w;
                if constexpr (WREG) w = wr[s];
                else w = *reinterpret_cast<const float4*>(wg + col4 + s * c.hop);
                val.x = fmaf(f.x, w.x, val.x);
                val.y = fmaf(f.y, w.y, val.y);
                val.z = fmaf(f.z, w.z, val.z);
                val.w = fmaf(f.w, w.w, val.w);
            }
        }
        if (q < P) {
            const int u = c.base_u + u_rel;
            if (u >= c.own_lo && u < c.own_hi) {
                float4 o;
                if (u >= c.N - c.hop && u + 3 < c.T * c.hop) {
                    o = make_float4(val.x * ie.x, val.y * ie.y, val.z * ie.z, val.w * ie.w);
                } else {
                    o.x = val.x / (istft_env_edge(u, c.hop, c.N, c.T, wg) * inv_scale2);
                    o.y = val.y / (istft_env_edge(u + 1, c.hop, c.N, c.T, wg) * inv_scale2);
                    o.z = val.z / (istft_env_edge(u + 2, c.hop, c.N, c.T, wg) * inv_scale2);
                    o.w = val.w / (istft_env_edge(u + 3, c.hop, c.N, c.T, wg) * inv_scale2);
                }
                *reinterpret_cast<float4*>(out_clip + (u - c.M)) = o;
            }
            *reinterpret_cast<float4*>(acc + pos) = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
            *reinterpret_cast<float4*>(acc + pos) = val;
        }
    }
}
