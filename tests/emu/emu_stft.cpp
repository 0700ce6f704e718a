// Host emulation of the STFT / iSTFT kernels in maavss_b200/csrc/stft.cu: the same phase functions
// (stft_phases.cuh / fft_core.cuh) compiled with g++, threads run serially, group / CTA barriers become
// loop boundaries, the CTA / batch / ring structure is restated 1:1.  Built by tests/test_emulation.py.
#include "cuda_emu.h"
#include <vector>
// fft_core.cuh includes <cuda_runtime.h>; tests/emu/cuda_runtime.h is an empty stand-in on the include path
#include "../../maavss_b200/csrc/stft_phases.cuh"

// fft_group of stft.cu with the group barriers turned into loops over the frame's TPF threads.
template <int M, bool PADFINAL>
static void fft_group_emu(std::vector<float2>& vin /*TPF*16*/, const float2* ttab, float2* buf) {
    using S = FftShape<M>;
    constexpr int TPF = S::TPF;
    float2 v[16];
    std::vector<float2> nxt(fft_buf_len(M));
    auto load = [&](int i) { for (int m = 0; m < 16; ++m) v[m] = vin[i * 16 + m]; };
    auto reload_all = [&]() {
        for (int i = 0; i < TPF; ++i) {
            fft_reload<M>(v, i, buf);
            for (int m = 0; m < 16; ++m) vin[i * 16 + m] = v[m];
        }
    };
    for (int i = 0; i < TPF; ++i) { load(i); fft_pass_t<M, 16, 1, 0, true>(v, i, ttab, nxt.data()); }
    std::memcpy(buf, nxt.data(), sizeof(float2) * nxt.size());
    reload_all();
    if constexpr (S::R2 > 1) {
        for (int i = 0; i < TPF; ++i) { load(i); fft_pass_t<M, S::R1, 16, 0, true>(v, i, ttab, nxt.data()); }
        std::memcpy(buf, nxt.data(), sizeof(float2) * nxt.size());
        reload_all();
        for (int i = 0; i < TPF; ++i) { load(i); fft_pass_t<M, S::R2, 16 * S::R1, S::SLOTS1, PADFINAL>(v, i, ttab, nxt.data()); }
    } else {
        for (int i = 0; i < TPF; ++i) { load(i); fft_pass_t<M, S::R1, 16, 0, PADFINAL>(v, i, ttab, nxt.data()); }
    }
    std::memcpy(buf, nxt.data(), sizeof(float2) * nxt.size());
}

static std::vector<float2> make_tw(int M) {
    std::vector<float2> tw(M);
    for (int q = 0; q < M; ++q) {
        double a = -2.0 * M_PI * q / (2.0 * M);
        tw[q] = make_float2((float)cos(a), (float)sin(a));
    }
    return tw;
}

template <int M>
static void stft_emu(const float* clip, int n_samples, int hop, const float* window, int F_out, float scale, int polar,
                     float* out /*2,T_out,F_out*/) {
    constexpr int TPF = M / 16;
    const int T_out = n_samples / hop;
    auto tw = make_tw(M);
    std::vector<float2> ttab(fft_ttab_len(M)), wsm(M), buf(fft_buf_len(M)), vin(TPF * 16);
    fft_build_ttab(M, ttab.data());
    const float hs = 0.5f * scale;
    for (int q = 0; q < M; ++q) wsm[q] = make_float2(window[2 * q] * hs, window[2 * q + 1] * hs);
    for (int t = 0; t < T_out; ++t) {
        for (int i = 0; i < TPF; ++i) {
            float2 v[16];
            stft_load_frame<M>(v, i, clip, wsm.data(), t, hop, n_samples);
            for (int m = 0; m < 16; ++m) vin[i * 16 + m] = v[m];
        }
        fft_group_emu<M, true>(vin, ttab.data(), buf.data());
        float* row0 = out + (size_t)t * F_out;
        for (int i = 0; i < TPF; ++i)
            stft_post_frame<M>(i, buf.data(), tw.data(), row0, row0 + (size_t)T_out * F_out, F_out, polar);
    }
}

template <int M, int V, int R>
static void istft_emu(const float* planes, int T, int F_in, int hop, const float* window, const float* ienv, float scale,
                      int polar, int G, float* audio) {
    constexpr int TPF = M / 16;
    constexpr int N = 2 * M;
    constexpr int NT = (M >= 2048) ? 512 : 256;
    constexpr int P = NT / TPF;
    constexpr int FS = 2 * fft_buf_len(M);
    const int out_len = hop * (T - 1);
    auto tw = make_tw(M);
    std::vector<float2> ttab(fft_ttab_len(M));
    fft_build_ttab(M, ttab.data());
    const float* in0 = planes;
    const float* in1 = planes + (size_t)T * F_in;
    const float sc = 0.5f * scale;
    const float inv_scale2 = 1.0f / (sc * sc);
    std::vector<float4> wg4(N / 4);
    float* wg = reinterpret_cast<float*>(wg4.data());
    for (int q = 0; q < N; ++q) wg[q] = window[q] * ((q & 1) ? -sc : sc);
    const int slots = (M + out_len + hop - 1) / hop;
    for (int bx = 0; bx * G < slots; ++bx) {
        IstftGather c;
        c.hop = hop; c.N = N; c.M = M; c.T = T;
        c.PH = P * hop;
        c.L = (P - 1) * hop + N;
        c.frame_stride = FS;
        c.inv_hop = 1.0f / (float)hop;
        const int t_a = bx * G, t_b = t_a + G;
        c.own_lo = max(M, t_a * hop);
        c.own_hi = min(M + out_len, t_b * hop);
        const int t_lim = min(t_b, T);
        const int t_start = max(0, t_a - (N - 1) / hop);
        std::vector<float4> acc4(c.L / 4 + 1, float4{0, 0, 0, 0}), bufs4((size_t)P * FS / 4 + 1);
        float* acc = reinterpret_cast<float*>(acc4.data());
        float* bufs = reinterpret_cast<float*>(bufs4.data());
        for (size_t q = 0; q < (size_t)P * FS; ++q) bufs[q] = NAN;     // stale shared memory must never be read
        c.off = 0;
        for (int tb = t_start; tb * hop < c.own_hi; tb += P) {
            for (int fi = 0; fi < P; ++fi) {
                const int t = tb + fi;
                const int t_group0 = (TPF >= 32) ? t : tb + ((fi * TPF) & ~31) / TPF;
                if (t_group0 >= t_lim) continue;
                std::vector<float2> vin(TPF * 16);
                float2* buf = reinterpret_cast<float2*>(bufs + (size_t)fi * FS);
                for (int i = 0; i < TPF; ++i) {
                    IstftRaw x;
                    istft_fetch_raw<M>(x, i, in0 + (size_t)t * F_in, in1 + (size_t)t * F_in, F_in, t < t_lim);
                    istft_build_z<M>(x, i, tw.data(), polar, buf);
                }
                for (int i = 0; i < TPF; ++i) {
                    float2 v[16];
                    fft_reload<M>(v, i, buf);
                    for (int m = 0; m < 16; ++m) vin[i * 16 + m] = v[m];
                }
                fft_group_emu<M, false>(vin, ttab.data(), reinterpret_cast<float2*>(bufs + (size_t)fi * FS));
            }
            c.base_u = tb * hop;
            c.nvalid = max(0, min(P, t_lim - tb));
            if constexpr (R > 0) {
                int log2C = 0;
                while ((4 << log2C) < hop) ++log2C;
                for (int tid = 0; tid < NT; ++tid) {
                    const int col4 = (tid & ((hop >> 2) - 1)) * 4;
                    float4 wr[R];
                    for (int s2 = 0; s2 < R; ++s2) {
                        const float* w = window + col4 + s2 * hop;
                        wr[s2] = make_float4(w[0] * sc, -w[1] * sc, w[2] * sc, -w[3] * sc);
                    }
                    const float4 ie = *reinterpret_cast<const float4*>(ienv + col4);
                    istft_gather_cols<R, (R <= 4)>(tid, NT, log2C, P, c, acc, bufs, wg, wr, ie, inv_scale2, audio);
                }
            } else {
                for (int e = 0; e < c.L / V; ++e) istft_gather_elem<V>(e, c, acc, bufs, wg, ienv, inv_scale2, audio);
            }
            c.off += c.PH;
            if (c.off >= c.L) c.off -= c.L;
        }
    }
}

#define DISPATCH(M_, CALL) \
    switch (M_) { \
        case 32: { constexpr int MM = 32; CALL; } break; \
        case 64: { constexpr int MM = 64; CALL; } break; \
        case 128: { constexpr int MM = 128; CALL; } break; \
        case 256: { constexpr int MM = 256; CALL; } break; \
        case 512: { constexpr int MM = 512; CALL; } break; \
        case 1024: { constexpr int MM = 1024; CALL; } break; \
        case 2048: { constexpr int MM = 2048; CALL; } break; \
        default: return -1; \
    }

extern "C" int emu_stft(const float* clip, int n_samples, int n_fft, int hop, const float* window, int trim,
                        int normalized, int polar, float* out) {
    double s = 0;
    for (int j = 0; j < n_fft; ++j) s += (double)window[j] * window[j];
    const float scale = normalized ? (float)(1.0 / sqrt(s)) : 1.f;
    const int F_out = trim ? n_fft / 2 : n_fft / 2 + 1;
    DISPATCH(n_fft / 2, stft_emu<MM>(clip, n_samples, hop, window, F_out, scale, polar, out));
    return 0;
}

// audio must be 16-byte aligned (numpy allocations are).
extern "C" int emu_istft(const float* planes, int T, int n_fft, int hop, const float* window, int trim, int normalized,
                         int polar, int G, float* audio) {
    const int F_in = trim ? n_fft / 2 : n_fft / 2 + 1;
    const float scale = (float)((normalized ? sqrt((double)n_fft) : 1.0) / (double)(n_fft / 2));
    std::vector<float4> ie4(hop / 4 + 1);
    float* ienv = reinterpret_cast<float*>(ie4.data());
    for (int r = 0; r < hop; ++r) {
        double e = 0;
        for (int j = r; j < n_fft; j += hop) e += (double)window[j] * window[j];
        ienv[r] = (float)(1.0 / e);
    }
    const int V = (hop % 4 == 0) ? 4 : (hop % 2 == 0) ? 2 : 1;
    const bool pow2 = (hop & (hop - 1)) == 0;
    const int NT = (n_fft / 2 >= 2048) ? 512 : 256;
    int R = (pow2 && hop >= 4 && hop / 4 <= NT && n_fft % hop == 0 && n_fft >= 1024) ? n_fft / hop : 0;
    if (R > 4) R = 0;   // same dispatch rule as launch_istft in stft.cu
    if (R == 1) { DISPATCH(n_fft / 2, (istft_emu<MM, 4, 1>(planes, T, F_in, hop, window, ienv, scale, polar, G, audio))); }
    else if (R == 2) { DISPATCH(n_fft / 2, (istft_emu<MM, 4, 2>(planes, T, F_in, hop, window, ienv, scale, polar, G, audio))); }
    else if (R == 4) { DISPATCH(n_fft / 2, (istft_emu<MM, 4, 4>(planes, T, F_in, hop, window, ienv, scale, polar, G, audio))); }
    else if (V == 4) { DISPATCH(n_fft / 2, (istft_emu<MM, 4, 0>(planes, T, F_in, hop, window, ienv, scale, polar, G, audio))); }
    else if (V == 2) { DISPATCH(n_fft / 2, (istft_emu<MM, 2, 0>(planes, T, F_in, hop, window, ienv, scale, polar, G, audio))); }
    else { DISPATCH(n_fft / 2, (istft_emu<MM, 1, 0>(planes, T, F_in, hop, window, ienv, scale, polar, G, audio))); }
    return 0;
}
