// Host emulation of the STFT / iSTFT kernels: same phase functions, threads run serially,
// barriers become loop boundaries.  Built by tests/test_emulation.py with g++.
#include "cuda_emu.h"
#define cudaError_t int
#include <vector>
#define CUDA_RUNTIME_H_SHIM
// fft_core.cuh includes <cuda_runtime.h>; provide an empty stand-in via include path (tests/emu/cuda_runtime.h)
#include "../../maavss_b200/csrc/stft_phases.cuh"

template <int M>
static void fft_block_emu(std::vector<float2>& buf /*M*/, const std::vector<float2>& tw, std::vector<float2>& in_v /*TPF*16*/) {
    constexpr int TPF = M / 16;
    constexpr int R1 = (M / 16 >= 16) ? 16 : (M / 16);
    constexpr int R2 = M / (16 * R1);
    float2 v[16];
    std::vector<float2> nxt(fft_buf_len(M));
    for (int i = 0; i < TPF; ++i) {
        for (int m = 0; m < 16; ++m) v[m] = in_v[i * 16 + m];
        fft_pass<M, 16, 1>(v, i, tw.data(), nxt.data());
    }
    buf = nxt;
    for (int i = 0; i < TPF; ++i) {
        fft_reload<M>(v, i, buf.data());
        fft_pass<M, R1, 16>(v, i, tw.data(), nxt.data());
    }
    buf = nxt;
    if constexpr (R2 > 1) {
        for (int i = 0; i < TPF; ++i) {
            fft_reload<M>(v, i, buf.data());
            fft_pass<M, R2, 16 * R1>(v, i, tw.data(), nxt.data());
        }
        buf = nxt;
    }
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
    std::vector<float2> buf(fft_buf_len(M)), vin(TPF * 16);
    for (int t = 0; t < T_out; ++t) {
        for (int i = 0; i < TPF; ++i) {
            float2 v[16];
            stft_load_frame<M>(v, i, clip, window, t, hop, n_samples);
            for (int m = 0; m < 16; ++m) vin[i * 16 + m] = v[m];
        }
        fft_block_emu<M>(buf, tw, vin);
        float* row0 = out + (size_t)t * F_out;
        for (int i = 0; i < TPF; ++i)
            stft_post_frame<M>(i, buf.data(), tw.data(), row0, row0 + (size_t)T_out * F_out, F_out, scale, polar);
    }
}

template <int M>
static void istft_emu(const float* planes, int T, int F_in, int hop, const float* window, float scale, int polar,
                      int G, float* audio) {
    constexpr int TPF = M / 16;
    constexpr int N = 2 * M;
    constexpr int P = 256 / TPF;
    const int out_len = hop * (T - 1);
    auto tw = make_tw(M);
    const float* in0 = planes;
    const float* in1 = planes + (size_t)T * F_in;
    const int strips = (out_len + G * hop - 1) / (G * hop);
    for (int g = 0; g < strips; ++g) {
        const int s0 = g * G * hop;
        const int S_out = min(G * hop, out_len - s0);
        const int u0 = s0 + M;
        int t_lo = (u0 - N + 1 > 0) ? (u0 - N + 1 + hop - 1) / hop : 0;
        int t_hi = min(T - 1, (u0 + S_out - 1) / hop);
        std::vector<float> acc(S_out, 0.f);
        std::vector<std::vector<float2>> bufs(P, std::vector<float2>(fft_buf_len(M)));
        for (int tb = t_lo; tb <= t_hi; tb += P) {
            const int t_last = min(tb + P - 1, t_hi);
            for (int fi = 0; fi < P && tb + fi <= t_hi; ++fi) {
                const int t = tb + fi;
                std::vector<float2> vin(TPF * 16);
                for (int i = 0; i < TPF; ++i) {
                    float2 v[16];
                    istft_load_frame<M>(v, i, in0 + (size_t)t * F_in, in1 + (size_t)t * F_in, F_in, tw.data(), polar);
                    for (int m = 0; m < 16; ++m) vin[i * 16 + m] = v[m];
                }
                fft_block_emu<M>(bufs[fi], tw, vin);
            }
            const int lo = max(u0, tb * hop);
            const int hi = min(u0 + S_out, t_last * hop + N);
            for (int u = lo; u < hi; ++u) {
                int tf = (u - N + 1 > 0) ? (u - N + 1 + hop - 1) / hop : 0;
                tf = max(tf, tb);
                const int tl = min(t_last, u / hop);
                float s = 0.f;
                for (int tt = tf; tt <= tl; ++tt) {
                    const int j = u - tt * hop;
                    s += istft_frame_sample(bufs[tt - tb].data(), j) * window[j];
                }
                acc[u - u0] += s;
            }
        }
        for (int q = 0; q < S_out; ++q) {
            const int u = u0 + q;
            const int tf = (u - N + 1 > 0) ? (u - N + 1 + hop - 1) / hop : 0;
            const int tl = min(T - 1, u / hop);
            float env = 0.f;
            for (int tt = tf; tt <= tl; ++tt) {
                const float w = window[u - tt * hop];
                env = fmaf(w, w, env);
            }
            audio[s0 + q] = acc[q] * scale / env;
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

extern "C" int emu_istft(const float* planes, int T, int n_fft, int hop, const float* window, int trim, int normalized,
                         int polar, int G, float* audio) {
    const int F_in = trim ? n_fft / 2 : n_fft / 2 + 1;
    const float scale = (float)((normalized ? sqrt((double)n_fft) : 1.0) / (double)(n_fft / 2));
    DISPATCH(n_fft / 2, istft_emu<MM>(planes, T, F_in, hop, window, scale, polar, G, audio));
    return 0;
}
