// Direct fp32 convolution kernels: forward, input gradient (= ConvTranspose forward)
// and weight gradient for any (KD,KH,KW)/(stride)/(padding) geometry, channels-first.
// These are the exact-fp32 SIMT kernels; the GEMM-shaped layers of the big model run on the tcgen05 path.
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

constexpr int kConvThreads = 128;

__device__ __forceinline__ void decode3(int p, int d1, int d2, int& a, int& b, int& c) {
    c = p % d2;
    const int q = p / d2;
    b = q % d1;
    a = q / d1;
}

// One thread = one output position, CT output channels.  Weights of the (cout tile, cin chunk) are staged in
// shared memory as [ci][tap][CT] so that the inner product reads them as broadcast vectors.
template <int CT>
__global__ void __launch_bounds__(kConvThreads) conv_fwd_kernel(maavss_conv_desc d, const float* __restrict__ x,
                                                                const float* __restrict__ w, const float* __restrict__ bias,
                                                                float* __restrict__ y, int CC) {
    extern __shared__ float ws[];
    const int KV = d.KD * d.KH * d.KW;
    const int O = d.OD * d.OH * d.OW;
    const int p = blockIdx.x * kConvThreads + threadIdx.x;
    const bool valid = p < O;
    int od, oh, ow;
    decode3(valid ? p : 0, d.OH, d.OW, od, oh, ow);
    const int co0 = blockIdx.y * CT, n = blockIdx.z;
    const int id0 = od * d.SD - d.PD, ih0 = oh * d.SH - d.PH, iw0 = ow * d.SW - d.PW;
    const size_t plane = (size_t)d.D * d.H * d.W;
    float acc[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) acc[c] = (bias && co0 + c < d.Cout) ? bias[co0 + c] : 0.f;
    for (int ci0 = 0; ci0 < d.Cin; ci0 += CC) {
        const int cc = min(CC, d.Cin - ci0);
        __syncthreads();
        for (int idx = threadIdx.x; idx < cc * KV * CT; idx += kConvThreads) {
            const int c = idx % CT, k = (idx / CT) % KV, ci = idx / (CT * KV);
            ws[idx] = (co0 + c < d.Cout) ? w[((size_t)(co0 + c) * d.Cin + ci0 + ci) * KV + k] : 0.f;
        }
        __syncthreads();
        if (!valid) continue;
        for (int ci = 0; ci < cc; ++ci) {
            const float* xb = x + ((size_t)n * d.Cin + ci0 + ci) * plane;
            const float* wc = ws + (size_t)ci * KV * CT;
            for (int kd = 0; kd < d.KD; ++kd) {
                const int id = id0 + kd;
                if (id < 0 || id >= d.D) continue;
                for (int kh = 0; kh < d.KH; ++kh) {
                    const int ih = ih0 + kh;
                    if (ih < 0 || ih >= d.H) continue;
                    const float* xr = xb + ((size_t)id * d.H + ih) * d.W;
                    const float* wr = wc + (size_t)((kd * d.KH + kh) * d.KW) * CT;
                    for (int kw = 0; kw < d.KW; ++kw) {
                        const int iw = iw0 + kw;
                        if (iw < 0 || iw >= d.W) continue;
                        const float xv = __ldg(xr + iw);
#pragma unroll
                        for (int c = 0; c < CT; ++c) acc[c] = fmaf(xv, wr[kw * CT + c], acc[c]);
                    }
                }
            }
        }
    }
    if (valid) {
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (co0 + c < d.Cout) y[((size_t)n * d.Cout + co0 + c) * O + p] = acc[c];
    }
}

// One thread = one INPUT position, CT input channels; gathers the output positions that touched it.
template <int CT>
__global__ void __launch_bounds__(kConvThreads) conv_dgrad_kernel(maavss_conv_desc d, const float* __restrict__ dy,
                                                                  const float* __restrict__ w, const float* __restrict__ bias,
                                                                  float* __restrict__ dx, int CC) {
    extern __shared__ float ws[];
    const int KV = d.KD * d.KH * d.KW;
    const int I = d.D * d.H * d.W;
    const size_t O = (size_t)d.OD * d.OH * d.OW;
    const int p = blockIdx.x * kConvThreads + threadIdx.x;
    const bool valid = p < I;
    int id, ih, iw;
    decode3(valid ? p : 0, d.H, d.W, id, ih, iw);
    const int ci0 = blockIdx.y * CT, n = blockIdx.z;
    float acc[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) acc[c] = (bias && ci0 + c < d.Cin) ? bias[ci0 + c] : 0.f;
    for (int co0 = 0; co0 < d.Cout; co0 += CC) {
        const int cc = min(CC, d.Cout - co0);
        __syncthreads();
        for (int idx = threadIdx.x; idx < cc * KV * CT; idx += kConvThreads) {
            const int c = idx % CT, k = (idx / CT) % KV, co = idx / (CT * KV);
            ws[idx] = (ci0 + c < d.Cin) ? w[((size_t)(co0 + co) * d.Cin + ci0 + c) * KV + k] : 0.f;
        }
        __syncthreads();
        if (!valid) continue;
        for (int co = 0; co < cc; ++co) {
            const float* yb = dy + ((size_t)n * d.Cout + co0 + co) * O;
            const float* wc = ws + (size_t)co * KV * CT;
            for (int kd = 0; kd < d.KD; ++kd) {
                const int td = id + d.PD - kd;
                if (td < 0 || td % d.SD) continue;
                const int od = td / d.SD;
                if (od >= d.OD) continue;
                for (int kh = 0; kh < d.KH; ++kh) {
                    const int th = ih + d.PH - kh;
                    if (th < 0 || th % d.SH) continue;
                    const int oh = th / d.SH;
                    if (oh >= d.OH) continue;
                    const float* yr = yb + ((size_t)od * d.OH + oh) * d.OW;
                    const float* wr = wc + (size_t)((kd * d.KH + kh) * d.KW) * CT;
                    for (int kw = 0; kw < d.KW; ++kw) {
                        const int tw = iw + d.PW - kw;
                        if (tw < 0 || tw % d.SW) continue;
                        const int ow = tw / d.SW;
                        if (ow >= d.OW) continue;
                        const float g = __ldg(yr + ow);
#pragma unroll
                        for (int c = 0; c < CT; ++c) acc[c] = fmaf(g, wr[kw * CT + c], acc[c]);
                    }
                }
            }
        }
    }
    if (valid) {
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (ci0 + c < d.Cin) dx[((size_t)n * d.Cin + ci0 + c) * I + p] = acc[c];
    }
}

// dw[co][ci][k] += sum_p dy[n][co][o] * x[n][ci][o*s - pad + k].  A CTA owns (CT couts) x (one cin) x (KT taps) and a
// slice of the (n, o) domain; per-thread register partials, warp-shuffle + shared reduction, one atomicAdd per weight.
template <int KT, int CT>
__global__ void __launch_bounds__(kConvThreads) conv_wgrad_kernel(maavss_conv_desc d, const float* __restrict__ x,
                                                                  const float* __restrict__ dy, float* __restrict__ dw,
                                                                  int tap_chunks, long long slice_len) {
    __shared__ int tap_d[KT], tap_h[KT], tap_w[KT];
    __shared__ float red[kConvThreads / 32][KT * CT];
    const int KV = d.KD * d.KH * d.KW;
    const int O = d.OD * d.OH * d.OW;
    const long long P = (long long)d.N * O;
    const int co0 = blockIdx.y * CT;
    const int ci = blockIdx.z / tap_chunks, k0 = (blockIdx.z % tap_chunks) * KT;
    if (threadIdx.x < KT) {
        const int k = k0 + threadIdx.x;
        if (k < KV) {
            tap_w[threadIdx.x] = k % d.KW;
            tap_h[threadIdx.x] = (k / d.KW) % d.KH;
            tap_d[threadIdx.x] = k / (d.KW * d.KH);
        } else {
            tap_d[threadIdx.x] = -(1 << 28);   // never in range
            tap_h[threadIdx.x] = 0;
            tap_w[threadIdx.x] = 0;
        }
    }
    __syncthreads();
    float acc[CT][KT];
#pragma unroll
    for (int c = 0; c < CT; ++c)
#pragma unroll
        for (int j = 0; j < KT; ++j) acc[c][j] = 0.f;
    const long long p0 = (long long)blockIdx.x * slice_len;
    const long long p1 = min(P, p0 + slice_len);
    const size_t plane = (size_t)d.D * d.H * d.W;
    for (long long p = p0 + threadIdx.x; p < p1; p += kConvThreads) {
        const int n = (int)(p / O), o = (int)(p - (long long)n * O);
        int od, oh, ow;
        decode3(o, d.OH, d.OW, od, oh, ow);
        float g[CT];
#pragma unroll
        for (int c = 0; c < CT; ++c) g[c] = (co0 + c < d.Cout) ? __ldg(dy + ((size_t)n * d.Cout + co0 + c) * O + o) : 0.f;
        const float* xb = x + ((size_t)n * d.Cin + ci) * plane;
        const int id0 = od * d.SD - d.PD, ih0 = oh * d.SH - d.PH, iw0 = ow * d.SW - d.PW;
#pragma unroll
        for (int j = 0; j < KT; ++j) {
            const int id = id0 + tap_d[j], ih = ih0 + tap_h[j], iw = iw0 + tap_w[j];
            float xv = 0.f;
            if (id >= 0 && id < d.D && ih >= 0 && ih < d.H && iw >= 0 && iw < d.W) xv = __ldg(xb + ((size_t)id * d.H + ih) * d.W + iw);
#pragma unroll
            for (int c = 0; c < CT; ++c) acc[c][j] = fmaf(g[c], xv, acc[c][j]);
        }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int c = 0; c < CT; ++c)
#pragma unroll
        for (int j = 0; j < KT; ++j) {
            const float s = warp_sum(acc[c][j]);
            if (lane == 0) red[warp][c * KT + j] = s;
        }
    __syncthreads();
    for (int q = threadIdx.x; q < KT * CT; q += kConvThreads) {
        const int c = q / KT, j = q % KT;
        if (co0 + c < d.Cout && k0 + j < KV) {
            float s = 0.f;
#pragma unroll
            for (int wv = 0; wv < kConvThreads / 32; ++wv) s += red[wv][q];
            atomicAdd(dw + ((size_t)(co0 + c) * d.Cin + ci) * KV + k0 + j, s);
        }
    }
}

// out[c] (+)= sum_{n,i} x[n,c,i]; two deterministic stages through a double workspace [C][S].
constexpr int kSumSplits = 16;
__global__ void channel_sum_partial(const float* __restrict__ x, int N, int C, long long inner, double* __restrict__ ws) {
    const int c = blockIdx.y, s = blockIdx.x;
    const long long total = (long long)N * inner;
    const long long chunk = (total + kSumSplits - 1) / kSumSplits;
    const long long e0 = s * chunk, e1 = min(total, e0 + chunk);
    float a = 0.f;
    for (long long e = e0 + threadIdx.x; e < e1; e += blockDim.x) {
        const long long n = e / inner, i = e - n * inner;
        a += x[((size_t)n * C + c) * inner + i];
    }
    double v = a;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __shared__ double sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int q = 0; q < (int)(blockDim.x >> 5); ++q) t += sm[q];
        ws[(size_t)c * kSumSplits + s] = t;
    }
}
__global__ void channel_sum_final(const double* __restrict__ ws, int C, float* __restrict__ out, int accumulate) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double t = 0;
    for (int s = 0; s < kSumSplits; ++s) t += ws[(size_t)c * kSumSplits + s];
    out[c] = accumulate ? out[c] + (float)t : (float)t;
}

int check_desc(const maavss_conv_desc* d, const char* who) {
    MAAVSS_CHECK_ARG(d, "%s: null descriptor", who);
    MAAVSS_CHECK_ARG(d->N >= 0 && d->Cin > 0 && d->Cout > 0 && d->D > 0 && d->H > 0 && d->W > 0, "%s: bad input dims", who);
    MAAVSS_CHECK_ARG(d->KD > 0 && d->KH > 0 && d->KW > 0 && d->SD > 0 && d->SH > 0 && d->SW > 0, "%s: bad kernel/stride", who);
    MAAVSS_CHECK_ARG(d->OD > 0 && d->OH > 0 && d->OW > 0, "%s: bad output dims", who);
    MAAVSS_CHECK_ARG(d->N <= 65535, "%s: batch %d > 65535", who, d->N);
    // every output position must read inside the padded input, and the input must not extend past what the taps reach + stride
    MAAVSS_CHECK_ARG((d->OD - 1) * d->SD - d->PD + d->KD - 1 < d->D + d->PD + d->SD && (d->OH - 1) * d->SH - d->PH + d->KH - 1 < d->H + d->PH + d->SH &&
                         (d->OW - 1) * d->SW - d->PW + d->KW - 1 < d->W + d->PW + d->SW,
                     "%s: output dims inconsistent with input/kernel/stride/padding", who);
    return MAAVSS_OK;
}

int pick_ct(int C) { return C <= 4 ? 4 : (C % 16 == 0 ? 16 : 8); }
int pick_cc(int KV, int CT, int Cred) {
    int cc = 4096 / (KV * CT);
    if (cc < 1) cc = 1;
    if (cc > Cred) cc = Cred;
    return cc;
}

}  // namespace

extern "C" int maavss_conv_fwd(const maavss_conv_desc* d, const float* x, const float* w, const float* bias, float* y, void* stream) {
    int rc = check_desc(d, "conv_fwd");
    if (rc) return rc;
    if (d->N == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && w && y, "conv_fwd: null argument");
    const int KV = d->KD * d->KH * d->KW, O = d->OD * d->OH * d->OW;
    const int CT = pick_ct(d->Cout), CC = pick_cc(KV, CT, d->Cin);
    dim3 grid(ceil_div(O, kConvThreads), ceil_div(d->Cout, CT), d->N);
    const size_t smem = sizeof(float) * CC * KV * CT;
    cudaStream_t st = (cudaStream_t)stream;
    if (CT == 4) conv_fwd_kernel<4><<<grid, kConvThreads, smem, st>>>(*d, x, w, bias, y, CC);
    else if (CT == 8) conv_fwd_kernel<8><<<grid, kConvThreads, smem, st>>>(*d, x, w, bias, y, CC);
    else conv_fwd_kernel<16><<<grid, kConvThreads, smem, st>>>(*d, x, w, bias, y, CC);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_conv_dgrad(const maavss_conv_desc* d, const float* dy, const float* w, const float* bias, float* dx, void* stream) {
    int rc = check_desc(d, "conv_dgrad");
    if (rc) return rc;
    if (d->N == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dy && w && dx, "conv_dgrad: null argument");
    const int KV = d->KD * d->KH * d->KW, I = d->D * d->H * d->W;
    const int CT = pick_ct(d->Cin), CC = pick_cc(KV, CT, d->Cout);
    dim3 grid(ceil_div(I, kConvThreads), ceil_div(d->Cin, CT), d->N);
    const size_t smem = sizeof(float) * CC * KV * CT;
    cudaStream_t st = (cudaStream_t)stream;
    if (CT == 4) conv_dgrad_kernel<4><<<grid, kConvThreads, smem, st>>>(*d, dy, w, bias, dx, CC);
    else if (CT == 8) conv_dgrad_kernel<8><<<grid, kConvThreads, smem, st>>>(*d, dy, w, bias, dx, CC);
    else conv_dgrad_kernel<16><<<grid, kConvThreads, smem, st>>>(*d, dy, w, bias, dx, CC);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_conv_wgrad(const maavss_conv_desc* d, const float* x, const float* dy, float* dw, int accumulate, void* stream) {
    int rc = check_desc(d, "conv_wgrad");
    if (rc) return rc;
    MAAVSS_CHECK_ARG(x && dy && dw || d->N == 0, "conv_wgrad: null argument");
    const int KV = d->KD * d->KH * d->KW, O = d->OD * d->OH * d->OW;
    cudaStream_t st = (cudaStream_t)stream;
    if (!accumulate) MAAVSS_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)d->Cout * d->Cin * KV, st));
    if (d->N == 0) return MAAVSS_OK;
    const long long P = (long long)d->N * O;
    const bool small_k = (KV % 9 == 0 && KV % 25 != 0) || KV < 16;
    const int KT = small_k ? 9 : 25;
    const int tap_chunks = ceil_div(KV, KT);
    constexpr int CT = 4;
    const long long base_ctas = (long long)ceil_div(d->Cout, CT) * d->Cin * tap_chunks;
    MAAVSS_CHECK_ARG((long long)d->Cin * tap_chunks <= 65535, "conv_wgrad: Cin * tap chunks too large");
    // enough slices to fill the GPU (>= 4 waves of 148 SMs), but at least 8 positions per thread
    long long slices = (148LL * 16 + base_ctas - 1) / base_ctas;
    const long long max_slices = (P + kConvThreads * 8 - 1) / (kConvThreads * 8);
    if (slices > max_slices) slices = max_slices;
    if (slices < 1) slices = 1;
    const long long slice_len = (P + slices - 1) / slices;
    dim3 grid((unsigned)ceil_div(P, slice_len), ceil_div(d->Cout, CT), d->Cin * tap_chunks);
    if (KT == 9) conv_wgrad_kernel<9, CT><<<grid, kConvThreads, 0, st>>>(*d, x, dy, dw, tap_chunks, slice_len);
    else conv_wgrad_kernel<25, CT><<<grid, kConvThreads, 0, st>>>(*d, x, dy, dw, tap_chunks, slice_len);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_reduce_workspace_bytes(int C) { return (int)(sizeof(double) * kSumSplits * (C > 0 ? C : 1)) + 64; }

extern "C" int maavss_channel_sum(const float* x, int N, int C, long long inner, float* out, int accumulate, void* workspace,
                                  void* stream) {
    if (C == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(out && workspace, "channel_sum: null argument");
    MAAVSS_CHECK_ARG(C <= 65535, "channel_sum: C too large");
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid(kSumSplits, C);
    channel_sum_partial<<<grid, 256, 0, st>>>(x, N, C, inner, (double*)workspace);
    MAAVSS_LAUNCH_CHECK();
    channel_sum_final<<<ceil_div(C, 128), 128, 0, st>>>((const double*)workspace, C, out, accumulate);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
