// Spectral layout helpers of the generator API and the polar option
// (/root/reference/generator.py:60-133, av_dataset.py:176-178,187-191).
#include <math.h>
#include <stdlib.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

struct maavss_dft_plan {
    int n;            // transform length
    float2* table;    // device, n entries exp(-2*pi*i*q/n), computed in double
};

namespace {

// planes: (outer, 2, plane) ; 16 B per bin (8 in, 8 out)
__global__ void magphase_split_kernel(const float* __restrict__ in, float* __restrict__ out, long long outer, long long plane) {
    const long long total = outer * plane;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const long long o = q / plane, r = q - o * plane;
        const size_t b = (size_t)o * 2 * plane + r;
        const float re = in[b], im = in[b + plane];
        out[b] = sqrtf(re * re + im * im);
        out[b + plane] = atan2f(im, re);
    }
}

__global__ void magphase_combine_kernel(const float* __restrict__ in, float* __restrict__ out, long long outer, long long plane) {
    const long long total = outer * plane;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const long long o = q / plane, r = q - o * plane;
        const size_t b = (size_t)o * 2 * plane + r;
        const float mag = in[b], ph = in[b + plane];
        float s, c;
        sincosf(ph, &s, &c);
        out[b] = mag * c;
        out[b + plane] = mag * s;
    }
}

// center_fft_bins (generator.py:117-122): out = cat(flip(x[..., ::2]), x[..., 1::2]) along the last axis.
// For output position p: p < ne (= ceil(K/2)) -> x[2*(ne-1-p)], else x[2*(p-ne)+1].
__global__ void bins_center_kernel(const float* __restrict__ in, float* __restrict__ out, long long rows, int K) {
    const int ne = (K + 1) / 2;
    const long long total = rows * K;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const long long r = q / K;
        const int p = (int)(q - r * K);
        const int src = (p < ne) ? 2 * (ne - 1 - p) : 2 * (p - ne) + 1;
        out[q] = in[r * K + src];
    }
}

// decenter_fft_bins (generator.py:126-133): half = K//2; out[::2] = flip(x[:half]); out[1::2] = x[half:].
// (For odd K the reference assignment would raise; the C ABI rejects odd K.)
__global__ void bins_decenter_kernel(const float* __restrict__ in, float* __restrict__ out, long long rows, int K) {
    const int half = K / 2;
    const long long total = rows * K;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const long long r = q / K;
        const int p = (int)(q - r * K);
        const int src = (p & 1) ? half + (p >> 1) : half - 1 - (p >> 1);
        out[q] = in[r * K + src];
    }
}

// Direct DFT of a real clip: out[b][k] = scale * sum_n x[n] * exp(-2*pi*i*k*n/N), k in [0, K).
// Exact integer phase (k*n mod N) into a double-precision-built table; x and the table live in
// shared memory.  One thread per output bin, 4 independent accumulators.
__global__ void dft_real_fwd_kernel(const float* __restrict__ x, float2* __restrict__ out, int N, int K, float scale,
                                    const float2* __restrict__ table) {
    extern __shared__ float2 sm[];
    float2* tab = sm;
    float* xs = reinterpret_cast<float*>(sm + N);
    const float* xb = x + (size_t)blockIdx.y * N;
    for (int q = threadIdx.x; q < N; q += blockDim.x) {
        tab[q] = table[q];
        xs[q] = xb[q];
    }
    __syncthreads();
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    float re = 0.f, im = 0.f;
    int ph = 0;
    for (int n = 0; n < N; ++n) {
        const float2 w = tab[ph];
        const float v = xs[n];
        re = fmaf(v, w.x, re);
        im = fmaf(v, w.y, im);
        ph += k;
        if (ph >= N) ph -= N;
    }
    out[(size_t)blockIdx.y * K + k] = make_float2(re * scale, im * scale);
}

// generator.ifft (generator.py:71-81): zero-extend K bins to N = 2K, inverse DFT (sign +), real part:
// out[b][m] = scale * sum_{k<K} (Xr[k]*cos(2*pi*k*m/N) - Xi[k]*sin(2*pi*k*m/N)).   table = exp(-i..) so sin flips.
__global__ void dft_half_inv_real_kernel(const float2* __restrict__ X, float* __restrict__ out, int K, float scale,
                                         const float2* __restrict__ table) {
    const int N = 2 * K;
    extern __shared__ float2 sm[];
    float2* tab = sm;
    float2* xs = sm + N;
    const float2* xb = X + (size_t)blockIdx.y * K;
    for (int q = threadIdx.x; q < N; q += blockDim.x) tab[q] = table[q];
    for (int q = threadIdx.x; q < K; q += blockDim.x) xs[q] = xb[q];
    __syncthreads();
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= N) return;
    float acc = 0.f;
    int ph = 0;
    for (int k = 0; k < K; ++k) {
        const float2 w = tab[ph];     // (cos, -sin)
        const float2 v = xs[k];
        acc = fmaf(v.x, w.x, acc);
        acc = fmaf(v.y, w.y, acc);    // -Xi*sin = Xi*w.y
        ph += m;
        if (ph >= N) ph -= N;
    }
    out[(size_t)blockIdx.y * N + m] = acc * scale;
}

inline int ew_blocks(long long n) {
    long long b = (n + 256 * 4 - 1) / (256 * 4);
    if (b > 148 * 16) b = 148 * 16;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace

extern "C" int maavss_magphase_split(const float* in, float* out, long long outer, long long plane, void* stream) {
    MAAVSS_CHECK_ARG(in && out, "magphase_split: null argument");
    if (outer * plane == 0) return MAAVSS_OK;
    magphase_split_kernel<<<ew_blocks(outer * plane), 256, 0, (cudaStream_t)stream>>>(in, out, outer, plane);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_magphase_combine(const float* in, float* out, long long outer, long long plane, void* stream) {
    MAAVSS_CHECK_ARG(in && out, "magphase_combine: null argument");
    if (outer * plane == 0) return MAAVSS_OK;
    magphase_combine_kernel<<<ew_blocks(outer * plane), 256, 0, (cudaStream_t)stream>>>(in, out, outer, plane);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bins_center(const float* in, float* out, long long rows, int K, void* stream) {
    MAAVSS_CHECK_ARG(in && out && in != out, "bins_center: null or aliased argument");
    if (rows * K == 0) return MAAVSS_OK;
    bins_center_kernel<<<ew_blocks(rows * K), 256, 0, (cudaStream_t)stream>>>(in, out, rows, K);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bins_decenter(const float* in, float* out, long long rows, int K, void* stream) {
    MAAVSS_CHECK_ARG(in && out && in != out, "bins_decenter: null or aliased argument");
    MAAVSS_CHECK_ARG((K & 1) == 0, "bins_decenter: odd K=%d (the reference's strided assignment raises for odd K)", K);
    if (rows * K == 0) return MAAVSS_OK;
    bins_decenter_kernel<<<ew_blocks(rows * K), 256, 0, (cudaStream_t)stream>>>(in, out, rows, K);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_dft_plan_create(maavss_dft_plan** out, int n) {
    MAAVSS_CHECK_ARG(out, "dft_plan_create: null argument");
    MAAVSS_CHECK_ARG(n >= 2 && n <= 8192, "dft_plan_create: n=%d out of range [2, 8192]", n);
    maavss_dft_plan* p = (maavss_dft_plan*)calloc(1, sizeof(maavss_dft_plan));
    p->n = n;
    float2* h = (float2*)malloc(sizeof(float2) * n);
    for (int q = 0; q < n; ++q) {
        const double ang = -2.0 * M_PI * (double)q / (double)n;
        h[q] = make_float2((float)cos(ang), (float)sin(ang));
    }
    cudaError_t e = cudaMalloc(&p->table, sizeof(float2) * n);
    if (e == cudaSuccess) e = cudaMemcpy(p->table, h, sizeof(float2) * n, cudaMemcpyHostToDevice);
    free(h);
    if (e != cudaSuccess) {
        maavss_set_error("dft_plan_create: %s", cudaGetErrorString(e));
        maavss_dft_plan_destroy(p);
        return MAAVSS_ERR_CUDA;
    }
    *out = p;
    return MAAVSS_OK;
}

extern "C" void maavss_dft_plan_destroy(maavss_dft_plan* p) {
    if (!p) return;
    if (p->table) cudaFree(p->table);
    free(p);
}

extern "C" int maavss_dft_fwd(const maavss_dft_plan* p, const float* audio, float* out_complex, int batch, int n_keep,
                              int norm_forward, void* stream) {
    MAAVSS_CHECK_ARG(p && audio && out_complex, "dft_fwd: null argument");
    MAAVSS_CHECK_ARG(n_keep >= 0 && n_keep <= p->n, "dft_fwd: n_keep=%d out of range", n_keep);
    if (batch == 0 || n_keep == 0) return MAAVSS_OK;
    const size_t smem = (sizeof(float2) + sizeof(float)) * p->n;
    MAAVSS_CUDA(cudaFuncSetAttribute(dft_real_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(ceil_div(n_keep, 128), batch);
    dft_real_fwd_kernel<<<grid, 128, smem, (cudaStream_t)stream>>>(audio, (float2*)out_complex, p->n, n_keep,
                                                                   norm_forward ? 1.0f / (float)p->n : 1.0f, p->table);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_dft_half_inv(const maavss_dft_plan* p, const float* in_complex, float* audio, int batch,
                                   int norm_forward, void* stream) {
    MAAVSS_CHECK_ARG(p && in_complex && audio, "dft_half_inv: null argument");
    MAAVSS_CHECK_ARG((p->n & 1) == 0, "dft_half_inv: plan length %d must be 2*K", p->n);
    if (batch == 0) return MAAVSS_OK;
    const int K = p->n / 2;
    const size_t smem = sizeof(float2) * (p->n + K);
    MAAVSS_CUDA(cudaFuncSetAttribute(dft_half_inv_real_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(ceil_div(p->n, 128), batch);
    // norm="forward" puts no factor on the inverse; norm=None (backward) puts 1/N on it.
    dft_half_inv_real_kernel<<<grid, 128, smem, (cudaStream_t)stream>>>((const float2*)in_complex, audio, K,
                                                                        norm_forward ? 1.0f : 1.0f / (float)p->n, p->table);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
