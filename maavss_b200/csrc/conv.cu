// Per-channel sums (bias gradients) for the convolution layers; the convolutions themselves are in conv_igemm.cu.
#include <stdint.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

// out[c] (+)= sum_{n,i} x[n,c,i]; two deterministic stages through a double workspace [C][S].
// Stage 1: CTA (s, c) walks the rows n = s, s + S, ... of channel c with float4 loads and four independent fp32 partial
// sums per thread (no index division in the loop), then reduces in double.
constexpr int kSumSplits = 16;
__global__ void __launch_bounds__(256) channel_sum_partial(const float* __restrict__ x, int N, int C, long long inner,
                                                           double* __restrict__ ws) {
    const int c = blockIdx.y, s = blockIdx.x;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    const bool vec = (inner & 3) == 0 && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
    // the 16 splits cover (row phase) x (inner range): few rows -> the inner axis is split instead
    int rs = 1;
    while (rs < kSumSplits && rs * 2 <= N) rs *= 2;
    const int is = kSumSplits / rs, sn = s % rs, si = s / rs;
    const long long chunk = (((inner + is - 1) / is) + 3) & ~3ll;
    const long long i0 = min(inner, si * chunk), i1 = min(inner, i0 + chunk);
    for (int n = sn; n < N; n += rs) {
        const float* row = x + ((size_t)n * C + c) * inner;
        if (vec) {
            const float4* r4 = reinterpret_cast<const float4*>(row);
            for (long long i = (i0 >> 2) + threadIdx.x; i < (i1 >> 2); i += 256) {
                const float4 v = r4[i];
                a0 += v.x; a1 += v.y; a2 += v.z; a3 += v.w;
            }
        } else {
            for (long long i = i0 + threadIdx.x; i < i1; i += 256) a0 += row[i];
        }
    }
    double v = ((double)a0 + (double)a1) + ((double)a2 + (double)a3);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __shared__ double sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int q = 0; q < 8; ++q) t += sm[q];
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

}  // namespace

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

// ------------------------------------------------------------------------------------------------------------------
// Allocation-free variant of maavss_channel_sum (no counterpart in the reference):
// the [C][16] double partial sums live in a file-scope device array instead of a caller workspace, so the call
// needs no cudaMalloc'd scratch -- for callers under CUDA graph capture, tight serving loops or harnesses that forbid
// allocation.  512 KB of device memory per GPU that loads the library, held until exit.  Every call shares the array:
// stage 1 overwrites all 16 * C entries that stage 2 reads (nothing has to start at zero), but two calls that overlap on
// the GPU (different streams, no ordering) race on it.
#include <mutex>

constexpr int kStaticSumMaxC = 4096;
static __device__ __align__(256) double g_channel_sum_scratch[(size_t)kSumSplits * kStaticSumMaxC];

namespace {

// Device address of g_channel_sum_scratch on the current device.  In host code the array's name is only a handle for the
// runtime, not a device pointer: the address comes from cudaGetSymbolAddress, looked up once per device with that device
// current (no allocation; it may load the module, so the first call per device belongs outside stream capture).
int static_scratch_ptr(double** out) {
    constexpr int kMaxDev = 64;
    static double* ptrs[kMaxDev] = {};
    static std::mutex mu;
    int dev = -1;
    MAAVSS_CUDA(cudaGetDevice(&dev));
    MAAVSS_CHECK_ARG(dev >= 0 && dev < kMaxDev, "channel_sum_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        double* p = nullptr;
        MAAVSS_CUDA(cudaGetSymbolAddress((void**)&p, g_channel_sum_scratch));
        MAAVSS_CHECK_ARG(p != nullptr, "channel_sum_static: cudaGetSymbolAddress returned a null address");
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return MAAVSS_OK;
}

}  // namespace

extern "C" int maavss_channel_sum_static_max_channels(void) { return kStaticSumMaxC; }

extern "C" int maavss_channel_sum_static(const float* x, int N, int C, long long inner, float* out, int accumulate, void* stream) {
    if (C == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && out, "channel_sum_static: null argument");
    MAAVSS_CHECK_ARG(N >= 0 && C > 0 && inner >= 0, "channel_sum_static: bad shape N=%d C=%d inner=%lld", N, C, inner);
    MAAVSS_CHECK_ARG(C <= kStaticSumMaxC,
                     "channel_sum_static: C=%d exceeds the static scratch limit of %d channels (use maavss_channel_sum with a workspace)", C,
                     kStaticSumMaxC);
    double* ws = nullptr;
    int rc = static_scratch_ptr(&ws);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid(kSumSplits, C);
    channel_sum_partial<<<grid, 256, 0, st>>>(x, N, C, inner, ws);
    MAAVSS_LAUNCH_CHECK();
    channel_sum_final<<<ceil_div(C, 128), 128, 0, st>>>(ws, C, out, accumulate);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
