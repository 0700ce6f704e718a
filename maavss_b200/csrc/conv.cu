// Per-channel sums (bias gradients) for the convolution layers; the convolutions themselves are in conv_igemm.cu.
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

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
