// Attention-map post-processing kernels (SURVEY 8a rows a6 tail, a8):
//   * head-sum + nearest x patch upsample + per-frame max-normalise (video_attention.py:80-96)
//   * temporal frame difference + per-clip max-normalise           (av_dataset.py:323-328)
// HBM-bound elementwise/reduction work: coalesced float4 traffic, per-clip signed max through an
// order-preserving uint atomicMax.
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

__device__ __forceinline__ unsigned int float_to_ordered(float f) {
    const unsigned int b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(unsigned int k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__device__ __forceinline__ float diff_at(const float* __restrict__ clip, long long hw, int t, long long r, int diff) {
    if (!diff) return clip[(size_t)t * hw + r];
    if (t == 0) return 0.f;                                   // the zeros_like(attn[0:1]) pad
    return clip[(size_t)t * hw + r] - clip[(size_t)(t - 1) * hw + r];
}

// grid (x, clips): signed max over the (optionally differenced) clip.
__global__ void diff_max_kernel(const float* __restrict__ in, int T, long long hw, int diff, unsigned int* __restrict__ ws) {
    const float* clip = in + (size_t)blockIdx.y * T * hw;
    const long long total = (long long)T * hw;
    float m = -INFINITY;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(q / hw);
        m = fmaxf(m, diff_at(clip, hw, t, q - (long long)t * hw, diff));
    }
    m = warp_max(m);
    __shared__ float sm[32];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = (threadIdx.x < (blockDim.x >> 5)) ? sm[threadIdx.x] : -INFINITY;
        m = warp_max(m);
        if (threadIdx.x == 0) atomicMax(ws + blockIdx.y, float_to_ordered(m));
    }
}

__global__ void diff_scale_kernel(const float* __restrict__ in, float* __restrict__ out, int T, long long hw, int diff,
                                  const unsigned int* __restrict__ ws) {
    const float* clip = in + (size_t)blockIdx.y * T * hw;
    float* o = out + (size_t)blockIdx.y * T * hw;
    const float inv = 1.0f / ordered_to_float(ws[blockIdx.y]);   // signed max, as in `attn *= 1/torch.max(attn)`
    const long long total = (long long)T * hw;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(q / hw);
        o[q] = diff_at(clip, hw, t, q - (long long)t * hw, diff) * inv;
    }
}

// One CTA per frame.  cls: (frames, heads, cls_stride) softmax rows of the CLS query; entries
// [1, 1+gh*gw) are the patch tokens.  out: (frames, gh*p, gw*p).
__global__ void headsum_upsample_kernel(const float* __restrict__ cls, int heads, int cls_stride, int gh, int gw, int p,
                                        float* __restrict__ out) {
    extern __shared__ float s[];   // gh*gw summed map
    const int np = gh * gw;
    const float* c = cls + (size_t)blockIdx.x * heads * cls_stride + 1;
    float m = -INFINITY;
    for (int q = threadIdx.x; q < np; q += blockDim.x) {
        float a = 0.f;
        for (int h = 0; h < heads; ++h) a += c[(size_t)h * cls_stride + q];   // torch.sum over heads, ascending
        s[q] = a;
        m = fmaxf(m, a);
    }
    __shared__ float red[32];
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : -INFINITY;
        m = warp_max(m);
        if (threadIdx.x == 0) red[0] = m;
    }
    __syncthreads();
    const float inv = 1.0f / red[0];
    const int W = gw * p, H = gh * p;
    float* o = out + (size_t)blockIdx.x * H * W;
    for (int q = threadIdx.x; q < H * W; q += blockDim.x) {
        const int y = q / W, x = q - y * W;
        o[q] = s[(y / p) * gw + (x / p)] * inv;
    }
}

inline int grid_x(long long n) {
    long long b = (n + 256 * 8 - 1) / (256 * 8);
    if (b > 148 * 4) b = 148 * 4;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace

extern "C" int maavss_attn_diff_normalise(const float* in, float* out, int clips, int frames, long long hw, int diff,
                                          void* workspace, void* stream) {
    MAAVSS_CHECK_ARG(in && out && workspace, "attn_diff_normalise: null argument");
    MAAVSS_CHECK_ARG(clips >= 0 && clips <= 65535, "attn_diff_normalise: clips=%d out of range", clips);
    if (clips == 0 || frames == 0 || hw == 0) return MAAVSS_OK;
    cudaStream_t st = (cudaStream_t)stream;
    MAAVSS_CUDA(cudaMemsetAsync(workspace, 0, sizeof(unsigned int) * clips, st));   // 0 orders below every float
    dim3 grid(grid_x((long long)frames * hw), clips);
    diff_max_kernel<<<grid, 256, 0, st>>>(in, frames, hw, diff, (unsigned int*)workspace);
    MAAVSS_LAUNCH_CHECK();
    diff_scale_kernel<<<grid, 256, 0, st>>>(in, out, frames, hw, diff, (const unsigned int*)workspace);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_attn_headsum_upsample(const float* cls_rows, float* out, int frames, int heads, int cls_stride,
                                            int grid_h, int grid_w, int patch, void* stream) {
    MAAVSS_CHECK_ARG(cls_rows && out, "attn_headsum_upsample: null argument");
    MAAVSS_CHECK_ARG(cls_stride >= 1 + grid_h * grid_w, "attn_headsum_upsample: cls_stride too small");
    if (frames == 0) return MAAVSS_OK;
    const size_t smem = sizeof(float) * grid_h * grid_w;
    headsum_upsample_kernel<<<frames, 256, smem, (cudaStream_t)stream>>>(cls_rows, heads, cls_stride, grid_h, grid_w, patch, out);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
