// Layout glue of the train steps that PyTorch would run as ATen copy kernels (torch.cat / slice copies): kept inside the
// library so that every launch of a step is one of ours (DESIGN.md "Execution model").  All HBM / L2 streaming copies.
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

// out[(j*B + b), c, t, f] = x[b, c, j*hop + t, f]: every (j, b, c) is one contiguous run of span*F floats.
template <typename V>
__global__ void __launch_bounds__(256) stack_windows_kernel(const V* __restrict__ x, V* __restrict__ out, int B, int C,
                                                            long long plane_in, long long run, long long hop_elems,
                                                            int chunks_per_run) {
    const long long blk = blockIdx.x / chunks_per_run;      // (j, b, c)
    const int chunk = blockIdx.x % chunks_per_run;
    const int c = (int)(blk % C);
    const int b = (int)((blk / C) % B);
    const int j = (int)(blk / ((long long)C * B));
    const V* src = x + ((long long)b * C + c) * plane_in + (long long)j * hop_elems;
    V* dst = out + blk * run;
    for (long long q = (long long)chunk * 256 + threadIdx.x; q < run; q += (long long)chunks_per_run * 256) dst[q] = __ldg(src + q);
}

// out[c] (+)= sum_r x[r * cols + c]: one thread per column walks the rows (coalesced across the warp); rows are few (the 64
// stacked window rows of a Linear layer's output gradient).
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ x, int rows, long long cols, float* __restrict__ out,
                                                     int accumulate) {
    const long long c = (long long)blockIdx.x * 256 + threadIdx.x;
    if (c >= cols) return;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int r = 0;
    for (; r + 3 < rows; r += 4) {
        s0 += __ldg(x + (size_t)r * cols + c);
        s1 += __ldg(x + (size_t)(r + 1) * cols + c);
        s2 += __ldg(x + (size_t)(r + 2) * cols + c);
        s3 += __ldg(x + (size_t)(r + 3) * cols + c);
    }
    for (; r < rows; ++r) s0 += __ldg(x + (size_t)r * cols + c);
    const float s = (s0 + s1) + (s2 + s3);
    out[c] = accumulate ? out[c] + s : s;
}

// Decoded video frames (T, H, W, 3) uint8 -> network input (T, 3, H, W) fp32: x / 255, then (x - mean[c]) / std[c]
// (av_dataset.py:312-317: permute, /255., Normalize).  One thread per pixel reads its 3 interleaved bytes and writes the three
// planes (coalesced fp32 stores; the byte loads of a warp cover 96 contiguous bytes).  Unfused division and subtraction so
// that the result is bit-identical to the torch expression (video / 255. - mean) / std.
__global__ void __launch_bounds__(256) frames_u8_kernel(const unsigned char* __restrict__ in, float* __restrict__ out, long long frames,
                                                        long long hw, float m0, float m1, float m2, float s0, float s1, float s2) {
    const long long n = frames * hw;
    for (long long p = (long long)blockIdx.x * 256 + threadIdx.x; p < n; p += (long long)gridDim.x * 256) {
        const long long f = p / hw, q = p - f * hw;
        const unsigned char* src = in + p * 3;
        float* dst = out + f * 3 * hw + q;
        dst[0] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)src[0], 255.f), m0), s0);
        dst[hw] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)src[1], 255.f), m1), s1);
        dst[2 * hw] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)src[2], 255.f), m2), s2);
    }
}

}  // namespace

extern "C" int maavss_frames_u8_normalize(const unsigned char* frames_hwc, float* out_chw, long long frames, int H, int W,
                                          const float* mean3_host, const float* std3_host, void* stream) {
    if (frames == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(frames_hwc && out_chw && mean3_host && std3_host && H > 0 && W > 0, "frames_u8_normalize: bad argument");
    const long long n = frames * H * W;
    int blocks = (int)((n + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    frames_u8_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(frames_hwc, out_chw, frames, (long long)H * W, mean3_host[0], mean3_host[1],
                                                                mean3_host[2], std3_host[0], std3_host[1], std3_host[2]);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_colsum(const float* x, int rows, long long cols, float* out, int accumulate, void* stream) {
    if (cols == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && out && rows >= 0, "colsum: bad argument");
    colsum_kernel<<<(unsigned)((cols + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, rows, cols, out, accumulate);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_stack_windows(const float* x, float* out, int B, int C, int T, int F, int G, int hop, int span, void* stream) {
    if (B == 0 || G == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && out, "stack_windows: null argument");
    MAAVSS_CHECK_ARG(B > 0 && C > 0 && G > 0 && span > 0 && hop >= 0 && (long long)(G - 1) * hop + span <= T,
                     "stack_windows: %d windows of %d rows (hop %d) do not fit %d rows", G, span, hop, T);
    cudaStream_t st = (cudaStream_t)stream;
    const long long plane = (long long)T * F, run = (long long)span * F, hop_e = (long long)hop * F;
    const long long runs = (long long)G * B * C;
    const bool v4 = plane % 4 == 0 && run % 4 == 0 && hop_e % 4 == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)out & 15) == 0;
    const long long units = v4 ? run / 4 : run;
    int cpr = (int)((units + 1023) / 1024);
    if (cpr < 1) cpr = 1;
    MAAVSS_CHECK_ARG(runs * cpr < (1ll << 31), "stack_windows: grid too large");
    if (v4)
        stack_windows_kernel<float4><<<(unsigned)(runs * cpr), 256, 0, st>>>(reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(out),
                                                                            B, C, plane / 4, run / 4, hop_e / 4, cpr);
    else
        stack_windows_kernel<float><<<(unsigned)(runs * cpr), 256, 0, st>>>(x, out, B, C, plane, run, hop_e, cpr);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
