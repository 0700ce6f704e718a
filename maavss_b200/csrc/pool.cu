// nn.MaxPool3d((1, ph, pw)) of the AV_Fusion_Model_Frames visual encoder (avse_model_final.py:36,41,46,51,57):
// stride = kernel, no padding, floor mode, over the trailing (H, W) plane of every (n, c, d).
// Forward keeps the arg-max (first maximum in row-major window order, as ATen) in one byte per output so that the
// backward pass is a gather: windows are disjoint, every input element belongs to at most one window.
#include <stdint.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

__global__ void maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, unsigned char* __restrict__ idx,
                                   long long planes, int H, int W, int ph, int pw, int OH, int OW) {
    const long long total = planes * OH * OW;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const int ow = (int)(e % OW);
        const long long q = e / OW;
        const int oh = (int)(q % OH);
        const long long pl = q / OH;
        const float* src = x + (pl * H + (long long)oh * ph) * W + (long long)ow * pw;
        float best = src[0];
        int arg = 0;
        for (int i = 0; i < ph; ++i)
            for (int j = 0; j < pw; ++j) {
                const float v = __ldg(src + (long long)i * W + j);
                if (v > best || (v != v && best == best)) {      // first maximum wins; NaN propagates like ATen
                    best = v;
                    arg = i * pw + j;
                }
            }
        y[e] = best;
        idx[e] = (unsigned char)arg;
    }
}

__global__ void maxpool_bwd_kernel(const float* __restrict__ dy, const unsigned char* __restrict__ idx, float* __restrict__ dx,
                                   long long planes, int H, int W, int ph, int pw, int OH, int OW) {
    const long long total = planes * H * W;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const int w = (int)(e % W);
        const long long q = e / W;
        const int h = (int)(q % H);
        const long long pl = q / H;
        const int oh = h / ph, ow = w / pw;
        float g = 0.f;
        if (oh < OH && ow < OW) {
            const long long o = (pl * OH + oh) * OW + ow;
            if (idx[o] == (unsigned char)((h - oh * ph) * pw + (w - ow * pw))) g = dy[o];
        }
        dx[e] = g;
    }
}

// 2x2 windows (every pool of avse_model_final.py:36-57), W % 4 == 0: a thread writes one float4 of dx (two windows of one
// input row), 32-bit index arithmetic, the two dy / arg-max values it needs as one 8-byte / 2-byte load.
__global__ void __launch_bounds__(256) maxpool_bwd_2x2_kernel(const float* __restrict__ dy, const unsigned char* __restrict__ idx,
                                                              float* __restrict__ dx, unsigned rows, unsigned H, unsigned W4, unsigned OH,
                                                              unsigned OW) {
    const unsigned total4 = rows * W4;
    for (unsigned e4 = blockIdx.x * 256u + threadIdx.x; e4 < total4; e4 += gridDim.x * 256u) {
        const unsigned row = e4 / W4, w4 = e4 - row * W4;
        const unsigned pl = row / H, h = row - pl * H;
        const unsigned oh = h >> 1;
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        if (oh < OH) {
            const size_t o = ((size_t)pl * OH + oh) * OW + 2 * w4;          // even: the pair loads are aligned
            const float2 d = *reinterpret_cast<const float2*>(dy + o);
            const uchar2 c = *reinterpret_cast<const uchar2*>(idx + o);
            const unsigned char base = (unsigned char)((h & 1) * 2);
            g.x = c.x == base ? d.x : 0.f;
            g.y = c.x == base + 1 ? d.x : 0.f;
            g.z = c.y == base ? d.y : 0.f;
            g.w = c.y == base + 1 ? d.y : 0.f;
        }
        reinterpret_cast<float4*>(dx)[e4] = g;
    }
}

}  // namespace

extern "C" int maavss_maxpool_fwd(const float* x, float* y, unsigned char* idx, long long planes, int H, int W, int ph, int pw,
                                  void* stream) {
    MAAVSS_CHECK_ARG(ph > 0 && pw > 0 && ph * pw <= 255 && H >= ph && W >= pw, "maxpool_fwd: bad window %dx%d for %dx%d", ph, pw, H, W);
    if (planes == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && y && idx, "maxpool_fwd: null argument");
    const int OH = H / ph, OW = W / pw;
    const long long total = planes * OH * OW;
    const int blocks = (int)((total + 255) / 256 < 148LL * 32 ? (total + 255) / 256 : 148LL * 32);
    maxpool_fwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, y, idx, planes, H, W, ph, pw, OH, OW);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_maxpool_bwd(const float* dy, const unsigned char* idx, float* dx, long long planes, int H, int W, int ph,
                                  int pw, void* stream) {
    MAAVSS_CHECK_ARG(ph > 0 && pw > 0 && ph * pw <= 255 && H >= ph && W >= pw, "maxpool_bwd: bad window");
    if (planes == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dy && idx && dx, "maxpool_bwd: null argument");
    const int OH = H / ph, OW = W / pw;
    const long long total = planes * H * W;
    if (ph == 2 && pw == 2 && W % 4 == 0 && total / 4 < (1LL << 31) && (((uintptr_t)dy | (uintptr_t)dx) & 15) == 0 && ((uintptr_t)idx & 1) == 0) {
        const long long t4 = total / 4;
        const int blocks4 = (int)((t4 + 255) / 256 < 148LL * 32 ? (t4 + 255) / 256 : 148LL * 32);
        maxpool_bwd_2x2_kernel<<<blocks4, 256, 0, (cudaStream_t)stream>>>(dy, idx, dx, (unsigned)(planes * H), (unsigned)H, (unsigned)(W / 4),
                                                                         (unsigned)OH, (unsigned)OW);
        MAAVSS_LAUNCH_CHECK();
        return MAAVSS_OK;
    }
    const int blocks = (int)((total + 255) / 256 < 148LL * 32 ? (total + 255) / 256 : 148LL * 32);
    maxpool_bwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(dy, idx, dx, planes, H, W, ph, pw, OH, OW);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
