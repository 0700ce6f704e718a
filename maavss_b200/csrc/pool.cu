// nn.MaxPool3d((1, ph, pw)) of the AV_Fusion_Model_Frames visual encoder (avse_model_final.py:36,41,46,51,57):
// stride = kernel, no padding, floor mode, over the trailing (H, W) plane of every (n, c, d).
// Forward keeps the arg-max (first maximum in row-major window order, as ATen) in one byte per output so that the
// backward pass is a gather: windows are disjoint, every input element belongs to at most one window.
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
    const int blocks = (int)((total + 255) / 256 < 148LL * 32 ? (total + 255) / 256 : 148LL * 32);
    maxpool_bwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(dy, idx, dx, planes, H, W, ph, pw, OH, OW);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
