// Bandwidth kernels of the ViT-S/8 attention-map extractor (video_attention.py:38-103 over dino/vision_transformer.py):
// patch gather (PatchEmbed's 8x8/8 conv as a GEMM operand), token assembly (CLS + pos-embed), LayerNorm -> bf16,
// fp32 -> bf16 weight conversion, and the last block's CLS-query softmax row.
#include <cuda_bf16.h>

#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

// frames (T, C, H, W) fp32, cropped to (gh*P, gw*P) -> patches (T*gh*gw, C*P*P) bf16, column = (c, ky, kx)
__global__ void patchify_kernel(const float* __restrict__ frames, __nv_bfloat16* __restrict__ out, int T, int C, int H, int W,
                                int gh, int gw, int P) {
    const int Kc = C * P * P;
    const long long total = (long long)T * gh * gw * Kc;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const int col = (int)(e % Kc);
        const long long row = e / Kc;
        const int kx = col % P, ky = (col / P) % P, c = col / (P * P);
        const int px = (int)(row % gw), py = (int)((row / gw) % gh), t = (int)(row / ((long long)gw * gh));
        out[e] = __float2bfloat16(frames[(((size_t)t * C + c) * H + py * P + ky) * W + px * P + kx]);
    }
}

// x (T, 1 + np, D) = cat(cls, patch_tokens (T*np, D)) + pos (1 + np, D)
__global__ void assemble_tokens_kernel(const float* __restrict__ patch_tok, const float* __restrict__ cls, const float* __restrict__ pos,
                                       float* __restrict__ x, int T, int np, int D) {
    const long long total = (long long)T * (np + 1) * D;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const int d = (int)(e % D);
        const long long r = e / D;
        const int n = (int)(r % (np + 1)), t = (int)(r / (np + 1));
        const float v = n == 0 ? cls[d] : patch_tok[((size_t)t * np + n - 1) * D + d];
        x[e] = v + pos[(size_t)n * D + d];
    }
}

// LayerNorm over the last dim (D = 32 * VPL), one warp per row, two-pass in registers; output bf16
template <int VPL>
__global__ void __launch_bounds__(256) layernorm_bf16_kernel(const float* __restrict__ x, const float* __restrict__ g,
                                                             const float* __restrict__ b, __nv_bfloat16* __restrict__ y,
                                                             long long rows, float eps) {
    constexpr int D = 32 * VPL;
    const int lane = threadIdx.x & 31;
    const long long row = blockIdx.x * 8LL + (threadIdx.x >> 5);
    if (row >= rows) return;
    const float* xr = x + row * D;
    float v[VPL];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        v[i] = xr[lane + 32 * i];
        s += v[i];
    }
    const float mean = warp_sum(s) * (1.f / D);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const float d = v[i] - mean;
        q = fmaf(d, d, q);
    }
    const float inv = rsqrtf(warp_sum(q) * (1.f / D) + eps);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = lane + 32 * i;
        y[row * D + c] = __float2bfloat16((v[i] - mean) * inv * g[c] + b[c]);
    }
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x)
        out[e] = __float2bfloat16(in[e]);
}

// Last block, CLS query only (dino/vision_transformer.py:82-90 with the query row 0; video_attention.py:56):
// q (T, D) fp32 = CLS rows of the q projection, k (T*N, D) fp32 (the last projections are kept in fp32: their logits
// feed the output softmax directly); one warp per (frame, head): out (T, heads, N) = softmax_j(q_h . k_{j,h} * scale)
__global__ void __launch_bounds__(128) cls_attn_row_kernel(const float* __restrict__ q, long long ldq, const float* __restrict__ k,
                                                           long long ldk, float* __restrict__ out, int T, int heads, int N,
                                                           float scale) {
    const int lane = threadIdx.x & 31;
    const int wid = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (wid >= T * heads) return;
    const int t = wid / heads, h = wid % heads;
    // head_dim 64: lane holds q[2*lane], q[2*lane + 1]
    const float2 qv = *reinterpret_cast<const float2*>(q + (size_t)t * ldq + h * 64 + 2 * lane);
    float* o = out + (size_t)wid * N;
    float mx = -INFINITY;
    for (int j0 = 0; j0 < N; j0 += 4) {
        float s[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int j = min(j0 + u, N - 1);
            const float2 kv = *reinterpret_cast<const float2*>(k + ((size_t)t * N + j) * ldk + h * 64 + 2 * lane);
            s[u] = qv.x * kv.x + qv.y * kv.y;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float v = warp_sum(s[u]) * scale;
            if (j0 + u < N) {
                if (lane == 0) o[j0 + u] = v;
                mx = fmaxf(mx, v);
            }
        }
    }
    __syncwarp();
    float sum = 0.f;
    for (int j = lane; j < N; j += 32) {
        const float e = __expf(o[j] - mx);
        o[j] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int j = lane; j < N; j += 32) o[j] *= inv;
}

int grid_for(long long total) {
    const long long b = (total + 255) / 256;
    return (int)(b < 148LL * 16 ? (b > 0 ? b : 1) : 148LL * 16);
}

}  // namespace

extern "C" int maavss_vit_patchify(const float* frames, void* patches_bf16, int T, int C, int H, int W, int gh, int gw, int P,
                                   void* stream) {
    if (T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(frames && patches_bf16 && gh * P <= H && gw * P <= W && P > 0, "vit_patchify: bad argument");
    patchify_kernel<<<grid_for((long long)T * gh * gw * C * P * P), 256, 0, (cudaStream_t)stream>>>(
        frames, (__nv_bfloat16*)patches_bf16, T, C, H, W, gh, gw, P);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_vit_assemble_tokens(const float* patch_tok, const float* cls, const float* pos, float* x, int T, int np, int D,
                                          void* stream) {
    if (T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(patch_tok && cls && pos && x, "vit_assemble_tokens: null argument");
    assemble_tokens_kernel<<<grid_for((long long)T * (np + 1) * D), 256, 0, (cudaStream_t)stream>>>(patch_tok, cls, pos, x, T, np, D);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_layernorm_bf16(const float* x, const float* gamma, const float* beta, void* y_bf16, long long rows, int D,
                                     float eps, void* stream) {
    if (rows == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && gamma && beta && y_bf16, "layernorm_bf16: null argument");
    MAAVSS_CHECK_ARG(D % 32 == 0 && D >= 32 && D <= 1024, "layernorm_bf16: D=%d must be a multiple of 32 up to 1024", D);
    const int blocks = (int)((rows + 7) / 8);
    cudaStream_t st = (cudaStream_t)stream;
    __nv_bfloat16* y = (__nv_bfloat16*)y_bf16;
    switch (D / 32) {
#define LN_CASE(V) case V: layernorm_bf16_kernel<V><<<blocks, 256, 0, st>>>(x, gamma, beta, y, rows, eps); break;
        LN_CASE(1) LN_CASE(2) LN_CASE(3) LN_CASE(4) LN_CASE(6) LN_CASE(8) LN_CASE(12) LN_CASE(16) LN_CASE(24) LN_CASE(32)
#undef LN_CASE
        default:
            maavss_set_error("layernorm_bf16: D=%d not instantiated", D);
            return MAAVSS_ERR_UNSUPPORTED;
    }
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_f32_to_bf16(const float* in, void* out_bf16, long long n, void* stream) {
    if (n == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(in && out_bf16, "f32_to_bf16: null argument");
    f32_to_bf16_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(in, (__nv_bfloat16*)out_bf16, n);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_vit_cls_attn_row(const float* q, long long ldq, const float* k, long long ldk, float* out, int T, int heads,
                                       int N, float scale, void* stream) {
    if (T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(q && k && out && heads > 0 && N > 0 && ldq % 2 == 0 && ldk % 2 == 0, "vit_cls_attn_row: bad argument");
    cls_attn_row_kernel<<<ceil_div(T * heads, 4), 128, 0, (cudaStream_t)stream>>>(q, ldq, k, ldk, out, T, heads, N, scale);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
