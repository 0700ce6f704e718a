// One-call ViT attention-map extractor behind the C ABI (maavss_vit_create / maavss_vit_attn_maps):
// VideoAttention._inference (video_attention.py:38-103) over dino/vision_transformer.py get_last_selfattention (:216-223),
// optionally followed by the dataset's frame differencing + clip normalise (av_dataset.py:323-328), for all frames of a call
// as one batch.  The plan owns bf16 copies of the GEMM weights, fp32 copies of biases / LayerNorm parameters and the
// bicubically interpolated position embedding of each prepared input size (interpolate_pos_encoding, :174-194); activations
// live in a caller-owned workspace.  Every launch is one of the library's own kernels (tc_gemm.cu, tc_attn.cu, vit.cu,
// attnmap.cu); nothing is allocated or synchronised on the per-call path.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

struct maavss_vit_plan {
    int dim, depth, heads, patch, in_chans, pos_grid;
    float *cls, *pos, *patch_b;
    void* patch_w;                       // bf16 (dim, in_chans * P * P)
    struct Block {
        float *ln1_w, *ln1_b, *qkv_b, *proj_b, *ln2_w, *ln2_b, *fc1_b, *fc2_b;
        void *qkv_w, *proj_w, *fc1_w, *fc2_w;   // bf16
    }* blocks;
    // interpolated position embeddings, one per prepared (gh, gw)
    struct Pos { int gh, gw; float* data; } pos_cache[8];
    int n_pos;
};

namespace {

// torch upsample_bicubic2d, align_corners = False, scale_factor given: source = (dst + 0.5) / scale - 0.5, cubic convolution
// with A = -0.75, border indices clamped.  in (s * s, D) tokens -> out (gh * gw, D).
__device__ __forceinline__ void cubic_coeffs(float t, float (&c)[4]) {
    const float A = -0.75f;
    float x = t + 1.f;
    c[0] = ((A * x - 5.f * A) * x + 8.f * A) * x - 4.f * A;
    x = t;
    c[1] = ((A + 2.f) * x - (A + 3.f)) * x * x + 1.f;
    x = 1.f - t;
    c[2] = ((A + 2.f) * x - (A + 3.f)) * x * x + 1.f;
    x = 2.f - t;
    c[3] = ((A * x - 5.f * A) * x + 8.f * A) * x - 4.f * A;
}

__global__ void __launch_bounds__(128) pos_bicubic_kernel(const float* __restrict__ in, float* __restrict__ out, int s, int gh, int gw, int D,
                                                          float inv_sy, float inv_sx) {
    const int tok = blockIdx.x, oy = tok / gw, ox = tok % gw;
    const float ry = ((float)oy + 0.5f) * inv_sy - 0.5f, rx = ((float)ox + 0.5f) * inv_sx - 0.5f;
    const int iy = (int)floorf(ry), ix = (int)floorf(rx);
    float cy[4], cx[4];
    cubic_coeffs(ry - (float)iy, cy);
    cubic_coeffs(rx - (float)ix, cx);
    for (int d = threadIdx.x; d < D; d += 128) {
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int yy = min(max(iy - 1 + i, 0), s - 1);
            float row = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int xx = min(max(ix - 1 + j, 0), s - 1);
                row += in[((size_t)yy * s + xx) * D + d] * cx[j];
            }
            acc += row * cy[i];
        }
        out[(size_t)tok * D + d] = acc;
    }
}

int dev_copy(float** dst, const float* src, size_t n) {
    MAAVSS_CUDA(cudaMalloc(dst, n * sizeof(float)));
    MAAVSS_CUDA(cudaMemcpy(*dst, src, n * sizeof(float), cudaMemcpyDeviceToDevice));
    return MAAVSS_OK;
}

int dev_bf16(void** dst, const float* src, size_t n) {
    MAAVSS_CUDA(cudaMalloc(dst, n * 2));
    const int rc = maavss_f32_to_bf16(src, *dst, (long long)n, nullptr);
    if (rc) return rc;
    MAAVSS_CUDA(cudaStreamSynchronize(nullptr));
    return MAAVSS_OK;
}

inline size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

struct Layout {
    size_t patches, tok, x, y, qkv, o, hdn, k, q, rows, maps, mx, total;
};

Layout layout_for(const maavss_vit_plan* p, int T, int gh, int gw, int clips) {
    const size_t np = (size_t)gh * gw, N = np + 1, D = p->dim;
    Layout L;
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += align_up(bytes); return o; };
    L.patches = take(2 * T * np * (size_t)p->in_chans * p->patch * p->patch);
    L.tok = take(4 * T * np * D);
    L.x = take(4 * T * N * D);
    L.y = take(2 * T * N * D);
    L.qkv = take(2 * T * N * 3 * D);
    L.o = take(2 * T * N * D);
    L.hdn = take(2 * T * N * 4 * D);
    L.k = take(4 * T * N * D);
    L.q = take(4 * (size_t)T * D);
    L.rows = take(4 * (size_t)T * p->heads * N);
    L.maps = take(4 * (size_t)T * gh * p->patch * gw * p->patch);
    L.mx = take(4 * (size_t)(clips > 0 ? clips : 1));
    L.total = off;
    return L;
}

}  // namespace

extern "C" void maavss_vit_destroy(maavss_vit_plan* p) {
    if (!p) return;
    cudaFree(p->cls); cudaFree(p->pos); cudaFree(p->patch_b); cudaFree(p->patch_w);
    if (p->blocks) {
        for (int i = 0; i < p->depth; ++i) {
            maavss_vit_plan::Block& b = p->blocks[i];
            float* f[] = {b.ln1_w, b.ln1_b, b.qkv_b, b.proj_b, b.ln2_w, b.ln2_b, b.fc1_b, b.fc2_b};
            for (float* q : f) cudaFree(q);
            void* v[] = {b.qkv_w, b.proj_w, b.fc1_w, b.fc2_w};
            for (void* q : v) cudaFree(q);
        }
        free(p->blocks);
    }
    for (int i = 0; i < p->n_pos; ++i) cudaFree(p->pos_cache[i].data);
    free(p);
}

extern "C" int maavss_vit_create(maavss_vit_plan** out, int dim, int depth, int patch, int in_chans, int pos_grid, const float* cls_token,
                                 const float* pos_embed, const float* patch_w, const float* patch_b, const float* const* block_tensors) {
    MAAVSS_CHECK_ARG(out && cls_token && pos_embed && patch_w && patch_b && block_tensors, "vit_create: null argument");
    MAAVSS_CHECK_ARG(dim % 64 == 0 && dim >= 64 && depth >= 1 && patch >= 1 && in_chans >= 1 && pos_grid >= 1, "vit_create: bad geometry");
    MAAVSS_CHECK_ARG((in_chans * patch * patch) % 8 == 0, "vit_create: in_chans * patch^2 must be a multiple of 8 (TMA pitch)");
    maavss_vit_plan* p = (maavss_vit_plan*)calloc(1, sizeof(maavss_vit_plan));
    p->dim = dim; p->depth = depth; p->heads = dim / 64; p->patch = patch; p->in_chans = in_chans; p->pos_grid = pos_grid;
    p->blocks = (maavss_vit_plan::Block*)calloc(depth, sizeof(maavss_vit_plan::Block));
    const size_t D = dim;
    int rc = dev_copy(&p->cls, cls_token, D);
    if (!rc) rc = dev_copy(&p->pos, pos_embed, (1 + (size_t)pos_grid * pos_grid) * D);
    if (!rc) rc = dev_copy(&p->patch_b, patch_b, D);
    if (!rc) rc = dev_bf16(&p->patch_w, patch_w, D * in_chans * patch * patch);
    for (int i = 0; i < depth && !rc; ++i) {
        const float* const* t = block_tensors + 12 * i;   // ln1_w ln1_b qkv_w qkv_b proj_w proj_b ln2_w ln2_b fc1_w fc1_b fc2_w fc2_b
        maavss_vit_plan::Block& b = p->blocks[i];
        for (int q = 0; q < 12 && !rc; ++q)
            if (!t[q]) { maavss_set_error("vit_create: block %d tensor %d is null", i, q); rc = MAAVSS_ERR_INVALID; }
        if (!rc) rc = dev_copy(&b.ln1_w, t[0], D);
        if (!rc) rc = dev_copy(&b.ln1_b, t[1], D);
        if (!rc) rc = dev_bf16(&b.qkv_w, t[2], 3 * D * D);
        if (!rc) rc = dev_copy(&b.qkv_b, t[3], 3 * D);
        if (!rc) rc = dev_bf16(&b.proj_w, t[4], D * D);
        if (!rc) rc = dev_copy(&b.proj_b, t[5], D);
        if (!rc) rc = dev_copy(&b.ln2_w, t[6], D);
        if (!rc) rc = dev_copy(&b.ln2_b, t[7], D);
        if (!rc) rc = dev_bf16(&b.fc1_w, t[8], 4 * D * D);
        if (!rc) rc = dev_copy(&b.fc1_b, t[9], 4 * D);
        if (!rc) rc = dev_bf16(&b.fc2_w, t[10], 4 * D * D);
        if (!rc) rc = dev_copy(&b.fc2_b, t[11], D);
    }
    if (rc) {
        maavss_vit_destroy(p);
        return rc;
    }
    *out = p;
    return MAAVSS_OK;
}

// interpolate_pos_encoding for an (H, W) input, evaluated once per size (allocates inside the plan: not a per-call path).
extern "C" int maavss_vit_prepare(maavss_vit_plan* p, int H, int W) {
    MAAVSS_CHECK_ARG(p && H >= p->patch && W >= p->patch, "vit_prepare: bad argument");
    const int gh = H / p->patch, gw = W / p->patch;
    for (int i = 0; i < p->n_pos; ++i)
        if (p->pos_cache[i].gh == gh && p->pos_cache[i].gw == gw) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(p->n_pos < 8, "vit_prepare: more than 8 input sizes on one plan");
    const size_t D = p->dim, N = 1 + (size_t)gh * gw;
    float* data = nullptr;
    MAAVSS_CUDA(cudaMalloc(&data, N * D * sizeof(float)));
    MAAVSS_CUDA(cudaMemcpy(data, p->pos, D * sizeof(float), cudaMemcpyDeviceToDevice));          // class token's embedding
    if (gh == p->pos_grid && gw == p->pos_grid) {
        MAAVSS_CUDA(cudaMemcpy(data + D, p->pos + D, (N - 1) * D * sizeof(float), cudaMemcpyDeviceToDevice));
    } else {
        // dino adds 0.1 to the target grid "to avoid floating point error in the interpolation" (:186-190); torch's
        // scale_factor path maps destination to source coordinates with 1 / scale_factor
        const float inv_sy = (float)((double)p->pos_grid / ((double)gh + 0.1)), inv_sx = (float)((double)p->pos_grid / ((double)gw + 0.1));
        pos_bicubic_kernel<<<gh * gw, 128>>>(p->pos + D, data + D, p->pos_grid, gh, gw, p->dim, inv_sy, inv_sx);
        MAAVSS_LAUNCH_CHECK();
        MAAVSS_CUDA(cudaStreamSynchronize(nullptr));
    }
    p->pos_cache[p->n_pos++] = {gh, gw, data};
    return MAAVSS_OK;
}

extern "C" long long maavss_vit_workspace_bytes(const maavss_vit_plan* p, int T, int H, int W, int clips) {
    if (!p || T <= 0) return 0;
    return (long long)layout_for(p, T, H / p->patch, W / p->patch, clips).total + 256;
}

extern "C" int maavss_vit_pos_embed(const maavss_vit_plan* p, int H, int W, float* out) {
    MAAVSS_CHECK_ARG(p && out, "vit_pos_embed: null argument");
    const int gh = H / p->patch, gw = W / p->patch;
    for (int i = 0; i < p->n_pos; ++i)
        if (p->pos_cache[i].gh == gh && p->pos_cache[i].gw == gw) {
            MAAVSS_CUDA(cudaMemcpy(out, p->pos_cache[i].data, (1 + (size_t)gh * gw) * p->dim * sizeof(float), cudaMemcpyDeviceToDevice));
            return MAAVSS_OK;
        }
    maavss_set_error("vit_pos_embed: size %dx%d was not prepared", H, W);
    return MAAVSS_ERR_INVALID;
}

extern "C" int maavss_vit_attn_maps(const maavss_vit_plan* p, const float* frames, int T, int H, int W, float* out, int post, int clips,
                                    void* workspace, long long workspace_bytes, void* stream) {
    if (T == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(p && frames && out && workspace, "vit_attn_maps: null argument");
    MAAVSS_CHECK_ARG(post >= 0 && post <= 2 && (post == 0 || (clips >= 1 && T % clips == 0)), "vit_attn_maps: bad post / clips");
    const int P = p->patch, D = p->dim, gh = H / P, gw = W / P;
    const int np = gh * gw, N = np + 1;
    const float* pos = nullptr;
    for (int i = 0; i < p->n_pos; ++i)
        if (p->pos_cache[i].gh == gh && p->pos_cache[i].gw == gw) pos = p->pos_cache[i].data;
    MAAVSS_CHECK_ARG(pos, "vit_attn_maps: call maavss_vit_prepare(plan, %d, %d) first", H, W);
    const Layout L = layout_for(p, T, gh, gw, clips);
    uint8_t* ws = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    MAAVSS_CHECK_ARG((long long)(ws - (uint8_t*)workspace) + (long long)L.total <= workspace_bytes, "vit_attn_maps: workspace too small");
    void* patches = ws + L.patches;
    float* tok = (float*)(ws + L.tok);
    float* x = (float*)(ws + L.x);
    void* y = ws + L.y;
    void* qkv = ws + L.qkv;
    void* o = ws + L.o;
    void* hdn = ws + L.hdn;
    float* kbuf = (float*)(ws + L.k);
    float* qbuf = (float*)(ws + L.q);
    float* rows = (float*)(ws + L.rows);
    float* maps = post ? (float*)(ws + L.maps) : out;
    const long long TN = (long long)T * N;
    const int KP = p->in_chans * P * P;
    int rc;
#define VIT_DO(call) do { rc = (call); if (rc) return rc; } while (0)
    VIT_DO(maavss_vit_patchify(frames, patches, T, p->in_chans, H, W, gh, gw, P, stream));
    VIT_DO(maavss_tc_gemm_bf16(patches, p->patch_w, T * np, D, KP, KP, KP, p->patch_b, nullptr, 0, tok, D, 0, MAAVSS_ACT_NONE, stream));
    VIT_DO(maavss_vit_assemble_tokens(tok, p->cls, pos, x, T, np, D, stream));
    for (int i = 0; i + 1 < p->depth; ++i) {
        const maavss_vit_plan::Block& b = p->blocks[i];
        VIT_DO(maavss_layernorm_bf16(x, b.ln1_w, b.ln1_b, y, TN, D, 1e-6f, stream));
        VIT_DO(maavss_tc_gemm_bf16(y, b.qkv_w, (int)TN, 3 * D, D, D, D, b.qkv_b, nullptr, 0, qkv, 3 * D, 1, MAAVSS_ACT_NONE, stream));
        VIT_DO(maavss_vit_attention(qkv, o, T, N, p->heads, 0.125f, stream));
        VIT_DO(maavss_tc_gemm_bf16(o, b.proj_w, (int)TN, D, D, D, D, b.proj_b, x, D, x, D, 0, MAAVSS_ACT_NONE, stream));
        VIT_DO(maavss_layernorm_bf16(x, b.ln2_w, b.ln2_b, y, TN, D, 1e-6f, stream));
        VIT_DO(maavss_tc_gemm_bf16(y, b.fc1_w, (int)TN, 4 * D, D, D, D, b.fc1_b, nullptr, 0, hdn, 4 * D, 1, MAAVSS_ACT_GELU, stream));
        VIT_DO(maavss_tc_gemm_bf16(hdn, b.fc2_w, (int)TN, D, 4 * D, 4 * D, 4 * D, b.fc2_b, x, D, x, D, 0, MAAVSS_ACT_NONE, stream));
    }
    {   // last block: K of every token, Q of the class rows only (the attention row video_attention.py:56 keeps)
        const maavss_vit_plan::Block& b = p->blocks[p->depth - 1];
        VIT_DO(maavss_layernorm_bf16(x, b.ln1_w, b.ln1_b, y, TN, D, 1e-6f, stream));
        const uint16_t* wq = (const uint16_t*)b.qkv_w;
        VIT_DO(maavss_tc_gemm_bf16(y, wq + (size_t)D * D, (int)TN, D, D, D, D, b.qkv_b + D, nullptr, 0, kbuf, D, 0, MAAVSS_ACT_NONE, stream));
        VIT_DO(maavss_tc_gemm_bf16(y, wq, T, D, D, (long long)N * D, D, b.qkv_b, nullptr, 0, qbuf, D, 0, MAAVSS_ACT_NONE, stream));
        VIT_DO(maavss_vit_cls_attn_row(qbuf, D, kbuf, D, rows, T, p->heads, N, 0.125f, stream));
    }
    VIT_DO(maavss_attn_headsum_upsample(rows, maps, T, p->heads, N, gh, gw, P, stream));
    if (post) VIT_DO(maavss_attn_diff_normalise(maps, out, clips, T / clips, (long long)gh * P * gw * P, post == 2, ws + L.mx, stream));
#undef VIT_DO
    return MAAVSS_OK;
}
