// Multi-head self-attention of the ViT blocks (dino/vision_transformer.py:82-90) on tcgen05 tensor cores:
//     O = softmax(Q K^T * scale) V        per (frame, head), head_dim 64, N tokens (1025 for 256 x 256 frames)
// qkv is the (T, N, 3*H*64) bf16 output of the qkv GEMM; Q / K / V tiles of 128 tokens x 64 dims are fetched by TMA
// (3-D tensor map, token overrun zero-filled).  One CTA = one 128-query tile of one (frame, head):
//   pass 1:  S_j = Q K_j^T into TMEM (UMMA 128x128x16), softmax warps reduce the row maxima
//   pass 2:  S_j recomputed, P_j = exp2((S_j - m) * scale*log2e) written as bf16 into a SWIZZLE_128B smem tile,
//            O += P_j V_j (V consumed MN-major straight from the qkv layout), row sums accumulated in registers
//   finally: O / l -> bf16 -> global.  No O rescaling (the maxima are final before pass 2), S never touches HBM.
// Warp 0: TMA producer, warp 1: UMMA issuer + TMEM owner, warps 2-9: softmax / epilogue (two threads per query row,
// each owning 64 of the 128 kv columns of a tile).
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/maavss_b200.h"

namespace tca {

// MUFU.EX2 directly: the arguments are <= 0 (score minus row maximum), exp2f()'s range scaling is not needed
__device__ __forceinline__ float ex2_approx(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}


constexpr int BQ = 128, BKV = 128, HD = 64, KSTAGES = 3, VSTAGES = 2, THREADS = 320;   // 2 + 8 softmax warps
constexpr int TILE_BYTES = 128 * HD * 2;                 // 16 KB: 128 tokens x 64 dims bf16
constexpr int P_BYTES = 2 * TILE_BYTES;                  // 128 x 128 bf16 as two K-major halves of 64 kv
constexpr int OFF_Q = 0, OFF_K = TILE_BYTES, OFF_V = OFF_K + KSTAGES * TILE_BYTES, OFF_P = OFF_V + VSTAGES * TILE_BYTES;
constexpr int OFF_BAR = OFF_P + 2 * P_BYTES;       // P is double-buffered: softmax(j+1) overlaps the P V product of tile j
constexpr int OFF_XCH = OFF_BAR + 256;                   // 2 x 128 floats: row max / row sum exchange between column halves
constexpr int SMEM_TOTAL = OFF_XCH + 1024 + 1024;
constexpr int TMEM_COLS = 512;                           // S0 [0,128) S1 [128,256) O [256,320)

__global__ void __launch_bounds__(THREADS, 1) attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, __nv_bfloat16* __restrict__ out,
                                                              int N, int heads, float scale_log2e) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);    // keeps the shared address space
    uint64_t* bars = (uint64_t*)(smem + OFF_BAR);
    uint64_t* q_full = bars;            // 1
    uint64_t* k_full = bars + 1;        // KSTAGES
    uint64_t* k_empty = k_full + KSTAGES;
    uint64_t* v_full = k_empty + KSTAGES;   // VSTAGES
    uint64_t* v_empty = v_full + VSTAGES;
    uint64_t* s_full = v_empty + VSTAGES;   // 2
    uint64_t* s_empty = s_full + 2;         // 2
    uint64_t* p_full = s_empty + 2;         // 2
    uint64_t* p_empty = p_full + 2;         // 2
    uint64_t* o_full = p_empty + 2;         // 1
    uint32_t* tmem_slot = (uint32_t*)(o_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * BQ, h = blockIdx.y, f = blockIdx.z;
    const int D = heads * HD;
    const int nkv = (N + BKV - 1) / BKV;

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tmQKV);
        tc::mbar_init(q_full, 1);
        for (int s = 0; s < KSTAGES; ++s) { tc::mbar_init(&k_full[s], 1); tc::mbar_init(&k_empty[s], 1); }
        for (int s = 0; s < VSTAGES; ++s) { tc::mbar_init(&v_full[s], 1); tc::mbar_init(&v_empty[s], 1); }
        for (int b = 0; b < 2; ++b) { tc::mbar_init(&s_full[b], 1); tc::mbar_init(&s_empty[b], 8); }
        for (int b = 0; b < 2; ++b) { tc::mbar_init(&p_full[b], 8); tc::mbar_init(&p_empty[b], 1); }
        tc::mbar_init(o_full, 1);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            tc::mbar_expect_tx(q_full, TILE_BYTES);
            tc::tma_load_3d(smem + OFF_Q, &tmQKV, q_full, h * HD, q0, f);
            int ks = 0, vs = 0;
            uint32_t kph = 0, vph = 0;
            for (int g = 0; g < 2 * nkv; ++g) {
                const int j = g % nkv;
                tc::mbar_wait(&k_empty[ks], kph ^ 1);
                tc::mbar_expect_tx(&k_full[ks], TILE_BYTES);
                tc::tma_load_3d(smem + OFF_K + ks * TILE_BYTES, &tmQKV, &k_full[ks], D + h * HD, j * BKV, f);
                if (++ks == KSTAGES) { ks = 0; kph ^= 1; }
                if (g >= nkv) {
                    tc::mbar_wait(&v_empty[vs], vph ^ 1);
                    tc::mbar_expect_tx(&v_full[vs], TILE_BYTES);
                    tc::tma_load_3d(smem + OFF_V + vs * TILE_BYTES, &tmQKV, &v_full[vs], 2 * D + h * HD, j * BKV, f);
                    if (++vs == VSTAGES) { vs = 0; vph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = tc::make_idesc(1, BQ, BKV);                     // Q K^T: both K-major
            constexpr uint32_t idesc_o = tc::make_idesc(1, BQ, HD) | (1u << 16);          // P V: B (= V) MN-major
            const uint64_t qdesc = tc::smem_desc_k_sw128(tc::smem_u32(smem + OFF_Q));
            int ks = 0, vs = 0;
            uint32_t kph = 0, vph = 0;
            tc::mbar_wait(q_full, 0);
            auto issue_s = [&](int g) {
                const int b = g & 1;
                tc::mbar_wait(&k_full[ks], kph);
                tc::mbar_wait(&s_empty[b], ((g >> 1) & 1) ^ 1);
                tc::tc_fence_after();
                const uint64_t kdesc = tc::smem_desc_k_sw128(tc::smem_u32(smem + OFF_K + ks * TILE_BYTES));
#pragma unroll
                for (int k = 0; k < HD / 16; ++k) tc::umma_f16(tmem + b * BKV, qdesc + 2 * k, kdesc + 2 * k, idesc_s, k != 0);
                tc::umma_commit(&k_empty[ks]);
                tc::umma_commit(&s_full[b]);
                if (++ks == KSTAGES) { ks = 0; kph ^= 1; }
            };
            for (int g = 0; g < nkv; ++g) issue_s(g);                      // pass 1
            issue_s(nkv);                                                  // pass 2, software-pipelined by one tile
            for (int j = 0; j < nkv; ++j) {
                if (j + 1 < nkv) issue_s(nkv + j + 1);
                tc::mbar_wait(&p_full[j & 1], (j >> 1) & 1);
                tc::mbar_wait(&v_full[vs], vph);
                tc::tc_fence_after();
                const uint32_t p_addr = tc::smem_u32(smem + OFF_P + (j & 1) * P_BYTES);
                const uint32_t v_addr = tc::smem_u32(smem + OFF_V + vs * TILE_BYTES);
#pragma unroll
                for (int k = 0; k < BKV / 16; ++k) {
                    // A = P: K-major, halves of 64 kv; +32 B per 16 kv inside a half
                    const uint64_t pdesc = tc::smem_desc_k_sw128(p_addr + (k >> 2) * TILE_BYTES) + 2 * (k & 3);
                    // B = V tile [kv][64 d]: MN-major (d contiguous, one 128-byte swizzle row per kv), 16 kv = 2048 bytes
                    const uint64_t vdesc = tc::smem_desc_k_sw128(v_addr + k * 2048);
                    tc::umma_f16(tmem + 2 * BKV, pdesc, vdesc, idesc_o, (j | k) != 0);
                }
                tc::umma_commit(&p_empty[j & 1]);
                tc::umma_commit(&v_empty[vs]);
                if (++vs == VSTAGES) { vs = 0; vph ^= 1; }
            }
            tc::umma_commit(o_full);
        }
    } else {
        // 8 softmax warps: TMEM lane quadrant qd = warp % 4 (one query row per lane), column half hf = which 64 of the
        // 128 kv columns of every S tile (and which 32 of the 64 output dims) this warp owns.
        const int qd = warp & 3, hf = (warp - 2) >> 2;
        const int row = qd * 32 + lane;                           // query row inside the tile
        const uint32_t lane_off = (uint32_t)(qd * 32) << 16;
        float* xch = (float*)(smem + OFF_XCH);
        float m = -INFINITY;
        // ---- pass 1: row maxima of the raw scores
        for (int g = 0; g < nkv; ++g) {
            const int b = g & 1;
            tc::mbar_wait(&s_full[b], (g >> 1) & 1);
            tc::tc_fence_after();
            const int valid = N - g * BKV;                        // columns >= valid are beyond the last token
#pragma unroll
            for (int c2 = 0; c2 < 2; ++c2) {
                const int c = hf * 2 + c2;
                uint32_t r[32];
                tc::tmem_ld_32x32(tmem + lane_off + b * BKV + c * 32, r);
                tc::tmem_ld_wait();
                if (valid >= BKV) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) m = fmaxf(m, __uint_as_float(r[i]));
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c * 32 + i < valid) m = fmaxf(m, __uint_as_float(r[i]));
                }
            }
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&s_empty[b]);
        }
        xch[hf * 128 + row] = m;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        m = fmaxf(m, xch[(hf ^ 1) * 128 + row]);
        const float mb = m * scale_log2e;
        float l = 0.f;
        // ---- pass 2: P = exp2(S * scale*log2e - m * scale*log2e) -> bf16 -> swizzled smem tile
        for (int j = 0; j < nkv; ++j) {
            const int g = nkv + j, b = g & 1;
            tc::mbar_wait(&s_full[b], (g >> 1) & 1);
            tc::tc_fence_after();
            const int valid = N - j * BKV;
            // this warp's 64 kv columns are the K-major half `hf` of the P tile: one 128-byte row per query
            uint8_t* prow = smem + OFF_P + (j & 1) * P_BYTES + hf * TILE_BYTES + row * 128;
            // both 32-column chunks are requested up front: the second tcgen05.ld is in flight while the first is turned
            // into probabilities.  Only the last kv tile has columns beyond the sequence: the mask lives in its own
            // (warp-uniform) branch instead of costing a compare + select per element everywhere.
            uint32_t rr[2][32];
            tc::tmem_ld_32x32(tmem + lane_off + b * BKV + (hf * 2) * 32, rr[0]);
            tc::tmem_ld_32x32(tmem + lane_off + b * BKV + (hf * 2 + 1) * 32, rr[1]);
            tc::tmem_ld_wait();
            float l0 = 0.f, l1 = 0.f;
            uint32_t pk[2][16];
#pragma unroll
            for (int c2 = 0; c2 < 2; ++c2) {
                const int c = hf * 2 + c2;
                const uint32_t (&r)[32] = rr[c2];
                uint32_t (&packed)[16] = pk[c2];
                if (valid >= BKV) {
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        const float p0 = ex2_approx(fmaf(__uint_as_float(r[i]), scale_log2e, -mb));
                        const float p1 = ex2_approx(fmaf(__uint_as_float(r[i + 1]), scale_log2e, -mb));
                        l0 += p0;
                        l1 += p1;
                        const __nv_bfloat162 pb = __floats2bfloat162_rn(p0, p1);
                        packed[i >> 1] = *reinterpret_cast<const uint32_t*>(&pb);
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        float p0 = ex2_approx(fmaf(__uint_as_float(r[i]), scale_log2e, -mb));
                        float p1 = ex2_approx(fmaf(__uint_as_float(r[i + 1]), scale_log2e, -mb));
                        if (c * 32 + i >= valid) p0 = 0.f;
                        if (c * 32 + i + 1 >= valid) p1 = 0.f;
                        l0 += p0;
                        l1 += p1;
                        const __nv_bfloat162 pb = __floats2bfloat162_rn(p0, p1);
                        packed[i >> 1] = *reinterpret_cast<const uint32_t*>(&pb);
                    }
                }
            }
            tc::mbar_wait(&p_empty[j & 1], ((j >> 1) & 1) ^ 1);      // the P V product that last read this P buffer (tile j - 2)
#pragma unroll
            for (int c2 = 0; c2 < 2; ++c2) {
                const uint32_t (&packed)[16] = pk[c2];
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    const int chunk = c2 * 4 + ch;
                    *reinterpret_cast<uint4*>(prow + ((chunk ^ (row & 7)) << 4)) =
                        make_uint4(packed[ch * 4], packed[ch * 4 + 1], packed[ch * 4 + 2], packed[ch * 4 + 3]);
                }
            }
            l += l0 + l1;
            tc::tc_fence_before();
            tc::fence_proxy_async();          // generic-proxy smem writes -> visible to the tensor core (async proxy)
            __syncwarp();
            if (lane == 0) {
                tc::mbar_arrive(&s_empty[b]);
                tc::mbar_arrive(&p_full[j & 1]);
            }
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");        // both halves are done reading the maxima
        xch[hf * 128 + row] = l;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        l += xch[(hf ^ 1) * 128 + row];
        // ---- epilogue: O / l, this warp's 32 of the 64 output dims
        tc::mbar_wait(o_full, 0);
        tc::tc_fence_after();
        const float inv = 1.f / l;
        const int qrow = q0 + row;
        {
            uint32_t r[32];
            tc::tmem_ld_32x32(tmem + lane_off + 2 * BKV + hf * 32, r);
            tc::tmem_ld_wait();
            if (qrow < N) {
                __nv_bfloat16* dst = out + ((size_t)f * N + qrow) * D + h * HD + hf * 32;
#pragma unroll
                for (int i = 0; i < 32; i += 8) {
                    uint32_t w[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const __nv_bfloat162 pk = __floats2bfloat162_rn(__uint_as_float(r[i + 2 * u]) * inv, __uint_as_float(r[i + 2 * u + 1]) * inv);
                        w[u] = *reinterpret_cast<const uint32_t*>(&pk);
                    }
                    *reinterpret_cast<uint4*>(dst + i) = make_uint4(w[0], w[1], w[2], w[3]);
                }
            }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc(tmem, TMEM_COLS);
}

}  // namespace tca

typedef CUresult (*PFN_encodeTiled3)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

extern "C" int maavss_vit_attention(const void* qkv_bf16, void* out_bf16, int T, int N, int heads, float scale, void* stream) {
    if (T == 0 || N == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(qkv_bf16 && out_bf16 && heads > 0 && heads <= 65535 && T <= 65535, "vit_attention: bad argument");
    static PFN_encodeTiled3 enc = nullptr;
    if (!enc) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        MAAVSS_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
        MAAVSS_CHECK_ARG(qres == cudaDriverEntryPointSuccess && p, "cuTensorMapEncodeTiled is unavailable");
        enc = (PFN_encodeTiled3)p;
    }
    const int D3 = 3 * heads * tca::HD;
    CUtensorMap tm;
    cuuint64_t dims[3] = {(cuuint64_t)D3, (cuuint64_t)N, (cuuint64_t)T};
    cuuint64_t strides[2] = {(cuuint64_t)D3 * 2, (cuuint64_t)N * D3 * 2};
    cuuint32_t box[3] = {(cuuint32_t)tca::HD, 128, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    MAAVSS_CHECK_ARG(((uintptr_t)qkv_bf16 & 15) == 0, "vit_attention: qkv must be 16-byte aligned");
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(qkv_bf16), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MAAVSS_CHECK_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    static bool attr_set = false;
    if (!attr_set) {
        MAAVSS_CUDA(cudaFuncSetAttribute(tca::attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tca::SMEM_TOTAL));
        attr_set = true;
    }
    dim3 grid(ceil_div(N, tca::BQ), heads, T);
    tca::attn_fwd_kernel<<<grid, tca::THREADS, tca::SMEM_TOTAL, (cudaStream_t)stream>>>(tm, (__nv_bfloat16*)out_bf16, N, heads,
                                                                                       scale * 1.4426950408889634f);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
