// Persistent warp-specialised bf16 GEMM on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), used by the
// ViT-S/8 attention-map extractor (dino/vision_transformer.py:59-65 Mlp, :80-92 Attention qkv/proj, :116-131 PatchEmbed):
//     D[M, N] = A[M, K] . B[N, K]^T (+ bias[N]) (-> GELU) (+ residual[M, N])     A, B bf16 K-major, fp32 accumulate in TMEM
// One CTA per SM loops over 128 x BN output tiles.  Warp 0: TMA producer (SWIZZLE_128B boxes of 64 bf16 along K into a
// 4/5-stage mbarrier ring).  Warp 1: allocates TMEM and issues tcgen05.mma (UMMA 128 x BN x 16) into one of two
// accumulator stages.  Warps 2-5: epilogue, each owning its TMEM lane quadrant: tcgen05.ld -> smem transpose ->
// fused bias / GELU / residual -> coalesced stores, overlapped with the next tile's main loop.
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/maavss_b200.h"

namespace tcg {

constexpr int BM = 128, BK = 64, ACC_STAGES = 2, EPI_WARPS = 8, THREADS = 64 + 32 * EPI_WARPS;

template <int BN>
struct Smem {
    static constexpr int STAGES = BN > 128 ? 4 : 5;
    static constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int EPI_BYTES = EPI_WARPS * 32 * 32 * 4;
    static constexpr int BAR_OFF = STAGES * STAGE_BYTES + EPI_BYTES;
    static constexpr int TOTAL = BAR_OFF + 256 + 1024;    // barriers + tmem slot, + slack for the 1024-byte alignment
};

struct Epilogue {
    const float* bias;       // [N] or null
    const float* residual;   // [M, ldr] fp32 or null
    long long ldr;
    void* out;               // fp32 or bf16 [M, ldc]
    long long ldc;
    int out_bf16;
    int act;                 // MAAVSS_ACT_NONE or MAAVSS_ACT_GELU
};

// erf by Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7, the size of erff's own rounding): one MUFU.RCP, one MUFU.EX2 and
// 8 FMA-class instructions instead of erff's ~45 -- the fc1 epilogue was bound by erff on its few epilogue warps
// (ncu: 322 us for fc1 + GELU against 103 us for fc2 with the same FLOPs).
__device__ __forceinline__ float rcp_approx(float x) {   // MUFU.RCP (1 ulp); __frcp_rn is the IEEE routine with a CALL slow path
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float erf_as(float x) {
    const float ax = fabsf(x);
    const float t = rcp_approx(fmaf(0.3275911f, ax, 1.f));
    float p = fmaf(t, 1.061405429f, -1.453152027f);
    p = fmaf(t, p, 1.421413741f);
    p = fmaf(t, p, -0.284496736f);
    p = fmaf(t, p, 0.254829592f);
    const float r = fmaf(-p * t, __expf(-ax * ax), 1.f);
    return copysignf(r, x);
}
__device__ __forceinline__ float gelu_erf(float v) { return 0.5f * v * (1.f + erf_as(v * 0.70710678118654752f)); }

template <int BN>
__global__ void __launch_bounds__(THREADS, 1) tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA,
                                                             const __grid_constant__ CUtensorMap tmB, int M, int N, int K,
                                                             Epilogue ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);    // keeps the shared address space
    using S = Smem<BN>;
    constexpr int STAGES = S::STAGES;
    constexpr int TMEM_COLS = ACC_STAGES * BN <= 128 ? 128 : (ACC_STAGES * BN <= 256 ? 256 : 512);
    float* epi = (float*)(smem + STAGES * S::STAGE_BYTES);
    uint64_t* full = (uint64_t*)(smem + S::BAR_OFF);
    uint64_t* empty = full + STAGES;
    uint64_t* tfull = empty + STAGES;
    uint64_t* tempty = tfull + ACC_STAGES;
    uint32_t* tmem_slot = (uint32_t*)(tempty + ACC_STAGES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles = (N + BN - 1) / BN, m_tiles = (M + BM - 1) / BM;
    const int tiles = n_tiles * m_tiles;
    const int KB = (K + BK - 1) / BK;

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tmA);
        tc::prefetch_tmap(&tmB);
        for (int s = 0; s < STAGES; ++s) {
            tc::mbar_init(&full[s], 1);
            tc::mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < ACC_STAGES; ++a) {
            tc::mbar_init(&tfull[a], 1);
            tc::mbar_init(&tempty[a], EPI_WARPS);
        }
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
                const int m0 = (t / n_tiles) * BM, n0 = (t % n_tiles) * BN;
                for (int kb = 0; kb < KB; ++kb) {
                    tc::mbar_wait(&empty[s], ph ^ 1);
                    tc::mbar_expect_tx(&full[s], S::STAGE_BYTES);
                    uint8_t* a_dst = smem + s * S::STAGE_BYTES;
                    tc::tma_load_2d(a_dst, &tmA, &full[s], kb * BK, m0);
                    tc::tma_load_2d(a_dst + S::A_BYTES, &tmB, &full[s], kb * BK, n0);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = tc::make_idesc(1, BM, BN);
            int s = 0, it = 0;
            uint32_t ph = 0;
            for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++it) {
                const int a = it & 1;
                const uint32_t aph = (it >> 1) & 1;
                tc::mbar_wait(&tempty[a], aph ^ 1);
                tc::tc_fence_after();
                const uint32_t d_tmem = tmem_base + a * BN;
                for (int kb = 0; kb < KB; ++kb) {
                    tc::mbar_wait(&full[s], ph);
                    tc::tc_fence_after();
                    const uint32_t a_addr = tc::smem_u32(smem + s * S::STAGE_BYTES);
                    const uint64_t adesc = tc::smem_desc_k_sw128(a_addr);
                    const uint64_t bdesc = tc::smem_desc_k_sw128(a_addr + S::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)      // +32 bytes along K inside the swizzle atom = +2 in the address field
                        tc::umma_f16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                    tc::umma_commit(&empty[s]);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
                tc::umma_commit(&tfull[a]);
            }
        }
    } else {
        // 8 epilogue warps: warp w may access TMEM lanes 32 * (w % 4) .. + 31, so two warps share each lane quadrant and
        // take alternate 32-column chunks; the tcgen05.ld of a warp's next chunk is in flight while it transposes,
        // finishes and stores the current one.
        const int ew = warp - 2, q = warp & 3, half = ew >> 2;
        float* stage = epi + ew * (32 * 32);         // 32 rows x 32 fp32, 16-byte chunks XOR-swizzled by (row & 7)
        const int sub = lane >> 3, cc = (lane & 7) * 4;     // phase 2: 4 rows per pass, 8 lanes x float4 per row
        constexpr int CH = (BN / 32 + 1) / 2;        // chunks per warp (the second half gets one fewer when BN / 32 is odd)
        int it = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++it) {
            const int m0 = (t / n_tiles) * BM, n0 = (t % n_tiles) * BN;
            const int a = it & 1;
            const uint32_t aph = (it >> 1) & 1;
            tc::mbar_wait(&tfull[a], aph);
            tc::tc_fence_after();
            const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + a * BN;
            uint32_t rbuf[2][32];
            if (half * 32 < BN) tc::tmem_ld_32x32(t_addr + half * 32, rbuf[0]);
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                const int c = half + 2 * i;
                if (c * 32 >= BN) break;
                uint32_t (&r)[32] = rbuf[i & 1];
                tc::tmem_ld_wait();
                if ((c + 2) * 32 < BN) tc::tmem_ld_32x32(t_addr + (c + 2) * 32, rbuf[(i + 1) & 1]);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<uint4*>(&stage[lane * 32 + ((j ^ (lane & 7)) << 2)]) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                __syncwarp();
                const int col = n0 + c * 32 + cc;
                if (col < N) {                       // N % 4 == 0: the float4 is entirely inside or outside
                    float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (ep.bias) b4 = __ldg(reinterpret_cast<const float4*>(ep.bias + col));
                    float4 res[8];
                    if (ep.residual) {               // all residual loads in flight before the first store (out may alias residual)
#pragma unroll
                        for (int p = 0; p < 8; ++p) {
                            const int row = m0 + q * 32 + p * 4 + sub;
                            res[p] = row < M ? *reinterpret_cast<const float4*>(ep.residual + (size_t)row * ep.ldr + col)
                                             : make_float4(0.f, 0.f, 0.f, 0.f);
                        }
                    }
#pragma unroll
                    for (int p = 0; p < 8; ++p) {
                        const int rr = p * 4 + sub;
                        const int row = m0 + q * 32 + rr;
                        float4 v = *reinterpret_cast<const float4*>(&stage[rr * 32 + (((lane & 7) ^ (rr & 7)) << 2)]);
                        v.x += b4.x; v.y += b4.y; v.z += b4.z; v.w += b4.w;
                        if (ep.act == MAAVSS_ACT_GELU) { v.x = gelu_erf(v.x); v.y = gelu_erf(v.y); v.z = gelu_erf(v.z); v.w = gelu_erf(v.w); }
                        if (ep.residual) { v.x += res[p].x; v.y += res[p].y; v.z += res[p].z; v.w += res[p].w; }
                        if (row < M) {
                            if (ep.out_bf16) {
                                const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
                                *reinterpret_cast<uint2*>((__nv_bfloat16*)ep.out + (size_t)row * ep.ldc + col) =
                                    make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
                            } else {
                                *reinterpret_cast<float4*>((float*)ep.out + (size_t)row * ep.ldc + col) = v;
                            }
                        }
                    }
                }
                __syncwarp();
            }
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&tempty[a]);
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace tcg

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiled)p;
    }
    return fn;
}

// 2-D row-major (rows, cols) tensor of `esize`-byte elements with row pitch `ld` elements; box = (box_rows, box_cols)
int maavss_make_tmap_2d_swz(CUtensorMap* map, const void* base, int dtype_code, int esize, long long rows, long long cols, long long ld,
                            int box_rows, int box_cols, int swizzle_code) {
    PFN_encodeTiled enc = get_encode();
    MAAVSS_CHECK_ARG(enc, "cuTensorMapEncodeTiled is unavailable (driver too old?)");
    MAAVSS_CHECK_ARG(((uintptr_t)base & 15) == 0 && (ld * esize) % 16 == 0, "TMA operand must be 16-byte aligned with a 16-byte multiple pitch");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * esize};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, (CUtensorMapDataType)dtype_code, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swizzle_code, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MAAVSS_CHECK_ARG(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return MAAVSS_OK;
}

int maavss_make_tmap_2d(CUtensorMap* map, const void* base, int dtype_code, int esize, long long rows, long long cols, long long ld,
                        int box_rows, int box_cols) {
    return maavss_make_tmap_2d_swz(map, base, dtype_code, esize, rows, cols, ld, box_rows, box_cols, CU_TENSOR_MAP_SWIZZLE_128B);
}

template <int BN>
static int launch_tc_gemm(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb, tcg::Epilogue ep,
                          cudaStream_t st) {
    CUtensorMap tmA, tmB;
    int rc = maavss_make_tmap_2d(&tmA, A, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, M, K, lda, tcg::BM, tcg::BK);
    if (rc) return rc;
    rc = maavss_make_tmap_2d(&tmB, B, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, N, K, ldb, BN, tcg::BK);
    if (rc) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        MAAVSS_CUDA(cudaFuncSetAttribute(tcg::tc_gemm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, tcg::Smem<BN>::TOTAL));
        attr_set = true;
    }
    const int tiles = ceil_div(M, tcg::BM) * ceil_div(N, BN);
    const int grid = tiles < maavss_num_sms() ? tiles : maavss_num_sms();
    tcg::tc_gemm_kernel<BN><<<grid, tcg::THREADS, tcg::Smem<BN>::TOTAL, st>>>(tmA, tmB, M, N, K, ep);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_tc_gemm_bf16(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb, const float* bias,
                                   const float* residual, long long ldr, void* out, long long ldc, int out_bf16, int act,
                                   void* stream) {
    if (M == 0 || N == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(A && B && out && K > 0, "tc_gemm_bf16: bad argument");
    MAAVSS_CHECK_ARG(act == MAAVSS_ACT_NONE || act == MAAVSS_ACT_GELU, "tc_gemm_bf16: activation %d not supported", act);
    tcg::Epilogue ep{bias, residual, ldr, out, ldc, out_bf16, act};
    MAAVSS_CHECK_ARG(N % 4 == 0 && ldc % 4 == 0 && (residual == nullptr || ldr % 4 == 0) && ((uintptr_t)out & 15) == 0 &&
                         ((uintptr_t)residual & 15) == 0 && ((uintptr_t)bias & 15) == 0,
                     "tc_gemm_bf16: N and the output / residual pitches must be multiples of 4 and 16-byte aligned");
    if (N <= 64) return launch_tc_gemm<64>(A, B, M, N, K, lda, ldb, ep, (cudaStream_t)stream);
    if (N % 192 == 0) return launch_tc_gemm<192>(A, B, M, N, K, lda, ldb, ep, (cudaStream_t)stream);
    return launch_tc_gemm<128>(A, B, M, N, K, lda, ldb, ep, (cudaStream_t)stream);
}
