// Weight gradient of the stride-1 "same" Conv3d / Conv2d layers on tcgen05 (companion of tc_conv.cu;
// avse_model_final.py:33-59 backward):
//     dW[co, ci, tap] (+)= sum_pos dy[pos, co] * x[pos @ tap, ci]
// Per tap this is a GEMM whose reduction runs over positions, so both operands are consumed MN-major straight from
// the channels-last bf16 copies: a stage is 128 positions; A = the dy tile [128 pos][64 co] (one TMA box, loaded once
// per stage and reused by every tap of the CTA's tap group), B = the x tile [128 pos][Cin] at the tap's shifted
// coordinate (zero padding = TMA out-of-bounds fill).  Each tap owns Cin fp32 accumulator columns in TMEM
// (64 lanes x Cin), so a CTA carries 512 / Cin taps; taps beyond that go to other CTAs (blockIdx.y) and the position
// range is split over blockIdx.x, partial sums meeting in dW through atomicAdd.
//   warp 0: TMA producer   warp 1: TMEM owner + MMA issuer (M = 64, N = Cin, K = 16 positions per tcgen05.mma)
//   warps 2-5: epilogue (tcgen05.ld -> atomicAdd)
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/maavss_b200.h"

namespace tcw {

constexpr int THREADS = 192, MAX_B_STAGES = 32, A_STAGES = 2, A_BYTES = 128 * 128;
constexpr int SMEM_BUDGET = 150 * 1024;      // A ring (2 x 16 KB) + B ring (deeper rings measured no faster)

struct Params {
    int N, D, H, W;
    int Cin, Cout;             // Cout <= 64 (dy copy is zero-padded to 64 channels)
    int KD, KH, KW, PD, PH, PW, KV;
    int BH, TG;                // tile rows; taps per CTA
    int tiles, tiles_per_cta;
    int swz_b;                 // bytes of one x row = 2 * Cin
    int b_stages;              // depth of the B ring (tiles of 128 * swz_b bytes)
    int lane_layout;           // TMEM lane of accumulator row r for M = 64: 0 -> r, 1 -> 32 * (r / 16) + r % 16
    float* dw;                 // (Cout, Cin, KV) fp32, accumulated into
};

__device__ __forceinline__ void tma_load_5d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
        ::"r"(tc::smem_u32(smem_dst)), "l"((uint64_t)m), "r"(tc::smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}

// MN-major operand: rows (= reduction index, one position each) of `swz` bytes written by TMA with the matching swizzle;
// the MN extent is exactly one swizzle atom wide, 8-row groups follow each other every 8 * swz bytes.
__device__ __forceinline__ uint64_t smem_desc_mn(uint32_t smem_addr, uint32_t swz) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((8 * swz) >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(swz == 128 ? 2 : (swz == 64 ? 4 : 6)) << 61;
    return d;
}

__global__ void __launch_bounds__(THREADS, 1) tc_conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmDy,
                                                                   const __grid_constant__ CUtensorMap tmX, Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* a_buf = smem;                                   // A_STAGES x 16 KB
    uint8_t* b_buf = smem + A_STAGES * A_BYTES;              // b_stages x (128 * swz_b) bytes
    const int B_STAGES = p.b_stages;
    const int B_MAX = 128 * p.swz_b;
    uint64_t* a_full = (uint64_t*)(b_buf + B_STAGES * B_MAX);
    uint64_t* a_empty = a_full + A_STAGES;
    uint64_t* b_full = a_empty + A_STAGES;
    uint64_t* b_empty = b_full + MAX_B_STAGES;
    uint64_t* tfull = b_empty + MAX_B_STAGES;
    uint32_t* tmem_slot = (uint32_t*)(tfull + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tap0 = blockIdx.y * p.TG;
    const int ntap = min(p.TG, p.KV - tap0);
    const int t_begin = blockIdx.x * p.tiles_per_cta, t_end = min(p.tiles, t_begin + p.tiles_per_cta);
    const int h_tiles = p.H / p.BH;
    const uint32_t b_bytes = 128u * p.swz_b;

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tmDy);
        tc::prefetch_tmap(&tmX);
        for (int s = 0; s < A_STAGES; ++s) { tc::mbar_init(&a_full[s], 1); tc::mbar_init(&a_empty[s], 1); }
        for (int s = 0; s < B_STAGES; ++s) { tc::mbar_init(&b_full[s], 1); tc::mbar_init(&b_empty[s], 1); }
        tc::mbar_init(tfull, 1);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            int as = 0, bs = 0;
            uint32_t aph = 0, bph = 0;
            for (int t = t_begin; t < t_end; ++t) {
                const int ht = t % h_tiles, nd = t / h_tiles;
                const int dd = nd % p.D, n = nd / p.D, h0 = ht * p.BH;
                tc::mbar_wait(&a_empty[as], aph ^ 1);
                tc::mbar_expect_tx(&a_full[as], A_BYTES);
                tma_load_5d(a_buf + as * A_BYTES, &tmDy, &a_full[as], 0, 0, h0, dd, n);
                if (++as == A_STAGES) { as = 0; aph ^= 1; }
                for (int j = 0; j < ntap; ++j) {
                    const int tap = tap0 + j;
                    const int kd = tap / (p.KH * p.KW), r = tap - kd * p.KH * p.KW;
                    const int kh = r / p.KW, kw = r - kh * p.KW;
                    tc::mbar_wait(&b_empty[bs], bph ^ 1);
                    tc::mbar_expect_tx(&b_full[bs], b_bytes);
                    tma_load_5d(b_buf + bs * B_MAX, &tmX, &b_full[bs], 0, kw - p.PW, h0 + kh - p.PH, dd + kd - p.PD, n);
                    if (++bs == B_STAGES) { bs = 0; bph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // kind::f16, bf16 x bf16 -> fp32, A and B both MN-major (bits 15, 16), M = 64, N = Cin
            const uint32_t idesc = tc::make_idesc(1, 64, p.Cin) | (1u << 15) | (1u << 16);
            int as = 0, bs = 0;
            uint32_t aph = 0, bph = 0;
            for (int t = t_begin; t < t_end; ++t) {
                tc::mbar_wait(&a_full[as], aph);
                tc::tc_fence_after();
                const uint32_t a_addr = tc::smem_u32(a_buf + as * A_BYTES);
                for (int j = 0; j < ntap; ++j) {
                    tc::mbar_wait(&b_full[bs], bph);
                    tc::tc_fence_after();
                    const uint32_t b_addr = tc::smem_u32(b_buf + bs * B_MAX);
#pragma unroll
                    for (int k = 0; k < 8; ++k) {            // 16 positions per MMA: 16 rows of 128 B (A) / swz_b bytes (B)
                        const uint64_t adesc = smem_desc_mn(a_addr + k * 2048, 128);
                        const uint64_t bdesc = smem_desc_mn(b_addr + k * 16 * p.swz_b, p.swz_b);
                        tc::umma_f16(tmem_base + j * p.Cin, adesc, bdesc, idesc, (t != t_begin) || (k != 0));
                    }
                    tc::umma_commit(&b_empty[bs]);
                    if (++bs == B_STAGES) { bs = 0; bph ^= 1; }
                }
                tc::umma_commit(&a_empty[as]);
                if (++as == A_STAGES) { as = 0; aph ^= 1; }
            }
            tc::umma_commit(tfull);
        }
    } else if (t_end > t_begin) {
        const int q = warp & 3;
        tc::mbar_wait(tfull, 0);
        tc::tc_fence_after();
        // accumulator row (= output channel) held by this thread's TMEM lane, or -1
        const int tl = q * 32 + lane;
        int co;
        if (p.lane_layout == 0) co = tl < 64 ? tl : -1;
        else co = (lane < 16) ? q * 16 + lane : -1;
        const int ncols = ntap * p.Cin;
        for (int c = 0; c < ncols; c += 32) {
            uint32_t r[32];
            tc::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + c, r);
            tc::tmem_ld_wait();
            if (co >= 0 && co < p.Cout) {
#pragma unroll
                for (int jj = 0; jj < 32; ++jj) {
                    const int col = c + jj;
                    if (col < ncols) {
                        const int j = col / p.Cin, ci = col - j * p.Cin;
                        atomicAdd(p.dw + ((size_t)co * p.Cin + ci) * p.KV + tap0 + j, __uint_as_float(r[jj]));
                    }
                }
            }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc(tmem_base, 512);
}

}  // namespace tcw

// ------------------------------------------------------------------------------------------------ host side
int maavss_tc_to_channels_last(const float* x, void* out, int C, int CP, long long S, long long total, cudaStream_t st);

typedef CUresult (*PFN_encodeTiledW)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

namespace {

PFN_encodeTiledW get_encode_w() {
    static PFN_encodeTiledW fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiledW)p;
    }
    return fn;
}

bool wg_supported(const maavss_conv_desc* d) {
    if (d->SD != 1 || d->SH != 1 || d->SW != 1) return false;
    if (d->OD != d->D || d->OH != d->H || d->OW != d->W) return false;
    if (!(d->Cin == 16 || d->Cin == 32 || d->Cin == 64)) return false;
    if (!(d->Cout == 16 || d->Cout == 32 || d->Cout == 64)) return false;
    if (!(d->W == 32 || d->W == 64 || d->W == 128)) return false;
    if (d->H % (128 / d->W) != 0) return false;
    return true;
}

size_t align_up_w(size_t v, size_t a) { return (v + a - 1) / a * a; }

int make_act_map(CUtensorMap* m, void* base, int C, int W, int H, int D, int N, int BW, int BH) {
    PFN_encodeTiledW enc = get_encode_w();
    MAAVSS_CHECK_ARG(enc, "cuTensorMapEncodeTiled is unavailable (driver too old?)");
    const int swz = C * 2;
    const CUtensorMapSwizzle swz_e = swz == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (swz == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
    cuuint64_t dims[5] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)D, (cuuint64_t)N};
    cuuint64_t strides[4] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2, (cuuint64_t)D * H * W * C * 2};
    cuuint32_t box[5] = {(cuuint32_t)C, (cuuint32_t)BW, (cuuint32_t)BH, 1, 1};
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz_e,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MAAVSS_CHECK_ARG(r == CUDA_SUCCESS, "conv_wgrad_tc: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return MAAVSS_OK;
}

}  // namespace

extern "C" int maavss_conv_wgrad_tc_supported(const maavss_conv_desc* d) { return d && wg_supported(d) ? 1 : 0; }

extern "C" long long maavss_conv_wgrad_tc_workspace_bytes(const maavss_conv_desc* d) {
    if (!d || !wg_supported(d)) return 0;
    const size_t pos = (size_t)d->N * d->D * d->H * d->W;
    return (long long)(align_up_w(pos * d->Cin * 2, 1024) + align_up_w(pos * 64 * 2, 1024));
}

extern "C" int maavss_conv_wgrad_tc(const maavss_conv_desc* d, const float* x, const float* dy, float* dw, int accumulate,
                                    void* workspace, long long workspace_bytes, void* stream) {
    MAAVSS_CHECK_ARG(d && x && dy && dw && workspace, "conv_wgrad_tc: null argument");
    MAAVSS_CHECK_ARG(wg_supported(d), "conv_wgrad_tc: layer outside the tensor-core envelope");
    MAAVSS_CHECK_ARG(workspace_bytes >= maavss_conv_wgrad_tc_workspace_bytes(d) && ((uintptr_t)workspace & 1023) == 0,
                     "conv_wgrad_tc: workspace too small or not 1024-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const int KV = d->KD * d->KH * d->KW;
    if (!accumulate) MAAVSS_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)d->Cout * d->Cin * KV, st));
    if (d->N == 0) return MAAVSS_OK;
    const long long S = (long long)d->D * d->H * d->W, total = S * d->N;
    __nv_bfloat16* x_cl = (__nv_bfloat16*)workspace;
    __nv_bfloat16* dy_cl = (__nv_bfloat16*)((uint8_t*)workspace + align_up_w((size_t)total * d->Cin * 2, 1024));
    int rc = maavss_tc_to_channels_last(x, x_cl, d->Cin, d->Cin, S, total, st);
    if (rc) return rc;
    rc = maavss_tc_to_channels_last(dy, dy_cl, d->Cout, 64, S, total, st);
    if (rc) return rc;
    const int BW = d->W, BH = 128 / d->W;
    CUtensorMap tmDy, tmX;
    rc = make_act_map(&tmDy, dy_cl, 64, d->W, d->H, d->D, d->N, BW, BH);
    if (rc) return rc;
    rc = make_act_map(&tmX, x_cl, d->Cin, d->W, d->H, d->D, d->N, BW, BH);
    if (rc) return rc;
    tcw::Params p;
    p.N = d->N; p.D = d->D; p.H = d->H; p.W = d->W;
    p.Cin = d->Cin; p.Cout = d->Cout;
    p.KD = d->KD; p.KH = d->KH; p.KW = d->KW; p.PD = d->PD; p.PH = d->PH; p.PW = d->PW; p.KV = KV;
    p.BH = BH;
    p.TG = 512 / d->Cin < KV ? 512 / d->Cin : KV;
    p.tiles = d->N * d->D * (d->H / BH);
    const int groups = ceil_div(KV, p.TG);
    int chunks = ceil_div(2 * maavss_num_sms(), groups);
    if (chunks > p.tiles) chunks = p.tiles;
    if (chunks < 1) chunks = 1;
    p.tiles_per_cta = ceil_div(p.tiles, chunks);
    chunks = ceil_div(p.tiles, p.tiles_per_cta);
    p.swz_b = d->Cin * 2;
    p.lane_layout = 1;      // verified on B200: an M = 64 accumulator keeps row r in lane 32 * (r / 16) + r % 16
    p.dw = dw;
    p.b_stages = (tcw::SMEM_BUDGET - tcw::A_STAGES * tcw::A_BYTES) / (128 * p.swz_b);
    if (p.b_stages > tcw::MAX_B_STAGES) p.b_stages = tcw::MAX_B_STAGES;
    const int smem = tcw::A_STAGES * tcw::A_BYTES + p.b_stages * 128 * p.swz_b + 1024 + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        MAAVSS_CUDA(cudaFuncSetAttribute(tcw::tc_conv_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tcw::SMEM_BUDGET + 2048));
        attr_set = true;
    }
    tcw::tc_conv_wgrad_kernel<<<dim3(chunks, groups), tcw::THREADS, smem, st>>>(tmDy, tmX, p);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
