// Stride-1 Conv3d / Conv2d on the 5th-generation tensor cores (tcgen05 + TMEM + TMA): the visual encoder of
// AV_Fusion_Model_Frames (avse_model_final.py:33-59, Conv3d k(3,5,5), 99.5 % of the model's FLOPs) forward and
// data gradient.
//
// Implicit GEMM without an im2col gather: the activation is first rewritten channels-last in bf16
// (N, D, H, W, C), so that for ONE kernel tap the A operand of a 128-position output tile (BH rows x BW = W columns)
// is a plain 5-D TMA box (C, BW, BH, 1, 1) at the tap's shifted coordinate -- zero padding is the TMA's
// out-of-bounds fill -- landing in shared memory as 128 K-major rows of C bf16 with the swizzle that matches the
// row width (C = 16 / 32 / 64 -> SWIZZLE_32B / 64B / 128B).  The B operand of the tap is the (Cout, C) slice of the
// weights repacked as [tap][Cout][C].  A persistent CTA per SM walks output tiles; per tile the reduction runs over
// the KD*KH*KW taps, C/16 tcgen05.mma (128 x Cout x 16, fp32 accumulators in TMEM) per tap.
//   warp 0: TMA producer (6-stage mbarrier ring)      warp 1: TMEM owner + MMA issuer (2 accumulator stages)
//   warps 2-5: epilogue, tcgen05.ld -> fp32 NCDHW stores (a tile is 128 consecutive positions of every channel)
// The data gradient is the same kernel on dy with the weights transposed and tap-flipped and padding K-1-P.
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/maavss_b200.h"

namespace tcc {

constexpr int BM = 128, MAX_STAGES = 32, ACC_STAGES = 2, THREADS = 192, TMEM_COLS = 128;
constexpr int SMEM_BUDGET = 150 * 1024;      // operand ring (6 stages of 24 KB at Cin = Cout = 64; a 200 KB / 32-stage ring measured no faster: the kernel is bound by L2->SM operand traffic, not by TMA latency)

struct Params {
    int N, D, H, W;            // spatial extents (input == output: stride 1, "same" geometry)
    int Ca, Co;                // reduction channels (operand A / B rows of C) and output channels
    int KD, KH, KW, PD, PH, PW;
    int BH, log2BW;            // tile = BH rows x BW (= W) columns = 128 positions
    int a_bytes, stage_bytes, swz, stages;
    float* out;                // (N, Co, D, H, W) fp32
};

__device__ __forceinline__ void tma_load_5d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
        ::"r"(tc::smem_u32(smem_dst)), "l"((uint64_t)m), "r"(tc::smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}

// K-major operand tile written by TMA with SWIZZLE_{32,64,128}B: rows of `swz` bytes, 8-row atoms stacked along M/N.
__device__ __forceinline__ uint64_t smem_desc_k(uint32_t smem_addr, uint32_t swz) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((8 * swz) >> 4) << 32;             // stride byte offset: 8 rows
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(swz == 128 ? 2 : (swz == 64 ? 4 : 6)) << 61;
    return d;
}

__global__ void __launch_bounds__(THREADS, 1) tc_conv_kernel(const __grid_constant__ CUtensorMap tmA,
                                                             const __grid_constant__ CUtensorMap tmB, Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    const int STAGES = p.stages;
    uint64_t* full = (uint64_t*)(smem + STAGES * p.stage_bytes);
    uint64_t* empty = full + MAX_STAGES;
    uint64_t* tfull = empty + MAX_STAGES;
    uint64_t* tempty = tfull + ACC_STAGES;
    uint32_t* tmem_slot = (uint32_t*)(tempty + ACC_STAGES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h_tiles = p.H / p.BH;
    const int tiles = p.N * p.D * h_tiles;
    const int KV = p.KD * p.KH * p.KW;

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tmA);
        tc::prefetch_tmap(&tmB);
        for (int s = 0; s < STAGES; ++s) {
            tc::mbar_init(&full[s], 1);
            tc::mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < ACC_STAGES; ++a) {
            tc::mbar_init(&tfull[a], 1);
            tc::mbar_init(&tempty[a], 4);
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
                const int ht = t % h_tiles;
                const int nd = t / h_tiles;
                const int dd = nd % p.D, n = nd / p.D;
                const int h0 = ht * p.BH;
                int tap = 0;
                for (int kd = 0; kd < p.KD; ++kd)
                    for (int kh = 0; kh < p.KH; ++kh)
                        for (int kw = 0; kw < p.KW; ++kw, ++tap) {
                            tc::mbar_wait(&empty[s], ph ^ 1);
                            tc::mbar_expect_tx(&full[s], p.a_bytes + p.Co * p.swz);
                            uint8_t* a_dst = smem + s * p.stage_bytes;
                            tma_load_5d(a_dst, &tmA, &full[s], 0, kw - p.PW, h0 + kh - p.PH, dd + kd - p.PD, n);
                            tc::tma_load_2d(a_dst + p.a_bytes, &tmB, &full[s], 0, tap * p.Co);
                            if (++s == STAGES) { s = 0; ph ^= 1; }
                        }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = tc::make_idesc(1, BM, p.Co);
            const int ksteps = p.Ca / 16;
            int s = 0, it = 0;
            uint32_t ph = 0;
            for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++it) {
                const int a = it & 1;
                const uint32_t aph = (it >> 1) & 1;
                tc::mbar_wait(&tempty[a], aph ^ 1);
                tc::tc_fence_after();
                const uint32_t d_tmem = tmem_base + a * 64;
                for (int tap = 0; tap < KV; ++tap) {
                    tc::mbar_wait(&full[s], ph);
                    tc::tc_fence_after();
                    const uint32_t a_addr = tc::smem_u32(smem + s * p.stage_bytes);
                    const uint64_t adesc = smem_desc_k(a_addr, p.swz);
                    const uint64_t bdesc = smem_desc_k(a_addr + p.a_bytes, p.swz);
                    for (int k = 0; k < ksteps; ++k)     // +32 bytes along K inside the swizzled row = +2 in the address field
                        tc::umma_f16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (tap | k) != 0);
                    tc::umma_commit(&empty[s]);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
                tc::umma_commit(&tfull[a]);
            }
        }
    } else {
        const int q = warp & 3;                      // TMEM lane quadrant this warp may access
        const size_t plane = (size_t)p.H << p.log2BW;    // H * W
        int it = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++it) {
            const int ht = t % h_tiles;
            const int nd = t / h_tiles;
            const int dd = nd % p.D, n = nd / p.D;
            const int a = it & 1;
            const uint32_t aph = (it >> 1) & 1;
            tc::mbar_wait(&tfull[a], aph);
            tc::tc_fence_after();
            // row r of the tile = position (h0 + r / BW, r % BW): 128 consecutive floats of every channel plane
            float* o = p.out + ((size_t)n * p.Co * p.D + dd) * plane + (((size_t)ht * p.BH) << p.log2BW) + q * 32 + lane;
            for (int c = 0; c < p.Co; c += 32) {
                uint32_t r[32];
                tc::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + a * 64 + c, r);
                tc::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (c + j < p.Co) o[(size_t)(c + j) * p.D * plane] = __uint_as_float(r[j]);
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

// (N, C, S) fp32 -> (N, S, C) bf16.  A thread gathers the C channels of one position (coalesced across the warp for
// every channel) and writes them as C/8 16-byte vectors.
// CP >= C pads every position with zero channels (the weight-gradient kernel wants 64-channel dy rows).
template <int C, int CP = C>
__global__ void __launch_bounds__(256) to_channels_last_bf16(const float* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                             long long S, long long total) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    const long long n = e / S, s = e - n * S;
    const float* src = x + (size_t)n * C * S + s;
    uint4* dst = reinterpret_cast<uint4*>(out + (size_t)e * CP);
#pragma unroll
    for (int c8 = C / 8; c8 < CP / 8; ++c8) dst[c8] = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
    for (int c8 = 0; c8 < C / 8; ++c8) {
        uint32_t w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const __nv_bfloat162 v = __floats2bfloat162_rn(__ldg(src + (size_t)(c8 * 8 + 2 * j) * S), __ldg(src + (size_t)(c8 * 8 + 2 * j + 1) * S));
            w[j] = *reinterpret_cast<const uint32_t*>(&v);
        }
        dst[c8] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// weights (Cout, Cin, KV) fp32 -> bf16 [tap][Co][Ca]:
//   forward : Co = Cout, Ca = Cin,  out[tap][co][ci] = w[co][ci][tap]
//   dgrad   : Co = Cin,  Ca = Cout, out[tap][ci][co] = w[co][ci][KV - 1 - tap]     (flipped taps, transposed channels)
__global__ void repack_weights_bf16(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int Cout, int Cin, int KV, int dgrad) {
    const int total = Cout * Cin * KV;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
        const int Ca = dgrad ? Cout : Cin, Co = dgrad ? Cin : Cout;
        const int ca = e % Ca, r = e / Ca;
        const int co = r % Co, tap = r / Co;
        const float v = dgrad ? w[((size_t)ca * Cin + co) * KV + (KV - 1 - tap)] : w[((size_t)co * Cin + ca) * KV + tap];
        out[e] = __float2bfloat16_rn(v);
    }
}


}  // namespace tcc

// (N, C, S) fp32 -> (N, S, CP) bf16 with zero channels C..CP (shared with tc_conv_wgrad.cu)
int maavss_tc_to_channels_last(const float* x, void* out, int C, int CP, long long S, long long total, cudaStream_t st) {
    __nv_bfloat16* o = (__nv_bfloat16*)out;
    const int blocks = ceil_div(total, 256);
    if (C == 16 && CP == 16) tcc::to_channels_last_bf16<16, 16><<<blocks, 256, 0, st>>>(x, o, S, total);
    else if (C == 32 && CP == 32) tcc::to_channels_last_bf16<32, 32><<<blocks, 256, 0, st>>>(x, o, S, total);
    else if (C == 64 && CP == 64) tcc::to_channels_last_bf16<64, 64><<<blocks, 256, 0, st>>>(x, o, S, total);
    else if (C == 16 && CP == 64) tcc::to_channels_last_bf16<16, 64><<<blocks, 256, 0, st>>>(x, o, S, total);
    else if (C == 32 && CP == 64) tcc::to_channels_last_bf16<32, 64><<<blocks, 256, 0, st>>>(x, o, S, total);
    else {
        maavss_set_error("tc_to_channels_last: unsupported channel counts %d -> %d", C, CP);
        return MAAVSS_ERR_UNSUPPORTED;
    }
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_conv() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiled)p;
    }
    return fn;
}

namespace {

struct TcGeo {
    int N, D, H, W, Ca, Co, KD, KH, KW, PD, PH, PW;
};

// mode 0 = forward (x -> y), 1 = data gradient (dy -> dx).  Returns false when the layer is outside the envelope.
bool tc_geo(const maavss_conv_desc* d, int mode, TcGeo* g) {
    if (d->SD != 1 || d->SH != 1 || d->SW != 1) return false;
    if (d->OD != d->D || d->OH != d->H || d->OW != d->W) return false;          // "same" geometry only
    g->N = d->N; g->D = d->D; g->H = d->H; g->W = d->W;
    g->KD = d->KD; g->KH = d->KH; g->KW = d->KW;
    if (mode == 0) {
        g->Ca = d->Cin; g->Co = d->Cout; g->PD = d->PD; g->PH = d->PH; g->PW = d->PW;
    } else {
        g->Ca = d->Cout; g->Co = d->Cin; g->PD = d->KD - 1 - d->PD; g->PH = d->KH - 1 - d->PH; g->PW = d->KW - 1 - d->PW;
    }
    if (g->PD < 0 || g->PH < 0 || g->PW < 0) return false;
    if (!(g->Ca == 16 || g->Ca == 32 || g->Ca == 64)) return false;
    if (!(g->Co == 16 || g->Co == 32 || g->Co == 64)) return false;
    if (!(g->W == 32 || g->W == 64 || g->W == 128)) return false;
    if (g->H % (128 / g->W) != 0) return false;
    return true;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

extern "C" int maavss_conv_tc_supported(const maavss_conv_desc* d, int mode) {
    TcGeo g;
    return d && (mode == 0 || mode == 1) && tc_geo(d, mode, &g) ? 1 : 0;
}

extern "C" long long maavss_conv_tc_workspace_bytes(const maavss_conv_desc* d, int mode) {
    TcGeo g;
    if (!d || !tc_geo(d, mode, &g)) return 0;
    const size_t act = align_up((size_t)g.N * g.D * g.H * g.W * g.Ca * 2, 1024);
    const size_t wts = align_up((size_t)g.KD * g.KH * g.KW * g.Co * g.Ca * 2, 1024);
    return (long long)(act + wts);
}

extern "C" int maavss_conv_tc(const maavss_conv_desc* d, int mode, const float* act, const float* w, float* out, void* workspace,
                              long long workspace_bytes, void* stream) {
    MAAVSS_CHECK_ARG(d && act && w && out && workspace, "conv_tc: null argument");
    TcGeo g;
    MAAVSS_CHECK_ARG((mode == 0 || mode == 1) && tc_geo(d, mode, &g), "conv_tc: layer outside the tensor-core envelope "
                     "(stride 1, same geometry, channels in {16,32,64}, W in {32,64,128})");
    MAAVSS_CHECK_ARG(workspace_bytes >= maavss_conv_tc_workspace_bytes(d, mode) && ((uintptr_t)workspace & 1023) == 0,
                     "conv_tc: workspace too small or not 1024-byte aligned");
    if (g.N == 0) return MAAVSS_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int KV = g.KD * g.KH * g.KW;
    const long long S = (long long)g.D * g.H * g.W, total = S * g.N;
    __nv_bfloat16* act_cl = (__nv_bfloat16*)workspace;
    __nv_bfloat16* w_rp = (__nv_bfloat16*)((uint8_t*)workspace + align_up((size_t)total * g.Ca * 2, 1024));
    {
        const int rc = maavss_tc_to_channels_last(act, act_cl, g.Ca, g.Ca, S, total, st);
        if (rc) return rc;
    }
    tcc::repack_weights_bf16<<<ceil_div(d->Cout * d->Cin * KV, 256), 256, 0, st>>>(w, w_rp, d->Cout, d->Cin, KV, mode);
    MAAVSS_LAUNCH_CHECK();

    PFN_encodeTiled enc = get_encode_conv();
    MAAVSS_CHECK_ARG(enc, "cuTensorMapEncodeTiled is unavailable (driver too old?)");
    const int swz = g.Ca * 2;
    const CUtensorMapSwizzle swz_e = swz == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (swz == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
    const int BW = g.W, BH = 128 / g.W;
    CUtensorMap tmA, tmB;
    {
        cuuint64_t dims[5] = {(cuuint64_t)g.Ca, (cuuint64_t)g.W, (cuuint64_t)g.H, (cuuint64_t)g.D, (cuuint64_t)g.N};
        cuuint64_t strides[4] = {(cuuint64_t)g.Ca * 2, (cuuint64_t)g.W * g.Ca * 2, (cuuint64_t)g.H * g.W * g.Ca * 2,
                                 (cuuint64_t)g.D * g.H * g.W * g.Ca * 2};
        cuuint32_t box[5] = {(cuuint32_t)g.Ca, (cuuint32_t)BW, (cuuint32_t)BH, 1, 1};
        cuuint32_t estr[5] = {1, 1, 1, 1, 1};
        CUresult r = enc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, act_cl, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         swz_e, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        MAAVSS_CHECK_ARG(r == CUDA_SUCCESS, "conv_tc: cuTensorMapEncodeTiled (activation) failed (%d)", (int)r);
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)g.Ca, (cuuint64_t)KV * g.Co};
        cuuint64_t strides[1] = {(cuuint64_t)g.Ca * 2};
        cuuint32_t box[2] = {(cuuint32_t)g.Ca, (cuuint32_t)g.Co};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w_rp, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         swz_e, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        MAAVSS_CHECK_ARG(r == CUDA_SUCCESS, "conv_tc: cuTensorMapEncodeTiled (weights) failed (%d)", (int)r);
    }
    tcc::Params p;
    p.N = g.N; p.D = g.D; p.H = g.H; p.W = g.W;
    p.Ca = g.Ca; p.Co = g.Co;
    p.KD = g.KD; p.KH = g.KH; p.KW = g.KW; p.PD = g.PD; p.PH = g.PH; p.PW = g.PW;
    p.BH = BH;
    p.log2BW = BW == 128 ? 7 : (BW == 64 ? 6 : 5);
    p.swz = swz;
    p.a_bytes = 128 * swz;
    p.stage_bytes = (int)align_up((size_t)p.a_bytes + (size_t)g.Co * swz, 1024);
    p.out = out;
    p.stages = tcc::SMEM_BUDGET / p.stage_bytes;
    if (p.stages > tcc::MAX_STAGES) p.stages = tcc::MAX_STAGES;
    const int smem = p.stages * p.stage_bytes + 1024 + 1024;       // ring + barriers / TMEM slot + alignment slack
    static bool attr_set = false;
    if (!attr_set) {
        MAAVSS_CUDA(cudaFuncSetAttribute(tcc::tc_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tcc::SMEM_BUDGET + 2048));
        attr_set = true;
    }
    const int tiles = g.N * g.D * (g.H / BH);
    const int grid = tiles < maavss_num_sms() ? tiles : maavss_num_sms();
    tcc::tc_conv_kernel<<<grid, tcc::THREADS, smem, st>>>(tmA, tmB, p);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
