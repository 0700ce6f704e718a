// fp32-accurate GEMM on the 5th-generation tensor cores: 3xTF32 on tcgen05.mma.kind::tf32 (TMA + TMEM), for the
// GEMM-shaped layers of the fusion models whose parity bar is fp32 (<= 1e-5): nn.Linear fc1 / fc2 / a_fc1 / v_fc1 forward,
// data gradient and weight gradient with the accumulation windows stacked (M = 64 rows; avse_model.py:550,553,618-627),
// the Bi-LSTM input projections (avse_model.py:542-543).
//
//     D[P, Q] = A . B^T,   A: P x K,  B: Q x K   (fp32 in HBM, fp32 accumulate in TMEM)
//
// Each operand is K-major (stored [rows][K], the nn.Linear weight / activation layout) or MN-major (stored [K][rows]:
// the same buffers seen from the data-gradient and weight-gradient products), so no transposed copy of a weight matrix is
// ever made.  x = hi + lo with hi = x & 0xffffe000 (exactly a tf32 number) and lo = x - hi (exact in fp32; the tensor core
// reads its top 19 bits), and D = Ahi.Bhi + Alo.Bhi + Ahi.Blo: three MMAs per k-step, error ~2^-21 per product -- fp32 grade.
// kind::tf32 ignores the low 13 mantissa bits of its fp32 inputs (truncation; results are bit-identical with and without
// masking them first, measured on B200), so the landed tile itself serves as hi and only lo is written.
//
//   warp 0      TMA producer: fp32 boxes of 32 elements (128 bytes, SWIZZLE_128B) into a 2-4 stage mbarrier ring
//   warps 2-9   split: read each landed tile once from shared memory and write lo = x - trunc(x) beside it
//               (element-wise, so the swizzled layout is irrelevant), fence.proxy.async, arrive
//   warp 1      TMEM owner + MMA issuer: 128 x BN x 8 tcgen05.mma x 3 per k-step into one of two accumulator stages
//   warps 10-17 epilogue: tcgen05.ld -> bias / activation / accumulate -> coalesced stores (transposed output straight from
//               the registers, row-major output through a swizzled shared-memory transpose), overlapped with the next tile
// Accumulation: the tensor core adds into its TMEM accumulator with round-toward-zero, a bias that grows with the number of
// accumulated MMAs (measured on B200: -2.5e-6 relative at K = 512, -0.7e-6 at K = 32).  So the K loop runs in chunks of
// CHUNK k-blocks (24 MMAs) into alternating TMEM stages that start from zero, and the epilogue warps add the chunks into
// fp32 registers with round-to-nearest (Ootomo & Yokota's correction for 3xTF32).
// MN-major tf32 operands exist in one shared-memory layout only, SWIZZLE_128B_BASE32B (32-byte chunks XOR-swizzled over 4
// rows of 128 bytes; TMA mode CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B): the plain 128-byte swizzle silently yields zeros.
// Persistent CTAs (one per SM) walk (tile, k-split) work items; k-splits write partial tiles to a caller-owned workspace
// that a small finalize kernel reduces deterministically (no atomics).  These layers stream a weight matrix once against
// 64 rows, so the kernel is HBM bound by design: per SM a stage (24 KB of new operand bytes at BN = 64) costs ~1000 cycles
// of HBM time against ~400 of tensor pipe and ~600 of the shared-memory split.
#include <stdlib.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/maavss_b200.h"

int maavss_make_tmap_2d_swz(CUtensorMap* map, const void* base, int dtype_code, int esize, long long rows, long long cols, long long ld,
                            int box_rows, int box_cols, int swizzle_code);

namespace tg3 {

constexpr int BM = 128, BK = 32, ACC_STAGES = 2, CHUNK = 2;
constexpr int SPLIT_WARPS = 8, EPI_WARPS = 8, THREADS = 64 + 32 * (SPLIT_WARPS + EPI_WARPS);
constexpr int A_BYTES = BM * BK * 4;      // 16 KB, K-major ([128 rows][128 B]) or MN-major (4 boxes of [32 k][128 B])

template <int BN>
struct Smem {
    static constexpr int B_BYTES = BN * BK * 4;
    static constexpr int STAGE_BYTES = 2 * (A_BYTES + B_BYTES);          // A, B, A_lo, B_lo
    static constexpr int STAGES = BN <= 64 ? 4 : 3;
    static constexpr int EPI_BYTES = EPI_WARPS * 32 * 32 * 4;
    static constexpr int BAR_OFF = STAGES * STAGE_BYTES + EPI_BYTES;
    static constexpr int TOTAL = BAR_OFF + 256 + 1024;
};

struct Params {
    int P, Q, K;
    int a_mn, b_mn;            // operand major-ness: 0 K-major, 1 MN-major
    int splits, kb_total;
    int transposed;            // 0: out[p * ldc + q]   1: out[q * ldc + p]
    float* out;
    long long ldc;
    int pb;                    // transposed output in blocks of pb rows of p: out[(p / pb) * pb_stride + q * ldc + p % pb]  (0: plain)
    long long pb_stride;
    float* ws;                 // splits > 1: partial tiles, (splits, rows, cols) compact in the layout of out
    const float* bias;         // fused epilogue only (splits == 1)
    int bias_on_q, act, accumulate;
    float slope;
    long long* trace;          // dbg & 64: clock64() of every pipeline hop of CTA 0's first 48 stages, [stage][8]
    int dbg;                   // timing experiments only (MAAVSS_G3_DBG): 1 skip the hi/lo split, 2 skip the MMAs, 4 skip the chunk loads
};

__device__ __forceinline__ float apply_act(float v, int act, float slope) {
    switch (act) {
        case MAAVSS_ACT_TANH: return tanhf(v);
        case MAAVSS_ACT_LEAKY: return v > 0.f ? v : v * slope;
        case MAAVSS_ACT_SIGMOID: return 1.f / (1.f + __expf(-v));
        default: return v;
    }
}

// MN-major tf32 operand tile: 32-element (128-byte) column blocks of [32 k-rows][128 B] written by TMA with
// CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B (UMMA layout SWIZZLE_128B_BASE32B = 1: swizzle atoms of 4 k-rows).  Column blocks follow
// each other every BK * 128 bytes (LBO), 4-row k-groups every 512 bytes (SBO); one tcgen05.mma (K = 8) reads two k-groups.
__device__ __forceinline__ uint64_t smem_desc_mn_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)((BK * 128) >> 4) << 16;            // leading byte offset: next 32-element block along M / N
    d |= (uint64_t)(512 >> 4) << 32;                   // stride byte offset: next group of 4 k-rows
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)1 << 61;                            // SWIZZLE_128B_BASE32B
    return d;
}

template <int ACT>
__device__ __forceinline__ float act_t(float v, float slope) {
    if constexpr (ACT == MAAVSS_ACT_TANH) return tanhf(v);
    else if constexpr (ACT == MAAVSS_ACT_LEAKY) return v > 0.f ? v : v * slope;
    else if constexpr (ACT == MAAVSS_ACT_SIGMOID) return 1.f / (1.f + __expf(-v));
    else return v;
}

// Transposed epilogue of one accumulator row: o[j * ld] = act(acc[j] + bias) (+ o[j * ld]), j < nq.  The lanes of a warp
// hold consecutive rows, so every store (and accumulate load) of the warp is one 128-byte line.
template <int BN, int ACT>
__device__ __forceinline__ void store_col(float (&acc)[BN], float* __restrict__ o, long long ld, int nq, float bp,
                                          const float* __restrict__ bq, float slope, bool accum) {
    if (bq) {
        // per-column bias first, 16 loads in flight: a load between two stores of the loop below waits out its full latency
        // behind the store (measured: 32 x ~350 cycles per tile on the convolution output path)
#pragma unroll
        for (int j0 = 0; j0 < BN; j0 += 16) {
            float t[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) t[j] = (j0 + j < nq) ? __ldg(bq + j0 + j) : 0.f;
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j0 + j] += t[j];
        }
        bq = nullptr;
    }
    if (nq == BN && !accum) {
#pragma unroll
        for (int j = 0; j < BN; ++j) o[(size_t)j * ld] = act_t<ACT>(acc[j] + bp, slope);
        return;
    }
#pragma unroll
    for (int j0 = 0; j0 < BN; j0 += 16) {
        float old[16];
        if (accum) {
#pragma unroll
            for (int j = 0; j < 16; ++j) old[j] = (j0 + j < nq) ? o[(size_t)(j0 + j) * ld] : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            if (j0 + j < nq) {
                float v = act_t<ACT>(acc[j0 + j] + bp, slope);
                if (accum) v += old[j];
                o[(size_t)(j0 + j) * ld] = v;
            }
        }
    }
}

// Row-major epilogue of one 32 x 32 block after its transpose through shared memory: lane (sub = lane / 8, chunk = lane % 8)
// stores the float4 of rows sub, sub + 4, ... (st = stage + sub * 32; rows_left = P - first row).
// MODE 0: plain store   1: accumulate into the output   2: bias (per column bq4 / per row bp) + activation (+ accumulate)
template <int MODE>
__device__ __forceinline__ void store_rows(const float* __restrict__ st, int lane, float* __restrict__ o, long long ld, int rows_left,
                                           const float* __restrict__ bq4, const float* __restrict__ bp, int act_acc, float slope) {
    const int ch = lane & 7, sub = lane >> 3;
    float4 v[8];
#pragma unroll
    for (int ps = 0; ps < 8; ++ps) {
        const int rr = ps * 4 + sub;
        v[ps] = *reinterpret_cast<const float4*>(&st[ps * 128 + ((ch ^ (rr & 7)) << 2)]);
    }
    if (MODE == 1 || (MODE == 2 && (act_acc & 256))) {
        float4 old[8];
#pragma unroll
        for (int ps = 0; ps < 8; ++ps)                // all 8 loads in flight before the first store
            old[ps] = (4 * ps < rows_left) ? *reinterpret_cast<const float4*>(o + (size_t)(4 * ps) * ld) : make_float4(0.f, 0.f, 0.f, 0.f);
        if (MODE == 2) {
            const int act = act_acc & 255;
            float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (bq4) b4 = __ldg(reinterpret_cast<const float4*>(bq4));
#pragma unroll
            for (int ps = 0; ps < 8; ++ps) {
                const float b = (bp && 4 * ps < rows_left) ? __ldg(bp + 4 * ps) : 0.f;
                v[ps].x = apply_act(v[ps].x + b4.x + b, act, slope); v[ps].y = apply_act(v[ps].y + b4.y + b, act, slope);
                v[ps].z = apply_act(v[ps].z + b4.z + b, act, slope); v[ps].w = apply_act(v[ps].w + b4.w + b, act, slope);
            }
        }
#pragma unroll
        for (int ps = 0; ps < 8; ++ps) { v[ps].x += old[ps].x; v[ps].y += old[ps].y; v[ps].z += old[ps].z; v[ps].w += old[ps].w; }
    } else if (MODE == 2) {
        const int act = act_acc & 255;
        float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (bq4) b4 = __ldg(reinterpret_cast<const float4*>(bq4));
#pragma unroll
        for (int ps = 0; ps < 8; ++ps) {
            const float b = (bp && 4 * ps < rows_left) ? __ldg(bp + 4 * ps) : 0.f;
            v[ps].x = apply_act(v[ps].x + b4.x + b, act, slope); v[ps].y = apply_act(v[ps].y + b4.y + b, act, slope);
            v[ps].z = apply_act(v[ps].z + b4.z + b, act, slope); v[ps].w = apply_act(v[ps].w + b4.w + b, act, slope);
        }
    }
#pragma unroll
    for (int ps = 0; ps < 8; ++ps)
        if (4 * ps < rows_left) *reinterpret_cast<float4*>(o + (size_t)(4 * ps) * ld) = v[ps];
}

template <int BN>
__global__ void __launch_bounds__(THREADS, 1) gemm_tf32x3_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                 const __grid_constant__ CUtensorMap tmB, Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    using S = Smem<BN>;
    constexpr int STAGES = S::STAGES;
    constexpr int B_BYTES = S::B_BYTES;
    constexpr int LO_OFF = A_BYTES + B_BYTES;          // lo tiles mirror the hi tiles at this offset inside a stage
    constexpr int TMEM_COLS = ACC_STAGES * BN <= 128 ? 128 : (ACC_STAGES * BN <= 256 ? 256 : 512);
    float* epi = (float*)(smem + STAGES * S::STAGE_BYTES);
    uint64_t* full = (uint64_t*)(smem + S::BAR_OFF);   // TMA landed
    uint64_t* ready = full + STAGES;                   // hi / lo split done
    uint64_t* empty = ready + STAGES;                  // MMAs that read the stage retired
    uint64_t* tfull = empty + STAGES;
    uint64_t* tempty = tfull + ACC_STAGES;
    uint32_t* tmem_slot = (uint32_t*)(tempty + ACC_STAGES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles = (p.Q + BN - 1) / BN, m_tiles = (p.P + BM - 1) / BM;
    const int items = n_tiles * m_tiles * p.splits;
    const int kb_base = p.kb_total / p.splits, kb_rem = p.kb_total % p.splits;

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tmA);
        tc::prefetch_tmap(&tmB);
        for (int s = 0; s < STAGES; ++s) {
            tc::mbar_init(&full[s], 1);
            tc::mbar_init(&ready[s], SPLIT_WARPS);
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
            int s = 0, tr_p = 0;
            uint32_t ph = 0;
            for (int w = blockIdx.x; w < items; w += gridDim.x) {
                const int tile = w / p.splits, sp = w % p.splits;
                const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
                const int kb0 = sp * kb_base + min(sp, kb_rem), nkb = kb_base + (sp < kb_rem ? 1 : 0);
                // Every CTA starts its walk over the k-blocks at a different one.  With all CTAs in step, every request in
                // flight would address the same 128-byte column of a (K * 4)-byte row pitch -- a power of two for these
                // layers -- and camp on a fraction of the HBM channels (measured: 1.7 TB/s without the rotation).
                const int rot = (p.dbg & 8) ? 0 : (int)((blockIdx.x * 5u + (unsigned)tile) % (unsigned)nkb);
                for (int i = 0; i < nkb; ++i) {
                    const int kb = kb0 + (i + rot) % nkb;
                    tc::mbar_wait(&empty[s], ph ^ 1);
                    if (p.trace && blockIdx.x == 0 && tr_p < 48) p.trace[tr_p * 8 + 0] = clock64();
                    tc::mbar_expect_tx(&full[s], ((p.dbg & 32) ? 0 : A_BYTES) + ((p.dbg & 16) ? 0 : B_BYTES));
                    uint8_t* a_dst = smem + s * S::STAGE_BYTES;
                    uint8_t* b_dst = a_dst + A_BYTES;
                    if (p.dbg & 32) {
                    } else if (p.a_mn) {
#pragma unroll
                        for (int j = 0; j < BM / 32; ++j) tc::tma_load_2d(a_dst + j * (BK * 128), &tmA, &full[s], m0 + 32 * j, kb * BK);
                    } else {
                        tc::tma_load_2d(a_dst, &tmA, &full[s], kb * BK, m0);
                    }
                    if (p.dbg & 16) {
                    } else if (p.b_mn) {
#pragma unroll
                        for (int j = 0; j < BN / 32; ++j) tc::tma_load_2d(b_dst + j * (BK * 128), &tmB, &full[s], n0 + 32 * j, kb * BK);
                    } else {
                        tc::tma_load_2d(b_dst, &tmB, &full[s], kb * BK, n0);
                    }
                    if (p.trace && blockIdx.x == 0 && tr_p < 48) p.trace[tr_p++ * 8 + 1] = clock64();
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // The whole warp walks the loop (uniform control flow: descriptors and barrier waits live in uniform registers) and one
        // elected lane issues.  Inside an `if (lane == 0)` region every tcgen05.mma paid vector -> uniform register moves for
        // its descriptors: ~62 cycles per issue against the 33 the tensor pipe needs for a 128 x 64 x 8 tf32 MMA.
        {
            const uint32_t idesc = tc::make_idesc(2, BM, BN) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16);
            // Descriptors are built once: the address field is the low 14 bits (units of 16 bytes) and every tile of the ring
            // stays below 256 KB, so a stage / k-step / lo-tile offset is a plain add.  The single issuing thread has to feed a
            // 33-cycle MMA (128 x 64 x 8) every 33 cycles: building four descriptors per k-step from scratch took 1100-1500
            // cycles per stage against 400 of tensor time (clock64 trace of the first version).
            const uint32_t smem0 = tc::smem_u32(smem);
            const uint64_t a_desc0 = p.a_mn ? smem_desc_mn_sw128(smem0) : tc::smem_desc_k_sw128(smem0);
            const uint64_t b_desc0 = p.b_mn ? smem_desc_mn_sw128(smem0 + A_BYTES) : tc::smem_desc_k_sw128(smem0 + A_BYTES);
            // per k-step of 8: K-major +32 bytes inside the 128-byte row; MN-major +8 rows of 128 bytes
            const uint64_t a_step = (p.a_mn ? 1024u : 32u) >> 4, b_step = (p.b_mn ? 1024u : 32u) >> 4;
            constexpr uint64_t LO = LO_OFF >> 4, STG = S::STAGE_BYTES >> 4;
            int s = 0, it = 0, tr_m = 0;           // it counts accumulator hand-offs (chunks), not items
            uint32_t ph = 0;
            for (int w = blockIdx.x; w < items; w += gridDim.x) {
                const int sp = w % p.splits;
                const int nkb = kb_base + (sp < kb_rem ? 1 : 0);
                for (int kc = 0; kc < nkb; kc += CHUNK, ++it) {
                    const int a = it & 1;
                    const uint32_t aph = (it >> 1) & 1;
                    tc::mbar_wait(&tempty[a], aph ^ 1);
                    tc::tc_fence_after();
                    const uint32_t d_tmem = tmem_base + a * BN;
                    const int kend = min(nkb, kc + CHUNK);
                    for (int kb = kc; kb < kend; ++kb) {
                        tc::mbar_wait(&ready[s], ph);
                        tc::tc_fence_after();
                        if (p.trace && blockIdx.x == 0 && lane == 0 && tr_m < 48) p.trace[tr_m * 8 + 4] = clock64();
                        uint64_t ad = a_desc0 + (uint64_t)s * STG, bd = b_desc0 + (uint64_t)s * STG;
                        if (!(p.dbg & 2) && tc::elect_one()) {
#pragma unroll
                            for (int k = 0; k < BK / 8; ++k) {
                                tc::umma_tf32(d_tmem, ad + LO, bd, idesc, (kb != kc) || (k != 0));      // the two small terms first
                                tc::umma_tf32(d_tmem, ad, bd + LO, idesc, 1);
                                tc::umma_tf32(d_tmem, ad, bd, idesc, 1);
                                ad += a_step;
                                bd += b_step;
                            }
                        }
                        __syncwarp();
                        if (tc::elect_one()) tc::umma_commit(&empty[s]);
                        if (p.trace && blockIdx.x == 0 && lane == 0 && tr_m < 48) p.trace[tr_m * 8 + 5] = clock64();
                        ++tr_m;
                        if (++s == STAGES) { s = 0; ph ^= 1; }
                    }
                    if (tc::elect_one()) tc::umma_commit(&tfull[a]);
                }
            }
        }
    } else if (warp < 2 + SPLIT_WARPS) {
        // hi / lo split of every landed stage: 16-byte vectors, consecutive threads on consecutive vectors (conflict free)
        const int tid = threadIdx.x - 64;
        constexpr int NV = (A_BYTES + B_BYTES) / 16, PER = NV / (32 * SPLIT_WARPS);
        static_assert(NV % (32 * SPLIT_WARPS) == 0, "stage vectors must divide over the split warps");
        int s = 0, tr_s = 0;
        uint32_t ph = 0;
        for (int w = blockIdx.x; w < items; w += gridDim.x) {
            const int sp = w % p.splits;
            const int nkb = kb_base + (sp < kb_rem ? 1 : 0);
            for (int kb = 0; kb < nkb; ++kb) {
                tc::mbar_wait(&full[s], ph);
                if (p.trace && blockIdx.x == 0 && tid == 0 && tr_s < 48) p.trace[tr_s * 8 + 2] = clock64();
                if (p.dbg & 1) { __syncwarp(); if (lane == 0) tc::mbar_arrive(&ready[s]); if (p.trace && blockIdx.x == 0 && tid == 0 && tr_s < 48) p.trace[tr_s * 8 + 3] = clock64(); ++tr_s; if (++s == STAGES) { s = 0; ph ^= 1; } continue; }
                uint4* hi = reinterpret_cast<uint4*>(smem + s * S::STAGE_BYTES);
                uint4* lo = reinterpret_cast<uint4*>(smem + s * S::STAGE_BYTES + LO_OFF);
                uint4 v[PER];
#pragma unroll
                for (int j = 0; j < PER; ++j) v[j] = hi[tid + j * 32 * SPLIT_WARPS];
#pragma unroll
                for (int j = 0; j < PER; ++j) {
                    uint4 h = make_uint4(v[j].x & 0xffffe000u, v[j].y & 0xffffe000u, v[j].z & 0xffffe000u, v[j].w & 0xffffe000u);
                    uint4 l;
                    l.x = __float_as_uint(__uint_as_float(v[j].x) - __uint_as_float(h.x));
                    l.y = __float_as_uint(__uint_as_float(v[j].y) - __uint_as_float(h.y));
                    l.z = __float_as_uint(__uint_as_float(v[j].z) - __uint_as_float(h.z));
                    l.w = __float_as_uint(__uint_as_float(v[j].w) - __uint_as_float(h.w));
                    if (p.dbg & 128) hi[tid + j * 32 * SPLIT_WARPS] = h;     // not needed: the tensor core truncates (see header)
                    lo[tid + j * 32 * SPLIT_WARPS] = l;
                }
                tc::fence_proxy_async();           // generic-proxy writes -> visible to the tensor core's async-proxy reads
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&ready[s]);
                if (p.trace && blockIdx.x == 0 && tid == 0 && tr_s < 48) p.trace[tr_s * 8 + 3] = clock64();
                ++tr_s;
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
    } else {
        // epilogue: warp w may access TMEM lanes 32 * (w % 4) .. + 31; two warps share a lane quadrant and take 32 accumulator
        // columns each.  Lane i owns accumulator row 32 * q + i; the chunks of a work item are added in fp32 registers (round
        // to nearest), then bias / activation / store.
        static_assert(BN == 64 && EPI_WARPS == 8, "two epilogue warps x 32 columns per lane quadrant");
        const int ew = warp - (2 + SPLIT_WARPS), q = warp & 3, half = ew >> 2;
        float* stage = epi + ew * (32 * 32);
        const int sub = lane >> 3, cc = (lane & 7) * 4;
        int it = 0;
        for (int w = blockIdx.x; w < items; w += gridDim.x) {
            const int tile = w / p.splits, sp = w % p.splits;
            const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN + half * 32;
            const int nkb = kb_base + (sp < kb_rem ? 1 : 0);
            const bool fused = p.splits == 1;
            float* dst = fused ? p.out : p.ws + (size_t)sp * p.P * p.Q;
            const long long ld = fused ? p.ldc : (p.transposed ? p.P : p.Q);
            const int prow = m0 + q * 32 + lane;               // this lane's accumulator row
            float acc[32];
            for (int kc = 0; kc < nkb; kc += CHUNK, ++it) {
                const int a = it & 1;
                const uint32_t aph = (it >> 1) & 1;
                tc::mbar_wait(&tfull[a], aph);
                tc::tc_fence_after();
                const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + a * BN + half * 32;
                if (!(p.dbg & 4)) {
                    uint32_t r[32];
                    tc::tmem_ld_32x32(t_addr, r);
                    tc::tmem_ld_wait();
                    if (kc == 0) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) acc[j] = __uint_as_float(r[j]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) acc[j] += __uint_as_float(r[j]);
                    }
                } else if (kc == 0) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) acc[j] = 0.f;
                }
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&tempty[a]);
                if (p.trace && blockIdx.x == 0 && ew == 0 && lane == 0 && it < 48) p.trace[it * 8 + 6] = clock64();
            }
            if (p.transposed) {
                // out[q * ld + p]: for one q the 32 lanes hold 32 consecutive p -> one 128-byte line per store
                const int nq = min(32, p.Q - n0);
                if (prow < p.P && nq > 0) {
                    float* o = dst + (size_t)n0 * ld + prow;
                    if (fused && p.pb) o = dst + (size_t)(prow / p.pb) * p.pb_stride + (size_t)n0 * ld + prow % p.pb;
                    const float bp = (fused && p.bias && !p.bias_on_q) ? __ldg(p.bias + prow) : 0.f;
                    const float* bq = (fused && p.bias && p.bias_on_q) ? p.bias + n0 : nullptr;
                    const int act = fused ? p.act : MAAVSS_ACT_NONE;
                    const bool accum = fused && p.accumulate;
                    switch (act) {
                        case MAAVSS_ACT_TANH: store_col<32, MAAVSS_ACT_TANH>(acc, o, ld, nq, bp, bq, p.slope, accum); break;
                        case MAAVSS_ACT_LEAKY: store_col<32, MAAVSS_ACT_LEAKY>(acc, o, ld, nq, bp, bq, p.slope, accum); break;
                        case MAAVSS_ACT_SIGMOID: store_col<32, MAAVSS_ACT_SIGMOID>(acc, o, ld, nq, bp, bq, p.slope, accum); break;
                        default: store_col<32, MAAVSS_ACT_NONE>(acc, o, ld, nq, bp, bq, p.slope, accum); break;
                    }
                }
            } else {
                // out[p * ld + q]: transpose the 32 x 32 block through shared memory (16-byte chunks XOR-swizzled by row)
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<float4*>(&stage[lane * 32 + ((j ^ (lane & 7)) << 2)]) =
                        make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
                __syncwarp();
                const int col = n0 + cc;
                if (col < p.Q) {                         // Q % 4 == 0: the float4 is entirely inside or outside
                    const int row0 = m0 + q * 32 + sub;
                    float* o = dst + (size_t)row0 * ld + col;
                    const float* st = stage + sub * 32;
                    // three specialisations: the general one carries ~200 (mostly predicated-off) instructions per stored
                    // float4 -- measured 5800 cycles per tile on the convolution data-gradient product, which has none of it
                    if (!fused || (!p.bias && p.act == MAAVSS_ACT_NONE && !p.accumulate))
                        store_rows<0>(st, lane, o, ld, p.P - row0, nullptr, nullptr, 0, 0.f);
                    else if (!p.bias && p.act == MAAVSS_ACT_NONE)
                        store_rows<1>(st, lane, o, ld, p.P - row0, nullptr, nullptr, 0, 0.f);
                    else
                        store_rows<2>(st, lane, o, ld, p.P - row0, p.bias_on_q ? p.bias + col : nullptr, p.bias_on_q ? nullptr : p.bias + row0,
                                      p.act | (p.accumulate ? 256 : 0), p.slope);
                }
                __syncwarp();
            }
            if (p.trace && blockIdx.x == 0 && ew == 0 && lane == 0 && it - 1 < 48) p.trace[(it - 1) * 8 + 7] = clock64();
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc(tmem_base, TMEM_COLS);
}

// out = act(sum_s ws[s] + bias) (+ out), element order of `out` (rows x cols with pitch ldc)
__global__ void __launch_bounds__(256) finalize_kernel(const float* __restrict__ ws, int splits, long long rows, long long cols,
                                                       float* __restrict__ out, long long ldc, const float* __restrict__ bias,
                                                       int bias_on_cols, int act, float slope, int accumulate, int pb,
                                                       long long pb_stride) {
    const long long n = rows * cols;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        const long long r = i / cols, c = i - r * cols;
        // 8 partials in flight per thread (the sum is latency-bound otherwise: up to 37 dependent L2 loads); fixed order
        float v8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        int s = 0;
        for (; s + 8 <= splits; s += 8) {
#pragma unroll
            for (int u = 0; u < 8; ++u) v8[u] += __ldcs(ws + (size_t)(s + u) * n + i);
        }
        for (int u = 0; s < splits; ++s, ++u) v8[u] += __ldcs(ws + (size_t)s * n + i);
        float v = ((v8[0] + v8[1]) + (v8[2] + v8[3])) + ((v8[4] + v8[5]) + (v8[6] + v8[7]));
        if (bias) v += __ldg(bias + (bias_on_cols ? c : r));
        v = apply_act(v, act, slope);
        float* o = pb ? out + (c / pb) * pb_stride + r * ldc + c % pb : out + r * ldc + c;
        *o = accumulate ? *o + v : v;
    }
}

static int choose_splits(int P, int Q, int K, int BN) {
    const int tiles = ceil_div(P, BM) * ceil_div(Q, BN);
    const int kb = ceil_div(K, BK);
    const int sms = maavss_num_sms();
    if (tiles * 2 > sms) return 1;
    int s = sms / tiles;
    if (s > kb / 4) s = kb / 4;          // at least 4 k-blocks per work item
    return s < 1 ? 1 : s;
}

template <int BN>
static int launch(const float* A, int a_mn, long long lda, const float* B, int b_mn, long long ldb, Params p, cudaStream_t st) {
    CUtensorMap tmA, tmB;
    int rc;
    // K-major: tensor (rows, K), box (tile rows, 32);  MN-major: tensor (K, rows), box (32 k, 32 rows)
    rc = a_mn ? maavss_make_tmap_2d_swz(&tmA, A, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, p.K, p.P, lda, BK, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)
              : maavss_make_tmap_2d_swz(&tmA, A, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, p.P, p.K, lda, BM, BK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    rc = b_mn ? maavss_make_tmap_2d_swz(&tmB, B, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, p.K, p.Q, ldb, BK, 32, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)
              : maavss_make_tmap_2d_swz(&tmB, B, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, p.Q, p.K, ldb, BN, BK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        MAAVSS_CUDA(cudaFuncSetAttribute(gemm_tf32x3_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<BN>::TOTAL));
        attr_set = true;
    }
    const int items = ceil_div(p.P, BM) * ceil_div(p.Q, BN) * p.splits;
    const int grid = items < maavss_num_sms() ? items : maavss_num_sms();
    gemm_tf32x3_kernel<BN><<<grid, THREADS, Smem<BN>::TOTAL, st>>>(tmA, tmB, p);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

static int pick_bn(int Q, int K) { (void)Q; (void)K; return 64; }   // one accumulator row = 64 fp32 registers per epilogue thread

}  // namespace tg3

extern "C" long long maavss_gemm_tf32x3_workspace_bytes(int P, int Q, int K) {
    const int s = tg3::choose_splits(P, Q, K, tg3::pick_bn(Q, K));
    return s > 1 ? (long long)s * P * Q * 4 : 0;
}

extern "C" int maavss_gemm_tf32x3_supported(int P, int Q, int K, int a_mn_major, long long lda, int b_mn_major, long long ldb,
                                            long long ldc, int out_transposed) {
    if (P <= 0 || Q <= 0 || K <= 0) return 0;
    if (lda % 4 || ldb % 4) return 0;                         // TMA: 16-byte multiple pitches
    if (!out_transposed && (Q % 4 || ldc % 4)) return 0;      // float4 stores of the row-major epilogue
    (void)a_mn_major; (void)b_mn_major;                       // rows / columns past P, Q, K are zero-filled by TMA
    return 1;
}

extern "C" int maavss_gemm_tf32x3(const float* A, int a_mn_major, long long lda, const float* B, int b_mn_major, long long ldb,
                                  int P, int Q, int K, float* out, long long ldc, int out_transposed, const float* bias,
                                  int bias_on_q, int act, float slope, int accumulate, void* workspace,
                                  long long workspace_bytes, void* stream) {
    return maavss_gemm_tf32x3_blocked(A, a_mn_major, lda, B, b_mn_major, ldb, P, Q, K, out, ldc, out_transposed, 0, 0, bias, bias_on_q, act,
                                      slope, accumulate, workspace, workspace_bytes, stream);
}

extern "C" int maavss_gemm_tf32x3_blocked(const float* A, int a_mn_major, long long lda, const float* B, int b_mn_major, long long ldb,
                                          int P, int Q, int K, float* out, long long ldc, int out_transposed, int p_block,
                                          long long p_block_stride, const float* bias, int bias_on_q, int act, float slope,
                                          int accumulate, void* workspace, long long workspace_bytes, void* stream) {
    if (P == 0 || Q == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(p_block == 0 || (out_transposed && p_block % 32 == 0 && p_block > 0),
                     "gemm_tf32x3_blocked: p_block must be a multiple of 32 and needs the transposed output");
    MAAVSS_CHECK_ARG(A && B && out && K > 0, "gemm_tf32x3: bad argument");
    MAAVSS_CHECK_ARG(maavss_gemm_tf32x3_supported(P, Q, K, a_mn_major, lda, b_mn_major, ldb, ldc, out_transposed),
                     "gemm_tf32x3: pitches must be multiples of 4 floats (and Q, ldc for a row-major output)");
    MAAVSS_CHECK_ARG(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0 && ((uintptr_t)out & 15) == 0 && ((uintptr_t)bias & 15) == 0,
                     "gemm_tf32x3: operands must be 16-byte aligned");
    MAAVSS_CHECK_ARG(act == MAAVSS_ACT_NONE || act == MAAVSS_ACT_TANH || act == MAAVSS_ACT_LEAKY || act == MAAVSS_ACT_SIGMOID,
                     "gemm_tf32x3: activation %d not supported", act);
    cudaStream_t st = (cudaStream_t)stream;
    const int BN = tg3::pick_bn(Q, K);
    tg3::Params p{};
    p.P = P; p.Q = Q; p.K = K;
    p.a_mn = a_mn_major ? 1 : 0;
    p.b_mn = b_mn_major ? 1 : 0;
    p.kb_total = ceil_div(K, tg3::BK);
    p.splits = tg3::choose_splits(P, Q, K, BN);
    p.transposed = out_transposed ? 1 : 0;
    p.out = out; p.ldc = ldc;
    p.pb = p_block; p.pb_stride = p_block_stride;
    p.bias = bias; p.bias_on_q = bias_on_q; p.act = act; p.slope = slope; p.accumulate = accumulate;
    if (const char* e = getenv("MAAVSS_G3_DBG")) p.dbg = atoi(e);
    static long long* trace_buf = nullptr;
    if (p.dbg & 64) {
        if (!trace_buf) cudaMalloc(&trace_buf, sizeof(long long) * 48 * 8);
        cudaMemsetAsync(trace_buf, 0, sizeof(long long) * 48 * 8, st);
        p.trace = trace_buf;
    }
    if (p.splits > 1) {
        MAAVSS_CHECK_ARG(workspace && workspace_bytes >= (long long)p.splits * P * Q * 4,
                         "gemm_tf32x3: workspace too small (%lld bytes, need maavss_gemm_tf32x3_workspace_bytes)", workspace_bytes);
        p.ws = (float*)workspace;
    }
    int rc;
    rc = tg3::launch<64>(A, p.a_mn, lda, B, p.b_mn, ldb, p, st);
    if (rc) return rc;
    if (p.trace) {
        long long h[48 * 8];
        cudaStreamSynchronize(st);
        cudaMemcpy(h, p.trace, sizeof(h), cudaMemcpyDeviceToHost);
        long long t0 = h[0];
        fprintf(stderr, "stage: prod_empty_seen prod_tma_issued | split_full_seen split_arrived | mma_ready_seen mma_committed | (per chunk) epi_drained tile_stored   (cycles since first)\n");
        for (int i = 0; i < 40; ++i) {
            if (!h[i * 8] && !h[i * 8 + 6]) continue;
            fprintf(stderr, "%3d: %7lld %7lld | %7lld %7lld | %7lld %7lld | %7lld %7lld\n", i, h[i * 8] - t0, h[i * 8 + 1] - t0, h[i * 8 + 2] - t0,
                    h[i * 8 + 3] - t0, h[i * 8 + 4] - t0, h[i * 8 + 5] - t0, h[i * 8 + 6] ? h[i * 8 + 6] - t0 : 0, h[i * 8 + 7] ? h[i * 8 + 7] - t0 : 0);
        }
    }
    if (p.splits > 1) {
        const long long rows = out_transposed ? Q : P, cols = out_transposed ? P : Q;
        // bias index in terms of the stored (rows, cols): per p -> rows unless transposed
        const int bias_on_cols = out_transposed ? !bias_on_q : bias_on_q;
        int blocks = (int)((rows * cols + 255) / 256);       // one element per thread while the grid fits: the sum is a latency chain
        if (blocks > 148 * 8) blocks = 148 * 8;
        tg3::finalize_kernel<<<blocks, 256, 0, st>>>(p.ws, p.splits, rows, cols, out, ldc, bias, bias_on_cols, act, slope, accumulate, p.pb, p.pb_stride);
        MAAVSS_LAUNCH_CHECK();
    }
    return MAAVSS_OK;
}
