// Attention-map cache kernels: the raw uint8 / fp16 frame store that replaces the reference's
// per-frame JPEG cache (av_dataset.py:251-278 reader, save_attn_videos.py:10-19 writer).
//   * pack:  fp32 maps -> uint8 with torchvision.utils.save_image's rounding (x*255 + 0.5, clamp, truncate) or fp16 (rne)
//   * load:  stored frames -> fp32 (uint8 / 255 as ToTensor; fp16 widened), temporal diff with a zero first frame and
//            the per-clip signed-max normalise of av_dataset.py:323-328, without an fp32 intermediate in HBM.
// HBM-bound elementwise work: 4 pixels per thread (float4 / uchar4 / 2 x half2), grid-stride, one atomicMax per CTA;
// the fp16 load runs both passes in one launch, one thread-block cluster per clip (cache_clip_fused_kernel).
#include <cooperative_groups.h>
#include <cuda_fp16.h>

#include <cstdlib>

#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace cg = cooperative_groups;

namespace {

__device__ __forceinline__ unsigned int float_to_ordered(float f) {
    const unsigned int b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(unsigned int k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// save_image: grid.mul(255).add_(0.5).clamp_(0, 255).to(uint8) -- two separately rounded fp32 operations (no FMA
// contraction, or .5 ties land on the other side), then truncation.
__device__ __forceinline__ unsigned char pack_u8(float x) {
    float v = __fadd_rn(__fmul_rn(x, 255.0f), 0.5f);
    v = fminf(fmaxf(v, 0.0f), 255.0f);     // NaN -> 0 (fmaxf returns the non-NaN operand; torch's clamp would keep NaN)
    return (unsigned char)(int)v;
}

// ToTensor's k / 255 (correctly rounded) in three issue slots instead of the division's slow path: q = k * fl(1/255) is
// within one ulp, the exact remainder k - 255 q corrects it (Markstein).  Equal to __fdiv_rn(k, 255) for every byte --
// checked exhaustively on the host (tests/test_attn_cache_host.py) and on the device (test_load_matches_oracle).
// The byte becomes a float without I2F (a quarter-rate pipe: 2 conversions per pixel made the uint8 load issue bound):
// PRMT drops it into the mantissa of 2^23, one exact subtraction removes the 2^23.
__device__ __forceinline__ float u8_over_255(unsigned int word, unsigned int sel = 0x7540u) {
    const float r = 0.0039215688593685627f;
    const float k = __fsub_rn(__uint_as_float(__byte_perm(word, 0x4B000000u, sel)), 8388608.0f);
    const float q = __fmul_rn(k, r);
    return __fmaf_rn(__fmaf_rn(-q, 255.0f, k), r, q);
}

struct U8 {
    typedef unsigned char T;
    static __device__ __forceinline__ float ld(const T* p, size_t i) { return u8_over_255(p[i]); }
    static __device__ __forceinline__ float4 ld4(const T* p, size_t i) { return cvt4(ldraw(p, i)); }
    typedef uchar4 Raw;                                                       // 4 stored pixels
    static __device__ __forceinline__ Raw ldraw(const T* p, size_t i) { return *reinterpret_cast<const uchar4*>(p + i); }
    static __device__ __forceinline__ float4 cvt4(Raw v) {
        const unsigned int w = *reinterpret_cast<const unsigned int*>(&v);
        return make_float4(u8_over_255(w, 0x7540u), u8_over_255(w, 0x7541u), u8_over_255(w, 0x7542u), u8_over_255(w, 0x7543u));
    }
};
struct F16 {
    typedef __half T;
    static __device__ __forceinline__ float ld(const T* p, size_t i) { return __half2float(p[i]); }
    static __device__ __forceinline__ float4 ld4(const T* p, size_t i) { return cvt4(ldraw(p, i)); }
    typedef uint2 Raw;
    static __device__ __forceinline__ Raw ldraw(const T* p, size_t i) { return *reinterpret_cast<const uint2*>(p + i); }
    static __device__ __forceinline__ float4 cvt4(Raw raw) {
        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
        const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
        return make_float4(a.x, a.y, b.x, b.y);
    }
};

__global__ void pack_u8_kernel(const float* __restrict__ in, unsigned char* __restrict__ out, long long n) {
    const long long n4 = n >> 2, stride = (long long)gridDim.x * blockDim.x;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += stride) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(in) + q);
        reinterpret_cast<uchar4*>(out)[q] = make_uchar4(pack_u8(v.x), pack_u8(v.y), pack_u8(v.z), pack_u8(v.w));
    }
    for (long long q = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += stride) out[q] = pack_u8(in[q]);
}

__global__ void pack_f16_kernel(const float* __restrict__ in, __half* __restrict__ out, long long n) {
    const long long n4 = n >> 2, stride = (long long)gridDim.x * blockDim.x;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += stride) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(in) + q);
        const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
        uint2 raw;
        raw.x = *reinterpret_cast<const unsigned int*>(&a);
        raw.y = *reinterpret_cast<const unsigned int*>(&b);
        reinterpret_cast<uint2*>(out)[q] = raw;
    }
    for (long long q = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += stride)
        out[q] = __float2half_rn(in[q]);
}

// value of the (optionally differenced) clip at frame t, pixels [r, r+4)
template <class S>
__device__ __forceinline__ float4 diff4(const typename S::T* __restrict__ clip, long long hw, int t, long long r, int diff) {
    const float4 c = S::ld4(clip, (size_t)t * hw + r);
    if (!diff) return c;
    if (t == 0) return make_float4(0.f, 0.f, 0.f, 0.f);       // the zeros_like(attn[0:1]) pad
    const float4 p = S::ld4(clip, (size_t)(t - 1) * hw + r);
    return make_float4(c.x - p.x, c.y - p.y, c.z - p.z, c.w - p.w);
}
template <class S>
__device__ __forceinline__ float diff1(const typename S::T* __restrict__ clip, long long hw, int t, long long r, int diff) {
    const float c = S::ld(clip, (size_t)t * hw + r);
    if (!diff) return c;
    if (t == 0) return 0.f;
    return c - S::ld(clip, (size_t)(t - 1) * hw + r);
}

// grid (x, clips).  VEC: hw % 4 == 0 and the base pointers are 4-element aligned.
template <class S, bool VEC, bool WRITE>
__global__ void __launch_bounds__(256) cache_clip_kernel(const typename S::T* __restrict__ in, float* __restrict__ out, int T,
                                                         long long hw, int diff, unsigned int* __restrict__ ws) {
    const typename S::T* clip = in + (size_t)blockIdx.y * T * hw;
    float* o = WRITE ? out + (size_t)blockIdx.y * T * hw : nullptr;
    const float inv = WRITE ? 1.0f / ordered_to_float(ws[blockIdx.y]) : 0.f;   // `attn *= 1/torch.max(attn)`: signed max
    float m = -INFINITY;
    const long long stride = (long long)gridDim.x * blockDim.x, start = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC) {
        const long long hw4 = hw >> 2, total4 = (long long)T * hw4;
        for (long long q = start; q < total4; q += stride) {
            const int t = (int)(q / hw4);
            const long long r = (q - (long long)t * hw4) << 2;
            const float4 v = diff4<S>(clip, hw, t, r, diff);
            if (WRITE) *reinterpret_cast<float4*>(o + (size_t)t * hw + r) = make_float4(v.x * inv, v.y * inv, v.z * inv, v.w * inv);
            else m = fmaxf(fmaxf(m, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
        }
    } else {
        const long long total = (long long)T * hw;
        for (long long q = start; q < total; q += stride) {
            const int t = (int)(q / hw);
            const float v = diff1<S>(clip, hw, t, q - (long long)t * hw, diff);
            if (WRITE) o[q] = v * inv;
            else m = fmaxf(m, v);
        }
    }
    if (!WRITE) {
        m = warp_max(m);
        __shared__ float sm[8];
        if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
        __syncthreads();
        if (threadIdx.x < 32) {
            m = threadIdx.x < 8 ? sm[threadIdx.x] : -INFINITY;
            m = warp_max(m);
            if (threadIdx.x == 0) atomicMax(ws + blockIdx.y, float_to_ordered(m));
        }
    }
}

// One thread-block cluster per clip (grid (CL, clips), cluster (CL, 1, 1)): both passes in one launch.  A thread owns
// FUSED_U pixel quads and walks the frames with the previous frame's values in registers (one conversion per pixel and
// pass, the diff costs no second load); the CTA maxima meet through distributed shared memory, so the second read of
// the clip's stored bytes (0.8 / 1.6 MB) comes from L2 while it is still there -- HBM sees s + 4 bytes per pixel.
constexpr int FUSED_U = 4;
constexpr int FUSED_THREADS = 256;

template <class S, bool WRITE>
__device__ __forceinline__ float fused_pass(const typename S::T* __restrict__ clip, float* __restrict__ o, int T, long long hw,
                                            int diff, float inv, long long first, long long nthreads) {
    typedef typename S::Raw Raw;
    const long long hw4 = hw >> 2;
    float m = -INFINITY;
    for (long long q0 = first; q0 < hw4; q0 += nthreads * FUSED_U) {
        // stored bytes of frame t + 1 are in flight while frame t is converted; the previous frame is kept as stored
        // bytes too (converted again: the issue slots are free, the registers are not)
        Raw nxt[FUSED_U], prv[FUSED_U];
#pragma unroll
        for (int u = 0; u < FUSED_U; ++u) {
            const long long q = q0 + u * nthreads;
            if (q < hw4) nxt[u] = S::ldraw(clip, (size_t)(q << 2));
        }
        for (int t = 0; t < T; ++t) {
            Raw cur[FUSED_U];
#pragma unroll
            for (int u = 0; u < FUSED_U; ++u) {
                cur[u] = nxt[u];
                const long long q = q0 + u * nthreads;
                if (t + 1 < T && q < hw4) nxt[u] = S::ldraw(clip, (size_t)(t + 1) * hw + (q << 2));
            }
#pragma unroll
            for (int u = 0; u < FUSED_U; ++u) {
                const long long q = q0 + u * nthreads;
                if (q >= hw4) continue;
                float4 v = S::cvt4(cur[u]);
                if (diff) {
                    if (t == 0) {
                        v = make_float4(0.f, 0.f, 0.f, 0.f);               // the zeros_like(attn[0:1]) pad
                    } else {
                        const float4 p = S::cvt4(prv[u]);
                        v = make_float4(v.x - p.x, v.y - p.y, v.z - p.z, v.w - p.w);
                    }
                    prv[u] = cur[u];
                }
                if (WRITE)      // streaming store: the fp32 output must not evict other clips' stored bytes from L2
                    __stcs(reinterpret_cast<float4*>(o + (size_t)t * hw + (q << 2)), make_float4(v.x * inv, v.y * inv, v.z * inv, v.w * inv));
                else
                    m = fmaxf(fmaxf(m, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
            }
        }
    }
    return m;
}

template <class S>
__global__ void __launch_bounds__(FUSED_THREADS, 4) cache_clip_fused_kernel(const typename S::T* __restrict__ in, float* __restrict__ out,
                                                                         int T, long long hw, int diff) {
    cg::cluster_group cluster = cg::this_cluster();
    const int ncta = gridDim.x;                                   // = cluster size: blockIdx.x is the rank in the cluster
    const typename S::T* clip = in + (size_t)blockIdx.y * T * hw;
    float* o = out + (size_t)blockIdx.y * T * hw;
    const long long first = (long long)blockIdx.x * FUSED_THREADS + threadIdx.x, nthreads = (long long)ncta * FUSED_THREADS;
    __shared__ float warp_m[FUSED_THREADS / 32];
    __shared__ float cta_m;
    float m = warp_max(fused_pass<S, false>(clip, nullptr, T, hw, diff, 0.f, first, nthreads));
    if ((threadIdx.x & 31) == 0) warp_m[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        m = warp_m[0];
#pragma unroll
        for (int w = 1; w < FUSED_THREADS / 32; ++w) m = fmaxf(m, warp_m[w]);
        cta_m = m;
    }
    cluster.sync();
    m = -INFINITY;
    for (int r = 0; r < ncta; ++r) m = fmaxf(m, *cluster.map_shared_rank(&cta_m, r));
    cluster.sync();                                               // no CTA leaves while a peer still reads its cta_m
    const float inv = 1.0f / m;                                   // `attn *= 1/torch.max(attn)`: signed max
    fused_pass<S, true>(clip, o, T, hw, diff, inv, first, nthreads);
}

// The one-launch cluster kernel serves the fp16 store (0.29 ms against 0.46-0.51 ms for 256 clips); the uint8 store
// stays on the two-launch path below, which is issue bound on the dequantiser either way and measured 0.48 / 0.56 ms
// against 0.52 / 0.61 ms fused.  MAAVSS_ATTN_CACHE_FUSED=1 / =0 forces one path for both formats (the parity tests run
// under each); clips whose frames are not a multiple of 4 pixels or are misaligned always take the two-launch path.
int fused_mode() {
    static const int mode = [] { const char* e = getenv("MAAVSS_ATTN_CACHE_FUSED"); return e ? (atoi(e) != 0) : -1; }();
    return mode;
}

int grid_for(long long items, int clips) {
    long long want = (items + 255) / 256;
    const long long cap = max(1, (maavss_num_sms() * 8 + clips - 1) / max(clips, 1));   // ~8 CTAs per SM over all clips
    if (want > cap) want = cap;
    return (int)max(1LL, want);
}

template <class S>
int launch_load(const void* stored, float* out, int clips, int frames, long long hw, int diff, void* workspace, cudaStream_t st) {
    const typename S::T* in = static_cast<const typename S::T*>(stored);
    const bool vec = (hw & 3) == 0 && (reinterpret_cast<uintptr_t>(in) % (4 * sizeof(typename S::T))) == 0 &&
                     (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    if (vec && (fused_mode() == 1 || (fused_mode() < 0 && sizeof(typename S::T) == 2))) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(8, clips);
        cfg.blockDim = dim3(FUSED_THREADS);
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 8;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        MAAVSS_CUDA(cudaLaunchKernelEx(&cfg, cache_clip_fused_kernel<S>, in, out, frames, hw, diff));
        maavss_count_launch();
        return MAAVSS_OK;
    }
    MAAVSS_CUDA(cudaMemsetAsync(workspace, 0, sizeof(unsigned int) * clips, st));   // 0 orders below every float
    const long long items = (long long)frames * (vec ? hw >> 2 : hw);
    dim3 grid(grid_for(items, clips), clips);
    unsigned int* ws = (unsigned int*)workspace;
    if (vec) {
        cache_clip_kernel<S, true, false><<<grid, 256, 0, st>>>(in, out, frames, hw, diff, ws);
        MAAVSS_LAUNCH_CHECK();
        cache_clip_kernel<S, true, true><<<grid, 256, 0, st>>>(in, out, frames, hw, diff, ws);
    } else {
        cache_clip_kernel<S, false, false><<<grid, 256, 0, st>>>(in, out, frames, hw, diff, ws);
        MAAVSS_LAUNCH_CHECK();
        cache_clip_kernel<S, false, true><<<grid, 256, 0, st>>>(in, out, frames, hw, diff, ws);
    }
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

}  // namespace

extern "C" int maavss_attn_cache_pack(const float* maps, void* stored, long long n, int format, void* stream) {
    MAAVSS_CHECK_ARG(format == MAAVSS_CACHE_U8 || format == MAAVSS_CACHE_F16, "attn_cache_pack: unknown format %d", format);
    MAAVSS_CHECK_ARG(n >= 0, "attn_cache_pack: n=%lld", n);
    if (n == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(maps && stored, "attn_cache_pack: null argument");
    MAAVSS_CHECK_ARG((reinterpret_cast<uintptr_t>(maps) & 15) == 0 && (reinterpret_cast<uintptr_t>(stored) & 7) == 0,
                     "attn_cache_pack: maps must be 16-byte and stored 8-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = grid_for((n + 3) >> 2, 1);
    if (format == MAAVSS_CACHE_U8) pack_u8_kernel<<<grid, 256, 0, st>>>(maps, (unsigned char*)stored, n);
    else pack_f16_kernel<<<grid, 256, 0, st>>>(maps, (__half*)stored, n);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_attn_cache_load(const void* stored, int format, float* out, int clips, int frames, long long hw, int diff,
                                      void* workspace, void* stream) {
    MAAVSS_CHECK_ARG(format == MAAVSS_CACHE_U8 || format == MAAVSS_CACHE_F16, "attn_cache_load: unknown format %d", format);
    MAAVSS_CHECK_ARG(clips >= 0 && clips <= 65535, "attn_cache_load: clips=%d out of range", clips);
    MAAVSS_CHECK_ARG(frames >= 0 && hw >= 0, "attn_cache_load: negative size");
    if (clips == 0 || frames == 0 || hw == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(stored && out && workspace, "attn_cache_load: null argument");
    cudaStream_t st = (cudaStream_t)stream;
    return format == MAAVSS_CACHE_U8 ? launch_load<U8>(stored, out, clips, frames, hw, diff, workspace, st)
                                     : launch_load<F16>(stored, out, clips, frames, hw, diff, workspace, st);
}
