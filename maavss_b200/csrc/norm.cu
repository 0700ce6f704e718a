// Train-mode BatchNorm (2-D / 3-D) fused with the activation that follows it, forward and backward
// (nn.BatchNorm2d + nn.Tanh, avse_model.py:440-441,498-499; avse_model_final.py:103-104).
// HBM-bound: x is read once for the statistics and once for the apply; the statistics are split in two halves
// so that a data-parallel caller can all-reduce the per-channel partial sums in between.
#include <stdint.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

constexpr int kBnThreads = 256;

__device__ __forceinline__ float act_apply(float v, int act, float slope) {
    switch (act) {
        case MAAVSS_ACT_TANH: return tanhf(v);
        case MAAVSS_ACT_LEAKY: return v > 0.f ? v : v * slope;
        case MAAVSS_ACT_SIGMOID: return 1.f / (1.f + expf(-v));
        default: return v;
    }
}
// derivative of the activation expressed through its OUTPUT y
__device__ __forceinline__ float act_grad(float y, int act, float slope) {
    switch (act) {
        case MAAVSS_ACT_TANH: return 1.f - y * y;
        case MAAVSS_ACT_LEAKY: return y > 0.f ? 1.f : slope;
        case MAAVSS_ACT_SIGMOID: return y * (1.f - y);
        default: return 1.f;
    }
}

__device__ __forceinline__ void block_sum2(double a, double b, double* out2) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    __shared__ double sa[kBnThreads / 32], sb[kBnThreads / 32];
    if ((threadIdx.x & 31) == 0) {
        sa[threadIdx.x >> 5] = a;
        sb[threadIdx.x >> 5] = b;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ta = 0, tb = 0;
        for (int q = 0; q < kBnThreads / 32; ++q) {
            ta += sa[q];
            tb += sb[q];
        }
        out2[0] = ta;
        out2[1] = tb;
    }
}

// grid (S, C): split s covers elements [s*chunk, (s+1)*chunk) of the flattened (n, inner) domain of channel c.
__global__ void __launch_bounds__(kBnThreads) bn_reduce_kernel(const float* __restrict__ x, int N, int C, long long inner,
                                                               double* __restrict__ partial) {
    const int c = blockIdx.y, S = gridDim.x;
    const long long total = (long long)N * inner;
    const long long chunk = (total + S - 1) / S;
    const long long e0 = (long long)blockIdx.x * chunk, e1 = min(total, e0 + chunk);
    float s1 = 0.f, s2 = 0.f;
    for (long long e = e0 + threadIdx.x; e < e1; e += kBnThreads) {
        const long long n = e / inner, i = e - n * inner;
        const float v = x[((size_t)n * C + c) * inner + i];
        s1 += v;
        s2 = fmaf(v, v, s2);
    }
    block_sum2((double)s1, (double)s2, partial + ((size_t)c * S + blockIdx.x) * 2);
}

__global__ void bn_finalize_kernel(const double* __restrict__ partial, int S, int C, double count, float eps, float momentum,
                                   float* __restrict__ running_mean, float* __restrict__ running_var,
                                   long long* __restrict__ nbt, int use_running, float* __restrict__ stats) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c == 0 && nbt && !use_running) nbt[0] += 1;
    if (c >= C) return;
    if (use_running) {
        stats[2 * c] = running_mean[c];
        stats[2 * c + 1] = (float)(1.0 / sqrt((double)running_var[c] + (double)eps));
        return;
    }
    double s1 = 0, s2 = 0;
    for (int s = 0; s < S; ++s) {
        s1 += partial[((size_t)c * S + s) * 2];
        s2 += partial[((size_t)c * S + s) * 2 + 1];
    }
    const double mean = s1 / count;
    double var = s2 / count - mean * mean;
    if (var < 0) var = 0;
    stats[2 * c] = (float)mean;
    stats[2 * c + 1] = (float)(1.0 / sqrt(var + (double)eps));
    if (running_mean) {
        const double unbiased = count > 1 ? var * count / (count - 1) : var;
        running_mean[c] = (float)((1.0 - momentum) * running_mean[c] + momentum * mean);
        running_var[c] = (float)((1.0 - momentum) * running_var[c] + momentum * unbiased);
    }
}

// grid (chunks, N*C)
__global__ void __launch_bounds__(kBnThreads) bn_apply_act_kernel(const float* __restrict__ x, const float* __restrict__ stats,
                                                                  const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                  int C, long long inner, int act, float slope,
                                                                  float* __restrict__ y) {
    const int nc = blockIdx.y, c = nc % C;
    const float mean = stats[2 * c], invstd = stats[2 * c + 1];
    const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
    const float* xp = x + (size_t)nc * inner;
    float* yp = y + (size_t)nc * inner;
    for (long long i = (long long)blockIdx.x * kBnThreads + threadIdx.x; i < inner; i += (long long)gridDim.x * kBnThreads)
        yp[i] = act_apply(fmaf((xp[i] - mean) * invstd, g, b), act, slope);
}

__global__ void __launch_bounds__(kBnThreads) bn_bwd_reduce_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                                                   const float* __restrict__ x, const float* __restrict__ stats,
                                                                   int N, int C, long long inner, int act, float slope,
                                                                   double* __restrict__ partial) {
    const int c = blockIdx.y, S = gridDim.x;
    const float mean = stats[2 * c], invstd = stats[2 * c + 1];
    const long long total = (long long)N * inner;
    const long long chunk = (total + S - 1) / S;
    const long long e0 = (long long)blockIdx.x * chunk, e1 = min(total, e0 + chunk);
    float s1 = 0.f, s2 = 0.f;
    for (long long e = e0 + threadIdx.x; e < e1; e += kBnThreads) {
        const long long n = e / inner, i = e - n * inner;
        const size_t q = ((size_t)n * C + c) * inner + i;
        const float dz = dy[q] * act_grad(y[q], act, slope);
        s1 += dz;
        s2 = fmaf(dz, (x[q] - mean) * invstd, s2);
    }
    block_sum2((double)s1, (double)s2, partial + ((size_t)c * S + blockIdx.x) * 2);
}

__global__ void bn_bwd_finalize_kernel(const double* __restrict__ partial, int S, int C, double count, int eval_mode,
                                       float* __restrict__ coef, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                       int accumulate) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double s1 = 0, s2 = 0;
    for (int s = 0; s < S; ++s) {
        s1 += partial[((size_t)c * S + s) * 2];
        s2 += partial[((size_t)c * S + s) * 2 + 1];
    }
    coef[2 * c] = eval_mode ? 0.f : (float)(s1 / count);
    coef[2 * c + 1] = eval_mode ? 0.f : (float)(s2 / count);
    if (dgamma) dgamma[c] = accumulate ? dgamma[c] + (float)s2 : (float)s2;
    if (dbeta) dbeta[c] = accumulate ? dbeta[c] + (float)s1 : (float)s1;
}

__global__ void __launch_bounds__(kBnThreads) bn_bwd_apply_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                                                  const float* __restrict__ x, const float* __restrict__ stats,
                                                                  const float* __restrict__ gamma, const float* __restrict__ coef,
                                                                  int C, long long inner, int act, float slope,
                                                                  float* __restrict__ dx) {
    const int nc = blockIdx.y, c = nc % C;
    const float mean = stats[2 * c], invstd = stats[2 * c + 1];
    const float scale = (gamma ? gamma[c] : 1.f) * invstd;
    const float c0 = coef[2 * c], c1 = coef[2 * c + 1];
    const size_t base = (size_t)nc * inner;
    for (long long i = (long long)blockIdx.x * kBnThreads + threadIdx.x; i < inner; i += (long long)gridDim.x * kBnThreads) {
        const float dz = dy[base + i] * act_grad(y[base + i], act, slope);
        const float xh = (x[base + i] - mean) * invstd;
        dx[base + i] = scale * (dz - c0 - xh * c1);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Grouped, two-launch BatchNorm: the batch is G independent groups of N/G samples with their own statistics (the
// num_seq accumulation windows of one optimizer step batched into one pass, train.py:136-159: the weights do not
// change between them).  Launch 1 = partial sums per (group, channel, split); launch 2 = apply, every CTA re-deriving
// its (group, channel) statistics from the few partials (no separate finalize launch); one designated CTA per
// channel publishes stats and applies the G running-statistics updates in window order.
// Walks this CTA's chunk of the flattened (n, i) index space of one (group, channel) in units of V floats, calling
// f(element offset).  One division per thread; afterwards (n, i) advance by the constant (q, r) = divmod(256, inner / V).
template <int V, class F>
__device__ __forceinline__ void bn_walk(int Ng, long long inner, size_t row0, int C, F&& f) {
    const long long innerV = inner / V, total = (long long)Ng * innerV;
    const long long chunk = (total + gridDim.x - 1) / gridDim.x;
    const long long e0 = (long long)blockIdx.x * chunk, e1 = min(total, e0 + chunk);
    long long e = e0 + threadIdx.x;
    if (e >= e1) return;
    long long n = e / innerV, i = e - n * innerV;
    const long long q = kBnThreads / innerV, r = kBnThreads - q * innerV;
    for (; e < e1; e += kBnThreads) {
        f((row0 + (size_t)n * C) * (size_t)inner + (size_t)i * V);
        i += r;
        n += q;
        if (i >= innerV) {
            i -= innerV;
            ++n;
        }
    }
}

__device__ __forceinline__ bool bn_vec_ok(long long inner, const void* a, const void* b, const void* c) {
    return (inner & 3) == 0 && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c)) & 15) == 0;
}

__global__ void __launch_bounds__(kBnThreads) bn_reduce_g_kernel(const float* __restrict__ x, int Ng, int C, long long inner,
                                                                 double* __restrict__ partial) {
    const int c = blockIdx.y, S = gridDim.x, g = blockIdx.z;
    const size_t row0 = (size_t)g * Ng * C + c;
    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
    if (bn_vec_ok(inner, x, x, x)) {
        bn_walk<4>(Ng, inner, row0, C, [&](size_t off) {
            const float4 v = *reinterpret_cast<const float4*>(x + off);
            s1[0] += v.x; s1[1] += v.y; s1[2] += v.z; s1[3] += v.w;
            s2[0] = fmaf(v.x, v.x, s2[0]); s2[1] = fmaf(v.y, v.y, s2[1]); s2[2] = fmaf(v.z, v.z, s2[2]); s2[3] = fmaf(v.w, v.w, s2[3]);
        });
    } else {
        bn_walk<1>(Ng, inner, row0, C, [&](size_t off) {
            const float v = x[off];
            s1[0] += v;
            s2[0] = fmaf(v, v, s2[0]);
        });
    }
    block_sum2(((double)s1[0] + (double)s1[1]) + ((double)s1[2] + (double)s1[3]),
               ((double)s2[0] + (double)s2[1]) + ((double)s2[2] + (double)s2[3]), partial + (((size_t)g * C + c) * S + blockIdx.x) * 2);
}

// Sum of the S (sum, sum-of-squares) partial pairs of one (group, channel) by the calling warp: every lane returns the totals.
__device__ __forceinline__ void warp_sum_partials(const double* __restrict__ pp, int S, double& s1, double& s2) {
    double a = 0, b = 0;
    for (int s = threadIdx.x & 31; s < S; s += 32) {
        a += pp[2 * s];
        b += pp[2 * s + 1];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    s1 = a;
    s2 = b;
}

// grid (chunks, C, G): CTA (k, c, g) normalises chunk k of the flattened (sample-in-group, inner) domain of channel c in
// group g -- the same walk as the reduce kernel, so the deep layers (inner = 32..512 floats per row) still run full CTAs.
// Warp 0 re-derives the (group, channel) statistics from the partials; CTA (0, c, 0) also applies the G running-statistics
// updates of its channel in window order.
__global__ void __launch_bounds__(kBnThreads) bn_apply_fused_kernel(const float* __restrict__ x, const double* __restrict__ partial,
                                                                    int S, int Ng, int G, double count, const float* __restrict__ gamma,
                                                                    const float* __restrict__ beta, float* __restrict__ running_mean,
                                                                    float* __restrict__ running_var, long long* __restrict__ nbt,
                                                                    float momentum, float eps, int C, long long inner, int act,
                                                                    float slope, float* __restrict__ stats, float* __restrict__ y) {
    __shared__ float sh[2];
    const int c = blockIdx.y, g = blockIdx.z;
    if (threadIdx.x < 32) {
        double s1, s2;
        warp_sum_partials(partial + ((size_t)g * C + c) * S * 2, S, s1, s2);
        const double mean = s1 / count;
        double var = s2 / count - mean * mean;
        if (var < 0) var = 0;
        if (threadIdx.x == 0) {
            sh[0] = (float)mean;
            sh[1] = (float)(1.0 / sqrt(var + (double)eps));
            if (blockIdx.x == 0) {
                stats[((size_t)g * C + c) * 2] = sh[0];
                stats[((size_t)g * C + c) * 2 + 1] = sh[1];
            }
        }
        if (blockIdx.x == 0 && g == 0) {
            if (running_mean) {
                double rm = running_mean[c], rv = running_var[c];
                for (int q = 0; q < G; ++q) {
                    warp_sum_partials(partial + ((size_t)q * C + c) * S * 2, S, s1, s2);
                    const double m2 = s1 / count;
                    double v2 = s2 / count - m2 * m2;
                    if (v2 < 0) v2 = 0;
                    const double unbiased = count > 1 ? v2 * count / (count - 1) : v2;
                    // fp32 rounding after every window, like G sequential reference updates
                    rm = (double)(float)((1.0 - momentum) * rm + momentum * m2);
                    rv = (double)(float)((1.0 - momentum) * rv + momentum * unbiased);
                }
                if (threadIdx.x == 0) {
                    running_mean[c] = (float)rm;
                    running_var[c] = (float)rv;
                }
            }
            if (threadIdx.x == 0 && c == 0 && nbt) nbt[0] += G;
        }
    }
    __syncthreads();
    const float mean = sh[0], invstd = sh[1];
    const float ga = gamma ? gamma[c] : 1.f, be = beta ? beta[c] : 0.f;
    const size_t row0 = (size_t)g * Ng * C + c;
    if (bn_vec_ok(inner, x, y, y)) {
        bn_walk<4>(Ng, inner, row0, C, [&](size_t off) {
            const float4 v = *reinterpret_cast<const float4*>(x + off);
            float4 o;
            o.x = act_apply(fmaf((v.x - mean) * invstd, ga, be), act, slope);
            o.y = act_apply(fmaf((v.y - mean) * invstd, ga, be), act, slope);
            o.z = act_apply(fmaf((v.z - mean) * invstd, ga, be), act, slope);
            o.w = act_apply(fmaf((v.w - mean) * invstd, ga, be), act, slope);
            *reinterpret_cast<float4*>(y + off) = o;
        });
    } else {
        bn_walk<1>(Ng, inner, row0, C, [&](size_t off) { y[off] = act_apply(fmaf((x[off] - mean) * invstd, ga, be), act, slope); });
    }
}

__global__ void __launch_bounds__(kBnThreads) bn_bwd_reduce_g_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                                                     const float* __restrict__ x, const float* __restrict__ stats,
                                                                     int Ng, int C, long long inner, int act, float slope,
                                                                     double* __restrict__ partial) {
    const int c = blockIdx.y, S = gridDim.x, g = blockIdx.z;
    const float mean = stats[((size_t)g * C + c) * 2], invstd = stats[((size_t)g * C + c) * 2 + 1];
    const size_t row0 = (size_t)g * Ng * C + c;
    float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
    if (bn_vec_ok(inner, dy, y, x)) {
        bn_walk<4>(Ng, inner, row0, C, [&](size_t off) {
            const float4 d4 = *reinterpret_cast<const float4*>(dy + off), y4 = *reinterpret_cast<const float4*>(y + off),
                         x4 = *reinterpret_cast<const float4*>(x + off);
            const float dv[4] = {d4.x, d4.y, d4.z, d4.w}, yv[4] = {y4.x, y4.y, y4.z, y4.w}, xv[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float dz = dv[k] * act_grad(yv[k], act, slope);
                s1[k] += dz;
                s2[k] = fmaf(dz, (xv[k] - mean) * invstd, s2[k]);
            }
        });
    } else {
        bn_walk<1>(Ng, inner, row0, C, [&](size_t off) {
            const float dz = dy[off] * act_grad(y[off], act, slope);
            s1[0] += dz;
            s2[0] = fmaf(dz, (x[off] - mean) * invstd, s2[0]);
        });
    }
    block_sum2(((double)s1[0] + (double)s1[1]) + ((double)s1[2] + (double)s1[3]),
               ((double)s2[0] + (double)s2[1]) + ((double)s2[2] + (double)s2[3]), partial + (((size_t)g * C + c) * S + blockIdx.x) * 2);
}

// grid (chunks, C, G), walk as bn_apply_fused_kernel.  CTA (0, c, 0) also emits dgamma / dbeta of its channel.
__global__ void __launch_bounds__(kBnThreads) bn_bwd_apply_fused_kernel(const float* __restrict__ dy, const float* __restrict__ y,
                                                                        const float* __restrict__ x, const float* __restrict__ stats,
                                                                        const float* __restrict__ gamma, const double* __restrict__ partial,
                                                                        const double* __restrict__ partial_local, int S, int Ng, int G,
                                                                        double count, int C, long long inner, int act,
                                                                        float slope, float* __restrict__ dgamma,
                                                                        float* __restrict__ dbeta, int accumulate, float* __restrict__ dx) {
    __shared__ float sh[2];
    const int c = blockIdx.y, g = blockIdx.z;
    if (threadIdx.x < 32) {
        double s1, s2;
        warp_sum_partials(partial + ((size_t)g * C + c) * S * 2, S, s1, s2);
        if (threadIdx.x == 0) {
            sh[0] = (float)(s1 / count);
            sh[1] = (float)(s2 / count);
        }
        if (blockIdx.x == 0 && g == 0 && (dgamma || dbeta)) {
            double t1 = 0, t2 = 0;
            for (int q = 0; q < G; ++q) {            // window order, fp32 accumulation like G sequential backward passes
                warp_sum_partials(partial_local + ((size_t)q * C + c) * S * 2, S, s1, s2);
                t1 = (double)(float)(t1 + (float)s1);
                t2 = (double)(float)(t2 + (float)s2);
            }
            if (threadIdx.x == 0) {
                if (dgamma) dgamma[c] = accumulate ? dgamma[c] + (float)t2 : (float)t2;
                if (dbeta) dbeta[c] = accumulate ? dbeta[c] + (float)t1 : (float)t1;
            }
        }
    }
    __syncthreads();
    if (!dx) return;
    const float mean = stats[((size_t)g * C + c) * 2], invstd = stats[((size_t)g * C + c) * 2 + 1];
    const float scale = (gamma ? gamma[c] : 1.f) * invstd;
    const float c0 = sh[0], c1 = sh[1];
    const size_t row0 = (size_t)g * Ng * C + c;
    if (bn_vec_ok(inner, dy, y, x) && bn_vec_ok(inner, dx, dx, dx)) {
        bn_walk<4>(Ng, inner, row0, C, [&](size_t off) {
            const float4 d4 = *reinterpret_cast<const float4*>(dy + off), y4 = *reinterpret_cast<const float4*>(y + off),
                         x4 = *reinterpret_cast<const float4*>(x + off);
            const float dv[4] = {d4.x, d4.y, d4.z, d4.w}, yv[4] = {y4.x, y4.y, y4.z, y4.w}, xv[4] = {x4.x, x4.y, x4.z, x4.w};
            float o[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float dz = dv[k] * act_grad(yv[k], act, slope);
                const float xh = (xv[k] - mean) * invstd;
                o[k] = scale * (dz - c0 - xh * c1);
            }
            *reinterpret_cast<float4*>(dx + off) = make_float4(o[0], o[1], o[2], o[3]);
        });
    } else {
        bn_walk<1>(Ng, inner, row0, C, [&](size_t off) {
            const float dz = dy[off] * act_grad(y[off], act, slope);
            const float xh = (x[off] - mean) * invstd;
            dx[off] = scale * (dz - c0 - xh * c1);
        });
    }
}

__global__ void act_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long n, int act, float slope) {
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x)
        y[q] = act_apply(x[q], act, slope);
}
__global__ void act_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ dz, long long n,
                               int act, float slope) {
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x)
        dz[q] = dy[q] * act_grad(y[q], act, slope);
}

inline int ew_grid(long long n) {
    long long b = (n + 1023) / 1024;
    if (b > 148 * 8) b = 148 * 8;
    return (int)(b < 1 ? 1 : b);
}
// chunks of one (group, channel) for the grouped apply kernels: >= 2048 elements per CTA, at most ~8 CTAs per SM over the grid
inline int group_chunks(long long per_gc, long long gc) {
    long long b = (per_gc + 2047) / 2048;
    const long long cap = (148LL * 8 + gc - 1) / gc;
    if (b > cap) b = cap;
    return (int)(b < 1 ? 1 : b);
}
inline int inner_chunks(long long inner, long long rows) {
    long long b = (inner + kBnThreads * 4 - 1) / (kBnThreads * 4);
    const long long cap = (148LL * 8 + rows - 1) / rows;
    if (b > cap) b = cap;
    return (int)(b < 1 ? 1 : b);
}

}  // namespace

extern "C" int maavss_bn_splits(int N, int C, long long inner) {
    if (C <= 0) return 1;
    long long s = (296 + C - 1) / C;
    const long long cap = ((long long)N * inner + 2047) / 2048;
    if (s > cap) s = cap;
    if (s > 64) s = 64;
    return (int)(s < 1 ? 1 : s);
}

extern "C" int maavss_bn_reduce(const float* x, int N, int C, long long inner, double* partial, void* stream) {
    MAAVSS_CHECK_ARG(partial && C > 0 && C <= 65535, "bn_reduce: bad argument");
    dim3 grid(maavss_bn_splits(N, C, inner), C);
    bn_reduce_kernel<<<grid, kBnThreads, 0, (cudaStream_t)stream>>>(x, N, C, inner, partial);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bn_finalize(const double* partial, int S, int C, double count, float eps, float momentum,
                                  float* running_mean, float* running_var, long long* num_batches_tracked, int use_running,
                                  float* stats, void* stream) {
    MAAVSS_CHECK_ARG(stats && C > 0, "bn_finalize: bad argument");
    MAAVSS_CHECK_ARG(use_running ? (running_mean && running_var) : (partial != nullptr), "bn_finalize: missing statistics source");
    bn_finalize_kernel<<<ceil_div(C, 128), 128, 0, (cudaStream_t)stream>>>(partial, S, C, count, eps, momentum, running_mean,
                                                                          running_var, num_batches_tracked, use_running, stats);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bn_apply_act(const float* x, const float* stats, const float* gamma, const float* beta, int N, int C,
                                   long long inner, int act, float slope, float* y, void* stream) {
    if ((long long)N * C * inner == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && stats && y, "bn_apply_act: null argument");
    MAAVSS_CHECK_ARG((long long)N * C <= 65535, "bn_apply_act: N*C > 65535");
    dim3 grid(inner_chunks(inner, (long long)N * C), N * C);
    bn_apply_act_kernel<<<grid, kBnThreads, 0, (cudaStream_t)stream>>>(x, stats, gamma, beta, C, inner, act, slope, y);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bn_bwd_reduce(const float* dy, const float* y, const float* x, const float* stats, int N, int C,
                                    long long inner, int act, float slope, double* partial, void* stream) {
    MAAVSS_CHECK_ARG(dy && y && x && stats && partial && C > 0 && C <= 65535, "bn_bwd_reduce: bad argument");
    dim3 grid(maavss_bn_splits(N, C, inner), C);
    bn_bwd_reduce_kernel<<<grid, kBnThreads, 0, (cudaStream_t)stream>>>(dy, y, x, stats, N, C, inner, act, slope, partial);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bn_bwd_finalize(const double* partial, int S, int C, double count, int eval_mode, float* coef,
                                      float* dgamma, float* dbeta, int accumulate, void* stream) {
    MAAVSS_CHECK_ARG(partial && coef && C > 0, "bn_bwd_finalize: bad argument");
    bn_bwd_finalize_kernel<<<ceil_div(C, 128), 128, 0, (cudaStream_t)stream>>>(partial, S, C, count, eval_mode, coef, dgamma,
                                                                              dbeta, accumulate);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bn_bwd_apply(const float* dy, const float* y, const float* x, const float* stats, const float* gamma,
                                   const float* coef, int N, int C, long long inner, int act, float slope, float* dx,
                                   void* stream) {
    if ((long long)N * C * inner == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dy && y && x && stats && coef && dx, "bn_bwd_apply: null argument");
    MAAVSS_CHECK_ARG((long long)N * C <= 65535, "bn_bwd_apply: N*C > 65535");
    dim3 grid(inner_chunks(inner, (long long)N * C), N * C);
    bn_bwd_apply_kernel<<<grid, kBnThreads, 0, (cudaStream_t)stream>>>(dy, y, x, stats, gamma, coef, C, inner, act, slope, dx);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bn_reduce_grouped(const float* x, int N, int C, long long inner, int groups, double* partial, void* stream) {
    if ((long long)N * C * inner == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && partial && C > 0 && C <= 65535, "bn_reduce_grouped: bad argument");
    MAAVSS_CHECK_ARG(groups >= 1 && groups <= 65535 && N % groups == 0, "bn_reduce_grouped: batch %d is not a multiple of %d groups", N, groups);
    const int Ng = N / groups, S = maavss_bn_splits(Ng, C, inner);
    bn_reduce_g_kernel<<<dim3(S, C, groups), kBnThreads, 0, (cudaStream_t)stream>>>(x, Ng, C, inner, partial);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bn_apply_grouped(const float* x, const double* partial, int N, int C, long long inner, int groups, double count,
                                       const float* gamma, const float* beta, float* running_mean, float* running_var,
                                       long long* num_batches_tracked, float momentum, float eps, int act, float slope, float* stats,
                                       float* y, void* stream) {
    if ((long long)N * C * inner == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && partial && stats && y && C > 0, "bn_apply_grouped: bad argument");
    MAAVSS_CHECK_ARG(groups >= 1 && groups <= 65535 && C <= 65535 && N % groups == 0 && count > 0, "bn_apply_grouped: bad batch / groups");
    const int Ng = N / groups, S = maavss_bn_splits(Ng, C, inner);
    dim3 grid(group_chunks((long long)Ng * inner, (long long)C * groups), C, groups);
    bn_apply_fused_kernel<<<grid, kBnThreads, 0, (cudaStream_t)stream>>>(x, partial, S, Ng, groups, count, gamma, beta, running_mean,
                                                                        running_var, num_batches_tracked, momentum, eps, C, inner,
                                                                        act, slope, stats, y);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bn_bwd_reduce_grouped(const float* dy, const float* y, const float* x, const float* stats, int N, int C,
                                            long long inner, int groups, int act, float slope, double* partial, void* stream) {
    if ((long long)N * C * inner == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dy && y && x && stats && partial && C > 0 && C <= 65535, "bn_bwd_reduce_grouped: bad argument");
    MAAVSS_CHECK_ARG(groups >= 1 && groups <= 65535 && N % groups == 0, "bn_bwd_reduce_grouped: bad batch / groups");
    const int Ng = N / groups, S = maavss_bn_splits(Ng, C, inner);
    bn_bwd_reduce_g_kernel<<<dim3(S, C, groups), kBnThreads, 0, (cudaStream_t)stream>>>(dy, y, x, stats, Ng, C, inner, act, slope, partial);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_bn_bwd_apply_grouped(const float* dy, const float* y, const float* x, const float* stats, const float* gamma,
                                           const double* partial, const double* partial_local, int N, int C, long long inner,
                                           int groups, double count, int act, float slope, float* dgamma, float* dbeta, int accumulate,
                                           float* dx, void* stream) {
    if ((long long)N * C * inner == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dy && y && x && stats && partial && partial_local && C > 0, "bn_bwd_apply_grouped: bad argument");
    MAAVSS_CHECK_ARG(groups >= 1 && groups <= 65535 && C <= 65535 && N % groups == 0 && count > 0, "bn_bwd_apply_grouped: bad batch / groups");
    const int Ng = N / groups, S = maavss_bn_splits(Ng, C, inner);
    dim3 grid(group_chunks((long long)Ng * inner, (long long)C * groups), C, groups);
    bn_bwd_apply_fused_kernel<<<grid, kBnThreads, 0, (cudaStream_t)stream>>>(dy, y, x, stats, gamma, partial, partial_local, S, Ng, groups,
                                                                            count, C, inner, act, slope, dgamma, dbeta, accumulate, dx);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_act_fwd(const float* x, float* y, long long n, int act, float slope, void* stream) {
    if (n == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && y, "act_fwd: null argument");
    act_fwd_kernel<<<ew_grid(n), 256, 0, (cudaStream_t)stream>>>(x, y, n, act, slope);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_act_bwd(const float* dy, const float* y, float* dz, long long n, int act, float slope, void* stream) {
    if (n == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dy && y && dz, "act_bwd: null argument");
    act_bwd_kernel<<<ew_grid(n), 256, 0, (cudaStream_t)stream>>>(dy, y, dz, n, act, slope);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
