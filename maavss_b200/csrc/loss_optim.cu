// nn.MSELoss forward/backward and the Adam step over a flat fp32 parameter arena
// (train.py:154-162, train_avse_frames.py:166-181).  All HBM-bound streaming kernels.
#include <math.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

constexpr int kMseBlocks = 128;   // partial sums: the workspace is maavss_reduce_workspace_bytes(8) = 128 doubles

__global__ void __launch_bounds__(256) mse_partial_kernel(const float* __restrict__ a, const float* __restrict__ b, long long n,
                                                          double* __restrict__ ws) {
    float s = 0.f;
    for (long long q = (long long)blockIdx.x * 256 + threadIdx.x; q < n; q += (long long)gridDim.x * 256) {
        const float d = a[q] - b[q];
        s = fmaf(d, d, s);
    }
    double v = s;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __shared__ double sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int q = 0; q < 8; ++q) t += sm[q];
        ws[blockIdx.x] = t;
    }
}

__global__ void mse_final_kernel(const double* __restrict__ ws, int blocks, long long n, float* __restrict__ loss) {
    if (threadIdx.x == 0) {
        double t = 0;
        for (int q = 0; q < blocks; ++q) t += ws[q];
        loss[0] = (float)(t / (double)n);
    }
}

__global__ void __launch_bounds__(256) mse_bwd_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                      const float* __restrict__ gout, long long n, float* __restrict__ da,
                                                      int accumulate) {
    const float k = gout[0] * (2.0f / (float)n);
    for (long long q = (long long)blockIdx.x * 256 + threadIdx.x; q < n; q += (long long)gridDim.x * 256) {
        const float g = k * (a[q] - b[q]);
        da[q] = accumulate ? da[q] + g : g;
    }
}

// Step counter and bias corrections live on the device so that a captured CUDA graph of the whole train step
// advances them on every replay.
struct AdamState {
    long long step;
    float step_size, inv_bc2_sqrt;
};

__global__ void adam_prepare_kernel(AdamState* st, float lr, float beta1, float beta2) {
    const long long step = st->step + 1;
    st->step = step;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    st->step_size = (float)((double)lr / bc1);
    st->inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
}

// torch.optim.Adam (no amsgrad, coupled weight decay): 4 reads + 3 writes of 4 B per parameter.
__global__ void __launch_bounds__(256) adam_kernel(float4* __restrict__ p, const float4* __restrict__ g, float4* __restrict__ m,
                                                   float4* __restrict__ v, long long n4, float beta1, float beta2, float eps,
                                                   float wd, const AdamState* __restrict__ st, float gscale, int zero_grad) {
    const float step_size = st->step_size, inv_bc2_sqrt = st->inv_bc2_sqrt;
    float4* gw = const_cast<float4*>(g);
    for (long long q = (long long)blockIdx.x * 256 + threadIdx.x; q < n4; q += (long long)gridDim.x * 256) {
        // streaming accesses (evict-first): 32 B/param pass through once per step and must not push the activations of the
        // concurrently running backward pass out of L2
        float4 pp = __ldcs(p + q), gg = __ldcs(g + q), mm = __ldcs(m + q), vv = __ldcs(v + q);
        float* P = &pp.x;
        float* G = &gg.x;
        float* Mv = &mm.x;
        float* V = &vv.x;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float gr = G[c] * gscale;
            if (wd != 0.f) gr = fmaf(wd, P[c], gr);
            Mv[c] = fmaf(beta1, Mv[c], (1.f - beta1) * gr);          // exp_avg.lerp_(grad, 1 - beta1)
            V[c] = fmaf(beta2, V[c], (1.f - beta2) * gr * gr);        // exp_avg_sq.mul_(beta2).addcmul_(g, g, 1 - beta2)
            const float denom = sqrtf(V[c]) * inv_bc2_sqrt + eps;
            P[c] -= step_size * (Mv[c] / denom);
        }
        __stcs(p + q, pp);
        __stcs(m + q, mm);
        __stcs(v + q, vv);
        if (zero_grad) __stcs(gw + q, make_float4(0.f, 0.f, 0.f, 0.f));       // optimizer.zero_grad() fused into the same pass
    }
}

__global__ void adam_tail_kernel(float* p, const float* g, float* m, float* v, long long start, long long n, float beta1,
                                 float beta2, float eps, float wd, const AdamState* __restrict__ st, float gscale, int zero_grad) {
    const float step_size = st->step_size, inv_bc2_sqrt = st->inv_bc2_sqrt;
    const long long q = start + threadIdx.x;
    if (q >= n) return;
    float gr = g[q] * gscale;
    if (wd != 0.f) gr = fmaf(wd, p[q], gr);
    const float mm = fmaf(beta1, m[q], (1.f - beta1) * gr);
    const float vv = fmaf(beta2, v[q], (1.f - beta2) * gr * gr);
    m[q] = mm;
    v[q] = vv;
    p[q] -= step_size * (mm / (sqrtf(vv) * inv_bc2_sqrt + eps));
    if (zero_grad) const_cast<float*>(g)[q] = 0.f;
}

__global__ void fill_kernel(float* __restrict__ x, long long n, float value) {
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) x[q] = value;
}

inline int stream_grid(long long n, int per_thread) {
    long long b = (n + 256LL * per_thread - 1) / (256LL * per_thread);
    if (b > 148 * 8) b = 148 * 8;
    return (int)(b < 1 ? 1 : b);
}

}  // namespace

extern "C" int maavss_mse_fwd(const float* a, const float* b, long long n, float* loss, void* workspace, void* stream) {
    MAAVSS_CHECK_ARG(a && b && loss && workspace && n > 0, "mse_fwd: bad argument");
    int blocks = stream_grid(n, 8);
    if (blocks > kMseBlocks) blocks = kMseBlocks;
    cudaStream_t st = (cudaStream_t)stream;
    mse_partial_kernel<<<blocks, 256, 0, st>>>(a, b, n, (double*)workspace);
    MAAVSS_LAUNCH_CHECK();
    mse_final_kernel<<<1, 32, 0, st>>>((const double*)workspace, blocks, n, loss);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_mse_bwd(const float* a, const float* b, const float* gout, long long n, float* da, int accumulate,
                              void* stream) {
    MAAVSS_CHECK_ARG(a && b && gout && da && n > 0, "mse_bwd: bad argument");
    mse_bwd_kernel<<<stream_grid(n, 4), 256, 0, (cudaStream_t)stream>>>(a, b, gout, n, da, accumulate);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_adam_prepare(void* state, float lr, float beta1, float beta2, void* stream) {
    MAAVSS_CHECK_ARG(state, "adam_prepare: null state");
    adam_prepare_kernel<<<1, 1, 0, (cudaStream_t)stream>>>((AdamState*)state, lr, beta1, beta2);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_adam_apply(float* p, float* g, float* m, float* v, long long n, float beta1, float beta2, float eps,
                                 float weight_decay, const void* state, float grad_scale, int zero_grad, int max_ctas, void* stream) {
    if (n == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(p && g && m && v && state, "adam_apply: bad argument");
    MAAVSS_CHECK_ARG(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) % 16 == 0, "adam_apply: arenas must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const AdamState* as = (const AdamState*)state;
    const long long n4 = n / 4;
    if (n4 > 0) {
        int grid = stream_grid(n4, 2);
        if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
        adam_kernel<<<grid, 256, 0, st>>>((float4*)p, (const float4*)g, (float4*)m, (float4*)v, n4, beta1, beta2, eps, weight_decay, as,
                                         grad_scale, zero_grad);
        MAAVSS_LAUNCH_CHECK();
    }
    if (n4 * 4 < n) {
        adam_tail_kernel<<<1, 32, 0, st>>>(p, g, m, v, n4 * 4, n, beta1, beta2, eps, weight_decay, as, grad_scale, zero_grad);
        MAAVSS_LAUNCH_CHECK();
    }
    return MAAVSS_OK;
}

extern "C" int maavss_adam_step(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1, float beta2,
                                float eps, float weight_decay, void* state, float grad_scale, void* stream) {
    if (n == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(p && g && m && v && state, "adam_step: bad argument");
    int rc = maavss_adam_prepare(state, lr, beta1, beta2, stream);
    if (rc) return rc;
    return maavss_adam_apply(p, const_cast<float*>(g), m, v, n, beta1, beta2, eps, weight_decay, state, grad_scale, 0, 0, stream);
}

extern "C" int maavss_fill(float* x, long long n, float value, void* stream) {
    if (n == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x, "fill: null argument");
    fill_kernel<<<stream_grid(n, 4), 256, 0, (cudaStream_t)stream>>>(x, n, value);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
