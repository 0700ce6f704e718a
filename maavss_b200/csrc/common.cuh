// Shared helpers for libmaavss_b200 (sm_100a only).  No torch types anywhere in this library.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define MAAVSS_OK 0
#define MAAVSS_ERR_INVALID (-1)
#define MAAVSS_ERR_CUDA (-2)
#define MAAVSS_ERR_UNSUPPORTED (-3)

void maavss_set_error(const char* fmt, ...);

#define MAAVSS_CHECK_ARG(cond, ...)                  \
    do {                                             \
        if (!(cond)) {                               \
            maavss_set_error(__VA_ARGS__);           \
            return MAAVSS_ERR_INVALID;               \
        }                                            \
    } while (0)

#define MAAVSS_CUDA(call)                                                              \
    do {                                                                               \
        cudaError_t e__ = (call);                                                      \
        if (e__ != cudaSuccess) {                                                      \
            maavss_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),  \
                             __FILE__, __LINE__);                                      \
            return MAAVSS_ERR_CUDA;                                                    \
        }                                                                              \
    } while (0)

// every kernel launch of the library goes through this macro: it also feeds maavss_launch_count()
void maavss_count_launch();
#define MAAVSS_LAUNCH_CHECK()          \
    do {                               \
        maavss_count_launch();         \
        MAAVSS_CUDA(cudaGetLastError()); \
    } while (0)

static inline int maavss_num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
