// Minimal host shim so the per-thread phase functions in maavss_b200/csrc/*_phases.cuh compile
// with g++ and can be driven thread-by-thread on the CPU (tests/test_emulation.py).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __global__
struct float2 { float x, y; };
struct alignas(16) float4 { float x, y, z, w; };
static inline float2 make_float2(float x, float y) { return float2{x, y}; }
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
static inline void sincosf_emu(float a, float* s, float* c) { *s = sinf(a); *c = cosf(a); }
#define sincosf sincosf_emu
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline void __syncthreads() {}
