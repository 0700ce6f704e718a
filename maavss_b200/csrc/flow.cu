// Sparse optical flow of the reference's flow.py: Shi-Tomasi corners (cv2.goodFeaturesToTrack,
// flow.py:11-14,28) tracked by the pyramidal Lucas-Kanade method (cv2.calcOpticalFlowPyrLK, flow.py:17-19,37) between two
// gray frames.  The arithmetic lives in OpenCV (4.x: imgproc/featureselect.cpp, corner.cpp, video/lkpyramid.cpp); it is
// restated here operation for operation -- integer where OpenCV is integer (gray conversion, pyrDown, Scharr, the 14-bit
// fixed-point bilinear patches), fp32 where it is fp32 -- so that the corner set is index-exact and the tracked points agree
// to a few 1e-3 px.  The cv2.imshow window of flow.py is a GUI: unavailable on this box, left out.
// Everything is small (256 x 256 frames, <= 100 points): plain global-memory kernels, one CTA per tracked point.
#include <math.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

__device__ __forceinline__ int r101(int i, int n) {      // BORDER_REFLECT_101
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

// cv2.cvtColor(BGR2GRAY) for uint8 (OpenCV 4.x, 15-bit fixed point): (B*3735 + G*19235 + R*9798 + 2^14) >> 15
__global__ void __launch_bounds__(256) bgr2gray_kernel(const unsigned char* __restrict__ bgr, unsigned char* __restrict__ gray, long long n) {
    for (long long p = (long long)blockIdx.x * 256 + threadIdx.x; p < n; p += (long long)gridDim.x * 256) {
        const unsigned char* s = bgr + 3 * p;
        gray[p] = (unsigned char)((s[0] * 3735 + s[1] * 19235 + s[2] * 9798 + (1 << 14)) >> 15);
    }
}

// cornerEigenValsVecs, first half: Dx, Dy = Sobel(ksize 3, scale 1 / (4 * block * 255)) with the scale folded into the
// smoothing kernel [1 2 1] as OpenCV does, then the products (Dx^2, Dx Dy, Dy^2).  Unfused multiplies / adds on purpose.
__global__ void __launch_bounds__(256) flow_cov_kernel(const unsigned char* __restrict__ g, float* __restrict__ cov, int H, int W, float k0) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= W || y >= H) return;
    const float k1 = __fmul_rn(2.f, k0);
    float p[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) p[i][j] = (float)g[(size_t)r101(y + i - 1, H) * W + r101(x + j - 1, W)];
    const float rd0 = p[0][2] - p[0][0], rd1 = p[1][2] - p[1][0], rd2 = p[2][2] - p[2][0];            // exact integers
    const float dx = __fadd_rn(__fmul_rn(rd1, k1), __fmul_rn(__fadd_rn(rd0, rd2), k0));
    float rs[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) rs[i] = __fadd_rn(__fmul_rn(p[i][1], k1), __fmul_rn(__fadd_rn(p[i][0], p[i][2]), k0));
    const float dy = __fsub_rn(rs[2], rs[0]);
    float* c = cov + ((size_t)y * W + x) * 3;
    c[0] = __fmul_rn(dx, dx);
    c[1] = __fmul_rn(dx, dy);
    c[2] = __fmul_rn(dy, dy);
}

// second half: un-normalised block x block box filter (reflect-101) of the products, then calcMinEigenVal
__global__ void __launch_bounds__(256) flow_eig_kernel(const float* __restrict__ cov, float* __restrict__ eig, int H, int W, int block,
                                                       int* __restrict__ max_bits) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    float e = 0.f;
    if (x < W && y < H) {
        const int r = block / 2;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f;
        for (int i = 0; i < block; ++i) {
            const int yy = r101(y + i - r, H);
            for (int j = 0; j < block; ++j) {
                const float* c = cov + ((size_t)yy * W + r101(x + j - r, W)) * 3;
                s0 = __fadd_rn(s0, c[0]);
                s1 = __fadd_rn(s1, c[1]);
                s2 = __fadd_rn(s2, c[2]);
            }
        }
        const float a = __fmul_rn(s0, 0.5f), b = s1, c = __fmul_rn(s2, 0.5f);
        const float d = __fsub_rn(a, c);
        e = __fsub_rn(__fadd_rn(a, c), __fsqrt_rn(__fadd_rn(__fmul_rn(d, d), __fmul_rn(b, b))));
        eig[(size_t)y * W + x] = e;
    }
    // block max -> one atomic (positive floats order like their bit patterns as signed ints; negatives never win)
    int bits = __float_as_int(e);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) bits = max(bits, __shfl_xor_sync(0xffffffffu, bits, o));
    if ((threadIdx.x & 31) == 0 && bits > 0) atomicMax(max_bits, bits);
}

// threshold(eig, max * quality, TOZERO), 3 x 3 dilate, keep interior points that equal their dilation: the candidate corners
constexpr int kMaxCand = 1 << 16;
__global__ void __launch_bounds__(256) flow_candidates_kernel(const float* __restrict__ eig, int H, int W, float quality,
                                                              const int* __restrict__ max_bits, unsigned long long* __restrict__ cand,
                                                              int* __restrict__ n_cand) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x < 1 || y < 1 || x >= W - 1 || y >= H - 1) return;
    const float thr = (float)((double)__int_as_float(*max_bits) * (double)quality);
    const float v = eig[(size_t)y * W + x];
    if (!(v > thr)) return;
    float m = v;
#pragma unroll
    for (int i = -1; i <= 1; ++i)
#pragma unroll
        for (int j = -1; j <= 1; ++j) {
            const float u = eig[(size_t)(y + i) * W + x + j];
            m = fmaxf(m, u > thr ? u : 0.f);
        }
    if (v != m) return;
    const int slot = atomicAdd(n_cand, 1);
    if (slot < kMaxCand) cand[slot] = ((unsigned long long)(unsigned)__float_as_int(v) << 32) | (unsigned)(y * W + x);
}

// One CTA: sort the candidates by (value, address) descending -- std::sort with greaterThanPtr -- then the sequential
// minimum-distance selection of featureselect.cpp.
__global__ void __launch_bounds__(1024) flow_select_kernel(unsigned long long* __restrict__ cand, const int* __restrict__ n_cand, int W,
                                                           int max_corners, float min_distance, float* __restrict__ corners,
                                                           int* __restrict__ count) {
    int n = min(*n_cand, kMaxCand);
    int np2 = 1;
    while (np2 < n) np2 <<= 1;
    for (int i = n + threadIdx.x; i < np2; i += 1024) cand[i] = 0ull;          // padding sorts to the end
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < np2; i += 1024) {
                const int l = i ^ j;
                if (l > i) {
                    const unsigned long long a = cand[i], b = cand[l];
                    const bool desc = (i & k) == 0;
                    if (desc ? a < b : a > b) { cand[i] = b; cand[l] = a; }
                }
            }
            __syncthreads();
        }
    if (threadIdx.x == 0) {
        int m = 0;
        const float md2 = min_distance * min_distance;
        for (int i = 0; i < n && (max_corners <= 0 || m < max_corners); ++i) {
            const unsigned idx = (unsigned)(cand[i] & 0xffffffffull);
            const float x = (float)(idx % (unsigned)W), y = (float)(idx / (unsigned)W);
            bool good = true;
            if (min_distance >= 1.f)
                for (int q = 0; q < m; ++q) {
                    const float dx = x - corners[2 * q], dy = y - corners[2 * q + 1];
                    if (dx * dx + dy * dy < md2) { good = false; break; }
                }
            if (good) { corners[2 * m] = x; corners[2 * m + 1] = y; ++m; }
        }
        *count = m;
    }
}

// cv2.pyrDown for uint8: 5 x 5 Gaussian [1 4 6 4 1]^2 / 256 with rounding, reflect-101, output ((W+1)/2, (H+1)/2)
__global__ void __launch_bounds__(256) pyrdown_kernel(const unsigned char* __restrict__ src, unsigned char* __restrict__ dst, int H, int W,
                                                      int OH, int OW) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= OW || y >= OH) return;
    const int k[5] = {1, 4, 6, 4, 1};
    int s = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const unsigned char* row = src + (size_t)r101(2 * y + i - 2, H) * W;
        int rs = 0;
#pragma unroll
        for (int j = 0; j < 5; ++j) rs += k[j] * row[r101(2 * x + j - 2, W)];
        s += k[i] * rs;
    }
    dst[(size_t)y * OW + x] = (unsigned char)((s + 128) >> 8);
}

// calcSharrDeriv: (dI/dx, dI/dy) as int16 pairs, Scharr [3 10 3] x [-1 0 1], reflect-101 at the image edge
__global__ void __launch_bounds__(256) scharr_kernel(const unsigned char* __restrict__ src, short2* __restrict__ d, int H, int W) {
    const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= W || y >= H) return;
    int p[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) p[i][j] = src[(size_t)r101(y + i - 1, H) * W + r101(x + j - 1, W)];
    int t0[3], t1[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        t0[j] = (p[0][j] + p[2][j]) * 3 + p[1][j] * 10;
        t1[j] = p[2][j] - p[0][j];
    }
    d[(size_t)y * W + x] = make_short2((short)(t0[2] - t0[0]), (short)((t1[2] + t1[0]) * 3 + t1[1] * 10));
}

}  // namespace

// ------------------------------------------------------------------------------------------ pyramidal Lucas-Kanade
namespace {

constexpr int kMaxLevels = 6, kWBits = 14;

struct LkLevels {
    const unsigned char* I[kMaxLevels];
    const unsigned char* J[kMaxLevels];
    const short2* dI[kMaxLevels];
    int H[kMaxLevels], W[kMaxLevels];
};

// image with its reflect-101 border of `win` pixels (OpenCV pads every pyramid level); derivatives are zero outside
__device__ __forceinline__ int pix(const unsigned char* img, int H, int W, int y, int x) { return img[(size_t)r101(y, H) * W + r101(x, W)]; }
__device__ __forceinline__ short2 dpix(const short2* d, int H, int W, int y, int x) {
    return (x < 0 || y < 0 || x >= W || y >= H) ? make_short2(0, 0) : d[(size_t)y * W + x];
}
__device__ __forceinline__ int descale(int v, int n) { return (v + (1 << (n - 1))) >> n; }

template <int NV>
__device__ __forceinline__ void block_sum(float (&v)[NV], float* red) {      // 256 threads -> every thread gets the totals
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int q = 0; q < NV; ++q) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
    }
    __syncthreads();
    if (lane == 0)
#pragma unroll
        for (int q = 0; q < NV; ++q) red[wid * NV + q] = v[q];
    __syncthreads();
#pragma unroll
    for (int q = 0; q < NV; ++q) {
        float s = 0.f;
        for (int w = 0; w < 8; ++w) s += red[w * NV + q];
        v[q] = s;
    }
}

// LKTrackerInvoker (video/lkpyramid.cpp) for one point: one CTA of 256 threads, thread t < win * win owns window pixel
// (t / win, t % win); the levels run from the top of the pyramid down inside the kernel.
__global__ void __launch_bounds__(256) lk_track_kernel(LkLevels L, const float* __restrict__ prev_pts, const int* __restrict__ count,
                                                       float* __restrict__ next_pts, unsigned char* __restrict__ status,
                                                       float* __restrict__ err, int win, int max_level, int max_iter, float eps2,
                                                       float min_eig_thr) {
    const int pt = blockIdx.x;
    if (count && pt >= *count) return;
    __shared__ float red[8 * 3];
    const int t = threadIdx.x;
    const bool act = t < win * win;
    const int wy = t / win, wx = t % win;
    const float half = (float)(win - 1) * 0.5f;
    const float px = prev_pts[2 * pt], py = prev_pts[2 * pt + 1];
    float nx = 0.f, ny = 0.f;                       // nextPts[pt], carried from level to level
    bool ok = true;
    float errv = 0.f;
    for (int level = max_level; level >= 0; --level) {
        const int H = L.H[level], W = L.W[level];
        const unsigned char* I = L.I[level];
        const unsigned char* J = L.J[level];
        const short2* dI = L.dI[level];
        const float sc = 1.f / (float)(1 << level);
        float ppx = px * sc, ppy = py * sc;
        if (level == max_level) { nx = ppx; ny = ppy; } else { nx *= 2.f; ny *= 2.f; }
        ppx -= half; ppy -= half;
        const int ipx = (int)floorf(ppx), ipy = (int)floorf(ppy);
        if (ipx < -win || ipx >= W || ipy < -win || ipy >= H) {
            if (level == 0) { ok = false; errv = 0.f; }
            continue;
        }
        float a = ppx - (float)ipx, b = ppy - (float)ipy;
        int iw00 = __float2int_rn((1.f - a) * (1.f - b) * (float)(1 << kWBits));
        int iw01 = __float2int_rn(a * (1.f - b) * (float)(1 << kWBits));
        int iw10 = __float2int_rn((1.f - a) * b * (float)(1 << kWBits));
        int iw11 = (1 << kWBits) - iw00 - iw01 - iw10;
        int Ival = 0, ix = 0, iy = 0;
        if (act) {
            const int y = ipy + wy, x = ipx + wx;
            Ival = descale(pix(I, H, W, y, x) * iw00 + pix(I, H, W, y, x + 1) * iw01 + pix(I, H, W, y + 1, x) * iw10 +
                           pix(I, H, W, y + 1, x + 1) * iw11, kWBits - 5);
            const short2 d00 = dpix(dI, H, W, y, x), d01 = dpix(dI, H, W, y, x + 1), d10 = dpix(dI, H, W, y + 1, x), d11 = dpix(dI, H, W, y + 1, x + 1);
            ix = descale(d00.x * iw00 + d01.x * iw01 + d10.x * iw10 + d11.x * iw11, kWBits);
            iy = descale(d00.y * iw00 + d01.y * iw01 + d10.y * iw10 + d11.y * iw11, kWBits);
            Ival = (short)Ival; ix = (short)ix; iy = (short)iy;
        }
        float A[3] = {(float)(ix * ix), (float)(ix * iy), (float)(iy * iy)};
        block_sum<3>(A, red);
        const float FLT_SCALE = 1.f / (float)(1 << 20);
        const float A11 = A[0] * FLT_SCALE, A12 = A[1] * FLT_SCALE, A22 = A[2] * FLT_SCALE;
        float D = A11 * A22 - A12 * A12;
        const float minEig = (A22 + A11 - sqrtf((A11 - A22) * (A11 - A22) + 4.f * A12 * A12)) / (float)(2 * win * win);
        if (minEig < min_eig_thr || D < 1.1920929e-07f) {
            if (level == 0) ok = false;
            continue;
        }
        D = 1.f / D;
        float qx = nx - half, qy = ny - half;        // nextPt (window corner)
        float pdx = 0.f, pdy = 0.f;
        for (int j = 0; j < max_iter; ++j) {
            const int inx = (int)floorf(qx), iny = (int)floorf(qy);
            if (inx < -win || inx >= W || iny < -win || iny >= H) {
                if (level == 0) ok = false;
                break;
            }
            a = qx - (float)inx; b = qy - (float)iny;
            iw00 = __float2int_rn((1.f - a) * (1.f - b) * (float)(1 << kWBits));
            iw01 = __float2int_rn(a * (1.f - b) * (float)(1 << kWBits));
            iw10 = __float2int_rn((1.f - a) * b * (float)(1 << kWBits));
            iw11 = (1 << kWBits) - iw00 - iw01 - iw10;
            float bb[2] = {0.f, 0.f};
            if (act) {
                const int y = iny + wy, x = inx + wx;
                const int diff = descale(pix(J, H, W, y, x) * iw00 + pix(J, H, W, y, x + 1) * iw01 + pix(J, H, W, y + 1, x) * iw10 +
                                         pix(J, H, W, y + 1, x + 1) * iw11, kWBits - 5) - Ival;
                bb[0] = (float)(diff * ix);
                bb[1] = (float)(diff * iy);
            }
            block_sum<2>(bb, red);
            const float b1 = bb[0] * FLT_SCALE, b2 = bb[1] * FLT_SCALE;
            const float dx = (A12 * b2 - A22 * b1) * D, dy = (A12 * b1 - A11 * b2) * D;
            qx += dx; qy += dy;
            nx = qx + half; ny = qy + half;
            if (dx * dx + dy * dy <= eps2) break;
            if (j > 0 && fabsf(dx + pdx) < 0.01f && fabsf(dy + pdy) < 0.01f) {
                nx -= dx * 0.5f; ny -= dy * 0.5f;
                break;
            }
            pdx = dx; pdy = dy;
        }
        if (ok && level == 0) {                      // the error measure: mean |J - I| over the window at the final position
            const float fx = nx - half, fy = ny - half;
            const int inx = (int)floorf(fx), iny = (int)floorf(fy);
            if (inx < -win || inx >= W || iny < -win || iny >= H) {
                ok = false;
            } else {
                a = fx - (float)inx; b = fy - (float)iny;
                iw00 = __float2int_rn((1.f - a) * (1.f - b) * (float)(1 << kWBits));
                iw01 = __float2int_rn(a * (1.f - b) * (float)(1 << kWBits));
                iw10 = __float2int_rn((1.f - a) * b * (float)(1 << kWBits));
                iw11 = (1 << kWBits) - iw00 - iw01 - iw10;
                float e[1] = {0.f};
                if (act) {
                    const int y = iny + wy, x = inx + wx;
                    const int diff = descale(pix(J, H, W, y, x) * iw00 + pix(J, H, W, y, x + 1) * iw01 + pix(J, H, W, y + 1, x) * iw10 +
                                             pix(J, H, W, y + 1, x + 1) * iw11, kWBits - 5) - Ival;
                    e[0] = fabsf((float)diff);
                }
                block_sum<1>(e, red);
                errv = e[0] * (1.f / (float)(32 * win * win));
            }
        }
    }
    if (t == 0) {
        next_pts[2 * pt] = nx;
        next_pts[2 * pt + 1] = ny;
        status[pt] = ok ? 1 : 0;
        if (err) err[pt] = ok ? errv : 0.f;
    }
}

inline size_t al(size_t v) { return (v + 255) & ~(size_t)255; }
inline dim3 grid2d(int H, int W) { return dim3((W + 31) / 32, (H + 7) / 8); }

}  // namespace

extern "C" int maavss_bgr_to_gray(const unsigned char* bgr, unsigned char* gray, long long pixels, void* stream) {
    if (pixels == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(bgr && gray, "bgr_to_gray: null argument");
    int blocks = (int)((pixels + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    bgr2gray_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(bgr, gray, pixels);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" long long maavss_flow_features_workspace_bytes(int H, int W) {
    return (long long)(al(sizeof(float) * 3 * (size_t)H * W) + al(sizeof(float) * (size_t)H * W) + al(8 * (size_t)kMaxCand) + 256 + 256);
}

extern "C" int maavss_flow_good_features(const unsigned char* gray, int H, int W, int max_corners, float quality_level, float min_distance,
                                         int block_size, float* corners_xy, int* count, void* workspace, long long workspace_bytes,
                                         void* stream) {
    MAAVSS_CHECK_ARG(gray && corners_xy && count && workspace, "flow_good_features: null argument");
    MAAVSS_CHECK_ARG(H >= 3 && W >= 3 && (long long)H * W < (1ll << 31) && block_size >= 1 && block_size % 2 == 1 && max_corners > 0 &&
                         quality_level > 0.f, "flow_good_features: bad geometry / parameters");
    MAAVSS_CHECK_ARG(workspace_bytes >= maavss_flow_features_workspace_bytes(H, W), "flow_good_features: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t* base = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    float* cov = (float*)base;
    float* eig = (float*)(base + al(sizeof(float) * 3 * (size_t)H * W));
    unsigned long long* cand = (unsigned long long*)((uint8_t*)eig + al(sizeof(float) * (size_t)H * W));
    int* scalars = (int*)((uint8_t*)cand + al(8 * (size_t)kMaxCand));       // [0] max bits, [1] candidate count
    MAAVSS_CUDA(cudaMemsetAsync(scalars, 0, 8, st));
    const float k0 = (float)(1.0 / (4.0 * block_size * 255.0));
    flow_cov_kernel<<<grid2d(H, W), 256, 0, st>>>(gray, cov, H, W, k0);
    MAAVSS_LAUNCH_CHECK();
    flow_eig_kernel<<<grid2d(H, W), 256, 0, st>>>(cov, eig, H, W, block_size, scalars);
    MAAVSS_LAUNCH_CHECK();
    flow_candidates_kernel<<<grid2d(H, W), 256, 0, st>>>(eig, H, W, quality_level, scalars, cand, scalars + 1);
    MAAVSS_LAUNCH_CHECK();
    flow_select_kernel<<<1, 1024, 0, st>>>(cand, scalars + 1, W, max_corners, min_distance, corners_xy, count);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" long long maavss_flow_lk_workspace_bytes(int H, int W, int max_level) {
    size_t total = 256;
    int h = H, w = W;
    for (int l = 0; l <= max_level; ++l) {
        if (l > 0) total += 2 * al((size_t)h * w);
        total += al(sizeof(short2) * (size_t)h * w);
        h = (h + 1) / 2; w = (w + 1) / 2;
    }
    return (long long)total;
}

extern "C" int maavss_flow_pyr_lk(const unsigned char* prev_gray, const unsigned char* next_gray, int H, int W, const float* prev_pts,
                                  int max_points, const int* count, float* next_pts, unsigned char* status, float* err, int win,
                                  int max_level, int max_iter, float epsilon, float min_eig_threshold, void* workspace,
                                  long long workspace_bytes, void* stream) {
    if (max_points == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(prev_gray && next_gray && prev_pts && next_pts && status && workspace, "flow_pyr_lk: null argument");
    MAAVSS_CHECK_ARG(win >= 3 && win <= 15 && max_level >= 0 && max_level < kMaxLevels && max_iter >= 1 && max_iter <= 100 && max_points > 0,
                     "flow_pyr_lk: window must be 3..15, max_level < %d, 1..100 iterations", kMaxLevels);
    MAAVSS_CHECK_ARG(workspace_bytes >= maavss_flow_lk_workspace_bytes(H, W, max_level), "flow_pyr_lk: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t* p = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
    LkLevels L;
    int h = H, w = W, levels = max_level;
    for (int l = 0; l <= max_level; ++l) {
        if (l > 0 && (w <= win || h <= win)) { levels = l - 1; break; }      // OpenCV stops building levels smaller than the window
        L.H[l] = h; L.W[l] = w;
        if (l == 0) {
            L.I[0] = prev_gray; L.J[0] = next_gray;
        } else {
            unsigned char* Il = p; p += al((size_t)h * w);
            unsigned char* Jl = p; p += al((size_t)h * w);
            pyrdown_kernel<<<grid2d(h, w), 256, 0, st>>>(L.I[l - 1], Il, L.H[l - 1], L.W[l - 1], h, w);
            MAAVSS_LAUNCH_CHECK();
            pyrdown_kernel<<<grid2d(h, w), 256, 0, st>>>(L.J[l - 1], Jl, L.H[l - 1], L.W[l - 1], h, w);
            MAAVSS_LAUNCH_CHECK();
            L.I[l] = Il; L.J[l] = Jl;
        }
        short2* d = (short2*)p; p += al(sizeof(short2) * (size_t)h * w);
        scharr_kernel<<<grid2d(h, w), 256, 0, st>>>(L.I[l], d, h, w);
        MAAVSS_LAUNCH_CHECK();
        L.dI[l] = d;
        h = (h + 1) / 2; w = (w + 1) / 2;
    }
    lk_track_kernel<<<max_points, 256, 0, st>>>(L, prev_pts, count, next_pts, status, err, win, levels, max_iter, epsilon * epsilon,
                                                min_eig_threshold);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}
