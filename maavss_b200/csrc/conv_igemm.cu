// Implicit-GEMM fp32 convolution kernels: the batch is folded into the position axis so that the many small layers
// of the fusion encoders (spatial extents down to 8 x 4, avse_model.py:427-446,474-502) still fill whole tiles.
//   FWD   : Y[p, co]      = sum_{ci,tap} X[p @ tap, ci] W[co, ci, tap]          p = (n, od, oh, ow)
//   DGRAD : dX[q, ci]     = sum_{co,tap} dY[(q + pad - tap) / stride, co] W[co, ci, tap]   q = (n, id, ih, iw)
//   WGRAD : dW[co, ci,tap] (+)= sum_p dY[p, co] X[p @ tap, ci]   (reduction over positions split across CTAs, red.add)
// Tiles TM x TN x 16 with 256 threads; gathered operands are loaded position-fastest (coalesced along W).
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace igemm {

constexpr int BK = 16;
enum { FWD = 0, DGRAD = 1, WGRAD = 2 };

struct Geo {
    maavss_conv_desc d;
    int KV, KHW, O, OHW, I, HW;
    int Mdim, Ndim, Kdim;     // GEMM extents
    int k_per_split;
};

constexpr int LT = 256;        // K-index decode table window (FWD / DGRAD)

__device__ __forceinline__ bool div_stride(int t, int S, int& q) {   // t / S when exact and t >= 0 (S is warp-uniform)
    if (t < 0) return false;
    if (S == 1) { q = t; return true; }
    if (S == 2) { q = t >> 1; return (t & 1) == 0; }
    q = t / S;
    return q * S == t;
}

template <int MODE, int TM, int TN, int TJ>
__global__ void __launch_bounds__(256) conv_igemm_kernel(Geo g, const float* __restrict__ act, const float* __restrict__ wgt_or_dy,
                                                         const float* __restrict__ bias, float* __restrict__ out, int accumulate) {
    // FWD:   act = x,  wgt_or_dy = w,  out = y          rows i = positions, cols j = cout, l = ci*KV + tap
    // DGRAD: act = dy, wgt_or_dy = w,  out = dx         rows i = input positions, cols j = cin, l = co*KV + tap
    // WGRAD: act = x,  wgt_or_dy = dy, out = dw         rows i = cout, cols j = ci*KV + tap, l = positions
    // All index decoding (div/mod) is hoisted out of the gather loops into small shared-memory tables:
    //   FWD/DGRAD: ltab[l] = {channel offset (+ tap offset for FWD), packed (kd, kh, kw)} for a window of LT reduction indices;
    //   WGRAD:     ptab[l] = {input base offset, packed (id0, ih0, iw0), dy base offset} for the BK positions of an iteration.
    __shared__ float As[BK][TM + 4];
    __shared__ float Bs[BK][TN + 4];
    __shared__ int4 ltab[MODE == WGRAD ? 1 : LT];
    __shared__ int4 ptab[MODE == WGRAD ? 2 * BK : 1];
    constexpr int TI = 256 / TJ;                        // thread grid TI x TJ
    constexpr int RM = TM / TI, RN = TN / TJ;          // per-thread micro tile
    static_assert(TM % TI == 0 && TN % TJ == 0 && 256 % TM == 0 && 256 % TN == 0, "bad tile");
    const maavss_conv_desc& d = g.d;
    const int i0 = blockIdx.x * TM, j0 = blockIdx.y * TN;     // rows on grid.x: N * positions can exceed the 65535 limit of grid.y
    const int ti = threadIdx.x % TI, tj = threadIdx.x / TI;
    const int l_begin = blockIdx.z * g.k_per_split;
    const int l_end = min(g.Kdim, l_begin + g.k_per_split);
    constexpr int A_PER = (TM * BK + 255) / 256, B_PER = (TN * BK + 255) / 256;

    // ---- per-thread fixed part of the gathered operand
    int p_base = 0, p_a = 0, p_b = 0, p_c = 0;          // FWD: input base offset and (id0, ih0, iw0); DGRAD: n*Cout*O and (td0, th0, tw0)
    bool pvalid = false;
    if (MODE != WGRAD) {
        const int p = i0 + (int)(threadIdx.x % TM);
        pvalid = p < g.Mdim;
        const int ext = (MODE == FWD) ? g.O : g.I;
        const int plane = (MODE == FWD) ? g.OHW : g.HW;
        const int wdt = (MODE == FWD) ? d.OW : d.W;
        const int pp = pvalid ? p : 0;
        const int n = pp / ext;
        int o = pp - n * ext;
        const int a = o / plane;
        o -= a * plane;
        const int b = o / wdt, c = o - b * wdt;
        if (MODE == FWD) {
            p_a = a * d.SD - d.PD; p_b = b * d.SH - d.PH; p_c = c * d.SW - d.PW;
            p_base = n * d.Cin * g.I + p_a * g.HW + p_b * d.W + p_c;
        } else {
            p_a = a + d.PD; p_b = b + d.PH; p_c = c + d.PW;
            p_base = n * d.Cout * g.O;
        }
    }
    int w_off = 0, w_kd = 0, w_kh = 0, w_kw = 0;        // WGRAD: this thread's column (ci, tap)
    bool wvalid = false;
    if (MODE == WGRAD) {
        const int j = j0 + (int)(threadIdx.x % TN);
        wvalid = j < g.Ndim;
        const int jj = wvalid ? j : 0;
        const int ci = jj / g.KV, tap = jj - ci * g.KV;
        w_kd = tap / g.KHW;
        const int r = tap - w_kd * g.KHW;
        w_kh = r / d.KW;
        w_kw = r - w_kh * d.KW;
        w_off = ci * g.I + w_kd * g.HW + w_kh * d.W + w_kw;
    }
    auto fill_ptab = [&](int l0, int buf) {             // WGRAD: decode the BK positions of the iteration starting at l0
        if (threadIdx.x < BK) {
            const int p = l0 + threadIdx.x;
            int4 e = make_int4(0, 0, -1, 0);
            if (p < l_end) {
                const int n = p / g.O;
                int o = p - n * g.O;
                const int a = o / g.OHW;
                const int r = o - a * g.OHW;
                const int b = r / d.OW, c = r - b * d.OW;
                const int id0 = a * d.SD - d.PD, ih0 = b * d.SH - d.PH, iw0 = c * d.SW - d.PW;
                e.x = n * d.Cin * g.I + id0 * g.HW + ih0 * d.W + iw0;
                e.y = (id0 + 64) | ((ih0 + 64) << 10);
                e.z = n * d.Cout * g.O + o;
                e.w = iw0;
            }
            ptab[buf * BK + threadIdx.x] = e;
        }
    };

    float acc[RM][RN];
#pragma unroll
    for (int u = 0; u < RM; ++u)
#pragma unroll
        for (int v = 0; v < RN; ++v) acc[u][v] = 0.f;

    // Operand tiles travel global -> registers (fetch, issued one iteration ahead so that the gathers overlap the FMAs
    // of the current tile) -> shared memory (stash).
    float ra[A_PER], rb[B_PER];
    auto build_ltab = [&](int l0) {                      // decode LT reduction indices starting at l0
        const int gl = l0 + threadIdx.x;
        int4 e = make_int4(0, 0, 0, 0);
        if (gl < l_end) {
            const int ch = gl / g.KV, tap = gl - ch * g.KV;
            const int kd = tap / g.KHW, r = tap - kd * g.KHW;
            const int kh = r / d.KW, kw = r - kh * d.KW;
            e.x = (MODE == FWD) ? ch * g.I + kd * g.HW + kh * d.W + kw : ch * g.O;
            e.y = kd | (kh << 8) | (kw << 16);
            e.z = ch * d.Cin * g.KV + tap;              // DGRAD: weight offset of (co, cin 0, tap)
        }
        ltab[threadIdx.x] = e;
    };
    auto fetch = [&](int l0, int iter) {
        const int tl = (iter % (LT / BK)) * BK;          // offset of this iteration inside the table window
        // ---------------- A tile
        if (MODE == WGRAD) {
            // A(i = cout, l = position) = dy[n, i, o]
#pragma unroll
            for (int q = 0; q < A_PER; ++q) {
                const int e = threadIdx.x + q * 256;
                if (TM * BK < 256 && e >= TM * BK) break;
                const int l = e % BK, r = e / BK;
                const int i = i0 + r;
                const int4 pe = ptab[(iter & 1) * BK + l];
                ra[q] = (pe.z >= 0 && i < g.Mdim) ? __ldg(wgt_or_dy + (size_t)pe.z + (size_t)i * g.O) : 0.f;
            }
        } else {
#pragma unroll
            for (int q = 0; q < A_PER; ++q) {
                const int e = threadIdx.x + q * 256;
                if (TM * BK < 256 && e >= TM * BK) break;
                const int l = e / TM;                   // row e % TM == threadIdx.x % TM for every q (256 % TM == 0)
                float v = 0.f;
                if (pvalid && l0 + l < l_end) {
                    const int4 le = ltab[tl + l];
                    const int kd = le.y & 255, kh = (le.y >> 8) & 255, kw = le.y >> 16;
                    if (MODE == FWD) {
                        if ((unsigned)(p_a + kd) < (unsigned)d.D && (unsigned)(p_b + kh) < (unsigned)d.H &&
                            (unsigned)(p_c + kw) < (unsigned)d.W)
                            v = __ldg(act + (p_base + le.x));
                    } else {
                        int od, oh, ow;
                        if (div_stride(p_a - kd, d.SD, od) && div_stride(p_b - kh, d.SH, oh) && div_stride(p_c - kw, d.SW, ow) &&
                            od < d.OD && oh < d.OH && ow < d.OW)
                            v = __ldg(act + (p_base + le.x + od * g.OHW + oh * d.OW + ow));
                    }
                }
                ra[q] = v;
            }
        }
        // ---------------- B tile
        if (MODE == FWD) {
#pragma unroll
            for (int q = 0; q < B_PER; ++q) {
                const int e = threadIdx.x + q * 256;
                if (TN * BK < 256 && e >= TN * BK) break;
                const int l = e % BK, c = e / BK;
                const int gl = l0 + l, j = j0 + c;
                rb[q] = (gl < l_end && j < g.Ndim) ? __ldg(wgt_or_dy + (size_t)j * g.Kdim + gl) : 0.f;
            }
        } else if (MODE == DGRAD) {
            // B(j = cin, l = co*KV + tap) = w[(co*Cin + j)*KV + tap]
#pragma unroll
            for (int q = 0; q < B_PER; ++q) {
                const int e = threadIdx.x + q * 256;
                if (TN * BK < 256 && e >= TN * BK) break;
                const int l = e % BK, c = e / BK;
                const int gl = l0 + l, j = j0 + c;
                float v = 0.f;
                if (gl < l_end && j < g.Ndim) v = __ldg(wgt_or_dy + (size_t)ltab[tl + l].z + (size_t)j * g.KV);
                rb[q] = v;
            }
        } else {
            // B(j = (ci, tap), l = position) = x gathered
#pragma unroll
            for (int q = 0; q < B_PER; ++q) {
                const int e = threadIdx.x + q * 256;
                if (TN * BK < 256 && e >= TN * BK) break;
                const int l = e / TN;
                const int4 pe = ptab[(iter & 1) * BK + l];
                float v = 0.f;
                if (wvalid && pe.z >= 0) {
                    const int id = (pe.y & 1023) - 64 + w_kd, ih = (pe.y >> 10) - 64 + w_kh, iw = pe.w + w_kw;
                    if ((unsigned)id < (unsigned)d.D && (unsigned)ih < (unsigned)d.H && (unsigned)iw < (unsigned)d.W)
                        v = __ldg(act + (pe.x + w_off));
                }
                rb[q] = v;
            }
        }
    };
    auto stash = [&]() {
#pragma unroll
        for (int q = 0; q < A_PER; ++q) {
            const int e = threadIdx.x + q * 256;
            if (TM * BK < 256 && e >= TM * BK) break;
            if (MODE == WGRAD) As[e % BK][e / BK] = ra[q];
            else As[e / TM][e % TM] = ra[q];
        }
#pragma unroll
        for (int q = 0; q < B_PER; ++q) {
            const int e = threadIdx.x + q * 256;
            if (TN * BK < 256 && e >= TN * BK) break;
            if (MODE == WGRAD) Bs[e / TN][e % TN] = rb[q];
            else Bs[e % BK][e / BK] = rb[q];
        }
    };

    if (MODE == WGRAD) fill_ptab(l_begin, 0);
    else build_ltab(l_begin);
    __syncthreads();
    fetch(l_begin, 0);
    int iter = 0;
    for (int l0 = l_begin; l0 < l_end; l0 += BK, ++iter) {
        // WGRAD: positions of the next iteration into the other table buffer (last read two barriers ago)
        if (MODE == WGRAD) fill_ptab(l0 + BK, (iter + 1) & 1);
        stash();
        __syncthreads();
        if (l0 + BK < l_end) {
            if (MODE != WGRAD && ((iter + 1) % (LT / BK)) == 0) {
                build_ltab(l0 + BK);                     // every fetch from the old window finished before the barrier above
                __syncthreads();
            }
            fetch(l0 + BK, iter + 1);
        }
#pragma unroll
        for (int l = 0; l < BK; ++l) {
            float ar[RM], br[RN];
#pragma unroll
            for (int u = 0; u < RM; ++u) ar[u] = As[l][ti + TI * u];
#pragma unroll
            for (int v = 0; v < RN; ++v) br[v] = Bs[l][tj + TJ * v];
#pragma unroll
            for (int u = 0; u < RM; ++u)
#pragma unroll
                for (int v = 0; v < RN; ++v) acc[u][v] = fmaf(ar[u], br[v], acc[u][v]);
        }
        __syncthreads();
    }

    const bool split = gridDim.z > 1;
#pragma unroll
    for (int u = 0; u < RM; ++u) {
        const int i = i0 + ti + TI * u;
        if (i >= g.Mdim) continue;
        size_t row_off;
        int ext = 0;
        if (MODE == WGRAD) {
            row_off = (size_t)i * g.Ndim;
        } else {
            ext = (MODE == FWD) ? g.O : g.I;
            const int n = i / ext, o = i - n * ext;
            row_off = (size_t)n * g.Ndim * ext + o;
        }
#pragma unroll
        for (int v = 0; v < RN; ++v) {
            const int j = j0 + tj + TJ * v;
            if (j >= g.Ndim) continue;
            float* dst = (MODE == WGRAD) ? out + row_off + j : out + row_off + (size_t)j * ext;
            if (MODE == WGRAD) {
                if (split || accumulate) atomicAdd(dst, acc[u][v]);
                else *dst = acc[u][v];
            } else {
                if (split) atomicAdd(dst, acc[u][v]);
                else *dst = acc[u][v] + (bias ? bias[j] : 0.f);
            }
        }
    }
}

// out[n, c, :] = bias[c] (or 0): initialises the destination of a split-reduction FWD / DGRAD
__global__ void fill_bias_kernel(float* __restrict__ out, const float* __restrict__ bias, int C, long long inner, long long total) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x)
        out[e] = bias ? bias[(e / inner) % C] : 0.f;
}

}  // namespace igemm

// ------------------------------------------------------------------------------------------------ C ABI
namespace {

int check_desc(const maavss_conv_desc* d, const char* who) {
    MAAVSS_CHECK_ARG(d, "%s: null descriptor", who);
    MAAVSS_CHECK_ARG(d->N >= 0 && d->Cin > 0 && d->Cout > 0 && d->D > 0 && d->H > 0 && d->W > 0, "%s: bad input dims", who);
    MAAVSS_CHECK_ARG(d->KD > 0 && d->KH > 0 && d->KW > 0 && d->SD > 0 && d->SH > 0 && d->SW > 0, "%s: bad kernel/stride", who);
    MAAVSS_CHECK_ARG(d->OD > 0 && d->OH > 0 && d->OW > 0, "%s: bad output dims", who);
    // every output position must read inside the padded input, and the input must not extend past what the taps reach + stride
    MAAVSS_CHECK_ARG((d->OD - 1) * d->SD - d->PD + d->KD - 1 < d->D + d->PD + d->SD && (d->OH - 1) * d->SH - d->PH + d->KH - 1 < d->H + d->PH + d->SH &&
                         (d->OW - 1) * d->SW - d->PW + d->KW - 1 < d->W + d->PW + d->SW,
                     "%s: output dims inconsistent with input/kernel/stride/padding", who);
    const long long pos = (long long)d->N * (d->OD * (long long)d->OH * d->OW), ipos = (long long)d->N * (d->D * (long long)d->H * d->W);
    MAAVSS_CHECK_ARG(pos * d->Cout < (1LL << 31) && ipos * d->Cin < (1LL << 31), "%s: tensors of 2^31 elements or more are not supported", who);
    MAAVSS_CHECK_ARG(d->KD < 256 && d->KH < 256 && d->KW < 256 && d->PD < 64 && d->PH < 64 && (long long)d->OD * d->SD < 900 &&
                         (long long)d->OH * d->SH < (1 << 20),
                     "%s: geometry outside the packed-index range", who);
    return MAAVSS_OK;
}

igemm::Geo make_geo(const maavss_conv_desc* d) {
    igemm::Geo g;
    g.d = *d;
    g.KV = d->KD * d->KH * d->KW;
    g.KHW = d->KH * d->KW;
    g.OHW = d->OH * d->OW;
    g.O = d->OD * g.OHW;
    g.HW = d->H * d->W;
    g.I = d->D * g.HW;
    g.Mdim = g.Ndim = g.Kdim = 0;
    g.k_per_split = 0;
    return g;
}

// pick the tile (64x64 or 32x32) and the reduction split so that the grid covers the SMs ~2x
template <int MODE>
int launch(igemm::Geo g, const float* a, const float* b, const float* bias, float* out, int accumulate, long long out_elems,
           int out_channels, long long out_inner, cudaStream_t st) {
    const int sms = maavss_num_sms();
    // tile shape: narrow tiles for the layers with very few channels on one GEMM side, 32x32 when 64x64 cannot fill the SMs
    int cfg, TMv, TNv;
    const long long t64 = (long long)ceil_div(g.Mdim, 64) * ceil_div(g.Ndim, 64);
    if (g.Ndim <= 8) { cfg = 2; TMv = 256; TNv = 8; }
    else if (g.Ndim <= 16) { cfg = 3; TMv = 128; TNv = 16; }
    else if (g.Mdim <= 8) { cfg = 4; TMv = 8; TNv = 256; }
    else if (g.Mdim <= 16) { cfg = 5; TMv = 16; TNv = 128; }
    // the weight gradient gets its parallelism from splitting the (huge) position reduction, not from small tiles:
    // 64 x 64 tiles keep 16 FMAs per 8 shared-memory loads (ncu: the 32 x 32 tiles ran at 31 % issue-active)
    else if (MODE == igemm::WGRAD && g.Mdim >= 48 && g.Ndim >= 48) { cfg = 0; TMv = 64; TNv = 64; }
    else if (t64 < sms) { cfg = 1; TMv = 32; TNv = 32; }
    else { cfg = 0; TMv = 64; TNv = 64; }
    const long long tiles = (long long)ceil_div(g.Mdim, TMv) * ceil_div(g.Ndim, TNv);
    int splits = 1;
    if (tiles < 2 * sms) {
        splits = (int)((2 * sms + tiles - 1) / tiles);
        const int max_splits = g.Kdim / (MODE == igemm::WGRAD ? 64 : 128);
        if (splits > max_splits) splits = max_splits;
        if (splits < 1) splits = 1;
    }
    g.k_per_split = ((g.Kdim + splits - 1) / splits + igemm::BK - 1) / igemm::BK * igemm::BK;
    splits = ceil_div(g.Kdim, g.k_per_split);
    if (MODE == igemm::WGRAD) {
        if (!accumulate && splits > 1) MAAVSS_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * (size_t)out_elems, st));
    } else if (splits > 1) {
        if (bias) {
            igemm::fill_bias_kernel<<<ceil_div(out_elems, 256 * 8), 256, 0, st>>>(out, bias, out_channels, out_inner, out_elems);
            MAAVSS_LAUNCH_CHECK();
        } else {
            MAAVSS_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * (size_t)out_elems, st));
        }
    }
    dim3 grid(ceil_div(g.Mdim, TMv), ceil_div(g.Ndim, TNv), splits);
    MAAVSS_CHECK_ARG(grid.y <= 65535 && grid.z <= 65535, "conv: grid too large");
    switch (cfg) {
        case 0: igemm::conv_igemm_kernel<MODE, 64, 64, 16><<<grid, 256, 0, st>>>(g, a, b, bias, out, accumulate); break;
        case 1: igemm::conv_igemm_kernel<MODE, 32, 32, 16><<<grid, 256, 0, st>>>(g, a, b, bias, out, accumulate); break;
        case 2: igemm::conv_igemm_kernel<MODE, 256, 8, 2><<<grid, 256, 0, st>>>(g, a, b, bias, out, accumulate); break;
        case 3: igemm::conv_igemm_kernel<MODE, 128, 16, 4><<<grid, 256, 0, st>>>(g, a, b, bias, out, accumulate); break;
        case 4: igemm::conv_igemm_kernel<MODE, 8, 256, 32><<<grid, 256, 0, st>>>(g, a, b, bias, out, accumulate); break;
        default: igemm::conv_igemm_kernel<MODE, 16, 128, 16><<<grid, 256, 0, st>>>(g, a, b, bias, out, accumulate); break;
    }
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

}  // namespace

int maavss_conv_wgrad_direct(const maavss_conv_desc*, const float*, const float*, float*, int, cudaStream_t);
int maavss_conv_dgrad_direct(const maavss_conv_desc*, const float*, const float*, const float*, float*, cudaStream_t);
int maavss_conv_fwd_direct(const maavss_conv_desc*, const float*, const float*, const float*, float*, cudaStream_t);

extern "C" int maavss_conv_fwd(const maavss_conv_desc* d, const float* x, const float* w, const float* bias, float* y, void* stream) {
    int rc = check_desc(d, "conv_fwd");
    if (rc) return rc;
    if (d->N == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(x && w && y, "conv_fwd: null argument");
    rc = maavss_conv_fwd_direct(d, x, w, bias, y, (cudaStream_t)stream);
    if (rc != 0) return rc < 0 ? rc : MAAVSS_OK;
    igemm::Geo g = make_geo(d);
    g.Mdim = d->N * g.O;
    g.Ndim = d->Cout;
    g.Kdim = d->Cin * g.KV;
    return launch<igemm::FWD>(g, x, w, bias, y, 0, (long long)g.Mdim * d->Cout, d->Cout, g.O, (cudaStream_t)stream);
}

extern "C" int maavss_conv_dgrad(const maavss_conv_desc* d, const float* dy, const float* w, const float* bias, float* dx, void* stream) {
    int rc = check_desc(d, "conv_dgrad");
    if (rc) return rc;
    if (d->N == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(dy && w && dx, "conv_dgrad: null argument");
    rc = maavss_conv_dgrad_direct(d, dy, w, bias, dx, (cudaStream_t)stream);
    if (rc != 0) return rc < 0 ? rc : MAAVSS_OK;
    igemm::Geo g = make_geo(d);
    g.Mdim = d->N * g.I;
    g.Ndim = d->Cin;
    g.Kdim = d->Cout * g.KV;
    return launch<igemm::DGRAD>(g, dy, w, bias, dx, 0, (long long)g.Mdim * d->Cin, d->Cin, g.I, (cudaStream_t)stream);
}

extern "C" int maavss_conv_wgrad(const maavss_conv_desc* d, const float* x, const float* dy, float* dw, int accumulate, void* stream) {
    int rc = check_desc(d, "conv_wgrad");
    if (rc) return rc;
    MAAVSS_CHECK_ARG((x && dy && dw) || d->N == 0, "conv_wgrad: null argument");
    igemm::Geo g = make_geo(d);
    const long long elems = (long long)d->Cout * d->Cin * g.KV;
    if (d->N == 0) {
        if (!accumulate) MAAVSS_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)elems, (cudaStream_t)stream));
        return MAAVSS_OK;
    }
    rc = maavss_conv_wgrad_direct(d, x, dy, dw, accumulate, (cudaStream_t)stream);
    if (rc != 0) return rc < 0 ? rc : MAAVSS_OK;
    g.Mdim = d->Cout;
    g.Ndim = d->Cin * g.KV;
    g.Kdim = d->N * g.O;
    return launch<igemm::WGRAD>(g, x, dy, nullptr, dw, accumulate, elems, 0, 0, (cudaStream_t)stream);
}
