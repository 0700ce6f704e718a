// Small-message sum all-reduce over NVLink peer memory: the train-mode BatchNorm partial sums that N data-parallel
// ranks exchange between bn_reduce and bn_apply (and between the two halves of the backward) -- 30 exchanges of <= 32 KB per
// optimizer step, each on the critical path.  As NCCL all-reduces they cost ~20 us apiece (launch + protocol latency); here
// one exchange is ONE single-CTA kernel per rank:
//   push   the rank stores its partial sums into row [seq % SLOTS][rank] of EVERY other rank's buffer (peer stores through
//          NVLink / NVSwitch; the buffers are cudaMalloc'ed once and opened by the peers through CUDA IPC), each value as 16
//          bytes that carry the sequence number as a tag
//   wait   it polls the rows of its OWN buffer until every value of every rank carries the tag (no fence, no flag word)
//   sum    it adds the world's contributions in rank order -- every rank computes the bit-identical sum -- in place
// The sequence counter lives in device memory and is advanced by the kernel, so a captured CUDA graph replays correctly.
// Exchanges of one channel must be issued in the same order on every rank; concurrent streams (the two encoder branches)
// use different channels (or one ordered stream).  A rank that waits longer than ~30 s sets an error word (maavss_peer_status) and gives up instead
// of hanging the GPU.
#include <stdint.h>
#include <string.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace px {

constexpr int MAXW = 16, SLOTS = 4, THREADS = 512;
// clock64 ticks (~30 s).  Long on purpose: like an NCCL collective, an exchange simply waits for a rank that is late (host-side
// skew between processes can be seconds at start-up); a rank that gave up early would run ahead, lap the 4 slots and leave
// every later exchange of every rank waiting out its limit.
constexpr long long WAIT_LIMIT = 60000000000LL;

struct State {
    unsigned long long seq;
    int error;
    int pad;
};

struct Layout {
    int channels, max_doubles;
    size_t state_off, flag_off, data_off, total;
};

struct Ptrs {
    uint8_t* p[MAXW];
};

static Layout make_layout(int channels, int max_doubles) {
    Layout L;
    L.channels = channels;
    L.max_doubles = max_doubles;
    L.state_off = 0;
    L.flag_off = 256 * (size_t)((channels * sizeof(State) + 255) / 256);
    L.data_off = L.flag_off;                                                    // (no flag words: the values carry their tag)
    L.total = L.data_off + (size_t)channels * SLOTS * MAXW * max_doubles * 16;   // 16 bytes per value
    return L;
}

// One value travels as 16 bytes {lo32, tag, hi32, tag} with tag = the low 32 bits of the sequence number (never 0): every
// 8-byte half validates itself, so the receiver needs no flag word and the sender no fence -- the exchange costs one launch
// and one one-way NVLink latency (NCCL's LL idea; the fence + flag version of this kernel measured 10-16 us per exchange).
__device__ __forceinline__ void st_vol_v4(uint4* p, uint4 v) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_vol_v4(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(THREADS) allreduce_kernel(Ptrs pp, Layout L, int rank, int world, int channel, double* __restrict__ data, int n) {
    __shared__ unsigned long long seq_s;
    uint8_t* local = pp.p[rank];
    State* st = reinterpret_cast<State*>(local + L.state_off) + channel;
    if (threadIdx.x == 0) seq_s = st->seq + 1;
    __syncthreads();
    const unsigned long long seq = seq_s;
    const unsigned tag = (unsigned)(seq & 0xffffffffull) | 0x80000000u;          // != 0 (the buffers start zeroed)
    const size_t slot = (size_t)channel * SLOTS + (size_t)(seq % SLOTS);
    // push: this rank's values into row [slot][rank] of every other rank's buffer
    for (int i = threadIdx.x; i < n; i += THREADS) {
        const unsigned long long bits = (unsigned long long)__double_as_longlong(data[i]);
        const uint4 v = make_uint4((unsigned)bits, tag, (unsigned)(bits >> 32), tag);
        for (int p = 0; p < world; ++p) {
            if (p == rank) continue;
            st_vol_v4(reinterpret_cast<uint4*>(pp.p[p] + L.data_off) + (slot * MAXW + rank) * L.max_doubles + i, v);
        }
    }
    // wait + sum in rank order (bit-identical on every rank); the own contribution comes from `data`
    const uint4* mine = reinterpret_cast<const uint4*>(local + L.data_off) + slot * MAXW * L.max_doubles;
    const long long t0 = clock64();
    bool failed = false;
    for (int i = threadIdx.x; i < n; i += THREADS) {
        double s = 0.0;
        for (int r = 0; r < world; ++r) {
            if (r == rank) {
                s += data[i];
                continue;
            }
            const uint4* src = mine + (size_t)r * L.max_doubles + i;
            uint4 v = ld_vol_v4(src);
            while ((v.y != tag || v.w != tag) && !failed) {
                if (clock64() - t0 > WAIT_LIMIT) failed = true;
                v = ld_vol_v4(src);
            }
            s += __longlong_as_double((long long)(((unsigned long long)v.z << 32) | v.x));
        }
        data[i] = s;
    }
    if (failed) st->error = 1;
    __syncthreads();
    if (threadIdx.x == 0) st->seq = seq;
}

}  // namespace px

struct maavss_peer {
    int rank, world, device;
    px::Layout L;
    uint8_t* local;
    px::Ptrs ptrs;
    bool opened[px::MAXW];
};

extern "C" int maavss_peer_create(maavss_peer** out, int rank, int world, int channels, int max_doubles, void* handle64) {
    MAAVSS_CHECK_ARG(out && handle64, "peer_create: null argument");
    MAAVSS_CHECK_ARG(world >= 1 && world <= px::MAXW && rank >= 0 && rank < world, "peer_create: rank %d / world %d (at most %d ranks)", rank,
                     world, px::MAXW);
    MAAVSS_CHECK_ARG(channels >= 1 && channels <= 8 && max_doubles >= 1 && max_doubles <= (1 << 20), "peer_create: bad channels / size");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle is 64 bytes");
    maavss_peer* p = new maavss_peer();
    p->rank = rank;
    p->world = world;
    p->L = px::make_layout(channels, max_doubles);
    MAAVSS_CUDA(cudaGetDevice(&p->device));
    MAAVSS_CUDA(cudaMalloc((void**)&p->local, p->L.total));
    MAAVSS_CUDA(cudaMemset(p->local, 0, p->L.total));
    MAAVSS_CUDA(cudaDeviceSynchronize());
    cudaIpcMemHandle_t h;
    MAAVSS_CUDA(cudaIpcGetMemHandle(&h, p->local));
    memcpy(handle64, &h, 64);
    memset(&p->ptrs, 0, sizeof(p->ptrs));
    memset(p->opened, 0, sizeof(p->opened));
    p->ptrs.p[rank] = p->local;
    *out = p;
    return MAAVSS_OK;
}

extern "C" int maavss_peer_connect(maavss_peer* p, const void* handles) {
    MAAVSS_CHECK_ARG(p && handles, "peer_connect: null argument");
    for (int r = 0; r < p->world; ++r) {
        if (r == p->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const uint8_t*)handles + 64 * (size_t)r, 64);
        void* ptr = nullptr;
        MAAVSS_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
        p->ptrs.p[r] = (uint8_t*)ptr;
        p->opened[r] = true;
    }
    return MAAVSS_OK;
}

extern "C" int maavss_peer_capacity(const maavss_peer* p) { return p ? p->L.max_doubles : 0; }

extern "C" int maavss_peer_allreduce_sum_f64(maavss_peer* p, int channel, double* data, long long n, void* stream) {
    if (n == 0) return MAAVSS_OK;
    MAAVSS_CHECK_ARG(p && data, "peer_allreduce_sum_f64: null argument");
    MAAVSS_CHECK_ARG(channel >= 0 && channel < p->L.channels, "peer_allreduce_sum_f64: channel %d of %d", channel, p->L.channels);
    MAAVSS_CHECK_ARG(n > 0 && n <= p->L.max_doubles, "peer_allreduce_sum_f64: %lld values exceed the buffer of %d", n, p->L.max_doubles);
    for (int r = 0; r < p->world; ++r) MAAVSS_CHECK_ARG(p->ptrs.p[r] != nullptr, "peer_allreduce_sum_f64: rank %d not connected (maavss_peer_connect)", r);
    px::allreduce_kernel<<<1, px::THREADS, 0, (cudaStream_t)stream>>>(p->ptrs, p->L, p->rank, p->world, channel, data, (int)n);
    MAAVSS_LAUNCH_CHECK();
    return MAAVSS_OK;
}

extern "C" int maavss_peer_status(maavss_peer* p, int* error_out) {
    MAAVSS_CHECK_ARG(p && error_out, "peer_status: null argument");
    int err = 0;
    for (int c = 0; c < p->L.channels; ++c) {
        px::State s;
        MAAVSS_CUDA(cudaMemcpy(&s, p->local + p->L.state_off + c * sizeof(px::State), sizeof(s), cudaMemcpyDeviceToHost));
        err |= s.error;
    }
    *error_out = err;
    return MAAVSS_OK;
}

extern "C" int maavss_peer_destroy(maavss_peer* p) {
    if (!p) return MAAVSS_OK;
    for (int r = 0; r < p->world; ++r)
        if (p->opened[r]) cudaIpcCloseMemHandle(p->ptrs.p[r]);
    cudaFree(p->local);
    delete p;
    return MAAVSS_OK;
}
