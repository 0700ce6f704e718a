// Data-parallel exchange behind the C ABI (maavss_comm_init / maavss_grad_allreduce): one NCCL
// communicator over the GPUs of the NVSwitch box, created from a caller-distributed ncclUniqueId, and the in-place sum
// all-reduce of the flat gradient arena (one collective per optimizer step; the 1/W average is folded into maavss_adam_step's
// grad_scale).  NCCL is resolved at run time (dlopen of the libnccl.so.2 the process already carries -- PyTorch's -- or the
// system one), so the library loads on machines without it and never links a second copy.
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>
#include "common.cuh"
#include "../../include/maavss_b200.h"

namespace {

typedef struct { char internal[128]; } NcclUniqueId;
typedef void* NcclComm;
typedef int (*PFN_GetUniqueId)(NcclUniqueId*);
typedef int (*PFN_CommInitRank)(NcclComm*, int, NcclUniqueId, int);
typedef int (*PFN_CommDestroy)(NcclComm);
typedef int (*PFN_AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef const char* (*PFN_GetErrorString)(int);

struct Nccl {
    void* handle;
    PFN_GetUniqueId get_id;
    PFN_CommInitRank init_rank;
    PFN_CommDestroy destroy;
    PFN_AllReduce all_reduce;
    PFN_GetErrorString err;
};

Nccl* nccl() {
    static Nccl n = {};
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (h) {
            n.handle = h;
            n.get_id = (PFN_GetUniqueId)dlsym(h, "ncclGetUniqueId");
            n.init_rank = (PFN_CommInitRank)dlsym(h, "ncclCommInitRank");
            n.destroy = (PFN_CommDestroy)dlsym(h, "ncclCommDestroy");
            n.all_reduce = (PFN_AllReduce)dlsym(h, "ncclAllReduce");
            n.err = (PFN_GetErrorString)dlsym(h, "ncclGetErrorString");
            if (!n.get_id || !n.init_rank || !n.destroy || !n.all_reduce) n.handle = nullptr;
        }
    }
    return n.handle ? &n : nullptr;
}

#define MAAVSS_NCCL(call)                                                                             \
    do {                                                                                              \
        int r__ = (call);                                                                             \
        if (r__ != 0) {                                                                               \
            maavss_set_error("%s failed: %s", #call, N->err ? N->err(r__) : "nccl error");           \
            return MAAVSS_ERR_CUDA;                                                                   \
        }                                                                                             \
    } while (0)

}  // namespace

struct maavss_comm {
    NcclComm comm;
    int rank, world;
};

extern "C" int maavss_comm_unique_id(void* id128) {
    MAAVSS_CHECK_ARG(id128, "comm_unique_id: null argument");
    Nccl* N = nccl();
    MAAVSS_CHECK_ARG(N, "comm_unique_id: libnccl.so.2 could not be loaded");
    NcclUniqueId id;
    MAAVSS_NCCL(N->get_id(&id));
    memcpy(id128, &id, sizeof(id));
    return MAAVSS_OK;
}

extern "C" int maavss_comm_init(maavss_comm** out, int rank, int world, const void* id128) {
    MAAVSS_CHECK_ARG(out && id128 && world >= 1 && rank >= 0 && rank < world, "comm_init: bad argument");
    Nccl* N = nccl();
    MAAVSS_CHECK_ARG(N, "comm_init: libnccl.so.2 could not be loaded");
    NcclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    maavss_comm* c = (maavss_comm*)calloc(1, sizeof(maavss_comm));
    c->rank = rank;
    c->world = world;
    const int r = N->init_rank(&c->comm, world, id, rank);
    if (r != 0) {
        maavss_set_error("ncclCommInitRank failed: %s", N->err ? N->err(r) : "nccl error");
        free(c);
        return MAAVSS_ERR_CUDA;
    }
    *out = c;
    return MAAVSS_OK;
}

extern "C" void maavss_comm_destroy(maavss_comm* c) {
    if (!c) return;
    Nccl* N = nccl();
    if (N && c->comm) N->destroy(c->comm);
    free(c);
}

extern "C" int maavss_grad_allreduce(maavss_comm* c, float* grad, long long n, void* stream) {
    MAAVSS_CHECK_ARG(c && grad && n >= 0, "grad_allreduce: bad argument");
    if (n == 0 || c->world == 1) return MAAVSS_OK;
    Nccl* N = nccl();
    MAAVSS_NCCL(N->all_reduce(grad, grad, (size_t)n, /*ncclFloat32*/ 7, /*ncclSum*/ 0, c->comm, (cudaStream_t)stream));
    maavss_count_launch();
    return MAAVSS_OK;
}

/* in-place MAX over ranks of `n` uint32 words (non-negative float bits order like integers): the window maxima of
 * utilities.video_phasegram (utilities.py:226-227 is a whole-batch max) */
extern "C" int maavss_allreduce_max_u32(maavss_comm* c, void* words, long long n, void* stream) {
    MAAVSS_CHECK_ARG(c && words && n >= 0, "allreduce_max_u32: bad argument");
    if (n == 0 || c->world == 1) return MAAVSS_OK;
    Nccl* N = nccl();
    MAAVSS_NCCL(N->all_reduce(words, words, (size_t)n, /*ncclUint32*/ 3, /*ncclMax*/ 2, c->comm, (cudaStream_t)stream));
    maavss_count_launch();
    return MAAVSS_OK;
}

/* in-place sum over ranks of `n` doubles: the BatchNorm partial sums between maavss_bn_reduce_grouped and
 * maavss_bn_apply_grouped */
extern "C" int maavss_allreduce_sum_f64(maavss_comm* c, double* v, long long n, void* stream) {
    MAAVSS_CHECK_ARG(c && v && n >= 0, "allreduce_sum_f64: bad argument");
    if (n == 0 || c->world == 1) return MAAVSS_OK;
    Nccl* N = nccl();
    MAAVSS_NCCL(N->all_reduce(v, v, (size_t)n, /*ncclFloat64*/ 8, /*ncclSum*/ 0, c->comm, (cudaStream_t)stream));
    maavss_count_launch();
    return MAAVSS_OK;
}
