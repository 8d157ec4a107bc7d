// K4 — gradient all-reduce across ranks (one process per GPU).
//
// The reference is single-process; its data-parallel precedent (GPT_SGLD_p.jl:70-108,194-218) splits the
// minibatch rows over Julia workers that write into SharedArrays.  Here each rank owns one B200, computes
// the gradient of its share of the minibatch rows and the [gradw | gradU] buffer (Q + n*r*D doubles) is
// summed with ncclAllReduce over NVLink 5 / NVSwitch, in place, on the handle's stream.
//
// NCCL is bound at run time (dlopen of libnccl.so.2 — inside a PyTorch process that resolves to the copy
// torch already loaded), so libgptb200.so itself has no link-time NCCL dependency and single-GPU users
// need no NCCL at all.
#include <dlfcn.h>
#include <string.h>

#include "common.cuh"

namespace {

typedef struct { char internal[128]; } ncclUniqueId_t;
typedef void *ncclComm_p;
typedef int ncclResult;
enum { kNcclSum = 0, kNcclFloat32 = 7, kNcclFloat64 = 8 };

struct NcclApi {
    void *lib = nullptr;
    ncclResult (*GetUniqueId)(ncclUniqueId_t *) = nullptr;
    ncclResult (*CommInitRank)(ncclComm_p *, int, ncclUniqueId_t, int) = nullptr;
    ncclResult (*CommInitAll)(ncclComm_p *, int, const int *) = nullptr;
    ncclResult (*CommDestroy)(ncclComm_p) = nullptr;
    ncclResult (*CommAbort)(ncclComm_p) = nullptr;
    ncclResult (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_p, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult) = nullptr;
    bool ok = false;
};

NcclApi &api() {
    static NcclApi a;
    if (a.lib) return a;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
        a.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (a.lib) break;
    }
    if (!a.lib) return a;
    a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(a.lib, "ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))dlsym(a.lib, "ncclCommInitRank");
    a.CommInitAll = (decltype(a.CommInitAll))dlsym(a.lib, "ncclCommInitAll");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(a.lib, "ncclCommDestroy");
    a.CommAbort = (decltype(a.CommAbort))dlsym(a.lib, "ncclCommAbort");
    a.AllReduce = (decltype(a.AllReduce))dlsym(a.lib, "ncclAllReduce");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(a.lib, "ncclGetErrorString");
    a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllReduce && a.GetErrorString;
    return a;
}

}  // namespace

extern "C" int gptb_comm_unique_id(void *id_out_128) {
    if (!id_out_128) return GPTB_ERR_ARG;
    NcclApi &a = api();
    if (!a.ok) {
        gptb_set_error(nullptr, "NCCL not found (dlopen libnccl.so.2 failed: %s)", dlerror());
        return GPTB_ERR_NCCL;
    }
    ncclUniqueId_t id;
    ncclResult r = a.GetUniqueId(&id);
    if (r != 0) {
        gptb_set_error(nullptr, "ncclGetUniqueId: %s", a.GetErrorString(r));
        return GPTB_ERR_NCCL;
    }
    memcpy(id_out_128, &id, 128);
    return GPTB_OK;
}

extern "C" int gptb_comm_init(gptb_handle *h, int32_t nranks, int32_t rank, const void *id_128) {
    if (!h || !id_128 || nranks < 1 || rank < 0 || rank >= nranks) return GPTB_ERR_ARG;
    if (h->nccl_comm) {
        gptb_set_error(h, "communicator already initialised");
        return GPTB_ERR_STATE;
    }
    if (nranks == 1) {
        h->nranks = 1;
        h->rank = 0;
        return GPTB_OK;
    }
    NcclApi &a = api();
    if (!a.ok) {
        gptb_set_error(h, "NCCL not found (dlopen libnccl.so.2 failed)");
        return GPTB_ERR_NCCL;
    }
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    ncclUniqueId_t id;
    memcpy(&id, id_128, 128);
    ncclComm_p comm = nullptr;
    ncclResult r = a.CommInitRank(&comm, nranks, id, rank);
    if (r != 0) {
        gptb_set_error(h, "ncclCommInitRank: %s", a.GetErrorString(r));
        return GPTB_ERR_NCCL;
    }
    h->nccl_comm = comm;
    h->nranks = nranks;
    h->rank = rank;
    return GPTB_OK;
}

// Single-process multi-GPU: the n handles (on n different devices of this process) become ranks 0..n-1 of one
// communicator.  Calls that contain a collective must then be made for all of them concurrently: multi.cu does that
// with one host thread per handle.
extern "C" int gptb_comm_init_local(gptb_handle *const *hs, int32_t n) {
    if (!hs || n < 1 || n > 64) return GPTB_ERR_ARG;
    for (int i = 0; i < n; ++i) {
        if (!hs[i]) return GPTB_ERR_ARG;
        if (hs[i]->nccl_comm) {
            gptb_set_error(hs[i], "communicator already initialised");
            return GPTB_ERR_STATE;
        }
        for (int j = 0; j < i; ++j)
            if (hs[j]->device == hs[i]->device) {
                gptb_set_error(hs[0], "gptb_comm_init_local: handles %d and %d share device %d", j, i, hs[i]->device);
                return GPTB_ERR_ARG;
            }
    }
    if (n == 1) {
        hs[0]->nranks = 1;
        hs[0]->rank = 0;
        return GPTB_OK;
    }
    NcclApi &a = api();
    if (!a.ok || !a.CommInitAll) {
        gptb_set_error(hs[0], "NCCL not found (dlopen libnccl.so.2 failed)");
        return GPTB_ERR_NCCL;
    }
    int devs[64];
    ncclComm_p comms[64];
    for (int i = 0; i < n; ++i) devs[i] = hs[i]->device;
    ncclResult r = a.CommInitAll(comms, n, devs);
    if (r != 0) {
        gptb_set_error(hs[0], "ncclCommInitAll: %s", a.GetErrorString(r));
        return GPTB_ERR_NCCL;
    }
    for (int i = 0; i < n; ++i) {
        hs[i]->nccl_comm = comms[i];
        hs[i]->nranks = n;
        hs[i]->rank = i;
    }
    return GPTB_OK;
}

// A rank of a single-process group failed outside a collective: abort every communicator of the group, so that the
// ranks waiting inside an all-reduce return instead of waiting for ever (multi.cu).  The handles fall back to one rank.
void gptb_comm_abort_group(gptb_handle *const *hs, int n) {
    NcclApi &a = api();
    for (int i = 0; i < n; ++i) {
        if (hs[i]->nccl_comm && a.CommAbort) a.CommAbort((ncclComm_p)hs[i]->nccl_comm);
        hs[i]->nccl_comm = nullptr;
        hs[i]->nranks = 1;
        hs[i]->rank = 0;
    }
}

extern "C" int gptb_comm_destroy(gptb_handle *h) {
    if (!h) return GPTB_ERR_ARG;
    if (h->nccl_comm) {
        api().CommDestroy((ncclComm_p)h->nccl_comm);
        h->nccl_comm = nullptr;
    }
    h->nranks = 1;
    h->rank = 0;
    return GPTB_OK;
}

// Sum the gradient buffer over ranks, in place (no-op for one rank).
int gptb_allreduce_grad(gptb_handle *h) {
    if (h->nranks <= 1) return GPTB_OK;
    NcclApi &a = api();
    size_t count = (size_t)h->cfg.Q + (size_t)h->cfg.n * h->cfg.r * h->cfg.D;
    ncclResult r = a.AllReduce(h->grad, h->grad, count, kNcclFloat64, kNcclSum, (ncclComm_p)h->nccl_comm, h->stream);
    if (r != 0) {
        gptb_set_error(h, "ncclAllReduce: %s", a.GetErrorString(r));
        return GPTB_ERR_NCCL;
    }
    h->launches++;
    return GPTB_OK;
}

// The same for a slice of the gradient buffer on a given stream (the dimension-group pipeline reduces each group's
// slice on its own NCCL stream, in the same order on every rank).
int gptb_allreduce_range(gptb_handle *h, double *buf, size_t count, cudaStream_t stream) {
    if (h->nranks <= 1 || count == 0) return GPTB_OK;
    NcclApi &a = api();
    ncclResult r = a.AllReduce(buf, buf, count, kNcclFloat64, kNcclSum, (ncclComm_p)h->nccl_comm, stream);
    if (r != 0) {
        gptb_set_error(h, "ncclAllReduce: %s", a.GetErrorString(r));
        return GPTB_ERR_NCCL;
    }
    h->launches++;
    return GPTB_OK;
}

int gptb_allreduce_range_f32(gptb_handle *h, float *buf, size_t count, cudaStream_t stream) {
    if (h->nranks <= 1 || count == 0) return GPTB_OK;
    NcclApi &a = api();
    ncclResult r = a.AllReduce(buf, buf, count, kNcclFloat32, kNcclSum, (ncclComm_p)h->nccl_comm, stream);
    if (r != 0) {
        gptb_set_error(h, "ncclAllReduce: %s", a.GetErrorString(r));
        return GPTB_ERR_NCCL;
    }
    h->launches++;
    return GPTB_OK;
}
