// comm.cu — data-parallel plumbing: a thin NCCL wrapper for the two collectives of the PPO update
// (per-minibatch gradient allreduce, per-epoch advantage-moment allreduce).  The
// reference has no collectives at all (single process, single device: PPO/PPO_Discrete.cpp:65).
//
// libnccl is resolved at run time with dlopen (the copy torch already loaded, else the system one),
// so libppo_b200.so has no link-time NCCL dependency and single-GPU use never touches it.
#include <dlfcn.h>
#include <string.h>

#include "common.cuh"

namespace ppo {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclFloat32 = 7, ncclFloat64 = 8 };  // nccl.h ncclDataType_t
enum { ncclSum = 0 };

struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
};

static NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.handle ? &api : nullptr;
    tried = true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
        api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) {
        set_error("NCCL not found: %s", dlerror());
        return nullptr;
    }
    api.GetUniqueId = (int (*)(ncclUniqueId*))dlsym(api.handle, "ncclGetUniqueId");
    api.CommInitRank = (int (*)(ncclComm_t*, int, ncclUniqueId, int))dlsym(api.handle, "ncclCommInitRank");
    api.CommDestroy = (int (*)(ncclComm_t))dlsym(api.handle, "ncclCommDestroy");
    api.AllReduce = (int (*)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t))dlsym(api.handle,
                                                                                                   "ncclAllReduce");
    api.GetErrorString = (const char* (*)(int))dlsym(api.handle, "ncclGetErrorString");
    api.GetVersion = (int (*)(int*))dlsym(api.handle, "ncclGetVersion");
    if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllReduce) {
        set_error("NCCL library lacks a required symbol");
        api.handle = nullptr;
        return nullptr;
    }
    return &api;
}

}  // namespace ppo

using namespace ppo;

struct ppo_comm {
    ncclComm_t comm;
    int rank, world;
};

extern "C" int ppo_comm_unique_id(char* h_id128) {
    NcclApi* api = nccl_api();
    if (!api) return PPO_ERR_COMM;
    PPO_REQUIRE(h_id128, "ppo_comm_unique_id: null output");
    ncclUniqueId id;
    const int rc = api->GetUniqueId(&id);
    if (rc != ncclSuccess) {
        set_error("ncclGetUniqueId failed: %s", api->GetErrorString ? api->GetErrorString(rc) : "?");
        return PPO_ERR_COMM;
    }
    memcpy(h_id128, id.internal, 128);
    return PPO_OK;
}

extern "C" ppo_comm* ppo_comm_init(const char* h_id128, int rank, int world_size) {
    NcclApi* api = nccl_api();
    if (!api) return nullptr;
    if (!h_id128 || rank < 0 || rank >= world_size) {
        set_error("ppo_comm_init: bad arguments (rank %d of %d)", rank, world_size);
        return nullptr;
    }
    ncclUniqueId id;
    memcpy(id.internal, h_id128, 128);
    ncclComm_t c = nullptr;
    const int rc = api->CommInitRank(&c, world_size, id, rank);
    if (rc != ncclSuccess) {
        set_error("ncclCommInitRank failed: %s", api->GetErrorString ? api->GetErrorString(rc) : "?");
        return nullptr;
    }
    ppo_comm* h = new ppo_comm{c, rank, world_size};
    return h;
}

extern "C" void ppo_comm_destroy(ppo_comm* c) {
    if (!c) return;
    NcclApi* api = nccl_api();
    if (api && c->comm) api->CommDestroy(c->comm);
    delete c;
}

extern "C" int ppo_comm_world_size(const ppo_comm* c) { return c ? c->world : 1; }
extern "C" int ppo_comm_rank(const ppo_comm* c) { return c ? c->rank : 0; }

static int allreduce(ppo_comm* c, void* buf, int64_t count, int dtype, ppo_stream_t stream) {
    PPO_REQUIRE(c && buf && count >= 0, "ppo_comm_allreduce: bad arguments");
    if (count == 0 || c->world == 1) return PPO_OK;
    NcclApi* api = nccl_api();
    if (!api) return PPO_ERR_COMM;
    const int rc = api->AllReduce(buf, buf, (size_t)count, dtype, ncclSum, c->comm, as_stream(stream));
    if (rc != ncclSuccess) {
        set_error("ncclAllReduce failed: %s", api->GetErrorString ? api->GetErrorString(rc) : "?");
        return PPO_ERR_COMM;
    }
    return PPO_OK;
}
extern "C" int ppo_comm_allreduce_f32(ppo_comm* c, float* buf, int64_t count, ppo_stream_t stream) {
    return allreduce(c, buf, count, ncclFloat32, stream);
}
extern "C" int ppo_comm_allreduce_f64(ppo_comm* c, double* buf, int64_t count, ppo_stream_t stream) {
    return allreduce(c, buf, count, ncclFloat64, stream);
}
