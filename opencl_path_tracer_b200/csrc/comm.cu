// comm.cu — the path's one real exchange step behind the C-ABI: SUM of the float accumulation buffers at a root over
// NCCL (NVLink 5 / NVSwitch), for hosts that run one PROCESS per GPU.  The reference has no counterpart
// (one device, one queue: main.cpp:466,476,506).  NCCL is bound at run time (dlopen of libnccl.so.2 — the copy a host
// framework such as torch has already loaded, else the system's), so the library has no link-time dependency on it and a
// single-GPU host never touches it.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>          // types and NCCL_UNIQUE_ID_BYTES only
#include <stdio.h>
#include <string.h>

#include "comm.h"

namespace {
struct Api {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
Api A;
ncclComm_t g_comm = nullptr;
int g_rank = 0, g_n = 1;
int *g_token = nullptr;            // one device int for the barrier's all-reduce

int load(char *err, size_t cap) {
    if (A.lib) return 0;
    static_assert(NCCL_UNIQUE_ID_BYTES == COMM_ID_BYTES, "NCCL unique id size");
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { snprintf(err, cap, "comm: cannot load libnccl.so.2 (%s)", dlerror()); return -1; }
#define SYM(field, name) do { *(void **)(&A.field) = dlsym(h, name); if (!A.field) { snprintf(err, cap, "comm: %s missing in libnccl", name); dlclose(h); return -1; } } while (0)
    SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank"); SYM(CommDestroy, "ncclCommDestroy");
    SYM(Reduce, "ncclReduce"); SYM(AllReduce, "ncclAllReduce"); SYM(Broadcast, "ncclBroadcast"); SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    A.lib = h;
    return 0;
}
#define NCCL_OK(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) { snprintf(err, cap, "comm: %s: %s", #call, A.GetErrorString(r_)); return -1; } } while (0)
}  // namespace

int comm_unique_id(void *out, char *err, size_t cap) {
    if (load(err, cap)) return -1;
    ncclUniqueId id;
    NCCL_OK(A.GetUniqueId(&id));
    memcpy(out, &id, sizeof id);
    return 0;
}
int comm_init(int rank, int nranks, const void *id_bytes, char *err, size_t cap) {
    if (load(err, cap)) return -1;
    if (g_comm) { snprintf(err, cap, "comm: already initialised (%d ranks)", g_n); return -1; }
    ncclUniqueId id;
    memcpy(&id, id_bytes, sizeof id);
    NCCL_OK(A.CommInitRank(&g_comm, nranks, id, rank));
    g_rank = rank; g_n = nranks;
    if (cudaMalloc((void **)&g_token, sizeof(int)) != cudaSuccess || cudaMemset(g_token, 0, sizeof(int)) != cudaSuccess) {
        snprintf(err, cap, "comm: cudaMalloc of the barrier token failed");
        return -1;
    }
    return 0;
}
int comm_ranks(void) { return g_comm ? g_n : 1; }
int comm_rank(void) { return g_comm ? g_rank : 0; }
int comm_reduce_sum(float *buf, size_t count, int root, cudaStream_t st, char *err, size_t cap) {
    if (!g_comm) { snprintf(err, cap, "comm: ptb_comm_init has not been called"); return -1; }
    NCCL_OK(A.Reduce(buf, buf, count, ncclFloat, ncclSum, root, g_comm, st));
    return 0;
}
int comm_broadcast(void *buf, size_t bytes, int root, cudaStream_t st, char *err, size_t cap) {
    if (!g_comm) { snprintf(err, cap, "comm: ptb_comm_init has not been called"); return -1; }
    NCCL_OK(A.Broadcast(buf, buf, bytes, ncclChar, root, g_comm, st));
    return 0;
}
int comm_barrier(cudaStream_t st, char *err, size_t cap) {
    if (!g_comm) return 0;
    NCCL_OK(A.AllReduce(g_token, g_token, 1, ncclInt, ncclSum, g_comm, st));
    if (cudaStreamSynchronize(st) != cudaSuccess) { snprintf(err, cap, "comm: barrier: %s", cudaGetErrorString(cudaGetLastError())); return -1; }
    return 0;
}
int comm_destroy(void) {
    if (g_comm) A.CommDestroy(g_comm);
    g_comm = nullptr; g_n = 1; g_rank = 0;
    cudaFree(g_token);
    g_token = nullptr;
    return 0;
}
