// NCCL is bound at run time (dlopen) so that the single-GPU library has no link-time dependency
// on it and so that, inside a torch process, the already-loaded libnccl.so.2 is the one used.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace rrl {

struct NcclUniqueId {
    char internal[128];
};

struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(NcclUniqueId *) = nullptr;
    int (*CommInitRank)(void **comm, int nranks, NcclUniqueId id, int rank) = nullptr;
    int (*CommDestroy)(void *comm) = nullptr;
    int (*AllReduce)(const void *send, void *recv, size_t count, int dtype, int op, void *comm, cudaStream_t s) = nullptr;
    int (*AllGather)(const void *send, void *recv, size_t sendcount, int dtype, void *comm, cudaStream_t s) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};

// ncclDataType_t / ncclRedOp_t values (nccl.h, stable since 2.x)
enum { NCCL_FLOAT32 = 7, NCCL_FLOAT64 = 8, NCCL_SUM = 0 };

// Returns nullptr (and a message) when libnccl.so.2 cannot be loaded.
NcclApi *nccl_api(const char **why);

}  // namespace rrl
