// Run-time binding to NCCL (dlopen) so libjcdf.so has no link-time NCCL dependency: inside a torchrun process it
// reuses the libnccl.so.2 that torch already mapped, in a Julia process it loads the system library.
#pragma once
#include <cstddef>
#include <cuda_runtime.h>

namespace jcdf {

struct NcclApi {
    bool ok = false;
    const char* why = "";
    // minimal ABI-stable subset of nccl.h
    typedef struct { char internal[128]; } UniqueId;
    int (*GetUniqueId)(UniqueId*) = nullptr;
    int (*CommInitRank)(void** comm, int nranks, UniqueId id, int rank) = nullptr;
    int (*CommInitAll)(void** comms, int ndev, const int* devlist) = nullptr;
    int (*CommDestroy)(void* comm) = nullptr;
    int (*AllReduce)(const void* send, void* recv, size_t count, int dtype, int op, void* comm, cudaStream_t st) = nullptr;
    int (*Broadcast)(const void* send, void* recv, size_t count, int dtype, int root, void* comm, cudaStream_t st) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
};
constexpr int NCCL_DOUBLE = 8;  // ncclFloat64
constexpr int NCCL_SUM = 0;     // ncclSum

const NcclApi& nccl();

}  // namespace jcdf
