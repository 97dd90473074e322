// nccl_dl.hpp — NCCL resolved at run time (dlopen) so that single-GPU users of libb2bfs.so need no
// NCCL at all, and so that inside a process that already loaded a libnccl.so.2 (PyTorch bundles one)
// the same copy is reused.  Replaces the reference's CUDA-aware Open MPI dependency
// (bfs_mpi.cu:2, README.md:19-22).
#pragma once
#include <nccl.h>

namespace b2 {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

// nullptr when no libnccl.so.2 can be loaded (err gets the dlerror text)
const NcclApi* nccl_api(const char** err);

}  // namespace b2
