// multi.cuh — multi-rank search driver (multi.cu).
#pragma once
#include "bfs.cuh"
#include "nccl_dl.hpp"

namespace b2 {

struct MultiContext {
  const NcclApi* nccl = nullptr;
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  int64_t n_global = 0;
  int64_t entries_global = 0;      // resident CSR entries summed over ranks
  int2* recv_buf = nullptr;        // (world - 1) * part pairs at most
  unsigned long long* stats_send = nullptr;  // 16 words
  unsigned long long* stats_all = nullptr;   // world * 16 words
  unsigned long long* stats_host = nullptr;  // pinned, world * 16 words
};

struct MultiResult {
  int levels = 0, td_levels = 0, bu_levels = 0;
  int64_t launches = 0;
  int64_t nccl_calls = 0;
  int64_t exchanged_bytes = 0;     // payload this rank sent in id exchanges + its share of bitmap allgathers
};

// src is a GLOBAL vertex id; every rank calls this with the same src
Status run_multi(const BfsParams& p, const MultiContext& mc, int64_t src, int grid, cudaStream_t stream,
                 MultiResult* out);

#define B2_NCCL(mc, expr)                                                                                     \
  do {                                                                                                        \
    ncclResult_t _r = (expr);                                                                                 \
    if (_r != ncclSuccess)                                                                                    \
      return b2::make_error(B2BFS_ERR_COMM, std::string(#expr) + ": " + (mc).nccl->GetErrorString(_r));       \
  } while (0)

}  // namespace b2
