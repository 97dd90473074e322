// bfs.cuh — search interface (bfs_kernels.cu).
#pragma once
#include "common.cuh"

namespace b2 {

struct BfsParams {
  GraphView g;
  SearchView s;
  Tuning tn;
};

Status search_grid_size(int device, int* grid_out);
Status run_fused(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, int64_t* launches);
Status run_stepwise(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, Control* pinned,
                    int64_t* launches);
Status run_reach_stats(const BfsParams& p, const uint32_t* orig_deg, cudaStream_t stream, int64_t* launches);

}  // namespace b2
