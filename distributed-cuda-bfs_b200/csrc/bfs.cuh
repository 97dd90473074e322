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
Status search_cluster_grid_size(int plain_grid, int* grid_out, int* size_out);
Status run_fused(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, int64_t* launches);
Status run_stepwise(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, Control* pinned,
                    int64_t* launches);
// launchers used by the multi-rank host loop (multi.cu)
void launch_init(const BfsParams& p, int64_t src_local, int grid, cudaStream_t s);
void launch_top_down_multi(const BfsParams& p, const LoopState& st, int grid, cudaStream_t s);
void launch_apply_remote(const BfsParams& p, const LoopState& st, const int2* recv, int64_t count, int grid,
                         cudaStream_t s);
void launch_bottom_up(const BfsParams& p, const LoopState& st, int grid, cudaStream_t s);
void launch_queue_to_bitmap(const BfsParams& p, const LoopState& st, int grid, cudaStream_t s);
void launch_bitmap_to_queue(const BfsParams& p, const LoopState& st, int grid, cudaStream_t s);
void launch_pack_stats(const BfsParams& p, int slot, unsigned long long* out, cudaStream_t s);

Status run_fused_multi(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, int64_t* launches);
Status multi_grid_size(int device, int* grid_out);

// the reference's alternative strategies as comparator engines (comparators.cu)
Status run_scan_engine(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, Control* pinned,
                       unsigned int* counts, unsigned int* block_totals, int64_t* launches);
Status run_pull_engine(const BfsParams& p, int64_t src, int grid, cudaStream_t stream, Control* pinned,
                       int* changed_dev, int64_t* launches);

Status run_reach_stats(const BfsParams& p, const uint32_t* orig_deg, cudaStream_t stream, int64_t* launches);

}  // namespace b2
