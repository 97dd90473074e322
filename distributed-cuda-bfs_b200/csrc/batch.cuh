// batch.cuh — bit-parallel multi-source search (batch_bfs.cu).
#pragma once
#include "common.cuh"

namespace b2 {

struct BatchState {
  uint64_t *seen = nullptr, *fa = nullptr, *fb = nullptr;  // n words each
  int32_t* depth = nullptr;                                // depth_rows x n
  int depth_rows = 0;
  int32_t* wide_list = nullptr;                            // vertices whose rows are >= kNarrow entries
  int2* rest = nullptr;                                    // (vertex, chunk >= 1) items of those rows
  int64_t n_wide = 0, n_rest = 0;
  unsigned long long* counters = nullptr;                  // device: [0..1] list build, [8..15] per-level finds, [16..143] V_r/D_r
  unsigned long long* host = nullptr;                      // page-locked mirror
  int64_t* roots_dev = nullptr;
};

void free_batch(BatchState* bs);
Status prepare_batch(BatchState* bs, int64_t n, int64_t wide_items_cap, const int64_t* rowptr, int nroots,
                     cudaStream_t stream, int64_t* launches);
Status run_batch(BatchState* bs, int64_t n, const int64_t* rowptr, const int32_t* col, int nroots,
                 const int64_t* roots_host, int grid, cudaStream_t stream, cudaEvent_t ev0, cudaEvent_t ev1,
                 int* levels_out, int64_t* launches);
Status batch_reach_stats(BatchState* bs, int64_t n, const uint32_t* orig_deg, cudaStream_t stream,
                         unsigned long long* vr_dr_host, int64_t* launches);

}  // namespace b2
