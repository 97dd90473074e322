// build.cuh — CSR builder interface (csr_build.cu).
#pragma once
#include "common.cuh"

namespace b2 {

struct BuiltCsr {
  int64_t nrows = 0;         // rows resident on this rank
  int64_t nwords = 0;        // bitmap words covering nrows (multiple of 4)
  int64_t entries = 0;       // resident col entries (after dedup)
  int64_t orig_entries = 0;  // directed entries before dedup (= the reference's numEdges for these rows)
  int64_t wide_items = 0;    // sum over rows with degree >= kNarrow of ceil(degree / kChunk)
  int64_t max_degree = 0;
  int64_t* rowptr = nullptr;        // nrows + 1
  int32_t* col = nullptr;           // entries (+64 padding ints); inside a row: hubs first (descending
                                    // floor(log2 degree)), then ascending id
  uint32_t* orig_deg = nullptr;     // nrows: the reference's edgesSize[]
  uint32_t* visited_init = nullptr; // nwords
  int32_t* first_nbr = nullptr;     // nwords * 32
};

Status generate_edges_device(int kind, int scale, int ef, uint64_t seed, int64_t first, int64_t count, int32_t* src,
                             int32_t* dst, cudaStream_t stream);

Status build_csr_device(const int32_t* src, const int32_t* dst, int64_t m, int64_t n, int64_t row_begin,
                        int64_t row_end, bool partitioned, int dedup, int hub_first, int directed,
                        int64_t nwords_override,
                        cudaStream_t stream,
                        BuiltCsr* out, int64_t* launches);

Status row_degrees_device(const int64_t* rowptr, int64_t nrows, int32_t* deg, cudaStream_t stream);

void free_csr(BuiltCsr* c);

}  // namespace b2
