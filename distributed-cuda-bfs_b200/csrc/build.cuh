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

// how a (row, neighbour) entry is packed into a 64-bit sort key: row << rshift | degree class << vbits | neighbour
struct KeySpec {
  int vbits, cbits, rshift;
};
// max_rows_per_rank: n on one GPU, `part` on several (every rank must derive the same layout)
KeySpec key_spec(int64_t n, int64_t max_rows_per_rank, int hub_first, int directed);

// one GPU: tuples -> CSR
Status build_csr_device(const int32_t* src, const int32_t* dst, int64_t m, int64_t n, int dedup, int hub_first, int directed,
                        cudaStream_t stream, BuiltCsr* out, int64_t* launches);

// several GPUs, step by step (api.cu drives the exchange between the steps):
Status count_degrees_device(const int32_t* src, const int32_t* dst, int64_t m, int64_t n, uint32_t* degc, cudaStream_t stream);
// write == false: cursors[d] += keys this rank's tuples yield for owner d; write == true: the keys go to keys[cursors[d]++]
Status owner_keys_device(bool write, const int32_t* src, const int32_t* dst, int64_t m, int64_t n, int64_t part, int world,
                         const KeySpec& ks, const uint32_t* degc, unsigned long long* cursors, uint64_t* keys, int* bad_flag,
                         cudaStream_t stream);
// N keys of rows [row_begin, row_end) (row relative to row_begin) -> CSR; takes ownership of ka
Status build_csr_from_keys(uint64_t* ka, int64_t N, int64_t n, int64_t row_begin, int64_t row_end, const KeySpec& ks, int dedup,
                           int directed, int64_t nwords_override, cudaStream_t stream, BuiltCsr* out, int64_t* launches);

Status row_degrees_device(const int64_t* rowptr, int64_t nrows, int32_t* deg, cudaStream_t stream);

void free_csr(BuiltCsr* c);

}  // namespace b2
