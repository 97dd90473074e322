// sort.cuh — device primitives shared by the CSR builder.
#pragma once
#include "common.cuh"

namespace b2 {

int sort_grid_size(int64_t n);
size_t sort_scratch_bytes(int64_t n);
Status radix_sort_u64(uint64_t* a, uint64_t* b, int64_t n, int key_bits, void* scratch, cudaStream_t stream,
                      uint64_t** sorted_out, int64_t* launches);

__global__ void scan_u64_single_block(uint64_t* data, int64_t len, uint64_t* total_out);

}  // namespace b2
