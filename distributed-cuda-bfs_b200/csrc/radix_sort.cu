// radix_sort.cu — hand-written stable LSD radix sort of 64-bit keys (8-bit digits) for the CSR
// builder.  Replaces the reference's per-vertex std::vector push_back + flatten
// (readGraphFromFile, bfs.cu:851-875) as the way neighbour lists get grouped by source vertex.
//
// Shape: a fixed grid of G CTAs, each owning one contiguous chunk of the input.  Per digit:
//   1. sort_histogram   per-CTA digit counts (warp-private shared-memory counters)
//   2. sort_scan        one CTA turns the digit-major [256][G] table into global offsets
//   3. sort_scatter     each CTA walks its chunk in order; stable in-tile ranks come from
//                       __match_any_sync (peers with the same digit) + per-warp digit counters
// HBM traffic per pass: 2 reads + 1 write of the key array (algorithmic 24 B/key/pass).
#include "common.cuh"
#include "sort.cuh"

namespace b2 {

constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;
static_assert(kRadix == kBlock, "one thread per digit in the scan steps");

__global__ void __launch_bounds__(kBlock) sort_histogram(const uint64_t* __restrict__ keys, int64_t n, int shift,
                                                         int64_t chunk, uint64_t* __restrict__ table, int G) {
  __shared__ unsigned int cnt[kWarpsPerBlock][kRadix];
  for (int w = 0; w < kWarpsPerBlock; ++w) cnt[w][threadIdx.x] = 0;
  __syncthreads();
  const int64_t begin = (int64_t)blockIdx.x * chunk;
  const int64_t end = min(begin + chunk, n);
  const int warp = threadIdx.x >> 5;
  for (int64_t i = begin + threadIdx.x; i < end; i += kBlock) {
    const unsigned d = (unsigned)(keys[i] >> shift) & (kRadix - 1);
    atomicAdd(&cnt[warp][d], 1u);
  }
  __syncthreads();
  unsigned int total = 0;
  for (int w = 0; w < kWarpsPerBlock; ++w) total += cnt[w][threadIdx.x];
  table[(int64_t)threadIdx.x * G + blockIdx.x] = total;
}

// exclusive scan of `len` uint64 values in place, one CTA of 1024 threads
__global__ void __launch_bounds__(1024) scan_u64_single_block(uint64_t* data, int64_t len, uint64_t* total_out) {
  __shared__ uint64_t warp_sum[32];
  __shared__ uint64_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t base = 0; base < len; base += 1024) {
    const int64_t i = base + threadIdx.x;
    const uint64_t x = i < len ? data[i] : 0;
    uint64_t incl = x;
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      uint64_t s = warp_sum[lane];
      uint64_t si = s;
      for (int o = 1; o < 32; o <<= 1) {
        const uint64_t y = __shfl_up_sync(0xffffffffu, si, o);
        if (lane >= o) si += y;
      }
      warp_sum[lane] = si - s;  // exclusive prefix of warp totals
    }
    __syncthreads();
    const uint64_t carry = carry_s;
    const uint64_t excl = carry + warp_sum[warp] + incl - x;
    if (i < len) data[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = excl + x;
    __syncthreads();
  }
  if (threadIdx.x == 0 && total_out) *total_out = carry_s;
}

__global__ void __launch_bounds__(kBlock) sort_scatter(const uint64_t* __restrict__ in, uint64_t* __restrict__ out,
                                                       int64_t n, int shift, int64_t chunk,
                                                       const uint64_t* __restrict__ table, int G) {
  __shared__ unsigned int wcnt[kWarpsPerBlock][kRadix];
  __shared__ uint64_t base[kRadix];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  base[threadIdx.x] = table[(int64_t)threadIdx.x * G + blockIdx.x];
  for (int w = 0; w < kWarpsPerBlock; ++w) wcnt[w][threadIdx.x] = 0;
  __syncthreads();
  const int64_t begin = (int64_t)blockIdx.x * chunk;
  const int64_t end = min(begin + chunk, n);
  for (int64_t tile = begin; tile < end; tile += kBlock) {
    const int64_t i = tile + threadIdx.x;
    const bool valid = i < end;
    const uint64_t key = valid ? in[i] : 0;
    const unsigned d = valid ? ((unsigned)(key >> shift) & (kRadix - 1)) : kRadix;  // invalid lanes form their own group
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    const unsigned rank = __popc(peers & lt);
    if (valid && rank == 0) wcnt[warp][d] = __popc(peers);
    __syncthreads();
    // thread t owns digit t: exclusive prefix over warps, tile total
    unsigned run = 0;
    for (int w = 0; w < kWarpsPerBlock; ++w) {
      const unsigned c = wcnt[w][threadIdx.x];
      wcnt[w][threadIdx.x] = run;
      run += c;
    }
    __syncthreads();
    if (valid) out[base[d] + wcnt[warp][d] + rank] = key;
    __syncthreads();
    base[threadIdx.x] += run;
    for (int w = 0; w < kWarpsPerBlock; ++w) wcnt[w][threadIdx.x] = 0;
    __syncthreads();
  }
}

int sort_grid_size(int64_t n) {
  const int64_t tiles = (n + kBlock - 1) / kBlock;
  const int64_t g = (int64_t)kSMs * 8;
  return (int)(tiles < g ? (tiles > 0 ? tiles : 1) : g);
}

size_t sort_scratch_bytes(int64_t n) { return (size_t)kRadix * sort_grid_size(n) * sizeof(uint64_t); }

// Sorts keys by bits [0, key_bits).  a holds the input; a/b are ping-ponged; *sorted_out points at
// whichever buffer holds the result.  scratch: sort_scratch_bytes(n) of device memory.
Status radix_sort_u64(uint64_t* a, uint64_t* b, int64_t n, int key_bits, void* scratch, cudaStream_t stream,
                      uint64_t** sorted_out, int64_t* launches) {
  *sorted_out = a;
  if (n <= 1 || key_bits <= 0) return Status();
  const int G = sort_grid_size(n);
  int64_t chunk = (n + G - 1) / G;
  chunk = (chunk + kBlock - 1) / kBlock * kBlock;
  uint64_t* table = reinterpret_cast<uint64_t*>(scratch);
  uint64_t* src = a;
  uint64_t* dst = b;
  for (int shift = 0; shift < key_bits; shift += kRadixBits) {
    sort_histogram<<<G, kBlock, 0, stream>>>(src, n, shift, chunk, table, G);
    scan_u64_single_block<<<1, 1024, 0, stream>>>(table, (int64_t)kRadix * G, nullptr);
    sort_scatter<<<G, kBlock, 0, stream>>>(src, dst, n, shift, chunk, table, G);
    if (launches) *launches += 3;
    uint64_t* t = src; src = dst; dst = t;
  }
  B2_CUDA(cudaGetLastError());
  *sorted_out = src;
  return Status();
}

}  // namespace b2
