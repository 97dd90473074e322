// noalloc.cu — b2bfs_search_noalloc: one top-down search over a CSR the caller already holds in
// device memory, with every byte of scratch (bitmaps, frontier work lists, level counters) taken from
// ONE array that is compiled into this library instead of being allocated.
//
// Who needs this: code that is not allowed to allocate device memory at the moment it wants a
// search — a callback running inside somebody else's memory-pool discipline, a real-time loop whose
// allocations were all done at start-up, a process whose allocator is owned by another framework
// that would see a foreign cudaMalloc as fragmentation (or fail it outright because it has already
// reserved the whole card).  b2bfs_run cannot serve them: loading a graph through a handle
// cudaMalloc's the CSR, the queues and the bitmaps.
//
// The price is a fixed ceiling (B2BFS_NOALLOC_MAX_VERTICES / _MAX_ENTRIES) and kNoallocBytes of
// device memory that every GPU holding this library keeps until the process exits, used or not.
//
// One thing the caller has to plan for: CUDA loads a library's device code lazily.  The FIRST call
// on a device is therefore the moment the driver places this file's kernels and the scratch array
// itself in device memory - a start-up step that takes memory once, exactly like loading any other
// CUDA library does.  Make that first call while the process is still setting up (a one-vertex graph
// is enough); every later call finds everything in place and takes nothing.
#include <mutex>

#include "bfs.cuh"

namespace {

using namespace b2;

constexpr int64_t kMaxN = B2BFS_NOALLOC_MAX_VERTICES;
constexpr int64_t kMaxE = B2BFS_NOALLOC_MAX_ENTRIES;
constexpr size_t up256(size_t x) { return (x + 255) / 256 * 256; }

constexpr size_t kWordBytes = up256((size_t)(kMaxN / 32) * sizeof(uint32_t));
constexpr size_t kQueueBytes = up256((size_t)kMaxN * sizeof(int32_t));
constexpr size_t kWideBytes = up256(((size_t)kMaxN + (size_t)kMaxE / kChunk + 1) * sizeof(int2));
constexpr size_t kOffCtl = 0;
constexpr size_t kOffInit = up256(sizeof(Control));
constexpr size_t kOffVisited = kOffInit + kWordBytes;
constexpr size_t kOffBitmap0 = kOffVisited + kWordBytes;
constexpr size_t kOffBitmap1 = kOffBitmap0 + kWordBytes;
constexpr size_t kOffQueue0 = kOffBitmap1 + kWordBytes;
constexpr size_t kOffQueue1 = kOffQueue0 + kQueueBytes;
constexpr size_t kOffWide0 = kOffQueue1 + kQueueBytes;
constexpr size_t kOffWide1 = kOffWide0 + kWideBytes;
constexpr size_t kNoallocBytes = kOffWide1 + kWideBytes;  // ~7.4 MiB
static_assert(kNoallocBytes < (size_t)8 << 20, "keep the resident scratch under 8 MiB");

// The scratch itself.  In HOST code this name is only a handle for cudaGetSymbolAddress; the device
// address is fetched once per device (below) and passed to the kernels like any other pointer.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocBytes];

// rows of degree 0 (and the padding bits past n) are never claimed nor scanned: same bitmap the
// handle-based path builds at load time (finish_rows_kernel), recomputed here on every call because
// the caller's graph may differ from the previous call's
__global__ void __launch_bounds__(kBlock) noalloc_prepare_kernel(int64_t n, int64_t nwords,
                                                                 const int64_t* __restrict__ rowptr,
                                                                 uint32_t* __restrict__ visited_init) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * kWarpsPerBlock;
  for (int64_t w = (int64_t)blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); w < nwords; w += warps) {
    const int64_t v = w * 32 + lane;
    const bool skip = v >= n || rowptr[v + 1] == rowptr[v];
    const unsigned m = __ballot_sync(0xffffffffu, skip);
    if (lane == 0) visited_init[w] = m;
  }
}

constexpr int kMaxDevices = 64;
struct DeviceSetup {
  bool ready = false;
  unsigned char* scratch = nullptr;  // device address of g_noalloc_scratch on this device
  int grid = 0;
};
DeviceSetup g_setup[kMaxDevices];
std::mutex g_mutex;
thread_local char g_error[256] = "";

int fail(int code, const char* what, const char* detail) {
  snprintf(g_error, sizeof(g_error), "%s%s%s", what, detail ? ": " : "", detail ? detail : "");
  return code;
}

}  // namespace

extern "C" {

const char* b2bfs_noalloc_last_error(void) { return g_error; }

int b2bfs_search_noalloc(int device, int64_t n, int64_t entries, const int64_t* rowptr_dev, const int32_t* col_dev,
                         int64_t src, int32_t* depth_dev, int32_t* parent_dev, void* cuda_stream,
                         int32_t* levels_out) {
  if (!rowptr_dev || !depth_dev || !parent_dev || (entries > 0 && !col_dev))
    return fail(B2BFS_ERR_ARG, "b2bfs_search_noalloc", "null device pointer");
  if (n < 1 || n > kMaxN || entries < 0 || entries > kMaxE) {
    snprintf(g_error, sizeof(g_error),
             "b2bfs_search_noalloc: graph of %lld vertices / %lld entries exceeds the built-in scratch "
             "(at most %lld vertices and %lld entries); use a b2bfs handle instead",
             (long long)n, (long long)entries, (long long)kMaxN, (long long)kMaxE);
    return B2BFS_ERR_RANGE;
  }
  if (src < 0 || src >= n) return fail(B2BFS_ERR_ARG, "b2bfs_search_noalloc", "source vertex outside [0, n)");
  if (((uintptr_t)col_dev & 15) || ((uintptr_t)depth_dev & 15) || ((uintptr_t)parent_dev & 15))
    return fail(B2BFS_ERR_ARG, "b2bfs_search_noalloc", "col, depth and parent must be 16-byte aligned");
  if (device < 0 || device >= kMaxDevices) return fail(B2BFS_ERR_ARG, "b2bfs_search_noalloc", "device index");

  std::lock_guard<std::mutex> lock(g_mutex);  // one shared array: calls are serialised on the host side
  int previous = -1;
  if (cudaGetDevice(&previous) != cudaSuccess) return fail(B2BFS_ERR_CUDA, "cudaGetDevice", "no CUDA device");
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return fail(B2BFS_ERR_CUDA, "cudaSetDevice", cudaGetErrorString(e));

  DeviceSetup& ds = g_setup[device];
  if (!ds.ready) {  // first call on this device: resolve the array's device address, size the grid
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) {
      cudaSetDevice(previous);
      return fail(B2BFS_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch)", cudaGetErrorString(e));
    }
    Status s = search_grid_size(device, &ds.grid);
    if (!s.ok()) {
      cudaSetDevice(previous);
      return fail(s.code, "search_grid_size", "fused kernel does not fit");
    }
    ds.scratch = static_cast<unsigned char*>(p);
    ds.ready = true;
  }

  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  unsigned char* base = ds.scratch;
  BfsParams p{};
  p.g.n = n;
  p.g.n_global = n;
  p.g.entries_global = (unsigned long long)entries;
  p.g.nwords = ((n + 31) / 32 + 3) / 4 * 4;
  p.g.entries = entries;
  p.g.vbase = 0;
  p.g.part = p.g.nwords * 32;
  p.g.rank = 0;
  p.g.world = 1;
  p.g.rowptr = rowptr_dev;
  p.g.col = col_dev;
  p.g.visited_init = reinterpret_cast<const uint32_t*>(base + kOffInit);
  p.g.first_nbr = nullptr;  // only the bottom-up phase reads it, and this entry point never leaves top-down
  p.s.depth = depth_dev;
  p.s.parent = parent_dev;
  p.s.visited = reinterpret_cast<uint32_t*>(base + kOffVisited);
  p.s.bitmap[0] = reinterpret_cast<uint32_t*>(base + kOffBitmap0);
  p.s.bitmap[1] = reinterpret_cast<uint32_t*>(base + kOffBitmap1);
  p.s.queue[0] = reinterpret_cast<int32_t*>(base + kOffQueue0);
  p.s.queue[1] = reinterpret_cast<int32_t*>(base + kOffQueue1);
  p.s.wideq[0] = reinterpret_cast<int2*>(base + kOffWide0);
  p.s.wideq[1] = reinterpret_cast<int2*>(base + kOffWide1);
  p.s.ctl = reinterpret_cast<Control*>(base + kOffCtl);
  p.s.gfrontier = nullptr;
  p.s.seen = nullptr;
  p.s.send_buf = nullptr;
  p.s.send_count = nullptr;
  p.s.send_cap = 0;
  p.tn.mode = B2BFS_MODE_TOP_DOWN;
  p.tn.alpha = 14;
  p.tn.beta = 24;
  p.tn.gamma = 0;
  p.tn.cluster = 0;
  p.tn.huge_min = 0;  // no hub-item lists in the built-in scratch: hub rows stay warp items
  p.tn.xfullpush = 0;

  int rc = B2BFS_OK;
  SearchResult result{};
  do {
    // the array still holds the previous call's counters, bitmaps and queues: the counters are
    // cleared here, the bitmaps by the kernel's own reset phase, the queues are write-before-read
    if ((e = cudaMemsetAsync(p.s.ctl, 0, sizeof(Control), stream)) != cudaSuccess) break;
    const int64_t blocks = (p.g.nwords + kWarpsPerBlock - 1) / kWarpsPerBlock;
    noalloc_prepare_kernel<<<(int)(blocks < 1184 ? blocks : 1184), kBlock, 0, stream>>>(
        n, p.g.nwords, rowptr_dev, reinterpret_cast<uint32_t*>(base + kOffInit));
    if ((e = cudaGetLastError()) != cudaSuccess) break;
    Status s = run_fused(p, src, ds.grid, stream, nullptr);
    if (!s.ok()) {
      rc = fail(s.code, "bfs_fused_kernel launch", "cooperative launch failed");
      break;
    }
    if ((e = cudaMemcpyAsync(&result, &p.s.ctl->result, sizeof(result), cudaMemcpyDeviceToHost, stream)) != cudaSuccess) break;
    e = cudaStreamSynchronize(stream);
  } while (false);
  if (rc == B2BFS_OK && e != cudaSuccess) rc = fail(B2BFS_ERR_CUDA, "b2bfs_search_noalloc", cudaGetErrorString(e));
  if (rc == B2BFS_OK && levels_out) *levels_out = result.levels;
  cudaSetDevice(previous);
  return rc;
}

}  // extern "C"
