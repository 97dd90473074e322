/*
 * b2bfs.h — C ABI of the B200-native level-synchronous BFS (libb2bfs.so).
 *
 * This is the drop-in boundary for the ONE hot path of xxcclong/Distributed-CUDA-BFS
 * (SURVEY.md §8): edge list -> CSR in HBM -> BFS from a source vertex -> per-vertex int32 depth
 * (INT_MAX = unreachable).  The reference has no FFI; the seam a maintainer would cut is the pair
 * initCuda2(Graph&) / runCudaQueueBfs(int, Graph&, vector<int>&, vector<int>&) that its main calls
 * (bfs.cu:804, 814).  Each entry point below names the reference code it stands in for.  All
 * citations are into the reference tree (bfs.cu / bfs_mpi.cu).
 *
 * Conventions: plain pointers and sizes only; every function returns 0 (B2BFS_OK) or a negative
 * B2BFS_ERR_* code and leaves a message retrievable with b2bfs_last_error().  A handle is not
 * thread-safe; it owns all device memory, streams and (multi-rank) the NCCL communicator.  Host
 * output buffers are caller-owned.  There is no CPU fallback: without a CUDA device every compute
 * call fails with B2BFS_ERR_CUDA.
 */
#ifndef B2BFS_H
#define B2BFS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2bfs b2bfs_t;

#define B2BFS_OK 0
#define B2BFS_ERR_ARG -1         /* null pointer, bad option, bad source vertex            */
#define B2BFS_ERR_IO -2          /* file cannot be opened / mapped                          */
#define B2BFS_ERR_FORMAT -3      /* malformed edge list (reference: undefined behaviour)   */
#define B2BFS_ERR_RANGE -4       /* vertex id >= vertex_num (reference: out-of-bounds)     */
#define B2BFS_ERR_CUDA -5        /* any CUDA runtime failure, incl. "no device"            */
#define B2BFS_ERR_NOMEM -6
#define B2BFS_ERR_STATE -7       /* e.g. run before a graph is resident                     */
#define B2BFS_ERR_COMM -8        /* NCCL failure                                            */
#define B2BFS_ERR_UNSUPPORTED -9

/* synthetic graph kinds (include/b2bfs_graphgen.h defines the edge streams) */
#define B2BFS_KIND_RMAT 0
#define B2BFS_KIND_KRON 1
#define B2BFS_KIND_GRID 2

/* option "mode" */
#define B2BFS_MODE_DIRECTION_OPT 0 /* top-down + bottom-up with the switch heuristic (default)     */
#define B2BFS_MODE_TOP_DOWN 1      /* top-down only: the reference's algorithm (queueBfs)           */
/* option "engine" */
#define B2BFS_ENGINE_FUSED 0       /* whole search in ONE persistent cooperative kernel (default)  */
#define B2BFS_ENGINE_STEPWISE 1    /* one launch per phase, host reads counters per level           */
/* comparators: the reference's two alternative strategies, dead code in its main
 * (bfs.cu:805-819), built here so the engines above can be measured against them; single GPU,
 * top-down by nature, option "mode" is ignored */
#define B2BFS_ENGINE_SCAN 2        /* runCudaScanBfs, bfs.cu:748-781: next queue by prefix sums, no tail atomic */
#define B2BFS_ENGINE_PULL 3        /* multiBfs, bfs.cu:101-130, 475-539: all vertices inspected every level     */

#define B2BFS_UNREACHED_DEPTH 2147483647 /* INT_MAX, bfs.cu:404, 793 */
#define B2BFS_UNREACHED_PARENT (-1)

typedef struct b2bfs_stats {
  int32_t levels;            /* BFS levels executed = eccentricity of the source + 1               */
  int32_t td_levels;         /* of which top-down                                                   */
  int32_t bu_levels;         /* of which bottom-up                                                  */
  int32_t reserved0;
  int64_t reached_vertices;  /* V_r = #{v : depth[v] != INT_MAX}                                    */
  int64_t reached_entries;   /* D_r = sum over reached v of the reference's edgesSize[v] = 2*m_r    */
  int64_t kernel_launches;   /* kernels this call launched                                          */
  int64_t h2d_bytes;         /* host->device bytes moved by this call                               */
  int64_t d2h_bytes;         /* device->host bytes moved by this call                               */
  double search_ms;          /* device time, CUDA events: state reset .. depth complete in HBM     */
  double total_ms;           /* host wall clock of the whole call incl. copies                      */
} b2bfs_stats;

/* Library version string. */
const char* b2bfs_version(void);

/* Create a single-GPU handle on CUDA device `device`.
 * Replaces the device set-up half of initCuda2 (bfs.cu:308-332: cudaSetDevice, peer enable). */
int b2bfs_create(b2bfs_t** h, int device);

/* Create one rank of a `world`-rank job (one process per GPU).  `nccl_id` is the 128-byte
 * ncclUniqueId every rank received from rank 0 (b2bfs_comm_unique_id), over any side channel.
 * Replaces MPI_Init / MPI_Comm_rank / MPI_Comm_size (bfs_mpi.cu:800-808) and the hard-wired
 * 2-device layout (`(i+1)%2`, bfs_mpi.cu:576, 615). */
int b2bfs_create_rank(b2bfs_t** h, int device, int rank, int world, const void* nccl_id);
int b2bfs_comm_unique_id(void* out128);

/* Options, set before loading a graph unless stated: "dedup" (1: drop duplicate entries and
 * self-loops while building the CSR — depths do not change; 0: keep the reference's multigraph,
 * bfs.cu:860-861), "hub_first" (1: inside a row neighbours are ordered by descending
 * floor(log2 degree), then ascending id, so that bottom-up meets a hub first; 0: ascending id),
 * "multi_engine" (multi-rank: 0 peer-memory kernel, 1 NCCL per level), "mode", "engine" (these
 * three may change between searches), "alpha", "beta" (switch
 * heuristic), "gamma" (a top-down level whose frontier degree sum times gamma exceeds the
 * unexplored entries / alpha only claims vertices and forces the switch to bottom-up; 0 = never),
 * "keep_edges" (1: keep the input tuples resident for b2bfs_copy_edges), "directed" (1: a tuple
 * "u v" is the single arc u -> v, as the reference's stdin reader builds it, bfs.cu:901, instead of
 * an undirected edge, bfs.cu:860-861; such graphs are searched top-down only, on one GPU),
 * "huge_min" (0, the default: off; >= 4096: rows of at least that many entries are expanded by whole
 * CTAs out of shared memory, filled by bulk asynchronous copies - a separate kernel instance),
 * "cluster" (1: graphs whose every row has fewer than 32 entries - grids, road-like meshes - run the
 * kernel instance that keeps small frontiers in distributed shared memory; 0: never).
 * Replaces the reference's compile-time switches (#define DeviceNum, commented call sites in main,
 * bfs.cu:19, 805-819). */
int b2bfs_set_option(b2bfs_t* h, const char* key, int64_t value);
int b2bfs_get_option(const b2bfs_t* h, const char* key, int64_t* value);

/* Load a graph: text file in the reference's format ("vertex_num edge_num" then "src dst" lines).
 * Replaces readGraphFromFile (bfs.cu:829-880) + the H2D half of initCuda2 (bfs.cu:333-351).
 * Unlike the reference, short files, ids >= vertex_num and unreadable paths are errors. */
int b2bfs_load_edge_file(b2bfs_t* h, const char* path);

/* Wider ingest.  The only dataset the reference's README names,
 * soc-LiveJournal1.mtx (README.md:22), cannot be read by the reference's own loader ('%' comments,
 * a 3-number header and 1-based ids, vs bfs.cu:845-861).  `format`:
 *   B2BFS_FORMAT_AUTO (-1)      sniff: cache magic, "%%MatrixMarket"/'%', leading '#', else reference
 *   B2BFS_FORMAT_REFERENCE (0)  "vertex_num edge_num" + "src dst" pairs, 0-based
 *   B2BFS_FORMAT_MTX (1)        MatrixMarket coordinate: '%' comments, "rows cols nnz", 1-based
 *                               "i j [value]" lines (values ignored); n = max(rows, cols)
 *   B2BFS_FORMAT_SNAP (2)       '#' comment lines, one 0-based "u v" per line; n = max id + 1
 *   B2BFS_FORMAT_CACHE (3)      binary edge cache written by b2bfs_save_edge_cache
 * All text formats are parsed by every host thread over a memory-mapped file. */
#define B2BFS_FORMAT_AUTO (-1)
#define B2BFS_FORMAT_REFERENCE 0
#define B2BFS_FORMAT_MTX 1
#define B2BFS_FORMAT_SNAP 2
#define B2BFS_FORMAT_CACHE 3
int b2bfs_load_graph_file(b2bfs_t* h, const char* path, int format);
/* Write the resident input tuples as a binary cache ("B2BFSEDG", n, m as int64, src[m], dst[m] as
 * int32) so that a multi-gigabyte text file is parsed once. */
int b2bfs_save_edge_cache(b2bfs_t* h, const char* path);
/* Host-only: parse a graph file into malloc'ed tuple arrays (no GPU involved; release them with
 * b2bfs_free_edges).  What b2bfs_load_graph_file feeds to b2bfs_load_edges. */
int b2bfs_read_graph_file(const char* path, int format, int64_t* n, int64_t* m, int32_t** src, int32_t** dst);
void b2bfs_free_edges(int32_t* src, int32_t* dst);

/* Load a graph from host arrays of m (src,dst) tuples over n vertices (0-based ids).
 * Replaces filling `struct Graph` (bfs.cu:21-27, 851-875) + initCuda2's cudaMalloc/cudaMemcpy. */
int b2bfs_load_edges(b2bfs_t* h, int64_t n, int64_t m, const int32_t* src, const int32_t* dst);

/* Generate a synthetic graph directly in HBM.  No live reference counterpart (its readGraph random
 * generator, bfs.cu:882-920, is never called); needed because BASELINE configs 2-5 are too large
 * for text files. */
int b2bfs_generate(b2bfs_t* h, int kind, int scale, int edgefactor, uint64_t seed);

/* Graph facts.  num_edges is what the reference prints as "Number of edges" (= 2m directed
 * entries, bfs.cu:790, 875); num_input_edges is m; csr_entries is what is resident after dedup. */
int b2bfs_num_vertices(const b2bfs_t* h, int64_t* n);
int b2bfs_num_edges(const b2bfs_t* h, int64_t* two_m);
int b2bfs_num_input_edges(const b2bfs_t* h, int64_t* m);
int b2bfs_csr_entries(const b2bfs_t* h, int64_t* entries);

/* Per-vertex degree in the reference's multigraph = its edgesSize[] (bfs.cu:868), n x int32. */
int b2bfs_copy_degrees(b2bfs_t* h, int32_t* deg_out);
/* Per-vertex degree in the RESIDENT CSR (after dedup), n x int32.  Multi-rank: collective. */
int b2bfs_copy_row_degrees(b2bfs_t* h, int32_t* deg_out);
/* Input tuples [first, first+count) back to the host (what the loader/generator produced). */
int b2bfs_copy_edges(b2bfs_t* h, int64_t first, int64_t count, int32_t* src_out, int32_t* dst_out);
/* Resident CSR back to the host: rowptr has n+1 int64 entries, col has csr_entries int32. */
int b2bfs_copy_csr(b2bfs_t* h, int64_t* rowptr_out, int32_t* col_out);

/* One search from `src`.  depth_out: n x int32 (INT_MAX = unreachable, depth[src] = 0).
 * parent_out (nullable): n x int32 VERTEX ids, parent[src] = src, -1 = unreachable (the reference
 * stores CSR edge indices and never copies them back, bfs.cu:147, 439).  stats nullable.
 * Replaces runCudaQueueBfs (bfs.cu:542-629; MPI: bfs_mpi.cu:549-643) with its
 * initializeCudaBfs2 (bfs.cu:402-422) and finalizeCudaBfs (bfs.cu:424-441; bfs_mpi.cu:424-448).
 * In a multi-rank job the call is collective: every rank calls it with the same src, the per-rank
 * slices are allgathered over NVLink and every rank that passed a buffer receives all n entries
 * (ranks that do not want them pass NULL; parents travel unless option "gather_parent" is 0 on all
 * ranks).  b2bfs_copy_degrees / b2bfs_copy_row_degrees are collective in the same way;
 * b2bfs_copy_csr returns this rank's rows. */
int b2bfs_run(b2bfs_t* h, int64_t src, int32_t* depth_out, int32_t* parent_out, b2bfs_stats* stats);

/* Multi-rank, for callers that are one process per GPU anyway: the same collective search, but every
 * rank receives only ITS OWN rows - depth_slice_out / parent_slice_out (nullable) hold `count`
 * entries for vertices [first_vertex, first_vertex + count) as b2bfs_local_range reports them - copied
 * over that GPU's own PCIe link, with no cross-GPU gather.  Ranks that point their slices into one
 * shared page-locked segment (b2bfs_host_register on a POSIX shared-memory mapping) assemble the
 * whole array with all links in parallel.  On a single-GPU handle it is b2bfs_run.
 * Replaces the rank-1 -> rank-0 MPI_Send of the full distance array and the host min-merge,
 * bfs_mpi.cu:424-448 (every rank there holds, sends and merges all n entries). */
int b2bfs_run_slice(b2bfs_t* h, int64_t src, int32_t* depth_slice_out, int32_t* parent_slice_out, b2bfs_stats* stats);
int b2bfs_local_range(const b2bfs_t* h, int64_t* first_vertex, int64_t* count);

/* Up to 64 searches over the resident graph in one pass per level (the reference
 * has no batched search - its metric's 64 roots would be 64 runs of runCudaQueueBfs, bfs.cu:542).
 * Bit-parallel multi-source BFS: one 64-bit word per vertex and role, one read of the adjacency
 * serves every search of the batch.  depth_out (nullable): nroots x n int32, row i = the depth array
 * b2bfs_run would return for roots[i] (INT_MAX = unreachable).  Depths only: no parent tree.
 * Single-GPU handles, undirected graphs.  The first call on a graph allocates 24 bytes per vertex
 * of state plus 4*n bytes per root. */
#define B2BFS_BATCH_MAX 64
typedef struct b2bfs_batch_stats {
  int32_t nroots;
  int32_t levels;              /* levels executed = largest eccentricity in the batch + 1 */
  int64_t kernel_launches;
  int64_t d2h_bytes;
  double search_ms;            /* device time, CUDA events: state reset .. all depth rows complete in HBM */
  double total_ms;
  int64_t reached_vertices[B2BFS_BATCH_MAX]; /* V_r per root */
  int64_t reached_entries[B2BFS_BATCH_MAX];  /* D_r per root */
} b2bfs_batch_stats;
int b2bfs_run_batch(b2bfs_t* h, int nroots, const int64_t* roots, int32_t* depth_out, b2bfs_batch_stats* stats);
int b2bfs_device_batch_depth(const b2bfs_t* h, const int32_t** depth_dev);

/* Same search, results left in HBM (no device->host copy). */
int b2bfs_run_device(b2bfs_t* h, int64_t src, b2bfs_stats* stats);
/* Device pointers to the last search's results (this rank's rows in a multi-rank job). */
int b2bfs_device_depth(const b2bfs_t* h, const int32_t** depth_dev);
int b2bfs_device_parent(const b2bfs_t* h, const int32_t** parent_dev);

/* Diagnostics: per-phase timestamps of the last fused search (level, phase code, microseconds since
 * the kernel started; phase codes: -1 start, 0 reset, 1 top-down, 2 bottom-up, 3 queue->bitmap,
 * 4 bitmap->queue, 5 a run of small-frontier levels).  Returns the number of records written
 * (<= cap) or a negative error.  Stands in for the commented-out per-level printf at bfs.cu:576. */
int b2bfs_last_trace(b2bfs_t* h, int32_t* level_out, int32_t* phase_out, double* us_out, int cap);

/* One search with NO device allocation (no counterpart in the reference).
 * For callers that must not allocate when they need a search: code running under another
 * framework's memory pool, loops whose allocations all happened at start-up, processes in which a
 * different allocator already owns the whole card.  The caller passes the graph as a CSR that is
 * ALREADY in device memory (rowptr: n+1 int64; col: `entries` int32, 16-byte aligned, neighbour
 * ids in [0,n)) and device buffers for the results (depth, parent: n int32 each, 16-byte aligned);
 * all scratch - bitmaps, frontier work lists, counters - lives in one array compiled into the
 * library (about 7.4 MiB per GPU, held until the process exits).  Limits: n <= B2BFS_NOALLOC_MAX_VERTICES
 * and entries <= B2BFS_NOALLOC_MAX_ENTRIES, otherwise B2BFS_ERR_RANGE.  Top-down only; results are
 * the same as b2bfs_run's (depth INT_MAX / parent -1 = unreachable).
 * `col` is read in 16-byte groups: when `entries` is not a multiple of 4 the last group reaches up
 * to 12 bytes past col[entries-1].  Those bytes are read and ignored, and they lie in the same
 * 16-byte unit as the last entry, but a caller that carves `col` out of a pool should size the
 * piece to a multiple of 4 entries so that a memory checker stays quiet.
 * The call blocks until the search is complete on `cuda_stream` (a cudaStream_t, or NULL for the
 * default stream); it synchronises that stream before it returns.
 * The first call on a device is a set-up call: it resolves the array's device address
 * (cudaGetSymbolAddress), sizes the launch, and - because CUDA loads device code lazily - is also
 * when the driver puts this library's kernels and the array itself on the device.  Make it during
 * start-up, with any small graph, and before starting any stream capture; later calls on that
 * device take no device memory at all.
 * All calls of a process share the one array (it still holds the previous call's contents; the
 * entry point clears what has to start at zero), so two searches overlapping on the GPU would
 * corrupt each other: the library takes a host lock and returns only when the search is complete,
 * which keeps its own calls apart. */
#define B2BFS_NOALLOC_MAX_VERTICES 262144   /* 2^18 */
#define B2BFS_NOALLOC_MAX_ENTRIES 8388608   /* 2^23 directed CSR entries */
int b2bfs_search_noalloc(int device, int64_t n, int64_t entries, const int64_t* rowptr_dev, const int32_t* col_dev,
                         int64_t src, int32_t* depth_dev, int32_t* parent_dev, void* cuda_stream,
                         int32_t* levels_out);
const char* b2bfs_noalloc_last_error(void);

/* Page-locked host memory for callers that want the copies in b2bfs_run at full PCIe speed. */
int b2bfs_host_alloc(void** p, int64_t bytes);
int b2bfs_host_free(void* p);
/* Page-lock memory the caller already owns (e.g. a shared-memory mapping several ranks write their
 * slices into); undo with b2bfs_host_unregister before unmapping it. */
int b2bfs_host_register(void* p, int64_t bytes);
int b2bfs_host_unregister(void* p);

/* Release everything the handle owns.  The reference never frees (finalizeCuda is commented out,
 * bfs.cu:362-372). */
int b2bfs_destroy(b2bfs_t* h);

/* Message of the last failure on this handle (h may be NULL for create-time failures). */
const char* b2bfs_last_error(const b2bfs_t* h);

#ifdef __cplusplus
}
#endif
#endif /* B2BFS_H */
