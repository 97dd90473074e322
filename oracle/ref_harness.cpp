// oracle/ref_harness.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Thin extern "C" door into the reference's OWN host code.  The reference translation unit
// /root/reference/bfs.cu is compiled where it lies, unmodified, with `-Dmain=ref_main` (see
// oracle/Makefile) and linked with this file into oracle/_ref/libref_oracle.so.  Nothing is copied:
// the declarations below only re-state the types and prototypes the reference defines at
// bfs.cu:21-28 (struct Graph, global G) and bfs.cu:36-38 (readGraphFromFile, bfsCPU) so that this
// file can call them.
//
// Used (a) to pin oracle/bfs_oracle.c and generate tests/golden/, (b) as the "reference" CPU baseline
// in bench.py.  It runs without a GPU: only the reference's host functions are called; its
// __managed__ globals are never touched.
#include <chrono>
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

struct Graph {                      // bfs.cu:21-27
  std::vector<int> adjacencyList;
  std::vector<int> edgesOffset;
  std::vector<int> edgesSize;
  int numVertices = 0;
  int numEdges = 0;
};
extern Graph G;                                              // bfs.cu:28
void readGraphFromFile(Graph& G, int argc, char** argv);     // bfs.cu:36, defined 829-880
void bfsCPU(int start, Graph& G, std::vector<int>& distance, std::vector<int>& parent,
            std::vector<bool>& visited);                     // bfs.cu:37-38, defined 923-945
void readGraph(Graph& G, int argc, char** argv);             // bfs.cu:35, defined 882-920 (dead code in the reference's main)

extern "C" {

// Load through the reference's own text loader (argv[2] = path, bfs.cu:834).
int ref_load_file(const char* path) {
  G = Graph();
  char a0[] = "ref", a1[] = "0";
  std::vector<char> p(path, path + strlen(path) + 1);
  char* argv[3] = {a0, a1, p.data()};
  fflush(stdout);
  readGraphFromFile(G, 3, argv);
  fflush(stdout);
  return G.numVertices;
}

// Fill the reference's Graph directly from a CSR in the reference's layout (what its loader would
// have built): used for inputs too large to round-trip through a text file.
int ref_set_graph(int n, int e, const int* offsets, const int* sizes, const int* adj) {
  G = Graph();
  G.edgesOffset.assign(offsets, offsets + n);
  G.edgesSize.assign(sizes, sizes + n);
  G.adjacencyList.assign(adj, adj + e);
  G.numVertices = n;
  G.numEdges = e;
  return 0;
}

// The reference's second reader in its stdin form (argc <= 2, bfs.cu:888): stdin is pointed at
// `path` for the duration of the call.  Directed: "u v" is pushed onto u's list only (bfs.cu:901).
int ref_read_graph_stdin(const char* path) {
  G = Graph();
  if (!freopen(path, "r", stdin)) return -1;
  char a0[] = "ref", a1[] = "0";
  char* argv[2] = {a0, a1};
  readGraph(G, 2, argv);
  return G.numVertices;
}

// ... and in its random form (bfs.cu:891-894): srand(12345), n = atoi(argv[2]), m = atoi(argv[3]).
int ref_read_graph_random(int n, int m) {
  G = Graph();
  char a0[] = "ref", a1[] = "0", a2[32], a3[32];
  snprintf(a2, sizeof(a2), "%d", n);
  snprintf(a3, sizeof(a3), "%d", m);
  char* argv[4] = {a0, a1, a2, a3};
  readGraph(G, 4, argv);
  return G.numVertices;
}

int ref_num_vertices() { return G.numVertices; }
int ref_num_edges() { return G.numEdges; }
void ref_get_csr(int* offsets, int* sizes, int* adj) {
  memcpy(offsets, G.edgesOffset.data(), G.edgesOffset.size() * sizeof(int));
  memcpy(sizes, G.edgesSize.data(), G.edgesSize.size() * sizeof(int));
  memcpy(adj, G.adjacencyList.data(), G.adjacencyList.size() * sizeof(int));
}

// One serial search exactly as main + runCpu drive it (bfs.cu:793-798, 211-219): vectors pre-filled
// with INT_MAX / false, steady_clock around bfsCPU alone.  Returns elapsed nanoseconds.
int64_t ref_bfs_cpu(int start, int* depth_out, int* parent_out) {
  std::vector<int> distance(G.numVertices, INT_MAX);
  std::vector<int> parent(G.numVertices, INT_MAX);
  std::vector<bool> visited(G.numVertices, false);
  auto t0 = std::chrono::steady_clock::now();
  bfsCPU(start, G, distance, parent, visited);
  auto t1 = std::chrono::steady_clock::now();
  if (depth_out) memcpy(depth_out, distance.data(), distance.size() * sizeof(int));
  if (parent_out) memcpy(parent_out, parent.data(), parent.size() * sizeof(int));
  return std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count();
}

}  // extern "C"
