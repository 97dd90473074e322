// cli_main.cpp — `b2bfs <src_vertex> <edge_file>`: the reference's command line (bfs.cu:783-823,
// README.md:13) over the C ABI.  Prints the reference's stdout lines (format strings at bfs.cu:830,
// 846, 848, 876, 789-790, 213, 218, 550, 626, 377-383) so scripts that grep
// "Elapsed time in milliseconds" / "Output OK!" keep working; exit status 0 on "Output OK!", 1 on
// a mismatch (bfs.cu:379) or any error.
//
// The search result always comes from the GPU.  Like the reference's main (bfs.cu:798-815) the
// binary then self-checks the depth array against a serial host BFS over the same tuples; that
// check is verification only (skip it with --no-check) — there is no CPU fallback.
//
// The reference's second reader, readGraph (bfs.cu:882-920; dead code in its main), is here too:
//   b2bfs <src_vertex> --stdin            "n m" + m pairs from standard input, every pair the single
//                                         arc u -> v (bfs.cu:899-901: NOT symmetrised; searched top-down)
//   b2bfs <src_vertex> --random N M       srand(12345); u = rand() % N; v = rand() % N per edge, both
//                                         directions (bfs.cu:891-906) - the C library's own rand() stream
// (`b2bfs <src_vertex>` alone and `b2bfs <src_vertex> N M` are accepted as the same two forms, which
// is how readGraph itself reads its argv: argc <= 2 means stdin, else argv[2] = n and argv[3] = m.)
//
// Extra, optional flags after the two positionals:
//   --no-check            skip the host self-check
//   --dump-depth PATH     write one decimal depth per line (2147483647 = unreachable)
//   --mode do|td          direction-optimising (default) or top-down only
//   --engine fused|stepwise
//   --no-dedup            keep duplicate entries and self-loops in the CSR, as the reference does
//   --gen KIND:SCALE:EF:SEED   ignore <edge_file>, generate rmat|kron|grid in HBM instead
//   --format auto|ref|mtx|snap|bin   how to read <edge_file> (default ref, the reference's format)
//   --save-cache PATH     also write the tuples as a binary cache for the next run
//   --device N
//   --gpus N              the reference's second binary, `mpirun -np 2 ./osu_bw <src> <file>` (README.md:19-22,
//                         bfs_mpi.cu:797-856), without MPI: the program starts N copies of itself, one per
//                         GPU (devices 0..N-1), which partition the rows and search together over NVLink;
//                         rank 0 prints the reference's lines and runs the self-check, as rank 0 does at
//                         bfs_mpi.cu:826-847 (where the check is vacuous; here it is real).  N in 2..8.
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include <chrono>
#include <climits>
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <queue>
#include <string>
#include <vector>

#include "../../include/b2bfs.h"
#include "../../include/b2bfs_graphgen.h"

static int die(b2bfs_t* h, const char* what, int rc) {
  fprintf(stderr, "b2bfs: %s failed (%d): %s\n", what, rc, b2bfs_last_error(h));
  if (h) b2bfs_destroy(h);
  return 1;
}

// serial host BFS for the self-check: same role as bfsCPU (bfs.cu:923-945), written for this file
static void host_check_bfs(int64_t n, int64_t m, const std::vector<int32_t>& s, const std::vector<int32_t>& d,
                           int64_t start, std::vector<int32_t>& depth, bool directed) {
  std::vector<int64_t> off(n + 1, 0);
  for (int64_t i = 0; i < m; ++i) { off[s[i] + 1]++; if (!directed) off[d[i] + 1]++; }
  for (int64_t v = 0; v < n; ++v) off[v + 1] += off[v];
  std::vector<int32_t> adj(directed ? m : 2 * m);
  std::vector<int64_t> cur(off.begin(), off.end() - 1);
  for (int64_t i = 0; i < m; ++i) { adj[cur[s[i]]++] = d[i]; if (!directed) adj[cur[d[i]]++] = s[i]; }
  depth.assign(n, INT_MAX);
  std::vector<int32_t> q;
  q.reserve(n);
  depth[start] = 0;
  q.push_back((int32_t)start);
  for (size_t head = 0; head < q.size(); ++head) {
    const int32_t u = q[head];
    for (int64_t i = off[u]; i < off[u + 1]; ++i) {
      const int32_t v = adj[i];
      if (depth[v] == INT_MAX) { depth[v] = depth[u] + 1; q.push_back(v); }
    }
  }
}

static bool all_digits(const char* s) {
  if (!*s) return false;
  for (; *s; ++s) if (*s < '0' || *s > '9') return false;
  return true;
}

static int run_rank(int argc, char** argv, int rank, int world, const char* id_file);

// --gpus N: start N copies of this program (ranks), hand them a file through which rank 0 publishes
// the NCCL id, wait for all of them; the exit status is the worst of theirs
static int launch_ranks(int argc, char** argv, int world) {
  char id_file[128];
  snprintf(id_file, sizeof(id_file), "/tmp/b2bfs_id_%d", (int)getpid());
  unlink(id_file);
  std::vector<pid_t> kids;
  fflush(stdout);
  for (int r = 0; r < world; ++r) {
    const pid_t pid = fork();  // before this process has touched CUDA
    if (pid < 0) { perror("b2bfs: fork"); break; }
    if (pid == 0) {
      const int rc = run_rank(argc, argv, r, world, id_file);
      fflush(stdout);
      fflush(stderr);
      _exit(rc);
    }
    kids.push_back(pid);
  }
  int worst = (int)kids.size() == world ? 0 : 1;
  for (size_t done = 0; done < kids.size(); ++done) {
    int st = 0;
    const pid_t pid = wait(&st);
    if (pid < 0) break;
    const int code = WIFEXITED(st) ? WEXITSTATUS(st) : 1;
    if (code != 0 && worst == 0) {  // one rank failed: the others would wait for it forever
      worst = code;
      for (pid_t k : kids) if (k != pid) kill(k, SIGTERM);
    }
  }
  unlink(id_file);
  unlink((std::string(id_file) + ".go").c_str());
  return worst;
}

int main(int argc, char** argv) {
  for (int i = 1; i + 1 < argc; ++i) {
    if (!strcmp(argv[i], "--gpus")) {
      const int world = atoi(argv[i + 1]);
      if (world == 1) break;
      if (world < 2 || world > 8) { fprintf(stderr, "b2bfs: --gpus takes 1..8\n"); return 2; }
      return launch_ranks(argc, argv, world);
    }
  }
  return run_rank(argc, argv, 0, 1, nullptr);
}

static int run_rank(int argc, char** argv, int rank, int world, const char* id_file) {
  const bool lead = rank == 0;  // the rank that talks (bfs_mpi.cu:826: `if (rank == 0)`)
  if (argc < 2) {
    fprintf(stderr, "usage: %s <src_vertex> <edge_file> | <src_vertex> --stdin | <src_vertex> --random N M\n"
                    "       [--no-check] [--dump-depth PATH] [--mode do|td] [--engine fused|stepwise] [--no-dedup] "
                    "[--gen KIND:SCALE:EF:SEED] [--format auto|ref|mtx|snap|bin] [--save-cache PATH] [--device N]\n", argv[0]);
    return 2;
  }
  const int64_t start = atoll(argv[1]);  // bfs.cu:786
  // which reader: readGraphFromFile (the reference's live one) or one of readGraph's two forms
  enum { kFile, kStdin, kRandom } source = kFile;
  long long rand_n = 0, rand_m = 0;
  int first_flag = 3;
  const char* path = argc > 2 ? argv[2] : "(stdin)";
  if (argc == 2 || !strcmp(argv[2], "--stdin")) {
    source = kStdin;
    path = "(stdin)";
    first_flag = argc == 2 ? 2 : 3;
  } else if (!strcmp(argv[2], "--random") && argc >= 5) {
    source = kRandom;
    rand_n = atoll(argv[3]);
    rand_m = atoll(argv[4]);
    first_flag = 5;
  } else if (argc >= 4 && all_digits(argv[2]) && all_digits(argv[3])) {  // readGraph's own argv layout
    source = kRandom;
    rand_n = atoll(argv[2]);
    rand_m = atoll(argv[3]);
    first_flag = 4;
  }
  bool check = true, dedup = true;
  const char* dump = nullptr;
  const char* gen = nullptr;
  const char* save_cache = nullptr;
  int format = B2BFS_FORMAT_REFERENCE;
  int mode = B2BFS_MODE_DIRECTION_OPT, engine = B2BFS_ENGINE_FUSED, device = 0;
  for (int i = first_flag; i < argc; ++i) {
    const std::string a(argv[i]);
    if (a == "--no-check") check = false;
    else if (a == "--no-dedup") dedup = false;
    else if (a == "--dump-depth" && i + 1 < argc) dump = argv[++i];
    else if (a == "--gen" && i + 1 < argc) gen = argv[++i];
    else if (a == "--device" && i + 1 < argc) device = atoi(argv[++i]);
    else if (a == "--gpus" && i + 1 < argc) ++i;  // handled by main
    else if (a == "--save-cache" && i + 1 < argc) save_cache = argv[++i];
    else if (a == "--format" && i + 1 < argc) {
      const std::string f(argv[++i]);
      format = f == "auto" ? B2BFS_FORMAT_AUTO : f == "mtx" ? B2BFS_FORMAT_MTX : f == "snap" ? B2BFS_FORMAT_SNAP
             : f == "bin" ? B2BFS_FORMAT_CACHE : B2BFS_FORMAT_REFERENCE;
    }
    else if (a == "--mode" && i + 1 < argc) mode = strcmp(argv[++i], "td") == 0 ? B2BFS_MODE_TOP_DOWN : B2BFS_MODE_DIRECTION_OPT;
    else if (a == "--engine" && i + 1 < argc) engine = strcmp(argv[++i], "stepwise") == 0 ? B2BFS_ENGINE_STEPWISE : B2BFS_ENGINE_FUSED;
    else { fprintf(stderr, "b2bfs: unknown argument %s\n", argv[i]); return 2; }
  }

  b2bfs_t* h = nullptr;
  int rc = 0;
  int gen_kind = 0, gen_scale = 0, gen_ef = 16;
  unsigned long long gen_seed = 1;
  if (world > 1) {
    if (source == kStdin) { fprintf(stderr, "b2bfs: --stdin (a directed graph) runs on one GPU\n"); return 2; }
    // rank 0 publishes the 128-byte NCCL id through a file, the others wait for it
    unsigned char id[128];
    const std::string tmp = std::string(id_file) + ".tmp";
    if (lead) {
      rc = b2bfs_comm_unique_id(id);
      if (rc) return die(nullptr, "b2bfs_comm_unique_id", rc);
      FILE* f = fopen(tmp.c_str(), "wb");
      if (!f || fwrite(id, 1, 128, f) != 128 || fclose(f) != 0 || rename(tmp.c_str(), id_file) != 0) {
        fprintf(stderr, "b2bfs: cannot publish the communicator id\n");
        return 1;
      }
    } else {
      FILE* f = nullptr;
      for (int tries = 0; tries < 1200 && !(f = fopen(id_file, "rb")); ++tries) usleep(50000);
      if (!f || fread(id, 1, 128, f) != 128) { fprintf(stderr, "b2bfs: rank %d never received the communicator id\n", rank); return 1; }
      fclose(f);
    }
    rc = b2bfs_create_rank(&h, rank, rank, world, id);
    if (rc) return die(nullptr, "b2bfs_create_rank", rc);
    if (!lead && !freopen("/dev/null", "w", stdout)) return 1;  // only rank 0 prints the reference's lines
  } else {
    rc = b2bfs_create(&h, device);
    if (rc) return die(nullptr, "b2bfs_create", rc);
  }
  b2bfs_set_option(h, "dedup", dedup);
  b2bfs_set_option(h, "mode", mode);
  b2bfs_set_option(h, "engine", engine);

  if (source == kFile) printf("%s\n", path);  // bfs.cu:830
  if (source == kStdin || source == kRandom) {
    long long n = rand_n, m = rand_m;
    std::vector<int32_t> s, d;
    if (source == kStdin) {
      if (scanf("%lld %lld", &n, &m) != 2 || n < 1 || m < 0) { fprintf(stderr, "b2bfs: stdin must start with \"n m\"\n"); return 2; }
      s.resize((size_t)m);
      d.resize((size_t)m);
      for (long long i = 0; i < m; ++i) {
        int u, v;
        if (scanf("%d %d", &u, &v) != 2) { fprintf(stderr, "b2bfs: stdin ends after %lld of %lld pairs\n", i, m); return 2; }
        s[(size_t)i] = u;
        d[(size_t)i] = v;
      }
      b2bfs_set_option(h, "directed", 1);  // bfs.cu:901 pushes v onto u's list only
    } else {
      if (n < 1 || n > INT_MAX || m < 0) { fprintf(stderr, "b2bfs: --random needs N >= 1 and M >= 0\n"); return 2; }
      s.resize((size_t)m);
      d.resize((size_t)m);
      srand(12345);  // bfs.cu:892
      for (long long i = 0; i < m; ++i) {
        s[(size_t)i] = (int32_t)(rand() % n);  // bfs.cu:903
        d[(size_t)i] = (int32_t)(rand() % n);  // bfs.cu:904
      }
    }
    rc = b2bfs_load_edges(h, n, m, s.data(), d.data());
    if (rc) return die(h, "b2bfs_load_edges", rc);
  } else if (gen) {
    char kind[16] = {0};
    int scale = 0, ef = 16;
    unsigned long long seed = 1;
    if (sscanf(gen, "%15[a-z]:%d:%d:%llu", kind, &scale, &ef, &seed) < 2) { fprintf(stderr, "b2bfs: bad --gen\n"); return 2; }
    const int k = !strcmp(kind, "kron") ? B2BFS_KIND_KRON : !strcmp(kind, "grid") ? B2BFS_KIND_GRID : B2BFS_KIND_RMAT;
    gen_kind = k; gen_scale = scale; gen_ef = ef; gen_seed = seed;
    rc = b2bfs_generate(h, k, scale, ef, seed);
    if (rc) return die(h, "b2bfs_generate", rc);
  } else {
    rc = b2bfs_load_graph_file(h, path, format);
    if (rc) {
      if (rc == B2BFS_ERR_IO) printf("not open %s\n", path);  // bfs.cu:878
      return die(h, "b2bfs_load_graph_file", rc);
    }
  }
  if (save_cache) {
    rc = b2bfs_save_edge_cache(h, save_cache);
    if (rc) return die(h, "b2bfs_save_edge_cache", rc);
  }
  int64_t n = 0, m = 0, e2 = 0;
  b2bfs_num_vertices(h, &n);
  b2bfs_num_input_edges(h, &m);
  b2bfs_num_edges(h, &e2);
  if (source == kFile) {  // readGraph prints nothing of its own
    printf("nodes num: %lld\n", (long long)n);  // bfs.cu:846
    printf("edge num: %lld\n", (long long)m);   // bfs.cu:848
    printf("finish load graph\n");              // bfs.cu:876
  }
  printf("Number of vertices %lld\n", (long long)n);    // bfs.cu:789
  printf("Number of edges %lld\n\n", (long long)e2);    // bfs.cu:790
  if (start < 0 || start >= n) {
    fprintf(stderr, "b2bfs: source vertex %lld outside [0, %lld)\n", (long long)start, (long long)n);
    b2bfs_destroy(h);
    return 1;
  }

  std::vector<int32_t> expected;
  check = check && lead;
  if (check) {
    std::vector<int32_t> s(m), d(m);
    if (world == 1) {
      rc = b2bfs_copy_edges(h, 0, m, s.data(), d.data());
      if (rc) return die(h, "b2bfs_copy_edges", rc);
    } else if (gen) {  // a rank holds only its share of the tuples on the device: the host draws the same stream
      for (int64_t i = 0; i < m; ++i) {
        uint32_t u, v;
        b2_gen_edge(gen_kind, gen_scale, gen_ef, gen_seed, (uint64_t)i, &u, &v);
        s[i] = (int32_t)u;
        d[i] = (int32_t)v;
      }
    } else if (source == kRandom) {
      srand(12345);
      for (int64_t i = 0; i < m; ++i) { s[i] = (int32_t)(rand() % n); d[i] = (int32_t)(rand() % n); }
    } else {
      int64_t fn = 0, fm = 0;
      int32_t *fs = nullptr, *fd = nullptr;
      rc = b2bfs_read_graph_file(path, format, &fn, &fm, &fs, &fd);
      if (rc || fm != m) return die(nullptr, "b2bfs_read_graph_file", rc);
      memcpy(s.data(), fs, (size_t)m * sizeof(int32_t));
      memcpy(d.data(), fd, (size_t)m * sizeof(int32_t));
      b2bfs_free_edges(fs, fd);
    }
    printf("Starting sequential bfs.\n");  // bfs.cu:213
    const auto t0 = std::chrono::steady_clock::now();
    host_check_bfs(n, m, s, d, start, expected, source == kStdin);
    const long ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
    printf("Elapsed time in milliseconds : %li ms.\n\n", ms);  // bfs.cu:218
  }

  if (world > 1) {
    // the ranks enter the collective search together: the others wait until rank 0 is through its
    // host-side check (inside the search a rank that waits too long for a peer gives up)
    const std::string go = std::string(id_file) + ".go";
    if (lead) {
      FILE* f = fopen(go.c_str(), "wb");
      if (f) fclose(f);
    } else {
      for (int tries = 0; tries < 72000 && access(go.c_str(), F_OK) != 0; ++tries) usleep(50000);
    }
  }
  std::vector<int32_t> depth(lead ? n : 0);
  b2bfs_stats st;
  printf("Starting queue parallel bfs.\n");  // bfs.cu:550
  if (world > 1) b2bfs_set_option(h, "gather_parent", 0);
  rc = b2bfs_run(h, start, lead ? depth.data() : nullptr, nullptr, &st);  // collective when world > 1
  if (rc) return die(h, "b2bfs_run", rc);
  printf("Elapsed time in milliseconds : %li ms.\n", (long)(st.search_ms + 0.5));  // bfs.cu:626
  if (lead) fprintf(stderr, "b2bfs: search %.3f ms on device, %d levels (%d top-down, %d bottom-up), %lld vertices reached, "
                  "%.3f GTEPS\n", st.search_ms, st.levels, st.td_levels, st.bu_levels, (long long)st.reached_vertices,
          st.search_ms > 0 ? (st.reached_entries / 2) / (st.search_ms * 1e6) : 0.0);

  if (dump && lead) {
    FILE* f = fopen(dump, "w");
    if (!f) { fprintf(stderr, "b2bfs: cannot write %s\n", dump); b2bfs_destroy(h); return 1; }
    for (int64_t i = 0; i < n; ++i) fprintf(f, "%d\n", depth[i]);
    fclose(f);
  }
  int status = 0;
  if (check) {  // checkOutput, bfs.cu:374-384
    for (int64_t i = 0; i < n; ++i) {
      if (depth[i] != expected[i]) {
        printf("%lld %d %d\n", (long long)i, depth[i], expected[i]);
        printf("Wrong output!\n");
        status = 1;
        break;
      }
    }
    if (!status) printf("Output OK!\n\n");
  }
  b2bfs_destroy(h);
  return status;
}
