// edge_file.cpp — mmap + multi-threaded parse of the reference's edge-list text format.
//
// Replaces the front half of readGraphFromFile (bfs.cu:829-863): `f >> n >> m` then m times
// `f >> u >> v` through std::ifstream on one core.  Same grammar — integers separated by any
// whitespace, no comments — but the file is mapped once, cut into byte ranges at whitespace, and
// every range is tokenised by its own thread straight into the (src, dst) arrays.  What the
// reference leaves undefined is an error here: a file with fewer than 2m integers after the header
// (bfs.cu:852-861 would push uninitialised values), an id outside [0, n) (bfs.cu:860 indexes out
// of range) and an unreadable path (bfs.cu:878 only prints "not open").  Tokens past the 2m-th
// are ignored, as the reference's loop never reads them.
#include "edge_file.hpp"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/b2bfs.h"

namespace b2 {

static inline bool is_space(char c) { return c == ' ' || c == '\n' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }

int map_file(const char* path, MappedFile* f, std::string* err) {
  f->fd = open(path, O_RDONLY);
  if (f->fd < 0) {
    *err = std::string("not open ") + path + ": " + strerror(errno);
    return B2BFS_ERR_IO;
  }
  struct stat st;
  if (fstat(f->fd, &st) != 0 || !S_ISREG(st.st_mode)) {
    *err = std::string("not a regular file: ") + path;
    close(f->fd);
    f->fd = -1;
    return B2BFS_ERR_IO;
  }
  f->size = (size_t)st.st_size;
  if (f->size == 0) {
    f->data = nullptr;
    return B2BFS_OK;
  }
  void* p = mmap(nullptr, f->size, PROT_READ, MAP_PRIVATE, f->fd, 0);
  if (p == MAP_FAILED) {
    *err = std::string("mmap failed: ") + strerror(errno);
    close(f->fd);
    f->fd = -1;
    return B2BFS_ERR_IO;
  }
  madvise(p, f->size, MADV_SEQUENTIAL);
  f->data = static_cast<const char*>(p);
  return B2BFS_OK;
}

void unmap_file(MappedFile* f) {
  if (f->data) munmap(const_cast<char*>(f->data), f->size);
  if (f->fd >= 0) close(f->fd);
  *f = MappedFile();
}

// one non-negative decimal integer starting at p (after optional '+'); false on anything else
static inline bool parse_uint(const char*& p, const char* end, int64_t* out) {
  if (p < end && *p == '+') ++p;
  if (p >= end || *p < '0' || *p > '9') return false;
  int64_t x = 0;
  while (p < end && *p >= '0' && *p <= '9') {
    x = x * 10 + (*p - '0');
    if (x > ((int64_t)1 << 40)) return false;
    ++p;
  }
  if (p < end && !is_space(*p)) return false;
  *out = x;
  return true;
}

int parse_header(const MappedFile& f, int64_t* n, int64_t* m, size_t* body_off, std::string* err) {
  const char* p = f.data;
  const char* end = f.data + f.size;
  int64_t v[2];
  for (int k = 0; k < 2; ++k) {
    while (p < end && is_space(*p)) ++p;
    if (!parse_uint(p, end, &v[k])) {
      *err = "edge list header must be two non-negative integers: vertex_num edge_num";
      return B2BFS_ERR_FORMAT;
    }
  }
  if (v[0] > 0x7fffffffLL) {
    *err = "vertex_num exceeds int32 ids";
    return B2BFS_ERR_RANGE;
  }
  *n = v[0];
  *m = v[1];
  *body_off = (size_t)(p - f.data);
  // every pair needs at least "d d " = 4 bytes: a header that promises more pairs than the rest of
  // the file can hold is a malformed file, not a reason to attempt a huge allocation
  if ((uint64_t)v[1] > (uint64_t)(f.size - *body_off) / 4 + 1) {
    *err = "edge list header declares " + std::to_string(v[1]) + " edges but the file is too short to hold them";
    return B2BFS_ERR_FORMAT;
  }
  return B2BFS_OK;
}

int parse_body(const MappedFile& f, size_t body_off, int64_t n, int64_t m, int32_t* src, int32_t* dst, int threads,
               std::string* err) {
  if (m == 0) return B2BFS_OK;
  const char* base = f.data;
  const size_t size = f.size;
  if (threads < 1) threads = 1;
  const size_t span = size - body_off;
  if (span / 64 < (size_t)threads) threads = (int)(span / 64) + 1;
  // a range owns the tokens that START in it
  std::vector<size_t> cut(threads + 1);
  for (int t = 0; t <= threads; ++t) {
    size_t b = body_off + span * (size_t)t / (size_t)threads;
    if (t == threads) b = size;
    else if (t > 0) while (b < size && b > body_off && !is_space(base[b - 1])) ++b;
    cut[t] = b;
  }
  std::vector<int64_t> ntok(threads, 0);
  std::atomic<int> status(B2BFS_OK);
  auto count = [&](int t) {
    const char* p = base + cut[t];
    const char* e = base + cut[t + 1];
    int64_t c = 0;
    while (p < e) {
      while (p < e && is_space(*p)) ++p;
      if (p >= e) break;
      ++c;
      while (p < base + size && !is_space(*p)) ++p;
    }
    ntok[t] = c;
  };
  {
    std::vector<std::thread> th;
    for (int t = 1; t < threads; ++t) th.emplace_back(count, t);
    count(0);
    for (auto& x : th) x.join();
  }
  std::vector<int64_t> first(threads + 1, 0);
  for (int t = 0; t < threads; ++t) first[t + 1] = first[t] + ntok[t];
  if (first[threads] < 2 * m) {
    *err = "edge list ends early: header promises " + std::to_string(m) + " pairs, file holds " +
           std::to_string(first[threads] / 2);
    return B2BFS_ERR_FORMAT;
  }
  auto parse = [&](int t) {
    const char* p = base + cut[t];
    const char* e = base + cut[t + 1];
    const char* fend = base + size;
    int64_t k = first[t];
    while (p < e && k < 2 * m) {
      while (p < e && is_space(*p)) ++p;
      if (p >= e) break;
      int64_t x;
      if (!parse_uint(p, fend, &x)) { status.store(B2BFS_ERR_FORMAT); return; }
      if (x >= n) { status.store(B2BFS_ERR_RANGE); return; }
      ((k & 1) ? dst : src)[k >> 1] = (int32_t)x;
      ++k;
    }
  };
  {
    std::vector<std::thread> th;
    for (int t = 1; t < threads; ++t) th.emplace_back(parse, t);
    parse(0);
    for (auto& x : th) x.join();
  }
  const int rc = status.load();
  if (rc == B2BFS_ERR_FORMAT) *err = "edge list holds a token that is not a non-negative integer";
  if (rc == B2BFS_ERR_RANGE) *err = "edge list holds a vertex id >= vertex_num";
  return rc;
}

}  // namespace b2

// ---------------------------------------------------------------------------------------------
// line-oriented formats (MatrixMarket coordinate files, SNAP edge lists)
// ---------------------------------------------------------------------------------------------
namespace b2 {

int detect_format(const MappedFile& f) {
  const char* p = f.data;
  const char* end = f.data + f.size;
  if (f.size >= 8 && memcmp(p, "B2BFSEDG", 8) == 0) return kFormatBinaryCache;
  if (f.size >= 14 && memcmp(p, "%%MatrixMarket", 14) == 0) return kFormatMatrixMarket;
  while (p < end && is_space(*p)) ++p;
  if (p < end && *p == '#') return kFormatSnap;
  if (p < end && *p == '%') return kFormatMatrixMarket;
  return kFormatReference;
}

static inline const char* line_end(const char* p, const char* end) {
  while (p < end && *p != '\n') ++p;
  return p;
}

// a data line = first non-blank character is neither the comment character nor a line break
static inline bool is_data_line(const char* p, const char* le, char comment) {
  while (p < le && (*p == ' ' || *p == '\t' || *p == '\r')) ++p;
  return p < le && *p != comment && *p != '%';
}

int parse_mtx_header(const MappedFile& f, int64_t* n, int64_t* nnz, size_t* body_off, std::string* err) {
  const char* p = f.data;
  const char* end = f.data + f.size;
  while (p < end) {
    const char* le = line_end(p, end);
    if (is_data_line(p, le, '%')) {
      int64_t v[3];
      const char* q = p;
      for (int k = 0; k < 3; ++k) {
        while (q < le && is_space(*q)) ++q;
        if (!parse_uint(q, le, &v[k])) {
          *err = "MatrixMarket size line must be: rows cols nonzeros";
          return B2BFS_ERR_FORMAT;
        }
      }
      *n = v[0] > v[1] ? v[0] : v[1];
      *nnz = v[2];
      if (*n > 0x7fffffffLL) { *err = "matrix dimension exceeds int32 ids"; return B2BFS_ERR_RANGE; }
      *body_off = (size_t)((le < end ? le + 1 : le) - f.data);
      return B2BFS_OK;
    }
    p = le < end ? le + 1 : le;
  }
  *err = "MatrixMarket file has no size line";
  return B2BFS_ERR_FORMAT;
}

int plan_lines(const MappedFile& f, size_t body_off, char comment, int64_t id_base, int threads, LinePlan* plan,
               std::string* err) {
  (void)err;
  const char* base = f.data;
  const size_t size = f.size;
  if (threads < 1) threads = 1;
  const size_t span = size > body_off ? size - body_off : 0;
  if (span / 4096 < (size_t)threads) threads = (int)(span / 4096) + 1;
  plan->comment = comment;
  plan->id_base = id_base;
  plan->cut.assign(threads + 1, size);
  for (int t = 0; t < threads; ++t) {
    size_t b = body_off + span * (size_t)t / (size_t)threads;
    if (t > 0) while (b < size && base[b - 1] != '\n') ++b;  // ranges start at line starts
    plan->cut[t] = b;
  }
  std::vector<int64_t> cnt(threads, 0);
  auto count = [&](int t) {
    const char* p = base + plan->cut[t];
    const char* e = base + plan->cut[t + 1];
    int64_t c = 0;
    while (p < e) {
      const char* le = line_end(p, base + size);
      c += is_data_line(p, le, comment);
      p = le + 1;
    }
    cnt[t] = c;
  };
  {
    std::vector<std::thread> th;
    for (int t = 1; t < threads; ++t) th.emplace_back(count, t);
    count(0);
    for (auto& x : th) x.join();
  }
  plan->first.assign(threads + 1, 0);
  for (int t = 0; t < threads; ++t) plan->first[t + 1] = plan->first[t] + cnt[t];
  plan->m = plan->first[threads];
  return B2BFS_OK;
}

int fill_lines(const MappedFile& f, const LinePlan& plan, int64_t n_limit, int32_t* src, int32_t* dst, int64_t* max_id,
               std::string* err) {
  const char* base = f.data;
  const size_t size = f.size;
  const int threads = (int)plan.cut.size() - 1;
  std::atomic<int> status(B2BFS_OK);
  std::vector<int64_t> mx(threads, -1);
  auto fill = [&](int t) {
    const char* p = base + plan.cut[t];
    const char* e = base + plan.cut[t + 1];
    int64_t k = plan.first[t], local_max = -1;
    while (p < e) {
      const char* le = line_end(p, base + size);
      if (is_data_line(p, le, plan.comment)) {
        int64_t v[2];
        const char* q = p;
        for (int j = 0; j < 2; ++j) {
          while (q < le && is_space(*q)) ++q;
          if (!parse_uint(q, le, &v[j])) { status.store(B2BFS_ERR_FORMAT); return; }
          v[j] -= plan.id_base;
          if (v[j] < 0 || v[j] > 0x7ffffffeLL || (n_limit > 0 && v[j] >= n_limit)) { status.store(B2BFS_ERR_RANGE); return; }
          if (v[j] > local_max) local_max = v[j];
        }
        src[k] = (int32_t)v[0];
        dst[k] = (int32_t)v[1];
        ++k;
      }
      p = le + 1;
    }
    mx[t] = local_max;
  };
  {
    std::vector<std::thread> th;
    for (int t = 1; t < threads; ++t) th.emplace_back(fill, t);
    fill(0);
    for (auto& x : th) x.join();
  }
  int64_t m = -1;
  for (int t = 0; t < threads; ++t) if (mx[t] > m) m = mx[t];
  if (max_id) *max_id = m;
  const int rc = status.load();
  if (rc == B2BFS_ERR_FORMAT) *err = "a data line does not start with two non-negative integers";
  if (rc == B2BFS_ERR_RANGE) *err = "a vertex id is outside the declared range";
  return rc;
}

}  // namespace b2
