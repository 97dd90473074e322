// edge_file.hpp — text edge-list reader (reference format, README.md:8-12, bfs.cu:845-861).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace b2 {

struct MappedFile {
  const char* data = nullptr;
  size_t size = 0;
  int fd = -1;
};

// 0 on success, B2BFS_ERR_* otherwise (err gets a message)
int map_file(const char* path, MappedFile* f, std::string* err);
void unmap_file(MappedFile* f);
// "vertex_num edge_num": the first two integers (bfs.cu:845).  body_off = first byte after them.
int parse_header(const MappedFile& f, int64_t* n, int64_t* m, size_t* body_off, std::string* err);
// the next 2*m integers as (src,dst) pairs, parsed by `threads` host threads; ids must be in [0,n)
int parse_body(const MappedFile& f, size_t body_off, int64_t n, int64_t m, int32_t* src, int32_t* dst, int threads,
               std::string* err);

}  // namespace b2

// ---- wider ingest: the dataset the reference's README names, soc-LiveJournal1.mtx
// (README.md:22), cannot be read by the reference's own loader (no '%' comments, 3-number header,
// 1-based ids: bfs.cu:845-861).  Line-oriented formats are parsed by all host threads as well.
namespace b2 {

enum GraphFileFormat {
  kFormatAuto = -1,
  kFormatReference = 0,     // "n m" then m "u v" pairs, 0-based (the reference's format)
  kFormatMatrixMarket = 1,  // %%MatrixMarket banner, '%' comments, "rows cols nnz", 1-based "i j [value]" lines
  kFormatSnap = 2,          // '#' comment lines, 0-based "u v" per line, vertex count = max id + 1
  kFormatBinaryCache = 3    // b2bfs edge cache: "B2BFSEDG", n, m (int64), src[m], dst[m] (int32)
};

int detect_format(const MappedFile& f);

struct LinePlan {
  std::vector<size_t> cut;      // byte ranges, one per thread, cut at line starts
  std::vector<int64_t> first;   // index of the first tuple of each range
  int64_t m = 0;                // data lines in all
  int64_t id_base = 0;          // subtracted from every id (1 for MatrixMarket)
  char comment = '#';
};
// pass 1: count the data lines (non-blank, not starting with the comment character) after body_off
int plan_lines(const MappedFile& f, size_t body_off, char comment, int64_t id_base, int threads, LinePlan* plan,
               std::string* err);
// pass 2: first two integers of every data line -> (src, dst); n_limit > 0 bounds the ids, max_id gets the largest id seen
int fill_lines(const MappedFile& f, const LinePlan& plan, int64_t n_limit, int32_t* src, int32_t* dst, int64_t* max_id,
               std::string* err);
// MatrixMarket: skips the banner/comments, reads "rows cols nnz"; body_off = first byte after that line
int parse_mtx_header(const MappedFile& f, int64_t* n, int64_t* nnz, size_t* body_off, std::string* err);

}  // namespace b2
