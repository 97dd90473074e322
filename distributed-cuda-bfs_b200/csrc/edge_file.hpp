// edge_file.hpp — text edge-list reader (reference format, README.md:8-12, bfs.cu:845-861).
#pragma once
#include <cstdint>
#include <string>

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
