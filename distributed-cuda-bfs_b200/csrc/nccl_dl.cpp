#include "nccl_dl.hpp"

#include <dlfcn.h>

#include <mutex>

namespace b2 {

static NcclApi g_api;
static bool g_ok = false;
static const char* g_err = "libnccl.so.2 was never looked up";
static std::once_flag g_once;

template <typename F>
static bool bind(void* lib, const char* name, F* out) {
  *out = reinterpret_cast<F>(dlsym(lib, name));
  return *out != nullptr;
}

const NcclApi* nccl_api(const char** err) {
  std::call_once(g_once, [] {
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) {
      g_err = "cannot load libnccl.so.2 (multi-rank handles need NCCL)";
      return;
    }
    g_ok = bind(lib, "ncclGetUniqueId", &g_api.GetUniqueId) && bind(lib, "ncclCommInitRank", &g_api.CommInitRank) &&
           bind(lib, "ncclCommDestroy", &g_api.CommDestroy) && bind(lib, "ncclAllGather", &g_api.AllGather) &&
           bind(lib, "ncclAllReduce", &g_api.AllReduce) && bind(lib, "ncclSend", &g_api.Send) &&
           bind(lib, "ncclRecv", &g_api.Recv) && bind(lib, "ncclGroupStart", &g_api.GroupStart) &&
           bind(lib, "ncclGroupEnd", &g_api.GroupEnd) && bind(lib, "ncclGetErrorString", &g_api.GetErrorString);
    if (!g_ok) g_err = "libnccl.so.2 lacks a required symbol";
  });
  if (!g_ok) {
    if (err) *err = g_err;
    return nullptr;
  }
  return &g_api;
}

}  // namespace b2
