// smmc_nccl.cpp -- run-time binding of the NCCL entry points the path needs (see smmc_nccl.h).
#include "smmc_nccl.h"

#include <dlfcn.h>

#include <cstdio>
#include <mutex>

namespace smmc {
namespace {

NcclApi g_api;
bool g_tried = false, g_ok = false;
char g_why[256] = "";
std::mutex g_mu;

template <typename F>
bool bind(void *h, const char *name, F *out) {
  void *p = dlsym(h, name);
  if (!p) {
    snprintf(g_why, sizeof g_why, "libnccl lacks %s", name);
    return false;
  }
  *out = reinterpret_cast<F>(p);
  return true;
}

}  // namespace

const NcclApi *nccl_api() {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_tried) return g_ok ? &g_api : nullptr;
  g_tried = true;
  void *h = nullptr;
  for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
    h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) {
    const char *e = dlerror();
    snprintf(g_why, sizeof g_why, "libnccl.so.2 not found (%s)", e ? e : "dlopen failed");
    return nullptr;
  }
  g_ok = bind(h, "ncclGetVersion", &g_api.GetVersion) && bind(h, "ncclGetUniqueId", &g_api.GetUniqueId) &&
         bind(h, "ncclCommInitRank", &g_api.CommInitRank) && bind(h, "ncclCommInitAll", &g_api.CommInitAll) &&
         bind(h, "ncclCommDestroy", &g_api.CommDestroy) && bind(h, "ncclCommCount", &g_api.CommCount) &&
         bind(h, "ncclCommUserRank", &g_api.CommUserRank) && bind(h, "ncclCommCuDevice", &g_api.CommCuDevice) &&
         bind(h, "ncclAllReduce", &g_api.AllReduce) && bind(h, "ncclGroupStart", &g_api.GroupStart) &&
         bind(h, "ncclGroupEnd", &g_api.GroupEnd) && bind(h, "ncclGetErrorString", &g_api.GetErrorString);
  return g_ok ? &g_api : nullptr;
}

const char *nccl_load_error() { return g_why; }

}  // namespace smmc
