// NCCL, bound at run time: libeqpcount.so does not link against libnccl, it dlopens "libnccl.so.2" on the first eqp_comm_* /
// eqp_reduce_* call (a process that already holds a copy of that SONAME — e.g. torch's bundled one — gets that copy, so the two
// never mix).  Only the handful of entry points the sample reduce needs are declared, with the public NCCL 2.x signatures.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <string>

namespace eqpb {

typedef struct ncclComm* nccl_comm_t;
struct nccl_unique_id { char internal[128]; };                       // ncclUniqueId
enum { NCCL_SUCCESS = 0, NCCL_UINT64 = 5, NCCL_SUM = 0 };             // ncclSuccess, ncclUint64, ncclSum

struct NcclApi {
  void* lib = nullptr; std::string error;
  int (*GetUniqueId)(nccl_unique_id*) = nullptr;
  int (*CommInitRank)(nccl_comm_t*, int, nccl_unique_id, int) = nullptr;
  int (*CommInitAll)(nccl_comm_t*, int, const int*) = nullptr;
  int (*Reduce)(const void*, void*, size_t, int, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  int (*CommDestroy)(nccl_comm_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;

  bool load() {
    if (lib) return true;
    if (!error.empty()) return false;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
    if (!lib) { error = std::string("cannot load libnccl.so.2: ") + dlerror(); return false; }
    bool ok = true;
    auto sym = [&](const char* n) { void* p = dlsym(lib, n); if (!p) { ok = false; error = std::string("symbol missing from libnccl: ") + n; } return p; };
    GetUniqueId = (decltype(GetUniqueId))sym("ncclGetUniqueId"); CommInitRank = (decltype(CommInitRank))sym("ncclCommInitRank");
    CommInitAll = (decltype(CommInitAll))sym("ncclCommInitAll"); Reduce = (decltype(Reduce))sym("ncclReduce");
    GroupStart = (decltype(GroupStart))sym("ncclGroupStart"); GroupEnd = (decltype(GroupEnd))sym("ncclGroupEnd");
    CommDestroy = (decltype(CommDestroy))sym("ncclCommDestroy"); GetErrorString = (decltype(GetErrorString))sym("ncclGetErrorString");
    if (!ok) { dlclose(lib); lib = nullptr; }
    return ok;
  }
  std::string text(int rc) const { return GetErrorString ? GetErrorString(rc) : "NCCL error " + std::to_string(rc); }
};

inline NcclApi& nccl_api() { static NcclApi api; return api; }

}  // namespace eqpb
