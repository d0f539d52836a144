// glp_nccl.inl -- run-time binding of NCCL (included by glprox.cu).
#include <dlfcn.h>

namespace {
struct NcclUniqueId { char internal[128]; };
typedef void* NcclComm;
enum { kNcclSum = 0, kNcclMax = 2, kNcclFloat64 = 8 };
typedef int (*PFN_ncclGetUniqueId)(NcclUniqueId*);
typedef int (*PFN_ncclCommInitRank)(NcclComm*, int, NcclUniqueId, int);
typedef int (*PFN_ncclCommDestroy)(NcclComm);
typedef int (*PFN_ncclAllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef const char* (*PFN_ncclGetErrorString)(int);

struct NcclApi {
  void* lib = nullptr;
  PFN_ncclGetUniqueId getUniqueId = nullptr;
  PFN_ncclCommInitRank commInitRank = nullptr;
  PFN_ncclCommDestroy commDestroy = nullptr;
  PFN_ncclAllReduce allReduce = nullptr;
  PFN_ncclGetErrorString errStr = nullptr;
};

NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.lib ? &api : nullptr;
  tried = true;
  // GLPROX_NCCL_LIB lets the host point at torch's bundled copy; otherwise the loader path is used
  // (a copy already mapped into the process by torch is found by its soname).
  const char* names[] = {getenv("GLPROX_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    if (nm == nullptr || nm[0] == 0) continue;
    api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  if (!api.lib) return nullptr;
  api.getUniqueId = reinterpret_cast<PFN_ncclGetUniqueId>(dlsym(api.lib, "ncclGetUniqueId"));
  api.commInitRank = reinterpret_cast<PFN_ncclCommInitRank>(dlsym(api.lib, "ncclCommInitRank"));
  api.commDestroy = reinterpret_cast<PFN_ncclCommDestroy>(dlsym(api.lib, "ncclCommDestroy"));
  api.allReduce = reinterpret_cast<PFN_ncclAllReduce>(dlsym(api.lib, "ncclAllReduce"));
  api.errStr = reinterpret_cast<PFN_ncclGetErrorString>(dlsym(api.lib, "ncclGetErrorString"));
  if (!api.getUniqueId || !api.commInitRank || !api.commDestroy || !api.allReduce) {
    dlclose(api.lib);
    api.lib = nullptr;
    return nullptr;
  }
  return &api;
}

int nccl_fail(NcclApi* api, const char* what, int rc) {
  return fail(GLP_ERR_NCCL, "%s failed: %s", what, api && api->errStr ? api->errStr(rc) : "nccl error");
}
}  // namespace

extern "C" int glp_nccl_unique_id(void* id128) {
  if (id128 == nullptr) return fail(GLP_ERR_ARG, "null id buffer");
  NcclApi* api = nccl_api();
  if (!api) return fail(GLP_ERR_NCCL, "libnccl.so.2 could not be loaded (%s)", dlerror());
  NcclUniqueId id;
  int rc = api->getUniqueId(&id);
  if (rc != 0) return nccl_fail(api, "ncclGetUniqueId", rc);
  memcpy(id128, &id, sizeof(id));
  return GLP_OK;
}

extern "C" int glp_nccl_comm_create(void** comm_out, const void* id128, int rank, int world, int device) {
  if (comm_out == nullptr || id128 == nullptr || world <= 0 || rank < 0 || rank >= world)
    return fail(GLP_ERR_ARG, "bad argument");
  NcclApi* api = nccl_api();
  if (!api) return fail(GLP_ERR_NCCL, "libnccl.so.2 could not be loaded (%s)", dlerror());
  CU_TRY(cudaSetDevice(device));
  NcclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  NcclComm comm = nullptr;
  int rc = api->commInitRank(&comm, world, id, rank);
  if (rc != 0) return nccl_fail(api, "ncclCommInitRank", rc);
  *comm_out = comm;
  return GLP_OK;
}

extern "C" int glp_nccl_comm_destroy(void* comm) {
  if (comm == nullptr) return GLP_OK;
  NcclApi* api = nccl_api();
  if (!api) return fail(GLP_ERR_NCCL, "libnccl.so.2 not loaded");
  int rc = api->commDestroy(comm);
  if (rc != 0) return nccl_fail(api, "ncclCommDestroy", rc);
  return GLP_OK;
}

static int glp_allreduce_f64(glp_handle* h, double* buf, size_t count, bool is_max) {
  NcclApi* api = nccl_api();
  if (!api || h->nccl_comm == nullptr) return fail(GLP_ERR_NCCL, "allreduce without an NCCL communicator");
  int rc = api->allReduce(buf, buf, count, kNcclFloat64, is_max ? kNcclMax : kNcclSum, h->nccl_comm, h->stream);
  if (rc != 0) return nccl_fail(api, "ncclAllReduce", rc);
  return GLP_OK;
}
