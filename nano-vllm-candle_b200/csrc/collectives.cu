// K9: tensor-parallel collectives.  The reference has none (src/tp.rs:1-70 is config only,
// src/layers/linear.rs:184-198 does no all-reduce); north_star (4) adds an all-reduce after
// each row-parallel linear.  NCCL is resolved at run time with dlopen so that the library
// loads (and TP=1 runs) without it; with torch in the process this binds to the same
// libnccl.so.2 torch already loaded.
#include "common.cuh"

#include <dlfcn.h>
#include <nccl.h>

struct nvc_nccl_api {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static nvc_nccl_api* load_nccl(std::string* err) {
  static nvc_nccl_api api;
  static bool tried = false, ok = false;
  if (tried) return ok ? &api : nullptr;
  tried = true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  if (!api.lib) {
    if (err) *err = std::string("dlopen(libnccl.so.2) failed: ") + dlerror();
    return nullptr;
  }
#define NVC_SYM(field, name)                                            \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.lib, name)); \
  if (!api.field) { if (err) *err = std::string("missing NCCL symbol ") + name; return nullptr; }
  NVC_SYM(GetUniqueId, "ncclGetUniqueId")
  NVC_SYM(CommInitRank, "ncclCommInitRank")
  NVC_SYM(CommDestroy, "ncclCommDestroy")
  NVC_SYM(AllReduce, "ncclAllReduce")
  NVC_SYM(AllGather, "ncclAllGather")
  NVC_SYM(GetErrorString, "ncclGetErrorString")
#undef NVC_SYM
  ok = true;
  return &api;
}

#define NVC_NCCL_CHECK(ctx, api, expr)                                                   \
  do {                                                                                   \
    ncclResult_t _r = (expr);                                                            \
    if (_r != ncclSuccess) {                                                             \
      nvc_set_error((ctx), std::string(#expr) + ": " + (api)->GetErrorString(_r));       \
      return NVC_NCCL_ERROR_BASE + (int)_r;                                              \
    }                                                                                    \
  } while (0)

void nvc_collectives_destroy(nvc_ctx* ctx) {
  if (ctx->nccl && ctx->nccl_comm) {
    ctx->nccl->CommDestroy(reinterpret_cast<ncclComm_t>(ctx->nccl_comm));
    ctx->nccl_comm = nullptr;
  }
}

extern "C" int nvc_nccl_unique_id(void* unique_id_128B) {
  if (!unique_id_128B) return NVC_EINVAL;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
  std::string err;
  nvc_nccl_api* api = load_nccl(&err);
  if (!api) { nvc_set_error(nullptr, err); return NVC_ENCCL; }
  ncclResult_t r = api->GetUniqueId(reinterpret_cast<ncclUniqueId*>(unique_id_128B));
  if (r != ncclSuccess) { nvc_set_error(nullptr, api->GetErrorString(r)); return NVC_NCCL_ERROR_BASE + (int)r; }
  return NVC_OK;
}

extern "C" int nvc_ctx_init_nccl(nvc_ctx* ctx, const void* unique_id_128B) {
  NVC_REQUIRE(ctx, ctx && unique_id_128B, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, ctx->nccl_comm == nullptr, NVC_ESTATE, "communicator already initialised");
  std::string err;
  nvc_nccl_api* api = load_nccl(&err);
  NVC_REQUIRE(ctx, api != nullptr, NVC_ENCCL, err);
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  ncclUniqueId id;
  memcpy(&id, unique_id_128B, sizeof(id));
  ncclComm_t comm;
  NVC_NCCL_CHECK(ctx, api, api->CommInitRank(&comm, ctx->tp_size, id, ctx->tp_rank));
  ctx->nccl = api;
  ctx->nccl_comm = comm;
  // a name every rank of this group derives identically (rendezvous of the NVLS set-up)
  uint64_t h = 1469598103934665603ull;
  for (size_t i = 0; i < sizeof(id); ++i) h = (h ^ (uint8_t)id.internal[i]) * 1099511628211ull;
  char nm[32];
  snprintf(nm, sizeof(nm), "%016llx", (unsigned long long)h);
  ctx->tp_group_name = nm;
  return NVC_OK;
}

static int allreduce(nvc_ctx* ctx, void* buf, size_t count, ncclDataType_t dt, void* stream) {
  NVC_REQUIRE(ctx, ctx && buf, NVC_EINVAL, "null pointer");
  if (ctx->tp_size == 1 || count == 0) return NVC_OK;
  NVC_REQUIRE(ctx, ctx->nccl_comm != nullptr, NVC_ESTATE, "nvc_ctx_init_nccl not called");
  NVC_NCCL_CHECK(ctx, ctx->nccl, ctx->nccl->AllReduce(buf, buf, count, dt, ncclSum,
                                                      reinterpret_cast<ncclComm_t>(ctx->nccl_comm), nvc_stream(stream)));
  return NVC_OK;
}

extern "C" int nvc_allreduce_bf16(nvc_ctx* ctx, nvc_bf16* buf, size_t count, void* stream) {
  return allreduce(ctx, buf, count, ncclBfloat16, stream);
}
extern "C" int nvc_allreduce_f32(nvc_ctx* ctx, float* buf, size_t count, void* stream) {
  return allreduce(ctx, buf, count, ncclFloat32, stream);
}

extern "C" int nvc_allgather_bytes(nvc_ctx* ctx, void* buf, size_t bytes_per_rank, void* stream) {
  NVC_REQUIRE(ctx, ctx && buf, NVC_EINVAL, "null pointer");
  if (ctx->tp_size == 1 || bytes_per_rank == 0) return NVC_OK;
  NVC_REQUIRE(ctx, ctx->nccl_comm != nullptr, NVC_ESTATE, "nvc_ctx_init_nccl not called");
  const char* send = reinterpret_cast<const char*>(buf) + (size_t)ctx->tp_rank * bytes_per_rank;
  NVC_NCCL_CHECK(ctx, ctx->nccl, ctx->nccl->AllGather(send, buf, bytes_per_rank, ncclInt8,
                                                      reinterpret_cast<ncclComm_t>(ctx->nccl_comm), nvc_stream(stream)));
  return NVC_OK;
}
