// Context lifetime, status strings, driver entry points.
#include "common.cuh"
#include "internal.h"

#include <cstdlib>
#include <cstring>

void nvc_collectives_destroy(nvc_ctx* ctx);   // collectives.cu
void nvc_p2p_destroy(nvc_ctx* ctx);           // p2p_allreduce.cu
void nvc_nvls_destroy(nvc_ctx* ctx);          // nvls_allreduce.cu

static thread_local std::string g_null_ctx_error;

void nvc_set_error(nvc_ctx* ctx, const std::string& msg) {
  if (ctx) ctx->last_error = msg;
  else g_null_ctx_error = msg;
}

extern "C" int nvc_abi_version(void) { return NVC_ABI_VERSION; }

extern "C" const char* nvc_status_string(int status) {
  switch (status) {
    case NVC_OK: return "ok";
    case NVC_EINVAL: return "invalid argument";
    case NVC_ESHAPE: return "unsupported shape";
    case NVC_ENOTIMPL: return "not implemented";
    case NVC_ESTATE: return "invalid state";
    case NVC_ENOMEM: return "workspace or KV pool too small";
    case NVC_EDEVICE: return "no usable sm_100 device";
    case NVC_ENCCL: return "NCCL unavailable";
    case NVC_ENVLS: return "NVLS multicast unavailable";
    default: break;
  }
  if (status >= NVC_NCCL_ERROR_BASE) return "NCCL error";
  if (status >= NVC_CUDA_ERROR_BASE) return cudaGetErrorString((cudaError_t)(status - NVC_CUDA_ERROR_BASE));
  return "unknown status";
}

extern "C" const char* nvc_last_error(const nvc_ctx* ctx) {
  return ctx ? ctx->last_error.c_str() : g_null_ctx_error.c_str();
}

extern "C" int nvc_ctx_create(int device, int tp_rank, int tp_size, nvc_ctx** out) {
  if (!out) return NVC_EINVAL;
  *out = nullptr;
  if (tp_size < 1 || tp_rank < 0 || tp_rank >= tp_size) {
    nvc_set_error(nullptr, "nvc_ctx_create: bad tp_rank/tp_size");
    return NVC_EINVAL;
  }
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || device < 0 || device >= count) {
    nvc_set_error(nullptr, std::string("nvc_ctx_create: no CUDA device ") + std::to_string(device) + " (" +
                               (e == cudaSuccess ? "out of range" : cudaGetErrorString(e)) + ")");
    return NVC_EDEVICE;
  }
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) return NVC_CUDA_ERROR_BASE + (int)e;
  if (prop.major != 10) {
    nvc_set_error(nullptr, "nvc_ctx_create: device is sm_" + std::to_string(prop.major * 10 + prop.minor) +
                               ", this library is built for sm_100a only");
    return NVC_EDEVICE;
  }
  e = cudaSetDevice(device);
  if (e != cudaSuccess) return NVC_CUDA_ERROR_BASE + (int)e;
  nvc_ctx* ctx = new nvc_ctx();
  ctx->device = device;
  ctx->tp_rank = tp_rank;
  ctx->tp_size = tp_size;
  ctx->num_sms = prop.multiProcessorCount;
  if (const char* e = getenv("NVC_PDL")) ctx->use_pdl = atoi(e) != 0;

  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
    delete ctx;
    nvc_set_error(nullptr, "nvc_ctx_create: cuTensorMapEncodeTiled not available from the driver");
    return NVC_EDEVICE;
  }
  ctx->encode_tiled = reinterpret_cast<nvc_encode_tiled_fn>(fn);

  ctx->gemm_ws_bytes = (size_t)96 << 20;
  ctx->gemm_num_counters = 8192;
  e = cudaMalloc(&ctx->gemm_ws, ctx->gemm_ws_bytes);
  if (e == cudaSuccess) e = cudaMalloc(&ctx->gemm_counters, ctx->gemm_num_counters * sizeof(int));
  if (e == cudaSuccess) e = cudaMemset(ctx->gemm_counters, 0, ctx->gemm_num_counters * sizeof(int));
  if (e != cudaSuccess) {
    if (ctx->gemm_ws) cudaFree(ctx->gemm_ws);
    if (ctx->gemm_counters) cudaFree(ctx->gemm_counters);
    delete ctx;
    return NVC_CUDA_ERROR_BASE + (int)e;
  }
  int rc = nvc_gemm_init(ctx);
  if (rc == NVC_OK) rc = nvc_gemm_pair_init(ctx);
  if (rc == NVC_OK) rc = nvc_prefill_attn_init(ctx);
  if (rc == NVC_OK) rc = nvc_paged_decode_init(ctx);
  if (rc != NVC_OK) {
    nvc_set_error(nullptr, "nvc_ctx_create: kernel attribute setup failed: " + ctx->last_error);
    cudaFree(ctx->gemm_ws);
    cudaFree(ctx->gemm_counters);
    if (ctx->pd_counters) cudaFree(ctx->pd_counters);
    delete ctx;
    return rc;
  }
  *out = ctx;
  return NVC_OK;
}

extern "C" uint64_t nvc_ctx_launch_count(const nvc_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" int nvc_ctx_destroy(nvc_ctx* ctx) {
  if (!ctx) return NVC_EINVAL;
  cudaSetDevice(ctx->device);
  nvc_nvls_destroy(ctx);
  nvc_p2p_destroy(ctx);
  nvc_collectives_destroy(ctx);
  if (ctx->gemm_ws) cudaFree(ctx->gemm_ws);
  if (ctx->gemm_counters) cudaFree(ctx->gemm_counters);
  if (ctx->pd_counters) cudaFree(ctx->pd_counters);
  delete ctx;
  return NVC_OK;
}

extern "C" int nvc_ctx_device(const nvc_ctx* ctx) { return ctx ? ctx->device : -1; }
extern "C" int nvc_ctx_tp_rank(const nvc_ctx* ctx) { return ctx ? ctx->tp_rank : -1; }
extern "C" int nvc_ctx_tp_size(const nvc_ctx* ctx) { return ctx ? ctx->tp_size : -1; }
extern "C" int nvc_ctx_num_sms(const nvc_ctx* ctx) { return ctx ? ctx->num_sms : -1; }
