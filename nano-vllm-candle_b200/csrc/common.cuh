// Shared declarations for the nvc_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>
#include <stdint.h>
#include <string>

#include "../../include/nvc_b200.h"
#include "../../include/nvc_b200_engine.h"

#define NVC_CUDA_CHECK(ctx, expr)                                               \
  do {                                                                          \
    cudaError_t _e = (expr);                                                    \
    if (_e != cudaSuccess) {                                                    \
      nvc_set_error((ctx), std::string(#expr) + ": " + cudaGetErrorString(_e)); \
      return NVC_CUDA_ERROR_BASE + (int)_e;                                     \
    }                                                                           \
  } while (0)

#define NVC_REQUIRE(ctx, cond, code, msg)                                       \
  do {                                                                          \
    if (!(cond)) {                                                              \
      nvc_set_error((ctx), std::string(__func__) + ": " + (msg));               \
      return (code);                                                            \
    }                                                                           \
  } while (0)

typedef CUresult (*nvc_encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                        const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                        const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct nvc_nccl_api;  // collectives.cu
struct nvc_nvls;      // nvls_allreduce.cu

struct nvc_ctx {
  int device = 0;
  int tp_rank = 0;
  int tp_size = 1;
  int num_sms = 148;
  std::string last_error;
  uint64_t launches = 0;   // kernels launched through this context (bench.py gpu_launches)
  int use_pdl = 1;         // programmatic dependent launch between the kernels of a step (NVC_PDL=0 disables)
  nvc_encode_tiled_fn encode_tiled = nullptr;
  // split-K scratch for the tcgen05 GEMM (fp32 partial tiles + per-tile arrival counters)
  float* gemm_ws = nullptr;
  size_t gemm_ws_bytes = 0;
  int* gemm_counters = nullptr;
  int gemm_num_counters = 0;
  int* pd_counters = nullptr;   // arrival tickets of the work-balanced paged decode kernel (zero between launches)
  int pair_clusters[5] = {0, 0, 0, 0, 0};   // co-resident clusters of 2S CTAs of the pair GEMM, S = 1..4
  // tensor-parallel communicator (NULL when tp_size == 1)
  nvc_nccl_api* nccl = nullptr;
  void* nccl_comm = nullptr;
  // peer-memory all-reduce over NVLink (p2p_allreduce.cu): one cudaMalloc'ed buffer per rank,
  // mapped into every peer with CUDA IPC.  Layout: [flags 4 KB][slot 0][slot 1]
  uint8_t* p2p_local = nullptr;          // this rank's buffer
  uint8_t* p2p_peer[8] = {nullptr};      // every rank's buffer as seen from this rank (own included)
  uint8_t** p2p_peer_d = nullptr;        // the same table in device memory
  uint32_t* p2p_state = nullptr;         // device: [0] epoch, [1] CTAs done
  size_t p2p_slot_bytes = 0;
  bool p2p_ready = false;
  // NVLS (NVSwitch multicast) all-reduce (nvls_allreduce.cu): one buffer per rank bound into a multicast object
  nvc_nvls* nvls = nullptr;
  bool nvls_ready = false;
  int nvls_generation = 0;               // set-ups so far on this context (part of the rendezvous name)
  std::string nvls_status = "not initialised";
  std::string tp_group_name;             // unique per TP group (from the NCCL unique id), same on every rank
};

void nvc_set_error(nvc_ctx* ctx, const std::string& msg);

static inline cudaStream_t nvc_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

#ifdef __CUDACC__
// Launch with programmatic stream serialization: the kernel may start (run its prologue up to
// pdl_wait()) while its predecessor in the stream is still draining.  Every kernel of this
// library calls pdl_trigger() at its top and pdl_wait() before it reads or writes anything another
// kernel of the step produces; operands no kernel of the step writes (weights, norm vectors, rope
// table, positions / slot_mapping / block_tables / context_lens, older cache rows) may be fetched
// ahead of the wait - the early-launch chain can be several kernels deep, so "written two kernels
// ago" is NOT safe before the wait.
template <typename... KArgs, typename... Args>
static inline cudaError_t nvc_launch_cluster3(const nvc_ctx* ctx, void (*kernel)(KArgs...), dim3 grid, dim3 block,
                                              size_t smem, cudaStream_t stream, dim3 cluster, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (ctx && ctx->use_pdl) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster.x * cluster.y * cluster.z > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster.x;
    attr[n].val.clusterDim.y = cluster.y;
    attr[n].val.clusterDim.z = cluster.z;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
static inline cudaError_t nvc_launch_cluster(const nvc_ctx* ctx, void (*kernel)(KArgs...), dim3 grid, dim3 block,
                                             size_t smem, cudaStream_t stream, int cluster_x, Args&&... args) {
  return nvc_launch_cluster3(ctx, kernel, grid, block, smem, stream, dim3(cluster_x > 1 ? cluster_x : 1, 1, 1),
                             static_cast<Args&&>(args)...);
}
template <typename... KArgs, typename... Args>
static inline cudaError_t nvc_launch(const nvc_ctx* ctx, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                     cudaStream_t stream, Args&&... args) {
  return nvc_launch_cluster(ctx, kernel, grid, block, smem, stream, 1, static_cast<Args&&>(args)...);
}
#endif

// ---------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------
#ifdef __CUDACC__

// Programmatic dependent launch (griddepcontrol): wait = all prerequisite grids have completed
// and their memory is visible; trigger = the next grid in the stream may be scheduled.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);   // .x = lo (low 16 bits), .y = hi
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  f[0] = bf16_lo(u.x); f[1] = bf16_hi(u.x); f[2] = bf16_lo(u.y); f[3] = bf16_hi(u.y);
  f[4] = bf16_lo(u.z); f[5] = bf16_hi(u.z); f[6] = bf16_lo(u.w); f[7] = bf16_hi(u.w);
}

__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 u;
  u.x = pack_bf16(f[0], f[1]); u.y = pack_bf16(f[2], f[3]);
  u.z = pack_bf16(f[4], f[5]); u.w = pack_bf16(f[6], f[7]);
  return u;
}

// q / k RMSNorm + neox RoPE of one 128-element head row held by 8 lanes: lane l owns ra = elements [8l, 8l + 8) and
// rb = [64 + 8l, 64 + 8l + 8) (the rotation pairs (d, d + 64) are lane-local); c0 c1 / s0 s1 = cos / sin of frequencies
// [8l, 8l + 8); mask = the 8 lanes of this unit.  has_w = false: rope only.  Every lane of the unit must call it (shuffles);
// inactive units pass zeros.  Shared by qk_norm_rope_store_kernel and the QKV GEMM epilogue, which are bit-identical.
// (src/models/qwen3.rs:206-234, src/layers/layernorm.rs:44-60, src/layers/rotary_embedding.rs:82-107)
__device__ __forceinline__ void qk_norm_rope8(const uint4& ra, const uint4& rb, unsigned mask, bool has_w, const uint4& w_lo,
                                              const uint4& w_hi, float eps, const float4& c0, const float4& c1,
                                              const float4& s0, const float4& s1, uint4& o1, uint4& o2) {
  float a[8], b[8];
  unpack8(ra, a);
  unpack8(rb, b);
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) ss += a[i] * a[i] + b[i] * b[i];
  ss += __shfl_xor_sync(mask, ss, 1);
  ss += __shfl_xor_sync(mask, ss, 2);
  ss += __shfl_xor_sync(mask, ss, 4);
  float rstd = 1.0f;
  float wa[8], wb[8];
  if (has_w) {
    rstd = 1.0f / sqrtf(ss * (1.0f / 128.0f) + eps);
    unpack8(w_lo, wa);
    unpack8(w_hi, wb);
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) wa[i] = wb[i] = 1.0f;
  }
  const float c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
  const float s[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
  float y1[8], y2[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float x1 = a[i] * rstd * wa[i];
    const float x2 = b[i] * rstd * wb[i];
    y1[i] = x1 * c[i] - x2 * s[i];
    y2[i] = x2 * c[i] + x1 * s[i];
  }
  o1 = pack8(y1);
  o2 = pack8(y2);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ uint4 ldg_nc_u4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

#endif  // __CUDACC__
