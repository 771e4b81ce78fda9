// K9 (B200-native): one-shot all-reduce over NVLink peer memory, fused with the residual add and
// RMSNorm that always follow a row-parallel linear (o_proj / down_proj).  The reference has no
// collective at all (src/layers/linear.rs:184-198); NCCL's all-reduce costs ~12 us for the 128 KB
// decode messages (56 of them per Qwen3-0.6B step), this kernel reads the peers' partial sums
// directly (P2P loads through NVSwitch) and replaces TWO kernels (all-reduce + fused_add_rmsnorm).
//
// Protocol (per call, epoch e = previous epoch + 1, all state lives in device memory so the call
// can be replayed from a CUDA graph):
//   1. the row-parallel GEMM of every rank wrote its partial [N, H] bf16 into its own slot s of
//      the IPC-mapped buffer (slots alternate between o_proj and down_proj, see below);
//   2. CTA 0 publishes "my partial is complete" by storing e into flag[my_rank] of EVERY rank's
//      buffer (system-scope release); every CTA then waits until all tp flags of its own buffer
//      reach e (acquire);
//   3. CTA = one token row: sum the tp partial rows in rank order (fp32, identical on every
//      rank), add the residual, store the new residual (fp32), RMSNorm, store bf16;
//   4. the last CTA to finish advances the epoch.
// No end barrier: a rank overwrites slot s again only two all-reduces later, and it cannot pass
// the barrier in between before every peer has entered it, i.e. after the peer finished reading.
#include "common.cuh"
#include "internal.h"

#include <cstdlib>
#include <cstring>

namespace {

constexpr size_t kFlagBytes = 4096;
constexpr unsigned kSpinLimit = 1u << 24;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_u4(const void* p) {     // peer memory: bypass the (non-coherent) L1
  uint4 r;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}

// x (the partial sums) is read from slot `slot` of every rank's buffer.  w == NULL: plain
// all-reduce (out = bf16(sum), no residual / norm).
template <int TP>
__global__ void __launch_bounds__(1024)
p2p_allreduce_rmsnorm_kernel(uint8_t* const* __restrict__ peers, uint32_t* __restrict__ state, int my_rank,
                             size_t slot_off, const float4* __restrict__ res_in, float4* __restrict__ res_out,
                             const uint4* __restrict__ w, float eps, uint4* __restrict__ out, int H) {
  __shared__ float red[32];
  __shared__ uint32_t s_epoch;
  pdl_trigger();
  pdl_wait();                                  // this rank's GEMM has completed and flushed
  if (threadIdx.x == 0) s_epoch = state[0] + 1;
  __syncthreads();
  const uint32_t e = s_epoch;
  if (blockIdx.x == 0 && threadIdx.x < TP) {
    // the partial was written by the previous kernel (complete and flushed: pdl_wait above), so a
    // system-scope release store is all the ordering the flag needs
    st_release_sys(reinterpret_cast<uint32_t*>(peers[threadIdx.x]) + my_rank, e);
  }
  if (threadIdx.x < TP) {
    const uint32_t* flag = reinterpret_cast<const uint32_t*>(peers[my_rank]) + threadIdx.x;
    unsigned spins = 0;
    while ((int32_t)(ld_acquire_sys(flag) - e) < 0) {
      if (++spins > kSpinLimit) {
        printf("nvc: p2p all-reduce timed out waiting for rank %d (epoch %u)\n", (int)threadIdx.x, e);
        __trap();
      }
    }
  }
  __syncthreads();

  const int row = blockIdx.x;
  const int nvec = H >> 3;
  const int v = threadIdx.x;
  float s[8];
  float ss = 0.f;
  if (v < nvec) {
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = 0.f;
    uint4 part[TP];
#pragma unroll
    for (int p = 0; p < TP; ++p)
      part[p] = ld_peer_u4(peers[p] + slot_off + ((size_t)row * nvec + v) * 16);
#pragma unroll
    for (int p = 0; p < TP; ++p) {             // fixed rank order: bit-identical on every rank
      float f[8];
      unpack8(part[p], f);
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] += f[i];
    }
    if (w != nullptr) {
      const int64_t o = ((int64_t)row * nvec + v) * 2;
      if (res_in) {
        const float4 a = res_in[o], b = res_in[o + 1];
        s[0] += a.x; s[1] += a.y; s[2] += a.z; s[3] += a.w;
        s[4] += b.x; s[5] += b.y; s[6] += b.z; s[7] += b.w;
      }
      if (res_out) {
        res_out[o] = make_float4(s[0], s[1], s[2], s[3]);
        res_out[o + 1] = make_float4(s[4], s[5], s[6], s[7]);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) ss += s[i] * s[i];
    }
  }
  if (w != nullptr) {
    ss = warp_sum(ss);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) red[wid] = ss;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    float tot = (lane < nw) ? red[lane] : 0.f;
    tot = warp_sum(tot);
    const float rstd = 1.0f / sqrtf(tot / (float)H + eps);
    if (v < nvec) {
      float wf[8];
      unpack8(__ldg(w + v), wf);
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] = s[i] * rstd * wf[i];
    }
  }
  if (v < nvec) out[(int64_t)row * nvec + v] = pack8(s);

  // the last CTA to finish publishes the epoch for the next call
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t done = atomicAdd(&state[1], 1u);
    if (done == gridDim.x - 1) {
      state[1] = 0;
      __threadfence();
      state[0] = e;
    }
  }
}

}  // namespace

void nvc_p2p_destroy(nvc_ctx* ctx) {
  if (!ctx->p2p_local) return;
  for (int r = 0; r < ctx->tp_size; ++r)
    if (r != ctx->tp_rank && ctx->p2p_peer[r]) cudaIpcCloseMemHandle(ctx->p2p_peer[r]);
  cudaFree(ctx->p2p_local);
  if (ctx->p2p_peer_d) cudaFree(ctx->p2p_peer_d);
  if (ctx->p2p_state) cudaFree(ctx->p2p_state);
  ctx->p2p_local = nullptr;
  ctx->p2p_ready = false;
}

extern "C" int nvc_ctx_p2p_export(nvc_ctx* ctx, size_t slot_bytes, void* handle_64B) {
  NVC_REQUIRE(ctx, ctx && handle_64B, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, ctx->tp_size > 1 && ctx->tp_size <= 8, NVC_EINVAL, "peer all-reduce needs 2..8 ranks");
  NVC_REQUIRE(ctx, ctx->p2p_local == nullptr, NVC_ESTATE, "peer buffer already exported");
  NVC_REQUIRE(ctx, slot_bytes > 0 && slot_bytes % 256 == 0, NVC_EINVAL, "slot_bytes must be a multiple of 256");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t size");
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  const size_t total = kFlagBytes + 2 * slot_bytes;      // flags | slot 0 | slot 1
  NVC_CUDA_CHECK(ctx, cudaMalloc(&ctx->p2p_local, total));
  NVC_CUDA_CHECK(ctx, cudaMemset(ctx->p2p_local, 0, total));
  NVC_CUDA_CHECK(ctx, cudaMalloc(&ctx->p2p_state, 4 * sizeof(uint32_t)));      // epoch, CTAs done, rows pushed
  NVC_CUDA_CHECK(ctx, cudaMemset(ctx->p2p_state, 0, 4 * sizeof(uint32_t)));
  NVC_CUDA_CHECK(ctx, cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  NVC_CUDA_CHECK(ctx, cudaIpcGetMemHandle(&h, ctx->p2p_local));
  memcpy(handle_64B, &h, sizeof(h));
  ctx->p2p_slot_bytes = slot_bytes;
  return NVC_OK;
}

extern "C" int nvc_ctx_p2p_import(nvc_ctx* ctx, const void* handles_tp_x_64B) {
  NVC_REQUIRE(ctx, ctx && handles_tp_x_64B, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, ctx->p2p_local != nullptr && !ctx->p2p_ready, NVC_ESTATE, "call nvc_ctx_p2p_export first, once");
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  const uint8_t* hs = reinterpret_cast<const uint8_t*>(handles_tp_x_64B);
  for (int r = 0; r < ctx->tp_size; ++r) {
    if (r == ctx->tp_rank) { ctx->p2p_peer[r] = ctx->p2p_local; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, hs + (size_t)r * 64, 64);
    void* p = nullptr;
    NVC_CUDA_CHECK(ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    ctx->p2p_peer[r] = reinterpret_cast<uint8_t*>(p);
  }
  NVC_CUDA_CHECK(ctx, cudaMalloc(&ctx->p2p_peer_d, 8 * sizeof(uint8_t*)));
  NVC_CUDA_CHECK(ctx, cudaMemcpy(ctx->p2p_peer_d, ctx->p2p_peer, 8 * sizeof(uint8_t*), cudaMemcpyHostToDevice));
  ctx->p2p_ready = true;
  return NVC_OK;
}

extern "C" int nvc_ctx_p2p_ready(const nvc_ctx* ctx) { return ctx && ctx->p2p_ready ? 1 : 0; }
extern "C" size_t nvc_ctx_p2p_slot_bytes(const nvc_ctx* ctx) { return ctx ? ctx->p2p_slot_bytes : 0; }

extern "C" void* nvc_ctx_p2p_slot(nvc_ctx* ctx, int slot) {
  if (!ctx || !ctx->p2p_local || slot < 0 || slot > 1) return nullptr;
  return ctx->p2p_local + kFlagBytes + (size_t)slot * ctx->p2p_slot_bytes;
}

extern "C" int nvc_p2p_allreduce_rmsnorm(nvc_ctx* ctx, int slot, const float* residual_in, float* residual_out,
                                         const nvc_bf16* w, float eps, nvc_bf16* out, int N, int H, void* stream) {
  NVC_REQUIRE(ctx, ctx && out, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, ctx->p2p_ready, NVC_ESTATE, "peer buffers not imported (nvc_ctx_p2p_import)");
  NVC_REQUIRE(ctx, slot == 0 || slot == 1, NVC_EINVAL, "slot must be 0 or 1");
  NVC_REQUIRE(ctx, N > 0 && H > 0 && H % 8 == 0 && H / 8 <= 1024, NVC_ESHAPE, "H must be a multiple of 8 and <= 8192");
  NVC_REQUIRE(ctx, (size_t)N * H * 2 <= ctx->p2p_slot_bytes, NVC_ENOMEM, "message larger than the peer slot");
  const size_t off = kFlagBytes + (size_t)slot * ctx->p2p_slot_bytes;
  const int threads = ((H / 8 + 31) / 32) * 32;
  cudaStream_t st = nvc_stream(stream);
  // (a two-shot variant - reduce-scatter by peer loads, all-gather by peer stores, second flag barrier - measured slower at
  // every size on 8 x B200, 44 vs 39 us for 2 MB, and was deleted in round 2; large messages go through the switch, nvls_allreduce.cu)
#define NVC_AR(TPV)                                                                                              \
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, p2p_allreduce_rmsnorm_kernel<TPV>, dim3(N), dim3(threads), 0, st,          \
                                 (uint8_t* const*)ctx->p2p_peer_d, ctx->p2p_state, ctx->tp_rank, off,            \
                                 reinterpret_cast<const float4*>(residual_in), reinterpret_cast<float4*>(residual_out), \
                                 reinterpret_cast<const uint4*>(w), eps, reinterpret_cast<uint4*>(out), H))
  switch (ctx->tp_size) {
    case 2: NVC_AR(2); break;
    case 4: NVC_AR(4); break;
    case 8: NVC_AR(8); break;
    default: NVC_REQUIRE(ctx, false, NVC_ESHAPE, "peer all-reduce supports tp_size 2, 4, 8");
  }
#undef NVC_AR
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}
