// K9 on NVLS: all-reduce through the NVSwitch multicast object of the TP group, fused with the residual
// add and RMSNorm that always follow a row-parallel linear (o_proj: src/layers/linear.rs:184-198; down_proj:
// src/models/qwen3.rs:326 - the reference has no collective there, north_star (4) adds one).
//
// Every rank owns one physical buffer (cuMemCreate) that is bound into ONE multicast object
// (cuMulticastCreate / AddDevice / BindMem) and mapped twice: at a unicast address (this rank's copy) and
// at the multicast address (a store there lands in every rank's copy, a multimem.ld_reduce there returns
// the sum over every rank's copy, added inside the switch).  Layout, identical on every rank:
//
//   [ flags 4 KB | slot (row-parallel GEMM output, the runner's x buffer) | LL buffer 0 | LL 1 | LL 2 ]
//
// Two kernels, chosen by message size (nvc_nvls_allreduce_rmsnorm):
//
//  * LL one-shot (small messages, latency-bound).  Rank r multimem.st's its partial rows into range r of LL
//    buffer (epoch % 3) - ONE store per 16 bytes reaches all ranks - and every rank then reads the tp ranges
//    from its OWN memory, spinning on the data itself: the buffers are kept filled with a sentinel word no
//    partial can contain (0xFFFFFFFF, two bf16 NaNs with a payload the sender never emits), so a vector is
//    there when none of its words is the sentinel.  No flag, no fence, no peer load: the critical path is
//    one NVLink traversal.  Sum in rank order (bit-identical on all ranks), residual add, RMSNorm, store.
//    Each call also refills buffer (epoch + 2) % 3 - the one the PREVIOUS call read - with sentinels; a
//    peer can be at most one call ahead (it needs this rank's data to finish a call), so the buffer being
//    refilled is never one a peer is writing.
//  * Two-shot in-switch (large messages, bandwidth-bound).  Barrier on a multicast flag
//    (multimem.red.add), then rank r multimem.ld_reduce's its 1/tp of the slot - the switch adds the tp
//    copies in flight, tp times fewer bytes enter each GPU than with peer loads - and multimem.st's the
//    sums back in place into every rank's slot; second flag barrier; residual add + RMSNorm of all rows
//    from local memory.  The optional LL variant broadcasts the sums into an LL buffer instead and skips
//    the second barrier (the store -> fence -> flag chain is two NVLink traversals, the LL read is one).
//
// All state (epochs, flags) lives in device memory, so the calls replay from a CUDA graph.
#include "common.cuh"
#include "internal.h"

#include <cerrno>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <thread>

#include <poll.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>

namespace {

constexpr size_t kNvlsFlagBytes = 4096;
constexpr unsigned kSpinLimit = 1u << 24;
constexpr uint32_t kSentinel = 0xFFFFFFFFu;
constexpr uint32_t kEpochMod = 3u << 28;     // the LL epoch counts modulo a multiple of 3

// ------------------------------------------------------------------------------------------ device
__device__ __forceinline__ void mm_st_u4(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 mm_ld_reduce_bf16x8(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
  return r;
}
__device__ __forceinline__ void mm_red_add_release(uint32_t* mc, uint32_t v) {
  asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_volatile_u4(const void* p) {     // written by peers through NVLink: never from L1
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ bool vec_ready(const uint4& v) {
  return v.x != kSentinel && v.y != kSentinel && v.z != kSentinel && v.w != kSentinel;
}
// a bf16 half of 0xFFFF (a NaN no arithmetic produces) would make a word look like "not arrived": send 0x7FFF instead
__device__ __forceinline__ uint32_t scrub(uint32_t w) {
  if ((w & 0xFFFFu) == 0xFFFFu) w &= ~0x8000u;
  if ((w >> 16) == 0xFFFFu) w &= ~0x80000000u;
  return w;
}

// state (this rank's device memory): [0] LL epoch, [1] CTAs finished (LL kernel), [2..4] vectors in use in LL
// buffer 0..2, [5] two-shot epoch, [6] CTAs finished (two-shot kernel), [7] CTAs arrived at barrier 2, [8] barrier-2 epoch
// (calls of the flag variant: the LL variant has no second barrier)
struct Epilogue {
  const float4* res_in;
  float4* res_out;
  const uint4* w;
  float eps;
  uint4* out;
};

// residual add + RMSNorm + store of one row held as 8 floats per thread (v < nvec)
__device__ __forceinline__ void row_epilogue(float (&s)[8], const Epilogue& ep, int row, int v, int nvec, int H, float* red) {
  float ss = 0.f;
  if (ep.w != nullptr && v < nvec) {
    const int64_t o = ((int64_t)row * nvec + v) * 2;
    if (ep.res_in) {
      const float4 a = ep.res_in[o], b = ep.res_in[o + 1];
      s[0] += a.x; s[1] += a.y; s[2] += a.z; s[3] += a.w;
      s[4] += b.x; s[5] += b.y; s[6] += b.z; s[7] += b.w;
    }
    if (ep.res_out) {
      ep.res_out[o] = make_float4(s[0], s[1], s[2], s[3]);
      ep.res_out[o + 1] = make_float4(s[4], s[5], s[6], s[7]);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) ss += s[i] * s[i];
  }
  if (ep.w != nullptr) {
    ss = warp_sum(ss);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();                               // red[] of the previous row has been read
    if (lane == 0) red[wid] = ss;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    float tot = (lane < nw) ? red[lane] : 0.f;
    tot = warp_sum(tot);
    const float rstd = 1.0f / sqrtf(tot / (float)H + ep.eps);
    if (v < nvec) {
      float wf[8];
      unpack8(__ldg(ep.w + v), wf);
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] = s[i] * rstd * wf[i];
    }
  }
  if (v < nvec) ep.out[(int64_t)row * nvec + v] = pack8(s);
}

// LL one-shot.  grid = N rows, block = H / 8 threads rounded up to a warp.  x = this rank's partial [N, H] bf16.
template <int TP>
__global__ void __launch_bounds__(1024)
nvls_ll_allreduce_rmsnorm_kernel(const uint4* __restrict__ x, uint8_t* __restrict__ uc_ll, uint8_t* __restrict__ mc_ll,
                                 size_t ll_buf_bytes, uint32_t* __restrict__ state, int my_rank, Epilogue ep, int H) {
  __shared__ float red[32];
  __shared__ uint32_t s_epoch, s_clear;
  pdl_trigger();
  pdl_wait();                                  // this rank's GEMM has completed; the previous all-reduce too
  const int nvec = H >> 3;
  const uint32_t T = gridDim.x * (uint32_t)nvec;          // vectors of one rank's message
  const int row = blockIdx.x, v = threadIdx.x;
  const uint32_t i = (uint32_t)row * nvec + v;
  uint4 p = make_uint4(0u, 0u, 0u, 0u);
  if (v < nvec) p = x[i];                      // in flight while thread 0 fetches the epoch
  if (threadIdx.x == 0) {                      // one L2 round trip; the epoch is published at the END of a call
    const volatile uint32_t* vs = state;
    const uint32_t e = (vs[0] + 1) % kEpochMod;
    s_epoch = e;
    s_clear = vs[2 + (e + 2) % 3];
  }
  __syncthreads();
  const uint32_t e = s_epoch;
  uint8_t* mine_mc = mc_ll + (size_t)(e % 3) * ll_buf_bytes;
  const uint8_t* cur = uc_ll + (size_t)(e % 3) * ll_buf_bytes;
  if (v < nvec) {                              // push: one store, tp destinations
    p.x = scrub(p.x); p.y = scrub(p.y); p.z = scrub(p.z); p.w = scrub(p.w);
    mm_st_u4(mine_mc + ((size_t)my_rank * T + i) * 16, p);
  }
  {                                            // refill the buffer the previous call read
    uint4* old = reinterpret_cast<uint4*>(uc_ll + (size_t)((e + 2) % 3) * ll_buf_bytes);
    const uint32_t used = s_clear;
    const uint4 sent = make_uint4(kSentinel, kSentinel, kSentinel, kSentinel);
    for (uint32_t j = (uint32_t)row * blockDim.x + v; j < used; j += gridDim.x * blockDim.x) old[j] = sent;
  }
  float s[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = 0.f;
  if (v < nvec) {
    uint4 part[TP];
    bool ok[TP];
#pragma unroll
    for (int p = 0; p < TP; ++p) {
      part[p] = ld_volatile_u4(cur + ((size_t)p * T + i) * 16);
      ok[p] = vec_ready(part[p]);
    }
    unsigned spins = 0;
    for (;;) {
      bool all = true;
#pragma unroll
      for (int p = 0; p < TP; ++p) {
        if (!ok[p]) {
          part[p] = ld_volatile_u4(cur + ((size_t)p * T + i) * 16);
          ok[p] = vec_ready(part[p]);
          all = all && ok[p];
        }
      }
      if (all) break;
      if (++spins > kSpinLimit) {
        printf("nvc: NVLS LL all-reduce timed out (rank %d, epoch %u, row %d)\n", my_rank, e, row);
        __trap();
      }
    }
#pragma unroll
    for (int p = 0; p < TP; ++p) {             // fixed rank order: bit-identical on every rank
      float f[8];
      unpack8(part[p], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) s[k] += f[k];
    }
  }
  row_epilogue(s, ep, row, v, nvec, H, red);
  // the LAST CTA TO FINISH publishes the epoch and this call's footprint (every CTA has read the old values; the
  // next call starts after this grid has completed)
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&state[1], 1u) == gridDim.x - 1) {
      state[1] = 0;
      state[2 + e % 3] = TP * T;
      state[0] = e;
    }
  }
}

// All-gather + merge of the per-rank (best logit, token) pairs that end a TP step (vocab-parallel lm_head,
// runner.cu run_head): rank r multimem.st's one 16-byte vector per sequence into range r of the current LL
// buffer, every rank reads the tp candidates of each sequence from its own memory and keeps the largest value,
// ties -> the larger token id (= the reference's "last maximum" over the full row, llm_engine.rs:135-142).
// Same epoch / refill protocol as the LL all-reduce, so the calls interleave freely.  CTAs beyond the first
// ceil(B / 128) only help refilling the previous buffer.
template <int TP>
__global__ void __launch_bounds__(128)
nvls_argmax_merge_kernel(const float* val, const int32_t* idx, uint8_t* __restrict__ uc_ll,
                         uint8_t* __restrict__ mc_ll, size_t ll_buf_bytes, uint32_t* __restrict__ state, int my_rank,
                         int32_t* out_tok, float* out_val, int B) {   // out_tok may alias idx

  __shared__ uint32_t s_epoch, s_clear;
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  uint4 p = make_uint4(0u, 0u, 0u, 0u);
  if (b < B) {
    uint32_t vb = __float_as_uint(val[b]);
    if (vb == kSentinel) vb = 0x7FFFFFFFu;       // a NaN either way; token ids are >= 0, never the sentinel
    p = make_uint4(vb, (uint32_t)idx[b], vb, (uint32_t)idx[b]);
  }
  if (threadIdx.x == 0) {
    const volatile uint32_t* vs = state;
    const uint32_t e = (vs[0] + 1) % kEpochMod;
    s_epoch = e;
    s_clear = vs[2 + (e + 2) % 3];
  }
  __syncthreads();
  const uint32_t e = s_epoch;
  const uint8_t* cur = uc_ll + (size_t)(e % 3) * ll_buf_bytes;
  if (b < B) mm_st_u4(mc_ll + (size_t)(e % 3) * ll_buf_bytes + ((size_t)my_rank * B + b) * 16, p);
  if (b < B) {
    float best_v = 0.f;
    int32_t best_i = -1;
#pragma unroll
    for (int r = 0; r < TP; ++r) {
      const void* src = cur + ((size_t)r * B + b) * 16;
      uint4 c = ld_volatile_u4(src);
      unsigned spins = 0;
      while (!vec_ready(c)) {
        c = ld_volatile_u4(src);
        if (++spins > kSpinLimit) {
          printf("nvc: NVLS argmax merge timed out (rank %d, epoch %u, seq %d)\n", my_rank, e, b);
          __trap();
        }
      }
      const float v = __uint_as_float(c.x);
      const int32_t i = (int32_t)c.y;
      if (best_i < 0 || v > best_v || (v == best_v && i > best_i)) { best_v = v; best_i = i; }
    }
    out_tok[b] = best_i;
    if (out_val) out_val[b] = best_v;
  }
  {                                            // refill the buffer the previous call read
    uint4* old = reinterpret_cast<uint4*>(uc_ll + (size_t)((e + 2) % 3) * ll_buf_bytes);
    const uint32_t used = s_clear;
    const uint4 sent = make_uint4(kSentinel, kSentinel, kSentinel, kSentinel);
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < used; j += gridDim.x * blockDim.x) old[j] = sent;
  }
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&state[1], 1u) == gridDim.x - 1) {
      state[1] = 0;
      state[2 + e % 3] = (uint32_t)TP * (uint32_t)B;
      state[0] = e;
    }
  }
}

// Two-shot in-switch.  Persistent grid (every CTA is resident: the flag waits cannot dead-lock), block = H / 8
// threads rounded up to a warp.  The partial sums are in the slot (uc_slot / mc_slot = its two mappings).
// LL = true: the sums are broadcast into LL buffer (epoch % 3) and read back by spinning on the data.
template <bool LL>
__global__ void __launch_bounds__(1024)
nvls_twoshot_allreduce_rmsnorm_kernel(uint8_t* __restrict__ uc_slot, uint8_t* __restrict__ mc_slot,
                                      uint32_t* __restrict__ uc_flags, uint32_t* __restrict__ mc_flags,
                                      uint8_t* __restrict__ uc_ll, uint8_t* __restrict__ mc_ll, size_t ll_buf_bytes,
                                      uint32_t* __restrict__ state, int my_rank, int tp, Epilogue ep, int H, int N) {
  __shared__ float red[32];
  __shared__ uint32_t s_epoch, s_ll_epoch, s_clear, s_epoch2;
  pdl_trigger();
  pdl_wait();                                  // this rank's GEMM has completed and flushed
  const int nvec = H >> 3;
  const uint32_t T = (uint32_t)N * nvec;
  if (threadIdx.x == 0) {
    // barrier 1: my partial is complete (previous kernel, flushed); one add reaches every rank's flag
    if (blockIdx.x == 0) mm_red_add_release(mc_flags, 1u);
    const volatile uint32_t* vs = state;
    const uint32_t e = vs[5] + 1;
    s_epoch = e;
    uint32_t le = 0;
    if (LL) {
      le = (vs[0] + 1) % kEpochMod;
      s_clear = vs[2 + (le + 2) % 3];
    }
    s_ll_epoch = le;
    const uint32_t e2 = vs[8] + 1;
    s_epoch2 = e2;
    const uint32_t want = e * (uint32_t)tp;
    unsigned spins = 0;
    while ((int32_t)(ld_acquire_sys_u32(uc_flags) - want) < 0) {
      if (++spins > kSpinLimit) {
        printf("nvc: NVLS two-shot all-reduce timed out at barrier 1 (rank %d, epoch %u)\n", my_rank, e);
        __trap();
      }
    }
  }
  __syncthreads();
  const uint32_t e = s_epoch, le = s_ll_epoch;
  // reduce-scatter + all-gather of my 1/tp of the vectors, both through the switch.  kUnroll independent
  // multimem.ld_reduce per thread are in flight before the first store: the loop is NVLink-latency-bound otherwise
  const uint32_t per = T / tp;                 // T % tp == 0: nvec is a multiple of 8
  uint8_t* dst_mc = LL ? mc_ll + (size_t)(le % 3) * ll_buf_bytes : mc_slot;
  constexpr int kUnroll = 8;
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t j0 = blockIdx.x * blockDim.x + threadIdx.x; j0 < per; j0 += stride * kUnroll) {
    uint4 sum[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const uint32_t j = j0 + u * stride;
      if (j < per) sum[u] = mm_ld_reduce_bf16x8(mc_slot + ((size_t)my_rank * per + j) * 16);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const uint32_t j = j0 + u * stride;
      if (j < per) {
        if (LL) { sum[u].x = scrub(sum[u].x); sum[u].y = scrub(sum[u].y); sum[u].z = scrub(sum[u].z); sum[u].w = scrub(sum[u].w); }
        mm_st_u4(dst_mc + ((size_t)my_rank * per + j) * 16, sum[u]);
      }
    }
  }
  const uint8_t* src = uc_slot;
  if (LL) {
    uint4* old = reinterpret_cast<uint4*>(uc_ll + (size_t)((le + 2) % 3) * ll_buf_bytes);
    const uint32_t used = s_clear;
    const uint4 sent = make_uint4(kSentinel, kSentinel, kSentinel, kSentinel);
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < used; j += gridDim.x * blockDim.x) old[j] = sent;
    src = uc_ll + (size_t)(le % 3) * ll_buf_bytes;
  } else {
    // barrier 2: my sums are in every rank's slot.  The CTAs of this rank meet on a local counter first; the last
    // one to arrive sends ONE add through the switch (a flag add per CTA would serialise tp x grid atomics).
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();                  // this CTA's multimem.st (ordered before by the barrier) ahead of the ticket
      if (atomicAdd(&state[7], 1u) == gridDim.x - 1) {
        state[7] = 0;
        __threadfence_system();                // every CTA's stores ahead of the flag
        mm_red_add_release(mc_flags + 32, 1u);
      }
      const uint32_t want = s_epoch2 * (uint32_t)tp;
      unsigned spins = 0;
      while ((int32_t)(ld_acquire_sys_u32(uc_flags + 32) - want) < 0) {
        if (++spins > kSpinLimit) {
          printf("nvc: NVLS two-shot all-reduce timed out at barrier 2 (rank %d, epoch %u)\n", my_rank, e);
          __trap();
        }
      }
    }
    __syncthreads();
  }
  const int v = threadIdx.x;
  for (int row = blockIdx.x; row < N; row += gridDim.x) {
    float s[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s[k] = 0.f;
    if (v < nvec) {
      const void* p = src + ((size_t)row * nvec + v) * 16;
      uint4 val = ld_volatile_u4(p);
      if (LL) {
        unsigned spins = 0;
        while (!vec_ready(val)) {
          val = ld_volatile_u4(p);
          if (++spins > kSpinLimit) {
            printf("nvc: NVLS two-shot LL all-reduce timed out (rank %d, epoch %u, row %d)\n", my_rank, e, row);
            __trap();
          }
        }
      }
      unpack8(val, s);
    }
    row_epilogue(s, ep, row, v, nvec, H, red);
  }
  if (threadIdx.x == 0) {                      // last CTA to finish: publish the epochs for the next call
    __threadfence();
    if (atomicAdd(&state[6], 1u) == gridDim.x - 1) {
      state[6] = 0;
      if (LL) { state[2 + le % 3] = T; state[0] = le; }
      else state[8] = s_epoch2;
      state[5] = e;
    }
  }
}

// ------------------------------------------------------------------------------------------ host: driver API
struct DriverApi {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*GetErrorString)(CUresult, const char**);
};

bool load_driver(DriverApi* d, std::string* err) {
  struct { const char* name; void** slot; } syms[] = {
      {"cuMemCreate", (void**)&d->MemCreate},
      {"cuMemRelease", (void**)&d->MemRelease},
      {"cuMemAddressReserve", (void**)&d->MemAddressReserve},
      {"cuMemAddressFree", (void**)&d->MemAddressFree},
      {"cuMemMap", (void**)&d->MemMap},
      {"cuMemUnmap", (void**)&d->MemUnmap},
      {"cuMemSetAccess", (void**)&d->MemSetAccess},
      {"cuMemGetAllocationGranularity", (void**)&d->MemGetAllocationGranularity},
      {"cuMemExportToShareableHandle", (void**)&d->MemExportToShareableHandle},
      {"cuMemImportFromShareableHandle", (void**)&d->MemImportFromShareableHandle},
      {"cuMulticastCreate", (void**)&d->MulticastCreate},
      {"cuMulticastAddDevice", (void**)&d->MulticastAddDevice},
      {"cuMulticastBindMem", (void**)&d->MulticastBindMem},
      {"cuMulticastUnbind", (void**)&d->MulticastUnbind},
      {"cuMulticastGetGranularity", (void**)&d->MulticastGetGranularity},
      {"cuDeviceGet", (void**)&d->DeviceGet},
      {"cuDeviceGetAttribute", (void**)&d->DeviceGetAttribute},
      {"cuGetErrorString", (void**)&d->GetErrorString},
  };
  for (auto& s : syms) {
    cudaDriverEntryPointQueryResult q;
    void* fn = nullptr;
    cudaError_t e = cudaGetDriverEntryPoint(s.name, &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
      *err = std::string("driver entry point ") + s.name + " not available";
      return false;
    }
    *s.slot = fn;
  }
  return true;
}

// ------------------------------------------------------------------------------------------ host: rendezvous
// Rank 0 listens on an abstract unix socket named after the group; the other ranks connect.  Carries three
// things the multicast set-up needs between the per-GPU processes: the POSIX file descriptor of the multicast
// object (SCM_RIGHTS), "every rank reached this point" barriers, and a go / no-go so that all ranks take
// the same path when one of them cannot proceed.
struct Rendezvous {
  int rank = 0, tp = 1;
  int lfd = -1;
  int fds[8] = {-1, -1, -1, -1, -1, -1, -1, -1};   // rank 0: socket to rank i; others: fds[0] = socket to rank 0
  std::string err;

  static bool io_all(int fd, void* buf, size_t n, bool wr, int timeout_ms) {
    uint8_t* p = reinterpret_cast<uint8_t*>(buf);
    while (n) {
      pollfd pf{fd, (short)(wr ? POLLOUT : POLLIN), 0};
      int pr = poll(&pf, 1, timeout_ms);
      if (pr <= 0) return false;
      ssize_t k = wr ? send(fd, p, n, MSG_NOSIGNAL) : recv(fd, p, n, 0);
      if (k <= 0) { if (k < 0 && (errno == EINTR || errno == EAGAIN)) continue; return false; }
      p += k; n -= (size_t)k;
    }
    return true;
  }

  bool open(int rank_, int tp_, const std::string& name, int timeout_ms) {
    rank = rank_; tp = tp_;
    sockaddr_un addr;
    memset(&addr, 0, sizeof(addr));
    addr.sun_family = AF_UNIX;
    const std::string nm = "nvc-nvls-" + name;
    const size_t len = std::min(nm.size(), sizeof(addr.sun_path) - 2);
    memcpy(addr.sun_path + 1, nm.data(), len);               // abstract namespace: sun_path[0] == 0
    const socklen_t alen = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + len);
    if (rank == 0) {
      lfd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
      if (lfd < 0 || bind(lfd, (sockaddr*)&addr, alen) != 0 || listen(lfd, 16) != 0) {
        err = std::string("rendezvous socket: ") + strerror(errno);
        return false;
      }
      for (int k = 1; k < tp; ++k) {
        pollfd pf{lfd, POLLIN, 0};
        if (poll(&pf, 1, timeout_ms) <= 0) { err = "rendezvous: a peer did not connect in time"; return false; }
        int c = accept4(lfd, nullptr, nullptr, SOCK_CLOEXEC);
        if (c < 0) { err = std::string("rendezvous accept: ") + strerror(errno); return false; }
        int32_t peer = -1;
        if (!io_all(c, &peer, 4, false, timeout_ms) || peer <= 0 || peer >= tp || fds[peer] >= 0) {
          ::close(c);
          err = "rendezvous: bad hello";
          return false;
        }
        fds[peer] = c;
      }
      return true;
    }
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
      int c = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
      if (c < 0) { err = std::string("rendezvous socket: ") + strerror(errno); return false; }
      if (connect(c, (sockaddr*)&addr, alen) == 0) {
        int32_t me = rank;
        if (!io_all(c, &me, 4, true, timeout_ms)) { ::close(c); err = "rendezvous: hello failed"; return false; }
        fds[0] = c;
        return true;
      }
      ::close(c);
      if (std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(timeout_ms)) {
        err = "rendezvous: rank 0 is not listening";
        return false;
      }
      std::this_thread::sleep_for(std::chrono::milliseconds(5));
    }
  }

  // all ranks pass their status; returns the first non-zero one (rank order), identical on every rank
  bool agree(int32_t mine, int32_t* out, int timeout_ms) {
    if (rank == 0) {
      int32_t agg = mine;
      for (int k = 1; k < tp; ++k) {
        int32_t s = 0;
        if (!io_all(fds[k], &s, 4, false, timeout_ms)) { err = "rendezvous: lost a peer"; return false; }
        if (agg == 0) agg = s;
      }
      for (int k = 1; k < tp; ++k)
        if (!io_all(fds[k], &agg, 4, true, timeout_ms)) { err = "rendezvous: lost a peer"; return false; }
      *out = agg;
      return true;
    }
    if (!io_all(fds[0], &mine, 4, true, timeout_ms) || !io_all(fds[0], out, 4, false, timeout_ms)) {
      err = "rendezvous: lost rank 0";
      return false;
    }
    return true;
  }

  bool send_fd(int to_sock, int fd) {
    char byte = 'F';
    iovec iov{&byte, 1};
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
    memset(ctrl, 0, sizeof(ctrl));
    msghdr msg{};
    msg.msg_iov = &iov; msg.msg_iovlen = 1;
    msg.msg_control = ctrl; msg.msg_controllen = sizeof(ctrl);
    cmsghdr* cm = CMSG_FIRSTHDR(&msg);
    cm->cmsg_level = SOL_SOCKET; cm->cmsg_type = SCM_RIGHTS; cm->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(cm), &fd, sizeof(int));
    return sendmsg(to_sock, &msg, MSG_NOSIGNAL) == 1;
  }
  int recv_fd(int from_sock, int timeout_ms) {
    pollfd pf{from_sock, POLLIN, 0};
    if (poll(&pf, 1, timeout_ms) <= 0) return -1;
    char byte = 0;
    iovec iov{&byte, 1};
    alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))];
    msghdr msg{};
    msg.msg_iov = &iov; msg.msg_iovlen = 1;
    msg.msg_control = ctrl; msg.msg_controllen = sizeof(ctrl);
    if (recvmsg(from_sock, &msg, MSG_CMSG_CLOEXEC) != 1) return -1;
    cmsghdr* cm = CMSG_FIRSTHDR(&msg);
    if (!cm || cm->cmsg_level != SOL_SOCKET || cm->cmsg_type != SCM_RIGHTS) return -1;
    int fd = -1;
    memcpy(&fd, CMSG_DATA(cm), sizeof(int));
    return fd;
  }
  void close_all() {
    for (int& f : fds) if (f >= 0) { ::close(f); f = -1; }
    if (lfd >= 0) { ::close(lfd); lfd = -1; }
  }
};

size_t round_up(size_t v, size_t g) { return (v + g - 1) / g * g; }

}  // namespace

struct nvc_nvls {
  DriverApi drv;
  CUmemGenericAllocationHandle mem = 0, mc = 0;
  bool have_mem = false, have_mc = false, bound = false, uc_mapped = false, mc_mapped = false;
  CUdeviceptr uc_va = 0, mc_va = 0;
  size_t total = 0, slot_bytes = 0, ll_buf_bytes = 0;
  CUdevice dev = 0;
  uint32_t* state = nullptr;
};

void nvc_nvls_destroy(nvc_ctx* ctx) {
  nvc_nvls* n = ctx->nvls;
  if (!n) return;
  cudaDeviceSynchronize();
  if (n->mc_mapped) { n->drv.MemUnmap(n->mc_va, n->total); }
  if (n->mc_va) n->drv.MemAddressFree(n->mc_va, n->total);
  if (n->bound) n->drv.MulticastUnbind(n->mc, n->dev, 0, n->total);
  if (n->uc_mapped) { n->drv.MemUnmap(n->uc_va, n->total); }
  if (n->uc_va) n->drv.MemAddressFree(n->uc_va, n->total);
  if (n->have_mc) n->drv.MemRelease(n->mc);
  if (n->have_mem) n->drv.MemRelease(n->mem);
  if (n->state) cudaFree(n->state);
  delete n;
  ctx->nvls = nullptr;
  ctx->nvls_ready = false;
}

extern "C" const char* nvc_ctx_nvls_status(const nvc_ctx* ctx) { return ctx ? ctx->nvls_status.c_str() : ""; }
extern "C" int nvc_ctx_nvls_ready(const nvc_ctx* ctx) { return ctx && ctx->nvls_ready ? 1 : 0; }
extern "C" size_t nvc_ctx_nvls_slot_bytes(const nvc_ctx* ctx) { return ctx && ctx->nvls_ready ? ctx->nvls->slot_bytes : 0; }
extern "C" size_t nvc_ctx_nvls_ll_bytes(const nvc_ctx* ctx) { return ctx && ctx->nvls_ready ? ctx->nvls->ll_buf_bytes : 0; }
extern "C" void* nvc_ctx_nvls_slot(nvc_ctx* ctx) {
  return ctx && ctx->nvls_ready ? reinterpret_cast<void*>(ctx->nvls->uc_va + kNvlsFlagBytes) : nullptr;
}

// Collective over the TP group: every rank calls it with the same sizes and the same group name.
extern "C" int nvc_ctx_nvls_init(nvc_ctx* ctx, size_t slot_bytes, size_t ll_buf_bytes, const char* group_name) {
  NVC_REQUIRE(ctx, ctx, NVC_EINVAL, "null pointer");
  if (!group_name) group_name = ctx->tp_group_name.c_str();
  NVC_REQUIRE(ctx, group_name[0] != 0, NVC_ESTATE, "no group name: call nvc_ctx_init_nccl first or pass one");
  NVC_REQUIRE(ctx, ctx->tp_size > 1 && ctx->tp_size <= 8, NVC_EINVAL, "NVLS all-reduce needs 2..8 ranks");
  NVC_REQUIRE(ctx, slot_bytes % 256 == 0 && ll_buf_bytes % 256 == 0 && slot_bytes > 0 && ll_buf_bytes > 0, NVC_EINVAL,
              "sizes must be positive multiples of 256");
  NVC_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
  if (ctx->nvls) nvc_nvls_destroy(ctx);
  const int tmo = 120000;
  const std::string name = std::string(group_name) + "-" + std::to_string(ctx->nvls_generation++);
  Rendezvous rv;
  if (!rv.open(ctx->tp_rank, ctx->tp_size, name, tmo)) {
    ctx->nvls_status = "unavailable: " + rv.err;
    rv.close_all();
    nvc_set_error(ctx, "nvc_ctx_nvls_init: " + ctx->nvls_status);
    return NVC_ENVLS;
  }
  nvc_nvls* n = new nvc_nvls();
  ctx->nvls = n;
  std::string why;
  auto cu_fail = [&](CUresult r, const char* what) {
    const char* s = nullptr;
    if (n->drv.GetErrorString) n->drv.GetErrorString(r, &s);
    why = std::string(what) + ": " + (s ? s : "CUDA driver error ") + " (" + std::to_string((int)r) + ")";
    return 1;
  };
  auto step = [&](int32_t mine) -> bool {      // false: some rank failed (or the rendezvous broke); all ranks bail out
    int32_t agg = 0;
    if (!rv.agree(mine, &agg, tmo)) { why = rv.err; return false; }
    if (agg != 0 && mine == 0) why = "a peer rank could not set up its side";
    return agg == 0;
  };
  auto bail = [&]() {
    ctx->nvls_status = "unavailable: " + why;
    rv.close_all();
    nvc_nvls_destroy(ctx);
    nvc_set_error(ctx, "nvc_ctx_nvls_init: " + ctx->nvls_status);
    return NVC_ENVLS;
  };

  // 1. driver entry points + multicast support of this device
  int32_t st = 0;
  if (!load_driver(&n->drv, &why)) st = 1;
  if (!st) {
    CUresult r = n->drv.DeviceGet(&n->dev, ctx->device);
    int sup = 0;
    if (r != CUDA_SUCCESS) st = cu_fail(r, "cuDeviceGet");
    else if ((r = n->drv.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, n->dev)) != CUDA_SUCCESS)
      st = cu_fail(r, "cuDeviceGetAttribute(MULTICAST_SUPPORTED)");
    else if (!sup) { why = "CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED is 0 on this device"; st = 1; }
  }
  if (!step(st)) return bail();

  // 2. sizes; rank 0 creates the multicast object and exports it
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)ctx->tp_size;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = ctx->device;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran_mc = 0, gran_mem = 0;
  const size_t want = kNvlsFlagBytes + slot_bytes + 3 * ll_buf_bytes;
  mp.size = want;
  {
    CUresult r = n->drv.MulticastGetGranularity(&gran_mc, &mp, CU_MULTICAST_GRANULARITY_MINIMUM);
    if (r != CUDA_SUCCESS) st = cu_fail(r, "cuMulticastGetGranularity");
    else if ((r = n->drv.MemGetAllocationGranularity(&gran_mem, &ap, CU_MEM_ALLOC_GRANULARITY_MINIMUM)) != CUDA_SUCCESS)
      st = cu_fail(r, "cuMemGetAllocationGranularity");
  }
  if (!st) {
    n->total = round_up(want, std::max(gran_mc, gran_mem));
    n->total = round_up(n->total, std::min(gran_mc, gran_mem));
    mp.size = n->total;
  }
  int mc_fd = -1;
  if (!st && ctx->tp_rank == 0) {
    CUresult r = n->drv.MulticastCreate(&n->mc, &mp);
    if (r != CUDA_SUCCESS) st = cu_fail(r, "cuMulticastCreate");
    else {
      n->have_mc = true;
      r = n->drv.MemExportToShareableHandle(&mc_fd, n->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
      if (r != CUDA_SUCCESS) st = cu_fail(r, "cuMemExportToShareableHandle(multicast)");
    }
  }
  if (!step(st)) { if (mc_fd >= 0) ::close(mc_fd); return bail(); }

  // 3. the descriptor travels to every peer; everyone adds its device
  if (ctx->tp_rank == 0) {
    for (int k = 1; k < ctx->tp_size && !st; ++k)
      if (!rv.send_fd(rv.fds[k], mc_fd)) { why = "sending the multicast descriptor failed"; st = 1; }
    ::close(mc_fd);
  } else {
    int fd = rv.recv_fd(rv.fds[0], tmo);
    if (fd < 0) { why = "receiving the multicast descriptor failed"; st = 1; }
    else {
      CUresult r = n->drv.MemImportFromShareableHandle(&n->mc, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
      ::close(fd);
      if (r != CUDA_SUCCESS) st = cu_fail(r, "cuMemImportFromShareableHandle(multicast)");
      else n->have_mc = true;
    }
  }
  if (!st) {
    CUresult r = n->drv.MulticastAddDevice(n->mc, n->dev);
    if (r != CUDA_SUCCESS) st = cu_fail(r, "cuMulticastAddDevice");
  }
  if (!step(st)) return bail();                // every device is in the group before anyone binds

  // 4. physical memory, bind, two mappings
  {
    CUresult r = n->drv.MemCreate(&n->mem, n->total, &ap, 0);
    if (r != CUDA_SUCCESS) st = cu_fail(r, "cuMemCreate");
    else n->have_mem = true;
    if (!st && (r = n->drv.MulticastBindMem(n->mc, 0, n->mem, 0, n->total, 0)) != CUDA_SUCCESS) st = cu_fail(r, "cuMulticastBindMem");
    else if (!st) n->bound = true;
    CUmemAccessDesc ad;
    memset(&ad, 0, sizeof(ad));
    ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ad.location.id = ctx->device;
    ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    const size_t align = std::max(gran_mc, gran_mem);
    if (!st && (r = n->drv.MemAddressReserve(&n->uc_va, n->total, align, 0, 0)) != CUDA_SUCCESS) st = cu_fail(r, "cuMemAddressReserve");
    if (!st && (r = n->drv.MemMap(n->uc_va, n->total, 0, n->mem, 0)) != CUDA_SUCCESS) st = cu_fail(r, "cuMemMap(unicast)");
    else if (!st) n->uc_mapped = true;
    if (!st && (r = n->drv.MemSetAccess(n->uc_va, n->total, &ad, 1)) != CUDA_SUCCESS) st = cu_fail(r, "cuMemSetAccess(unicast)");
    if (!st && (r = n->drv.MemAddressReserve(&n->mc_va, n->total, align, 0, 0)) != CUDA_SUCCESS) st = cu_fail(r, "cuMemAddressReserve(multicast)");
    if (!st && (r = n->drv.MemMap(n->mc_va, n->total, 0, n->mc, 0)) != CUDA_SUCCESS) st = cu_fail(r, "cuMemMap(multicast)");
    else if (!st) n->mc_mapped = true;
    if (!st && (r = n->drv.MemSetAccess(n->mc_va, n->total, &ad, 1)) != CUDA_SUCCESS) st = cu_fail(r, "cuMemSetAccess(multicast)");
    if (!st) {
      cudaError_t ce = cudaMemset(reinterpret_cast<void*>(n->uc_va), 0, kNvlsFlagBytes + slot_bytes);
      if (ce == cudaSuccess) ce = cudaMemset(reinterpret_cast<void*>(n->uc_va + kNvlsFlagBytes + slot_bytes), 0xFF, 3 * ll_buf_bytes);
      if (ce == cudaSuccess) ce = cudaMalloc(&n->state, 16 * sizeof(uint32_t));
      if (ce == cudaSuccess) ce = cudaMemset(n->state, 0, 16 * sizeof(uint32_t));
      if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
      if (ce != cudaSuccess) { why = std::string("initialising the buffer: ") + cudaGetErrorString(ce); st = 1; }
    }
  }
  if (!step(st)) return bail();                // every rank's copy is initialised before the first store arrives
  rv.close_all();
  n->slot_bytes = slot_bytes;
  n->ll_buf_bytes = ll_buf_bytes;
  ctx->nvls_ready = true;
  ctx->nvls_status = "ready: multicast object over " + std::to_string(ctx->tp_size) + " devices, " +
                     std::to_string(n->total >> 20) + " MiB per rank";
  return NVC_OK;
}

// mode 0 = choose by size, 1 = LL one-shot, 2 = two-shot (flag barriers), 3 = two-shot with LL broadcast
extern "C" int nvc_nvls_allreduce_rmsnorm(nvc_ctx* ctx, const nvc_bf16* x_partial, const float* residual_in,
                                          float* residual_out, const nvc_bf16* w, float eps, nvc_bf16* out, int N, int H,
                                          int mode, void* stream) {
  NVC_REQUIRE(ctx, ctx && out && x_partial, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, ctx->nvls_ready, NVC_ESTATE, "NVLS buffers not set up (nvc_ctx_nvls_init)");
  NVC_REQUIRE(ctx, N > 0 && H > 0 && H % 64 == 0 && H / 8 <= 1024, NVC_ESHAPE, "H must be a multiple of 64 and <= 8192");
  nvc_nvls* n = ctx->nvls;
  const int tp = ctx->tp_size;
  const size_t msg = (size_t)N * H * 2;
  static const int mode_env = getenv("NVC_NVLS_MODE") ? atoi(getenv("NVC_NVLS_MODE")) : 0;
  if (mode == 0) mode = mode_env;
  if (mode == 0) {
    // one-shot moves tp x msg into every GPU, two-shot ~1.1 x msg behind one more switch round trip
    if (msg * tp <= n->ll_buf_bytes && msg * (tp - 1) <= ((size_t)3 << 20)) mode = 1;
    else if (msg <= n->ll_buf_bytes) mode = 3;
    else mode = 2;
  }
  uint8_t* uc = reinterpret_cast<uint8_t*>(n->uc_va);
  uint8_t* mc = reinterpret_cast<uint8_t*>(n->mc_va);
  uint8_t* uc_slot = uc + kNvlsFlagBytes;
  uint8_t* mc_slot = mc + kNvlsFlagBytes;
  uint8_t* uc_ll = uc_slot + n->slot_bytes;
  uint8_t* mc_ll = mc_slot + n->slot_bytes;
  const int threads = ((H / 8 + 31) / 32) * 32;
  cudaStream_t st = nvc_stream(stream);
  Epilogue ep{reinterpret_cast<const float4*>(residual_in), reinterpret_cast<float4*>(residual_out),
              reinterpret_cast<const uint4*>(w), eps, reinterpret_cast<uint4*>(out)};
  if (mode == 1) {
    NVC_REQUIRE(ctx, msg * tp <= n->ll_buf_bytes, NVC_ENOMEM, "message larger than the LL buffer");
#define NVC_LL(TPV)                                                                                                    \
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, nvls_ll_allreduce_rmsnorm_kernel<TPV>, dim3(N), dim3(threads), 0, st,             \
                                 reinterpret_cast<const uint4*>(x_partial), uc_ll, mc_ll, n->ll_buf_bytes, n->state,   \
                                 ctx->tp_rank, ep, H))
    switch (tp) {
      case 2: NVC_LL(2); break;
      case 4: NVC_LL(4); break;
      case 8: NVC_LL(8); break;
      default: NVC_REQUIRE(ctx, false, NVC_ESHAPE, "NVLS all-reduce supports tp_size 2, 4, 8");
    }
#undef NVC_LL
  } else {
    NVC_REQUIRE(ctx, mode == 2 || mode == 3, NVC_EINVAL, "mode must be 0..3");
    NVC_REQUIRE(ctx, reinterpret_cast<const uint8_t*>(x_partial) == uc_slot, NVC_EINVAL,
                "the two-shot kernel reduces in the switch: the partial sums must be in nvc_ctx_nvls_slot()");
    NVC_REQUIRE(ctx, msg <= n->slot_bytes, NVC_ENOMEM, "message larger than the NVLS slot");
    NVC_REQUIRE(ctx, mode == 2 || msg <= n->ll_buf_bytes, NVC_ENOMEM, "message larger than the LL buffer");
    NVC_REQUIRE(ctx, tp == 2 || tp == 4 || tp == 8, NVC_ESHAPE, "NVLS all-reduce supports tp_size 2, 4, 8");
    // every CTA must be resident (flag waits): as many per SM as the occupancy calculator grants, at most 8
    int occ = 0;
    if (mode == 2) NVC_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, nvls_twoshot_allreduce_rmsnorm_kernel<false>, threads, 0));
    else NVC_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, nvls_twoshot_allreduce_rmsnorm_kernel<true>, threads, 0));
    const int per_sm = std::max(1, std::min(occ, 8));
    const int grid = std::min(N, ctx->num_sms * per_sm);
    uint32_t* ucf = reinterpret_cast<uint32_t*>(uc);
    uint32_t* mcf = reinterpret_cast<uint32_t*>(mc);
    if (mode == 2)
      NVC_CUDA_CHECK(ctx, nvc_launch(ctx, nvls_twoshot_allreduce_rmsnorm_kernel<false>, dim3(grid), dim3(threads), 0, st,
                                     uc_slot, mc_slot, ucf, mcf, uc_ll, mc_ll, n->ll_buf_bytes, n->state, ctx->tp_rank, tp, ep, H, N));
    else
      NVC_CUDA_CHECK(ctx, nvc_launch(ctx, nvls_twoshot_allreduce_rmsnorm_kernel<true>, dim3(grid), dim3(threads), 0, st,
                                     uc_slot, mc_slot, ucf, mcf, uc_ll, mc_ll, n->ll_buf_bytes, n->state, ctx->tp_rank, tp, ep, H, N));
  }
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// (value, token) candidates of every rank's vocab shard -> the winning token per sequence, on every rank alike
extern "C" int nvc_nvls_argmax_merge(nvc_ctx* ctx, const float* best_val, const int32_t* best_tok, int32_t* out_tok,
                                     float* out_val, int B, void* stream) {
  NVC_REQUIRE(ctx, ctx && best_val && best_tok && out_tok, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, ctx->nvls_ready, NVC_ESTATE, "NVLS buffers not set up (nvc_ctx_nvls_init)");
  nvc_nvls* n = ctx->nvls;
  const int tp = ctx->tp_size;
  NVC_REQUIRE(ctx, B > 0 && (size_t)B * tp * 16 <= n->ll_buf_bytes, NVC_ENOMEM, "batch larger than the LL buffer");
  uint8_t* uc_ll = reinterpret_cast<uint8_t*>(n->uc_va) + kNvlsFlagBytes + n->slot_bytes;
  uint8_t* mc_ll = reinterpret_cast<uint8_t*>(n->mc_va) + kNvlsFlagBytes + n->slot_bytes;
  const int grid = std::max((B + 127) / 128, 16);
  cudaStream_t st = nvc_stream(stream);
#define NVC_AM(TPV)                                                                                              \
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, nvls_argmax_merge_kernel<TPV>, dim3(grid), dim3(128), 0, st, best_val, best_tok, \
                                 uc_ll, mc_ll, n->ll_buf_bytes, n->state, ctx->tp_rank, out_tok, out_val, B))
  switch (tp) {
    case 2: NVC_AM(2); break;
    case 4: NVC_AM(4); break;
    case 8: NVC_AM(8); break;
    default: NVC_REQUIRE(ctx, false, NVC_ESHAPE, "NVLS merge supports tp_size 2, 4, 8");
  }
#undef NVC_AM
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}
