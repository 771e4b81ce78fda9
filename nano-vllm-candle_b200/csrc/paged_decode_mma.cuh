// K5 (tensor-core variant): paged decode attention where the whole GQA group (up to 16 query
// heads sharing one kv head) is the M dimension of mma.sync.m16n8k16, so every K / V row is read
// from HBM once and the per-row arithmetic no longer scales with the group size on the CUDA cores
// (the SIMT variant in paged_decode.cu is FMA/shuffle-bound from G = 4 up: 72 % of HBM peak at
// G = 4, 48 % at G = 5, 30 % at G = 8).
//
// This kernel is HBM-bound by construction (4 flop per byte at G = 8), so the legacy warp-level
// mma.sync path is the right tool: tcgen05 would need M >= 64 tiles, a TMEM round trip per
// 16-token tile and buys nothing when the tensor pipe is < 5 % busy.
//
// CTA = 4 independent warps on one (sequence, kv head, KV split).  A warp owns 16-token tiles
// (tokens t*64 + 16w ..): K and V rows of the tile are copied with cp.async (16 B per lane,
// zero-filled past the context length) into a private 3-stage ring in shared memory with a
// 16-byte-chunk XOR swizzle, read back with ldmatrix (K) / ldmatrix.trans (V); S = Q K^T and
// O += P V run on mma.sync with fp32 accumulators, online softmax in the log2 domain on the
// accumulator fragments.  Warps never synchronise with each other until the final merge.
#pragma once
#include "common.cuh"

namespace pdmma {

constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
constexpr int kTileTok = 16;                         // tokens per warp tile
constexpr int kCtaTok = kWarps * kTileTok;           // 64 tokens per CTA iteration
constexpr int kStages = 3;
constexpr int kStageBytes = 2 * kTileTok * 256;      // K tile + V tile, 256 B per token row
constexpr int kSmemBytes = kWarps * kStages * kStageBytes;   // 96 KB

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
// D (f32 16x8) += A (bf16 16x16, row) * B (bf16 16x8, col)
__device__ __forceinline__ void mma_bf16(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ int atomic_add_acq_rel(int* p, int v) {
  int old;
  asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}

// group: query heads per kv head handled by this CTA (<= 16); head0: first query head.
// num_splits > 1: every split parks its (m, l, o) partial; with `counters` (one zeroed int per (sequence, blockIdx.x))
// the CTA that takes the last ticket merges the partials in split order - no second kernel.
__global__ void __launch_bounds__(kThreads, 2)
paged_decode_mma_kernel(const nvc_bf16* __restrict__ q, int ldq, const nvc_bf16* __restrict__ k_cache,
                        const nvc_bf16* __restrict__ v_cache, const int32_t* __restrict__ block_tables, int max_blocks,
                        const int32_t* __restrict__ context_lens, int nh, int kv, int group_total, int gchunk,
                        int block_size, float scale_log2e, nvc_bf16* __restrict__ out, float* __restrict__ part_o,
                        float* __restrict__ part_ml, int num_splits, int* __restrict__ counters) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int ticket_sm;
  pdl_trigger();
  // The dependency wait comes AFTER the first K / V tiles are in flight (see below): context_lens and
  // block_tables are step inputs (copied before the step's first kernel), and the only cache rows a
  // kernel of this step writes are the newest token's (position ctx_len - 1, K2 of this layer).
  const int chunks = (group_total + gchunk - 1) / gchunk;
  const int kvh = blockIdx.x / chunks;
  const int chunk = blockIdx.x - kvh * chunks;
  const int head0 = kvh * group_total + chunk * gchunk;
  const int G = min(gchunk, group_total - chunk * gchunk);
  const int b = blockIdx.y, split = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quad = lane >> 2, tq = lane & 3;          // accumulator row (quad, quad + 8), column pair tq

  const int ctx_len = __ldg(context_lens + b);
  const int n_tiles = (ctx_len + kCtaTok - 1) / kCtaTok;
  const int tiles_per_split = (n_tiles + num_splits - 1) / num_splits;
  const int t_begin = split * tiles_per_split;
  const int t_end = min(n_tiles, t_begin + tiles_per_split);

  float o[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

  const uint32_t ring = smem_u32(smem) + warp * (kStages * kStageBytes);
  const int32_t* bt = block_tables + (int64_t)b * max_blocks;
  const int64_t row_stride = (int64_t)kv * 256;        // bytes between consecutive token rows of one kv head
  const int my_tiles = t_end - t_begin;                // CTA tiles; this warp's tile i covers tokens (t_begin+i)*64 + 16w

  // issue the loads of warp tile i into ring stage i % kStages (empty group if i is out of range)
  auto issue = [&](int i) {
    if (i < my_tiles) {
      const int tok0 = (t_begin + i) * kCtaTok + warp * kTileTok;
      if (tok0 < ctx_len) {
        const int blk = __ldg(bt + tok0 / block_size);
        const int64_t slot0 = (int64_t)blk * block_size + (tok0 % block_size);
        const uint32_t st = ring + (i % kStages) * kStageBytes;
        const int c = lane & 15;                       // 16-byte chunk of the 256-byte row
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
          const int r = rr * 2 + (lane >> 4);          // token row inside the tile
          const uint32_t ok = (tok0 + r < ctx_len) ? 16u : 0u;         // zero-fill past the context
          const int64_t goff = ((slot0 + (ok ? r : 0)) * kv + kvh) * 256 + c * 16;
          const uint32_t soff = r * 256 + ((c ^ (r & 7)) << 4);
          cp_async16(st + soff, reinterpret_cast<const uint8_t*>(k_cache) + goff, ok);
          cp_async16(st + kTileTok * 256 + soff, reinterpret_cast<const uint8_t*>(v_cache) + goff, ok);
        }
      }
    }
    cp_async_commit();
  };
  (void)row_stride;

  // Tiles that cannot hold the newest token are immutable during this step: start them while the
  // predecessor (q/k-norm + RoPE + KV store) is still running.  A warp's tiles ascend in position, so
  // "tile i is safe" is monotone in i; an out-of-range tile is an empty group.
  auto safe = [&](int i) {
    const int tok0 = (t_begin + i) * kCtaTok + warp * kTileTok;
    return i >= my_tiles || tok0 >= ctx_len || tok0 + kTileTok <= ctx_len - 1;
  };
  int pre = 0;
  if (safe(0)) {
    issue(0);
    pre = 1;
    if (safe(1)) {
      issue(1);
      pre = 2;
    }
  }
  pdl_wait();
  // ---- Q as A fragments: 8 k-steps of 16 dims; rows >= G are zero
  uint32_t qa[8][4];
  {
    const nvc_bf16* qrow0 = q + (int64_t)b * ldq + (int64_t)(head0 + quad) * 128;
    const nvc_bf16* qrow1 = qrow0 + 8 * 128;
    const bool r0 = quad < G, r1 = quad + 8 < G;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      const int d = ks * 16 + 2 * tq;
      qa[ks][0] = r0 ? __ldg(reinterpret_cast<const uint32_t*>(qrow0 + d)) : 0u;
      qa[ks][1] = r1 ? __ldg(reinterpret_cast<const uint32_t*>(qrow1 + d)) : 0u;
      qa[ks][2] = r0 ? __ldg(reinterpret_cast<const uint32_t*>(qrow0 + d + 8)) : 0u;
      qa[ks][3] = r1 ? __ldg(reinterpret_cast<const uint32_t*>(qrow1 + d + 8)) : 0u;
    }
  }
  if (pre < 1) issue(0);
  if (pre < 2) issue(1);
  for (int i = 0; i < my_tiles; ++i) {
    issue(i + 2);
    cp_async_wait<2>();                               // tile i has landed (two younger groups may be in flight)
    __syncwarp();
    const int tok0 = (t_begin + i) * kCtaTok + warp * kTileTok;
    if (tok0 < ctx_len) {                             // warp-uniform
      const uint32_t st = ring + (i % kStages) * kStageBytes;
      // ---- S = Q K^T: two n-tiles of 8 tokens, fp32
      float s[2][4];
#pragma unroll
      for (int j = 0; j < 2; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
      {
        // ldmatrix.x4: matrices (tok 0-7, d lo), (tok 0-7, d hi), (tok 8-15, d lo), (tok 8-15, d hi)
        const int r = (lane & 7) + ((lane >> 4) << 3);
        const int chi = (lane >> 3) & 1;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const int c = 2 * ks + chi;
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4(st + r * 256 + ((c ^ (r & 7)) << 4), b0, b1, b2, b3);
          mma_bf16(s[0], qa[ks], b0, b1);
          mma_bf16(s[1], qa[ks], b2, b3);
        }
      }
      // ---- online softmax on the fragments (rows quad and quad + 8; columns tok0 + 8j + 2tq + {0,1})
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int t0 = tok0 + 8 * j + 2 * tq;
        const bool v0 = t0 < ctx_len, v1 = t0 + 1 < ctx_len;
        s[j][0] = v0 ? s[j][0] * scale_log2e : -INFINITY;
        s[j][1] = v1 ? s[j][1] * scale_log2e : -INFINITY;
        s[j][2] = v0 ? s[j][2] * scale_log2e : -INFINITY;
        s[j][3] = v1 ? s[j][3] * scale_log2e : -INFINITY;
        mx0 = fmaxf(mx0, fmaxf(s[j][0], s[j][1]));
        mx1 = fmaxf(mx1, fmaxf(s[j][2], s[j][3]));
      }
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);      // tok0 < ctx_len: token tok0 is valid, so finite
      const float al0 = fast_exp2(m0 - mn0), al1 = fast_exp2(m1 - mn1);
      m0 = mn0;
      m1 = mn1;
      uint32_t pa[4];
      float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float p0 = fast_exp2(s[j][0] - mn0), p1 = fast_exp2(s[j][1] - mn0);
        const float p2 = fast_exp2(s[j][2] - mn1), p3 = fast_exp2(s[j][3] - mn1);
        ps0 += p0 + p1;
        ps1 += p2 + p3;
        pa[2 * j] = pack_bf16(p0, p1);                 // A fragment of P: a0/a2 = row quad, a1/a3 = row quad + 8
        pa[2 * j + 1] = pack_bf16(p2, p3);
      }
      l0 = l0 * al0 + ps0;
      l1 = l1 * al1 + ps1;
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
        o[nt][0] *= al0; o[nt][1] *= al0;
        o[nt][2] *= al1; o[nt][3] *= al1;
      }
      // ---- O += P V: ldmatrix.trans: matrices (tok 0-7, d-chunk nt), (tok 8-15, nt), (tok 0-7, nt+1), (tok 8-15, nt+1)
      {
        const uint32_t vs = st + kTileTok * 256;
        const int r = (lane & 7) + (((lane >> 3) & 1) << 3);
        const int cadd = lane >> 4;
#pragma unroll
        for (int np = 0; np < 8; ++np) {
          const int c = 2 * np + cadd;
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4_trans(vs + r * 256 + ((c ^ (r & 7)) << 4), b0, b1, b2, b3);
          mma_bf16(o[2 * np], pa, b0, b1);
          mma_bf16(o[2 * np + 1], pa, b2, b3);
        }
      }
    }
    __syncwarp();                                     // everyone is done with stage i % kStages before it is refilled
  }
  cp_async_wait<0>();
  // row sums live per lane of a quad: finish them
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);

  // ---- merge the 4 warps through shared memory (the rings are dead now)
  __syncthreads();
  float* sm_m = reinterpret_cast<float*>(smem);                    // [kWarps][16]
  float* sm_l = sm_m + kWarps * 16;                                // [kWarps][16]
  float* sm_o = sm_l + kWarps * 16;                                // [kWarps][16][128]
  if (tq == 0) {
    sm_m[warp * 16 + quad] = m0;
    sm_m[warp * 16 + quad + 8] = m1;
    sm_l[warp * 16 + quad] = l0;
    sm_l[warp * 16 + quad + 8] = l1;
  }
#pragma unroll
  for (int nt = 0; nt < 16; ++nt) {
    const int d = nt * 8 + 2 * tq;
    float* r0 = sm_o + (warp * 16 + quad) * 128 + d;
    float* r1 = sm_o + (warp * 16 + quad + 8) * 128 + d;
    r0[0] = o[nt][0]; r0[1] = o[nt][1];
    r1[0] = o[nt][2]; r1[1] = o[nt][3];
  }
  __syncthreads();
  const int d = threadIdx.x;                                       // 128 threads = 128 dims
  for (int g = 0; g < G; ++g) {
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) M = fmaxf(M, sm_m[w * 16 + g]);
    float L = 0.f, O = 0.f;
    if (M != -INFINITY) {
#pragma unroll
      for (int w = 0; w < kWarps; ++w) {
        const float wt = fast_exp2(sm_m[w * 16 + g] - M);
        L += sm_l[w * 16 + g] * wt;
        O += sm_o[(w * 16 + g) * 128 + d] * wt;
      }
    }
    const int head = head0 + g;
    if (num_splits == 1) {
      const float r = L > 0.f ? O / L : 0.f;
      out[((int64_t)b * nh + head) * 128 + d] = __bfloat16_as_ushort(__float2bfloat16_rn(r));
    } else {
      const int64_t p = ((int64_t)b * nh + head) * num_splits + split;
      __stcg(part_o + p * 128 + d, O);
      if (d == 0) { __stcg(part_ml + p * 2, M); __stcg(part_ml + p * 2 + 1, L); }
    }
  }
  if (num_splits > 1 && counters != nullptr) {
    // CTA barrier orders every thread's partial stores before thread 0's release; the acquire half of the same
    // atomic plus the second barrier orders the last CTA's loads after the other splits' stores.
    const int pair = b * gridDim.x + blockIdx.x;
    __syncthreads();
    if (threadIdx.x == 0) ticket_sm = atomic_add_acq_rel(counters + pair, 1);
    __syncthreads();
    if (ticket_sm == num_splits - 1) {
      for (int g = 0; g < G; ++g) {
        const int64_t p0 = ((int64_t)b * nh + head0 + g) * num_splits;
        float M = -INFINITY;
        for (int s = 0; s < num_splits; ++s) M = fmaxf(M, __ldcg(part_ml + (p0 + s) * 2));
        float L = 0.f, O = 0.f;
        if (M != -INFINITY) {
          for (int s = 0; s < num_splits; ++s) {                       // split order whoever merges: deterministic
            const float wt = fast_exp2(__ldcg(part_ml + (p0 + s) * 2) - M);
            L += __ldcg(part_ml + (p0 + s) * 2 + 1) * wt;
            O += __ldcg(part_o + (p0 + s) * 128 + d) * wt;
          }
        }
        const float r = L > 0.f ? O / L : 0.f;
        out[((int64_t)b * nh + head0 + g) * 128 + d] = __bfloat16_as_ushort(__float2bfloat16_rn(r));
      }
      if (threadIdx.x == 0) counters[pair] = 0;                        // re-arm for the next launch
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Work-balanced variant: the (sequence, 64-token tile) items of the whole batch are laid out flat
// and cut into equal contiguous ranges, one per GROUP of CTAs (one CTA per kv head, 2 CTAs per SM
// resident), so ragged context lengths and batch sizes that do
// not divide the SM count cost no tail wave (Qwen3-0.6B, 64 seqs x 8 kv heads = 512 CTAs of 10
// tiles on 296 slots ran 2 waves for 1.73 waves of work).  A range may cover several
// (kv head, sequence) pairs and a pair may be cut between CTAs: a CTA that holds a whole pair
// writes the output directly; otherwise it writes an (m, l, o) partial to slot pair + group (both
// grow monotonically along the flat order, so the sum is unique) and takes a ticket on the pair;
// the CTA that takes the last ticket merges the partials in CTA order (deterministic) - no
// second kernel.  Everything is derived from context_lens on the device: CUDA-graph replayable.
// ---------------------------------------------------------------------------------------------
constexpr int kMaxFlatBatch = 1024;

__global__ void __launch_bounds__(kThreads, 2)
paged_decode_flat_kernel(const nvc_bf16* __restrict__ q, int ldq, const nvc_bf16* __restrict__ k_cache,
                         const nvc_bf16* __restrict__ v_cache, const int32_t* __restrict__ block_tables, int max_blocks,
                         const int32_t* __restrict__ context_lens, int B, int nh, int kv, int group_total, int gchunk,
                         int block_size, float scale_log2e, nvc_bf16* __restrict__ out, float* __restrict__ part_o,
                         float* __restrict__ part_ml, int* __restrict__ counters) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int prefix[kMaxFlatBatch + 1];          // prefix[b] = tiles of sequences < b (per kv head)
  __shared__ int warp_tot[kWarps];
  __shared__ int ticket_sm;
  pdl_trigger();
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quad = lane >> 2, tq = lane & 3;          // accumulator row (quad, quad + 8), column pair tq
  const int chunks = (group_total + gchunk - 1) / gchunk;
  const int kvx_total = kv * chunks;

  // ---- tiles per sequence -> exclusive prefix sum (every CTA needs the whole map; B <= 1024)
  {
    const int per = (B + kThreads - 1) / kThreads;
    const int b0 = threadIdx.x * per;
    int sum = 0;
    for (int i = 0; i < per; ++i) {
      const int b = b0 + i;
      if (b < B) sum += (__ldg(context_lens + b) + kCtaTok - 1) / kCtaTok;
    }
    int incl = sum;
#pragma unroll
    for (int o2 = 1; o2 < 32; o2 <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o2);
      if (lane >= o2) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    int base = incl - sum;
    for (int w = 0; w < warp; ++w) base += warp_tot[w];
    for (int i = 0; i < per; ++i) {
      const int b = b0 + i;
      if (b < B) {
        prefix[b] = base;
        base += (__ldg(context_lens + b) + kCtaTok - 1) / kCtaTok;
      }
    }
    if (threadIdx.x == kThreads - 1) prefix[B] = base;   // the last thread's range ends at (or beyond) B
    __syncthreads();
  }
  // gridDim.x = groups * kvx_total: the kvx_total CTAs of a group walk the SAME (sequence, tile) range in
  // lockstep, one kv head each, so that at any time the group reads all kv heads of the same tokens
  // (2 KB contiguous per token): cutting a kv-head-major list instead cost 7 % of HBM bandwidth
  const int T = prefix[B];
  const long long W = T;
  const int P = gridDim.x / kvx_total;
  const int cta_group = blockIdx.x / kvx_total;
  // empty sequences own no item: their output rows are zero (the per-(sequence, head) kernel writes 0 too)
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    if (__ldg(context_lens + b) <= 0)
      for (int i = threadIdx.x; i < nh * 128; i += kThreads) out[(int64_t)b * nh * 128 + i] = 0;
  }
  const int chunk = W > 0 ? (int)((W + P - 1) / P) : 1;
  long long x = (long long)cta_group * chunk;
  const long long x1 = min(W, x + chunk);
  if (x >= x1) return;
  const int kvx = blockIdx.x - cta_group * kvx_total;
  int rem = (int)x;
  int b;
  {                                                   // largest b with prefix[b] <= rem
    int lo = 0, hi = B - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (prefix[mid] <= rem) lo = mid; else hi = mid - 1;
    }
    b = lo;
  }
  const uint32_t ring = smem_u32(smem) + warp * (kStages * kStageBytes);
  float* sm_m = reinterpret_cast<float*>(smem);                    // [kWarps][16]  (aliases the rings: used after them)
  float* sm_l = sm_m + kWarps * 16;                                // [kWarps][16]
  float* sm_o = sm_l + kWarps * 16;                                // [kWarps][16][128]

  while (x < x1) {
    while (prefix[b + 1] <= rem) ++b;                 // skip empty sequences (x < W: a non-empty one follows)
    const int n_tiles = prefix[b + 1] - prefix[b];
    const int t_begin = rem - prefix[b];
    const int my_tiles = min((int)(x1 - x), n_tiles - t_begin);
    const int kvh = kvx / chunks, gch = kvx - kvh * chunks;
    const int head0 = kvh * group_total + gch * gchunk;
    const int G = min(gchunk, group_total - gch * gchunk);
    const int ctx_len = __ldg(context_lens + b);

    // ---- Q as A fragments: 8 k-steps of 16 dims; rows >= G are zero
    uint32_t qa[8][4];
    {
      const nvc_bf16* qrow0 = q + (int64_t)b * ldq + (int64_t)(head0 + quad) * 128;
      const nvc_bf16* qrow1 = qrow0 + 8 * 128;
      const bool r0 = quad < G, r1 = quad + 8 < G;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const int d = ks * 16 + 2 * tq;
        qa[ks][0] = r0 ? __ldg(reinterpret_cast<const uint32_t*>(qrow0 + d)) : 0u;
        qa[ks][1] = r1 ? __ldg(reinterpret_cast<const uint32_t*>(qrow1 + d)) : 0u;
        qa[ks][2] = r0 ? __ldg(reinterpret_cast<const uint32_t*>(qrow0 + d + 8)) : 0u;
        qa[ks][3] = r1 ? __ldg(reinterpret_cast<const uint32_t*>(qrow1 + d + 8)) : 0u;
      }
    }
    float o[16][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    const int32_t* bt = block_tables + (int64_t)b * max_blocks;

    // loads of warp tile i into ring stage i % kStages (empty group if i is out of range)
    auto issue = [&](int i) {
      if (i < my_tiles) {
        const int tok0 = (t_begin + i) * kCtaTok + warp * kTileTok;
        if (tok0 < ctx_len) {
          const int blk = __ldg(bt + tok0 / block_size);
          const int64_t slot0 = (int64_t)blk * block_size + (tok0 % block_size);
          const uint32_t st = ring + (i % kStages) * kStageBytes;
          const int c = lane & 15;                       // 16-byte chunk of the 256-byte row
#pragma unroll
          for (int rr = 0; rr < 8; ++rr) {
            const int r = rr * 2 + (lane >> 4);          // token row inside the tile
            const uint32_t ok = (tok0 + r < ctx_len) ? 16u : 0u;         // zero-fill past the context
            const int64_t goff = ((slot0 + (ok ? r : 0)) * kv + kvh) * 256 + c * 16;
            const uint32_t soff = r * 256 + ((c ^ (r & 7)) << 4);
            cp_async16(st + soff, reinterpret_cast<const uint8_t*>(k_cache) + goff, ok);
            cp_async16(st + kTileTok * 256 + soff, reinterpret_cast<const uint8_t*>(v_cache) + goff, ok);
          }
        }
      }
      cp_async_commit();
    };

    issue(0);
    issue(1);
    for (int i = 0; i < my_tiles; ++i) {
      issue(i + 2);
      cp_async_wait<2>();                               // tile i has landed (two younger groups may be in flight)
      __syncwarp();
      const int tok0 = (t_begin + i) * kCtaTok + warp * kTileTok;
      if (tok0 < ctx_len) {                             // warp-uniform
        const uint32_t st = ring + (i % kStages) * kStageBytes;
        float s[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
        {
          const int r = (lane & 7) + ((lane >> 4) << 3);
          const int chi = (lane >> 3) & 1;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const int c = 2 * ks + chi;
            uint32_t b0, b1, b2, b3;
            ldmatrix_x4(st + r * 256 + ((c ^ (r & 7)) << 4), b0, b1, b2, b3);
            mma_bf16(s[0], qa[ks], b0, b1);
            mma_bf16(s[1], qa[ks], b2, b3);
          }
        }
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int t0 = tok0 + 8 * j + 2 * tq;
          const bool v0 = t0 < ctx_len, v1 = t0 + 1 < ctx_len;
          s[j][0] = v0 ? s[j][0] * scale_log2e : -INFINITY;
          s[j][1] = v1 ? s[j][1] * scale_log2e : -INFINITY;
          s[j][2] = v0 ? s[j][2] * scale_log2e : -INFINITY;
          s[j][3] = v1 ? s[j][3] * scale_log2e : -INFINITY;
          mx0 = fmaxf(mx0, fmaxf(s[j][0], s[j][1]));
          mx1 = fmaxf(mx1, fmaxf(s[j][2], s[j][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
        const float al0 = fast_exp2(m0 - mn0), al1 = fast_exp2(m1 - mn1);
        m0 = mn0;
        m1 = mn1;
        uint32_t pa[4];
        float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const float p0 = fast_exp2(s[j][0] - mn0), p1 = fast_exp2(s[j][1] - mn0);
          const float p2 = fast_exp2(s[j][2] - mn1), p3 = fast_exp2(s[j][3] - mn1);
          ps0 += p0 + p1;
          ps1 += p2 + p3;
          pa[2 * j] = pack_bf16(p0, p1);
          pa[2 * j + 1] = pack_bf16(p2, p3);
        }
        l0 = l0 * al0 + ps0;
        l1 = l1 * al1 + ps1;
#pragma unroll
        for (int nt = 0; nt < 16; ++nt) {
          o[nt][0] *= al0; o[nt][1] *= al0;
          o[nt][2] *= al1; o[nt][3] *= al1;
        }
        {
          const uint32_t vs = st + kTileTok * 256;
          const int r = (lane & 7) + (((lane >> 3) & 1) << 3);
          const int cadd = lane >> 4;
#pragma unroll
          for (int np = 0; np < 8; ++np) {
            const int c = 2 * np + cadd;
            uint32_t b0, b1, b2, b3;
            ldmatrix_x4_trans(vs + r * 256 + ((c ^ (r & 7)) << 4), b0, b1, b2, b3);
            mma_bf16(o[2 * np], pa, b0, b1);
            mma_bf16(o[2 * np + 1], pa, b2, b3);
          }
        }
      }
      __syncwarp();
    }
    cp_async_wait<0>();
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);

    // ---- merge the 4 warps through shared memory (the rings are idle now)
    __syncthreads();
    if (tq == 0) {
      sm_m[warp * 16 + quad] = m0;
      sm_m[warp * 16 + quad + 8] = m1;
      sm_l[warp * 16 + quad] = l0;
      sm_l[warp * 16 + quad + 8] = l1;
    }
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
      const int d = nt * 8 + 2 * tq;
      float* r0 = sm_o + (warp * 16 + quad) * 128 + d;
      float* r1 = sm_o + (warp * 16 + quad + 8) * 128 + d;
      r0[0] = o[nt][0]; r0[1] = o[nt][1];
      r1[0] = o[nt][2]; r1[1] = o[nt][3];
    }
    __syncthreads();
    // how many CTAs share this (kv head, sequence) pair: its items are [lo, hi) of the flat order
    const long long lo = prefix[b];
    const int i_lo = (int)(lo / chunk), i_hi = (int)((lo + n_tiles - 1) / chunk);
    const int nparts = i_hi - i_lo + 1;
    const int pair = kvx * B + b;                                    // ticket index
    const int slot0 = kvx * (B + P) + b;                             // + group: unique per kv head (both grow monotonically)
    const int d = threadIdx.x;                                       // 128 threads = 128 dims
    for (int g = 0; g < G; ++g) {
      float M = -INFINITY;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) M = fmaxf(M, sm_m[w * 16 + g]);
      float L = 0.f, O = 0.f;
      if (M != -INFINITY) {
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
          const float wt = fast_exp2(sm_m[w * 16 + g] - M);
          L += sm_l[w * 16 + g] * wt;
          O += sm_o[(w * 16 + g) * 128 + d] * wt;
        }
      }
      if (nparts == 1) {
        const float r = L > 0.f ? O / L : 0.f;
        out[((int64_t)b * nh + head0 + g) * 128 + d] = __bfloat16_as_ushort(__float2bfloat16_rn(r));
      } else {
        const int64_t e = (int64_t)(slot0 + cta_group) * gchunk + g;
        __stcg(part_o + e * 128 + d, O);
        if (d == 0) { __stcg(part_ml + e * 2, M); __stcg(part_ml + e * 2 + 1, L); }
      }
    }
    if (nparts > 1) {
      // CTA barrier orders every thread's partial stores before thread 0's release; the acquire half of the
      // same atomic plus the second barrier orders the last CTA's loads after all other CTAs' stores.
      __syncthreads();
      if (threadIdx.x == 0) ticket_sm = atomic_add_acq_rel(counters + pair, 1);
      __syncthreads();
      if (ticket_sm == nparts - 1) {
        for (int g = 0; g < G; ++g) {
          float M = -INFINITY;
          for (int i = i_lo; i <= i_hi; ++i) M = fmaxf(M, __ldcg(part_ml + ((int64_t)(slot0 + i) * gchunk + g) * 2));
          float L = 0.f, O = 0.f;
          if (M != -INFINITY) {
            for (int i = i_lo; i <= i_hi; ++i) {                       // CTA order whoever merges: deterministic
              const int64_t e = (int64_t)(slot0 + i) * gchunk + g;
              const float wt = fast_exp2(__ldcg(part_ml + e * 2) - M);
              L += __ldcg(part_ml + e * 2 + 1) * wt;
              O += __ldcg(part_o + e * 128 + d) * wt;
            }
          }
          const float r = L > 0.f ? O / L : 0.f;
          out[((int64_t)b * nh + head0 + g) * 128 + d] = __bfloat16_as_ushort(__float2bfloat16_rn(r));
        }
        if (threadIdx.x == 0) counters[pair] = 0;                      // re-arm for the next launch
      }
    }
    __syncthreads();                                                   // sm_m / sm_l / sm_o alias the rings of the next segment
    x += my_tiles;
    rem += my_tiles;
  }
}

}  // namespace pdmma
