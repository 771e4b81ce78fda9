// K5: paged decode attention where the whole GQA group (up to 16 query heads sharing one kv head)
// is the M dimension of mma.sync.m16n8k16, so every K / V row is read from HBM once and the per-row
// arithmetic does not scale with the group size on the CUDA cores.
//
// This kernel is HBM-bound by construction (4 flop per byte at G = 8), so the legacy warp-level
// mma.sync path is the right tool: tcgen05 would need M >= 64 tiles, a TMEM round trip per
// 16-token tile and buys nothing when the tensor pipe is < 5 % busy.
//
// CTA = kWarps independent warps on one (sequence, kv head, KV split).  A warp owns 16-token tiles
// (tokens t*16*kWarps + 16w ..): K and V rows of the tile are copied with cp.async (16 B per lane,
// zero-filled past the context length) into a private 3-stage ring in shared memory with a
// 16-byte-chunk XOR swizzle, read back with ldmatrix (K) / ldmatrix.trans (V); S = Q K^T and
// O += P V run on mma.sync with fp32 accumulators, online softmax in the log2 domain on the
// accumulator fragments.  Warps never synchronise with each other until the final merge.
//
// kWarps = 4: 96 KB of rings, 2 CTAs per SM (296 resident).  kWarps = 2: 48 KB, 4 CTAs per SM (592
// resident): the same bytes in flight per SM, but a grid of 297..592 CTAs (Qwen3-0.6B at batch 64:
// 512) is resident at once instead of running 1.7 waves with a synchronised bubble between them.
#pragma once
#include "common.cuh"

namespace pdmma {

constexpr int kTileTok = 16;                         // tokens per warp tile
constexpr int kStages = 3;
constexpr int kStageBytes = 2 * kTileTok * 256;      // K tile + V tile, 256 B per token row
template <int kWarps>
constexpr int smem_bytes() { return kWarps * kStages * kStageBytes; }

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
// group_total: query heads per kv head; gchunk: heads of the group handled by one CTA (<= 16).
// num_splits > 1: every split parks its (m, l, o) partial in part_o / part_ml for paged_decode_combine_kernel.
template <int kWarps>
__global__ void __launch_bounds__(kWarps * 32, kWarps == 4 ? 2 : 4)
paged_decode_mma_kernel(const nvc_bf16* __restrict__ q, int ldq, const nvc_bf16* __restrict__ k_cache,
                        const nvc_bf16* __restrict__ v_cache, const int32_t* __restrict__ block_tables, int max_blocks,
                        const int32_t* __restrict__ context_lens, int nh, int kv, int group_total, int gchunk,
                        int block_size, float scale_log2e, nvc_bf16* __restrict__ out, float* __restrict__ part_o,
                        float* __restrict__ part_ml, int num_splits) {
  constexpr int kThreads = kWarps * 32;
  constexpr int kCtaTok = kWarps * kTileTok;          // tokens per CTA iteration
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_trigger();
  // The dependency wait comes AFTER the first K / V tiles are in flight (see below): context_lens and
  // block_tables are step inputs (copied before the step's first kernel), and the only cache rows a
  // kernel of this step writes are the newest token's (position ctx_len - 1: K2 of this layer).
  const int chunks = (group_total + gchunk - 1) / gchunk;
  const int kvh = blockIdx.x / chunks;
  const int chunk = blockIdx.x - kvh * chunks;
  const int head0 = kvh * group_total + chunk * gchunk;
  const int G = min(gchunk, group_total - chunk * gchunk);
  const int b = blockIdx.y, split = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quad = lane >> 2, tq = lane & 3;          // accumulator row (quad, quad + 8), column pair tq

  // independent first loads, issued together: the context length and the first block id (the block of this warp's
  // first tile whenever the CTA starts at tile 0) - one load latency instead of two before the first cp.async
  const int32_t* bt = block_tables + (int64_t)b * max_blocks;
  const int blk_first = __ldg(bt);
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
  const int my_tiles = t_end - t_begin;                // CTA tiles; this warp's tile i covers tokens (t_begin+i)*kCtaTok + 16w

  // issue the loads of warp tile i into ring stage i % kStages (empty group if i is out of range); rows past the
  // context are zero-filled
  auto issue = [&](int i) {
    if (i < my_tiles) {
      const int tok0 = (t_begin + i) * kCtaTok + warp * kTileTok;
      if (tok0 < ctx_len) {
        const int bi = tok0 / block_size;
        const int blk = bi == 0 ? blk_first : __ldg(bt + bi);
        const int64_t slot0 = (int64_t)blk * block_size + (tok0 % block_size);
        const uint32_t st = ring + (i % kStages) * kStageBytes;
        const int c = lane & 15;                       // 16-byte chunk of the 256-byte row
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
          const int r = rr * 2 + (lane >> 4);          // token row inside the tile
          const uint32_t ok = (tok0 + r < ctx_len) ? 16u : 0u;
          const int64_t goff = ((slot0 + (ok ? r : 0)) * kv + kvh) * 256 + c * 16;
          const uint32_t soff = r * 256 + ((c ^ (r & 7)) << 4);
          cp_async16(st + soff, reinterpret_cast<const uint8_t*>(k_cache) + goff, ok);
          cp_async16(st + kTileTok * 256 + soff, reinterpret_cast<const uint8_t*>(v_cache) + goff, ok);
        }
      }
    }
    cp_async_commit();
  };

  // Tiles that cannot hold the newest token are immutable during this step: start them while the
  // predecessor is still running.  A warp's tiles ascend in position, so "tile i is safe" is monotone in i; an
  // out-of-range tile is an empty group.
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


  // ---- merge the warps through shared memory (the rings are dead now)
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
  for (int idx = threadIdx.x; idx < G * 128; idx += kThreads) {
    const int g = idx >> 7, d = idx & 127;
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
}

}  // namespace pdmma
