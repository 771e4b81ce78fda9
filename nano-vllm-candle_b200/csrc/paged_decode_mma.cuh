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
//
// kFused = true folds K2 (q/k RMSNorm + RoPE + KV-cache store of the step's new token,
// src/models/qwen3.rs:206-234, src/layers/rotary_embedding.rs:82-107) into the prologue: q is read raw
// from the fused qkv row, normalised and rotated into shared memory with the arithmetic (and summation
// order) of qk_norm_rope_store_kernel, and the warp whose tile holds position ctx_len - 1 computes the
// new k / v row, writes it to the cache and patches it into its shared-memory tile.  Every cache row
// the kernel LOADS is then older than this step, so all of them may be requested before the PDL wait.
#pragma once
#include "common.cuh"

namespace pdmma {

constexpr int kTileTok = 16;                         // tokens per warp tile
constexpr int kStages = 3;
constexpr int kStageBytes = 2 * kTileTok * 256;      // K tile + V tile, 256 B per token row
constexpr int kQBytes = 16 * 256;                    // fused: the group's normalised q rows (bf16, swizzled)
template <int kWarps, bool kFused>
constexpr int smem_bytes() { return kWarps * kStages * kStageBytes + (kFused ? kQBytes : 0); }

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
__device__ __forceinline__ void st_shared_u4(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// The step's new token of a fused launch (kFused): raw fused qkv rows in, normalised + rotated k and the raw v out
// to the cache.  positions / slot_mapping / rope / norm weights are step inputs or weights (no kernel of the step
// writes them).
struct FusedArgs {
  const nvc_bf16* q_norm_w;
  const nvc_bf16* k_norm_w;
  const float* cos_sin;            // [max_pos][128]: cos[64], sin[64]
  const int32_t* positions;        // [B]
  const int32_t* slot_mapping;     // [B]
  nvc_bf16* k_cache_w;             // the same pools the kernel reads, writable
  nvc_bf16* v_cache_w;
  float eps;
  int dbg;                         // TEMP bisect: 1 = skip the owner's k / v work, 2 = skip the q transform
};

// One 8-lane unit (lane l = 0..7 owns dims [8l, 8l+8) and [64+8l, 64+8l+8)) normalises and rotates one head row
// with the arithmetic and the summation order of qk_norm_rope_store_kernel (elementwise.cu): the fused and the
// stand-alone K2 give bit-identical q / k.  Inactive units contribute zeros and return zero rows.
__device__ __forceinline__ void norm_rope_unit(bool active, const uint4& ra, const uint4& rb, unsigned mask, const uint4& w_lo,
                                               const uint4& w_hi, const float4& c0, const float4& c1, const float4& s0,
                                               const float4& s1, float eps, uint4& o1, uint4& o2) {
  float a[8], b[8];                                    // ra / rb: this lane's two 16-byte pieces (zeros when inactive)
  unpack8(ra, a);
  unpack8(rb, b);
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) ss += a[i] * a[i] + b[i] * b[i];
  ss += __shfl_xor_sync(mask, ss, 1);
  ss += __shfl_xor_sync(mask, ss, 2);
  ss += __shfl_xor_sync(mask, ss, 4);
  const float rstd = 1.0f / sqrtf(ss * (1.0f / 128.0f) + eps);
  float wa[8], wb[8];
  unpack8(w_lo, wa);
  unpack8(w_hi, wb);
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
  if (!active) o1 = o2 = make_uint4(0u, 0u, 0u, 0u);
}

// group_total: query heads per kv head; gchunk: heads of the group handled by one CTA (<= 16).
// num_splits > 1: every split parks its (m, l, o) partial in part_o / part_ml for paged_decode_combine_kernel.
// kFused: q points at the fused qkv rows ([q heads | k heads | v heads], ldq elements per sequence).
template <int kWarps, bool kFused>
__global__ void __launch_bounds__(kWarps * 32, kWarps == 4 ? 2 : 4)
paged_decode_mma_kernel(const nvc_bf16* __restrict__ q, int ldq, const nvc_bf16* __restrict__ k_cache,
                        const nvc_bf16* __restrict__ v_cache, const int32_t* __restrict__ block_tables, int max_blocks,
                        const int32_t* __restrict__ context_lens, int nh, int kv, int group_total, int gchunk,
                        int block_size, float scale_log2e, nvc_bf16* __restrict__ out, float* __restrict__ part_o,
                        float* __restrict__ part_ml, int num_splits, FusedArgs fa) {
  constexpr int kThreads = kWarps * 32;
  constexpr int kCtaTok = kWarps * kTileTok;          // tokens per CTA iteration
  extern __shared__ __align__(128) uint8_t smem[];
  pdl_trigger();
  // The dependency wait comes AFTER the first K / V tiles are in flight (see below): context_lens and
  // block_tables are step inputs (copied before the step's first kernel), and the only cache rows a
  // kernel of this step writes are the newest token's (position ctx_len - 1: K2 of this layer, or, fused, this
  // kernel itself - which then never loads that row).
  const int chunks = (group_total + gchunk - 1) / gchunk;
  const int kvh = blockIdx.x / chunks;
  const int chunk = blockIdx.x - kvh * chunks;
  const int head0 = kvh * group_total + chunk * gchunk;
  const int G = min(gchunk, group_total - chunk * gchunk);
  const int b = blockIdx.y, split = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quad = lane >> 2, tq = lane & 3;          // accumulator row (quad, quad + 8), column pair tq

  // independent first loads, issued together: the context length, the first block id (the block of this warp's first
  // tile whenever the CTA starts at tile 0) and, fused, the position that selects the cos / sin row
  const int32_t* bt = block_tables + (int64_t)b * max_blocks;
  const int pos_in = kFused ? __ldg(fa.positions + b) : 0;
  const int blk_first = __ldg(bt);
  const int ctx_len = __ldg(context_lens + b);
  const int l8 = lane & 7;
  const unsigned mask8 = 0xffu << (lane & 24);
  uint4 qw_lo, qw_hi, kw_lo, kw_hi, k_new1, k_new2, v_new;
  float4 c0, c1, s0, s1;
  if (kFused) {
    // static operands of the q / k transform: in flight while the block table is walked
    const uint4* qw = reinterpret_cast<const uint4*>(fa.q_norm_w);
    const uint4* kw = reinterpret_cast<const uint4*>(fa.k_norm_w);
    qw_lo = __ldg(qw + l8); qw_hi = __ldg(qw + l8 + 8);
    kw_lo = __ldg(kw + l8); kw_hi = __ldg(kw + l8 + 8);
    const float4* cs = reinterpret_cast<const float4*>(fa.cos_sin) + (int64_t)pos_in * 32;
    c0 = __ldg(cs + 2 * l8); c1 = __ldg(cs + 2 * l8 + 1);
    s0 = __ldg(cs + 16 + 2 * l8); s1 = __ldg(cs + 16 + 2 * l8 + 1);
  }
  const int n_load = kFused ? ctx_len - 1 : ctx_len;  // cache rows [0, n_load) are loaded; fused: the last one is made here
  const int n_tiles = (ctx_len + kCtaTok - 1) / kCtaTok;
  const int tiles_per_split = (n_tiles + num_splits - 1) / num_splits;
  const int t_begin = split * tiles_per_split;
  const int t_end = min(n_tiles, t_begin + tiles_per_split);

  float o[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

  const uint32_t ring = smem_u32(smem) + warp * (kStages * kStageBytes);
  const uint32_t q_sm = smem_u32(smem) + kWarps * (kStages * kStageBytes);     // fused only
  const int my_tiles = t_end - t_begin;                // CTA tiles; this warp's tile i covers tokens (t_begin+i)*kCtaTok + 16w

  // issue the loads of warp tile i into ring stage i % kStages (empty group if i is out of range); rows at or past
  // n_load are zero-filled (a fused launch zero-fills the new token's row and patches it after the copy has landed)
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
          const uint32_t ok = (tok0 + r < n_load) ? 16u : 0u;
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
  // out-of-range tile is an empty group.  Fused: every row that is loaded is older than the step.
  auto safe = [&](int i) {
    if (kFused) return true;
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
  // fused: the new token's position / tile owner, and the static operands of the q / k transform
  const int pos_new = ctx_len - 1;
  const bool owner = kFused && ctx_len > 0 && pos_new / kCtaTok >= t_begin && pos_new / kCtaTok < t_end &&
                     (pos_new % kCtaTok) / kTileTok == warp;            // warp-uniform
  const int r_new = pos_new & (kTileTok - 1);
  pdl_wait();
  // ---- Q as A fragments: 8 k-steps of 16 dims; rows >= G are zero
  uint32_t qa[8][4];
  if (kFused) {
    // every load first: the raw q rows of this lane's units (8 lanes per head), the owner's new k / v rows and slot
    constexpr int kUnits = 16 / (kWarps * 4);         // head rows per 8-lane unit: 1 (4 warps) or 2 (2 warps)
    uint4 qa_raw[kUnits], qb_raw[kUnits], ka_raw = make_uint4(0, 0, 0, 0), kb_raw = ka_raw;
    int slot = -1;
#pragma unroll
    for (int i = 0; i < kUnits; ++i) {
      const int u = warp * 4 + (lane >> 3) + i * kWarps * 4;
      qa_raw[i] = qb_raw[i] = make_uint4(0, 0, 0, 0);
      if (u < G) {
        const uint4* row = reinterpret_cast<const uint4*>(q + (int64_t)b * ldq + (int64_t)(head0 + u) * 128);
        qa_raw[i] = row[l8];
        qb_raw[i] = row[l8 + 8];
      }
    }
    if (owner && !(fa.dbg & 1)) {
      const uint4* krow = reinterpret_cast<const uint4*>(q + (int64_t)b * ldq + (int64_t)(nh + kvh) * 128);
      const uint4* vrow = reinterpret_cast<const uint4*>(q + (int64_t)b * ldq + (int64_t)(nh + kv + kvh) * 128);
      if (lane < 8) { ka_raw = krow[l8]; kb_raw = krow[l8 + 8]; }
      if (lane >= 16) v_new = vrow[lane - 16];
      slot = __ldg(fa.slot_mapping + b);
    }
    // normalise + rotate into shared memory (16-byte chunk c of row u at chunk c ^ (u & 7)); rows >= G are zeros
#pragma unroll
    for (int i = 0; i < kUnits; ++i) {
      const int u = warp * 4 + (lane >> 3) + i * kWarps * 4;
      uint4 o1, o2;
      if (fa.dbg & 2) { o1 = qa_raw[i]; o2 = qb_raw[i]; }
      else norm_rope_unit(u < G, qa_raw[i], qb_raw[i], mask8, qw_lo, qw_hi, c0, c1, s0, s1, fa.eps, o1, o2);
      st_shared_u4(q_sm + u * 256 + ((l8 ^ (u & 7)) << 4), o1);
      st_shared_u4(q_sm + u * 256 + (((l8 + 8) ^ (u & 7)) << 4), o2);
    }
    if (owner && !(fa.dbg & 1)) {
      // the step's new k row (lanes 0..7, same unit arithmetic) and v row (lanes 16..31, one 16-byte chunk each):
      // kept in registers for the tile patch below; the CTA of the group's first chunk writes them to the cache
      norm_rope_unit(lane < 8, ka_raw, kb_raw, mask8, kw_lo, kw_hi, c0, c1, s0, s1, fa.eps, k_new1, k_new2);
      if (chunk == 0 && slot >= 0) {
        if (lane < 8) {
          uint4* dst = reinterpret_cast<uint4*>(fa.k_cache_w) + ((int64_t)slot * kv + kvh) * 16;
          dst[lane] = k_new1;
          dst[lane + 8] = k_new2;
        } else if (lane >= 16) {
          reinterpret_cast<uint4*>(fa.v_cache_w)[((int64_t)slot * kv + kvh) * 16 + (lane - 16)] = v_new;
        }
      }
    }
    __syncthreads();
    // ldmatrix.x4: matrices (rows 0-7, k lo), (rows 8-15, k lo), (rows 0-7, k hi), (rows 8-15, k hi) = a0..a3
    const int mi = lane >> 3;
    const int row = (lane & 7) + ((mi & 1) << 3);
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      const int c = 2 * ks + (mi >> 1);
      ldmatrix_x4(q_sm + row * 256 + ((c ^ (row & 7)) << 4), qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
    }
  } else {
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
      if (kFused && owner && tok0 + r_new == pos_new) {
        // the zero-filled row of the new token becomes the k / v computed above
        if (lane < 8) {
          st_shared_u4(st + r_new * 256 + ((lane ^ (r_new & 7)) << 4), k_new1);
          st_shared_u4(st + r_new * 256 + (((lane + 8) ^ (r_new & 7)) << 4), k_new2);
        } else if (lane >= 16) {
          st_shared_u4(st + kTileTok * 256 + r_new * 256 + (((lane - 16) ^ (r_new & 7)) << 4), v_new);
        }
        __syncwarp();
      }
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


  // ---- merge the warps through shared memory (the rings are dead now; the fused q rows sit behind them)
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
