// K5: paged decode attention over block_tables with GQA head grouping (N2 + the decode half
// of src/models/qwen3.rs:236-277; the reference recomputes dense attention over the whole
// sequence each step and has no KV cache, src/models/qwen3.rs:402).
//
// HBM-bound: every K and V row of a (sequence, kv head) is read exactly once and shared by
// the G = nh/kv query heads of the group.  Algorithmic bytes per (sequence, layer):
//   context_len * 2 * kv * 128 * 2  (K+V)  +  2 * nh * 128 * 2  (q in, out).
//
// CTA = 4 warps on one (sequence, kv head, group chunk, KV split).  A 64-token tile is split
// 16 tokens per warp; inside a warp the two 16-lane halves take alternating tokens and each
// lane owns 8 of the 128 head dims (one 16-byte load per row), so a warp issues 16 independent
// 16-byte loads per lane (K and V of 8 row pairs) before it touches the data.  Scores are
// reduced with 4 xor-shuffles inside the half warp, softmax is online in the log2 domain, and
// the 8 (warp, half) partial states are merged through shared memory.  With KV splits > 1
// the partials go to a workspace and a second kernel combines them.
#include "common.cuh"
#include "internal.h"
#include "paged_decode_mma.cuh"

#include <cstdlib>

extern int g_nvc_debug[8];

namespace {

constexpr int kTile = 64;          // tokens per CTA iteration
constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int G>
__global__ void __launch_bounds__(kThreads)
paged_decode_kernel(const nvc_bf16* __restrict__ q, int ldq, const uint4* __restrict__ k_cache,
                    const uint4* __restrict__ v_cache, const int32_t* __restrict__ block_tables, int max_blocks,
                    const int32_t* __restrict__ context_lens, int nh, int kv, int group, int block_size,
                    float scale_log2e, nvc_bf16* __restrict__ out, float* __restrict__ part_o,
                    float* __restrict__ part_ml, int num_splits) {
  __shared__ float sm_m[2 * kWarps][G];
  __shared__ float sm_l[2 * kWarps][G];
  __shared__ float sm_acc[2 * kWarps][G][128];
  pdl_trigger();
  pdl_wait();

  const int chunks = group / G;
  const int kvh = blockIdx.x / chunks;
  const int chunk = blockIdx.x - kvh * chunks;
  const int head0 = kvh * group + chunk * G;
  const int b = blockIdx.y;
  const int split = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int half = lane >> 4, l16 = lane & 15;

  const int ctx_len = __ldg(context_lens + b);
  const int n_tiles = (ctx_len + kTile - 1) / kTile;
  const int tiles_per_split = (n_tiles + num_splits - 1) / num_splits;
  const int t_begin = split * tiles_per_split;
  const int t_end = min(n_tiles, t_begin + tiles_per_split);

  // q (pre-scaled into the log2 domain): this lane's 8 dims of each of the G heads
  float qf[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(q + (int64_t)b * ldq + (head0 + g) * 128) + l16);
    unpack8(u, qf[g]);
#pragma unroll
    for (int j = 0; j < 8; ++j) qf[g][j] *= scale_log2e;
  }
  float m[G], l[G], acc[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    m[g] = -INFINITY;
    l[g] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[g][j] = 0.f;
  }

  const int32_t* bt = block_tables + (int64_t)b * max_blocks;
  for (int t = t_begin; t < t_end; ++t) {
    const int tok0 = t * kTile + warp * 16;
    if (tok0 >= ctx_len) break;                     // warp-uniform
    const int blk = __ldg(bt + tok0 / block_size);
    const int64_t slot0 = (int64_t)blk * block_size + (tok0 % block_size);
    uint4 kr[8], vr[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int tok = tok0 + 2 * i + half;
      const int64_t off = ((slot0 + 2 * i + half) * kv + kvh) * 16 + l16;
      if (tok < ctx_len) {
        kr[i] = ldg_nc_u4(k_cache + off);
        vr[i] = ldg_nc_u4(v_cache + off);
      } else {
        kr[i] = make_uint4(0, 0, 0, 0);
        vr[i] = make_uint4(0, 0, 0, 0);
      }
    }
    float s[G][8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float kf[8];
      unpack8(kr[i], kf);
#pragma unroll
      for (int g = 0; g < G; ++g) {
        float d = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) d += qf[g][j] * kf[j];
        s[g][i] = d;
      }
    }
#pragma unroll
    for (int g = 0; g < G; ++g)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float d = s[g][i];
        d += __shfl_xor_sync(0xffffffffu, d, 8);
        d += __shfl_xor_sync(0xffffffffu, d, 4);
        d += __shfl_xor_sync(0xffffffffu, d, 2);
        d += __shfl_xor_sync(0xffffffffu, d, 1);
        s[g][i] = (tok0 + 2 * i + half < ctx_len) ? d : -INFINITY;
      }
    float alpha[G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
      float mx = m[g];
#pragma unroll
      for (int i = 0; i < 8; ++i) mx = fmaxf(mx, s[g][i]);
      const float ms = (mx == -INFINITY) ? 0.f : mx;     // all-masked so far: keep everything at zero
      alpha[g] = fast_exp2(m[g] - ms);
      float ps = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s[g][i] = fast_exp2(s[g][i] - ms);
        ps += s[g][i];
      }
      l[g] = l[g] * alpha[g] + ps;
      m[g] = mx;
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[g][j] *= alpha[g];
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float vf[8];
      unpack8(vr[i], vf);
#pragma unroll
      for (int g = 0; g < G; ++g)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[g][j] += s[g][i] * vf[j];
    }
  }

  // ---- merge the 8 (warp, half) states
  const int st = warp * 2 + half;
#pragma unroll
  for (int g = 0; g < G; ++g) {
    if (l16 == 0) { sm_m[st][g] = m[g]; sm_l[st][g] = l[g]; }
#pragma unroll
    for (int j = 0; j < 8; ++j) sm_acc[st][g][8 * l16 + j] = acc[g][j];
  }
  __syncthreads();
  const int d = threadIdx.x;   // 128 threads = 128 dims
#pragma unroll
  for (int g = 0; g < G; ++g) {
    float M = -INFINITY;
#pragma unroll
    for (int s2 = 0; s2 < 2 * kWarps; ++s2) M = fmaxf(M, sm_m[s2][g]);
    float L = 0.f, O = 0.f;
    if (M != -INFINITY) {
#pragma unroll
      for (int s2 = 0; s2 < 2 * kWarps; ++s2) {
        const float w = fast_exp2(sm_m[s2][g] - M);
        L += sm_l[s2][g] * w;
        O += sm_acc[s2][g][d] * w;
      }
    }
    const int head = head0 + g;
    if (num_splits == 1) {
      const float o = L > 0.f ? O / L : 0.f;
      out[((int64_t)b * nh + head) * 128 + d] = __bfloat16_as_ushort(__float2bfloat16_rn(o));
    } else {
      const int64_t p = ((int64_t)b * nh + head) * num_splits + split;
      part_o[p * 128 + d] = O;
      if (d == 0) { part_ml[p * 2] = M; part_ml[p * 2 + 1] = L; }
    }
  }
}

__global__ void __launch_bounds__(128)
paged_decode_combine_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                            nvc_bf16* __restrict__ out, int nh, int num_splits) {
  pdl_trigger();
  pdl_wait();
  const int head = blockIdx.x, b = blockIdx.y, d = threadIdx.x;
  const int64_t p0 = ((int64_t)b * nh + head) * num_splits;
  float M = -INFINITY;
  for (int s = 0; s < num_splits; ++s) M = fmaxf(M, __ldg(part_ml + (p0 + s) * 2));
  float L = 0.f, O = 0.f;
  if (M != -INFINITY) {
    for (int s = 0; s < num_splits; ++s) {
      const float w = fast_exp2(__ldg(part_ml + (p0 + s) * 2) - M);
      L += __ldg(part_ml + (p0 + s) * 2 + 1) * w;
      O += __ldg(part_o + (p0 + s) * 128 + d) * w;
    }
  }
  const float o = L > 0.f ? O / L : 0.f;
  out[((int64_t)b * nh + head) * 128 + d] = __bfloat16_as_ushort(__float2bfloat16_rn(o));
}

constexpr int kFlatMaxPairs = 1 << 16;

int max_splits_for_batch(int B, int num_sms) {
  int s = (2 * num_sms + B - 1) / B;
  return s < 1 ? 1 : (s > 64 ? 64 : s);
}

}  // namespace

extern "C" int nvc_paged_decode_init(nvc_ctx* ctx) {
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(pdmma::paged_decode_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           pdmma::kSmemBytes));
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(pdmma::paged_decode_flat_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           pdmma::kSmemBytes));
  // per-(kv head, sequence) arrival tickets of the work-balanced kernel; zero between launches
  NVC_CUDA_CHECK(ctx, cudaMalloc(&ctx->pd_counters, (size_t)kFlatMaxPairs * sizeof(int)));
  NVC_CUDA_CHECK(ctx, cudaMemset(ctx->pd_counters, 0, (size_t)kFlatMaxPairs * sizeof(int)));
  return NVC_OK;
}

extern "C" size_t nvc_paged_decode_workspace_bytes(int B, int nh, int hd, int num_sms) {
  if (B <= 0 || nh <= 0 || hd <= 0) return 0;
  const int sms = num_sms > 0 ? num_sms : 148;
  const size_t S = (size_t)max_splits_for_batch(B, sms);
  const size_t split = (size_t)B * nh * S * (hd + 2) * sizeof(float);
  // work-balanced kernel: one (m, l, o) entry per head and slot, slots = pairs + CTAs (2 per SM); head groups
  // of more than 16 are chunked, which can pad nh by < 2x
  const size_t flat = (size_t)2 * nh * ((size_t)B + 2 * sms) * (hd + 2) * sizeof(float);
  return split > flat ? split : flat;
}

extern "C" int nvc_paged_decode_attn_ex(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                                        const nvc_bf16* v_cache, const int32_t* block_tables, int max_blocks,
                                        int max_blocks_hint, const int32_t* context_lens, int B, int nh, int kv,
                                        int hd, int block_size, float scale, nvc_bf16* out, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  NVC_REQUIRE(ctx, ctx && q && k_cache && v_cache && block_tables && context_lens && out, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, B >= 0 && nh > 0 && kv > 0 && max_blocks > 0 && max_blocks_hint > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, hd == 128, NVC_ESHAPE, "head_dim must be 128");
  NVC_REQUIRE(ctx, nh % kv == 0, NVC_ESHAPE, "nh must be a multiple of kv");
  NVC_REQUIRE(ctx, block_size > 0 && block_size % 16 == 0, NVC_ESHAPE, "block_size must be a multiple of 16");
  NVC_REQUIRE(ctx, ldq % 8 == 0 && ldq >= nh * hd, NVC_ESHAPE, "ldq must be >= nh*hd and a multiple of 8");
  if (B == 0) return NVC_OK;
  const int group = nh / kv;
  // default: tensor-core kernel (whole group = the M dimension of mma.sync); NVC_PD_SIMT=1 selects the
  // CUDA-core kernel (kept for A/B measurements, tools/decode_microbench.py)
  static const bool use_simt = getenv("NVC_PD_SIMT") != nullptr && atoi(getenv("NVC_PD_SIMT")) != 0;
  int G;
  if (!use_simt) G = group < 16 ? group : 16;
  else if (group % 4 == 0) G = 4;
  else if (group % 5 == 0) G = 5;
  else if (group % 2 == 0) G = 2;
  else G = 1;
  const int chunks = (group + G - 1) / G;
  const int max_tiles = (int)(((int64_t)max_blocks_hint * block_size + kTile - 1) / kTile);
  // KV splits: only when the (sequence, kv head) grid leaves most of the 2 x 148 resident slots empty, and never
  // beyond them (tools/decode_microbench.py --splits, profiles/r02_pd_split_sweep.txt: 64 CTAs -> 4 splits 10.3 ->
  // 7.6 us, 128 -> 2, 256 and up -> 1; rounding the split count UP to cover the slots cost 3 - 5 us on the
  // TP-sharded shapes through the second wave it created)
  int splits = 2 * ctx->num_sms / (B * kv * chunks);
  const int cap = max_splits_for_batch(B, ctx->num_sms);
  if (splits > cap) splits = cap;
  if (splits > max_tiles) splits = max_tiles;
  if (splits < 1) splits = 1;
  if (const char* e = getenv("NVC_PD_SPLITS")) {   // tuning override (tools/decode_microbench.py)
    const int f = atoi(e);
    if (f > 0) splits = f < cap ? f : cap;
  }
  float* part_o = nullptr;
  float* part_ml = nullptr;
  const float scale_log2e = scale * 1.4426950408889634f;
  cudaStream_t st = nvc_stream(stream);
  // Work-balanced kernel (flat item list cut into equal ranges, fused last-arriver merge): opt-in through
  // NVC_PD_FLAT=1 / nvc_debug_set(1, 1).  Measured on B200 it removes the tail wave but pays ~5 us of per-CTA
  // overhead (prefix sums, a pipeline restart per (head, sequence) segment, the ticket round trip): 34.0 us vs
  // 29.3 us on Qwen3-0.6B bs 64, 323 vs 307 us on 8B bs 256 ctx 2048 - so the per-(head, sequence) grid stays
  // the default until the segment boundaries are software-pipelined.
  static const bool flat_env = getenv("NVC_PD_FLAT") != nullptr && atoi(getenv("NVC_PD_FLAT")) != 0;
  const bool use_flat = flat_env || g_nvc_debug[1] == 1;
  if (!use_simt && use_flat && !getenv("NVC_PD_SPLITS") && B <= pdmma::kMaxFlatBatch && ctx->pd_counters != nullptr &&
      (int64_t)kv * chunks * B <= kFlatMaxPairs) {
    const int kvx = kv * chunks;
    const int groups = 2 * ctx->num_sms / kvx > 0 ? 2 * ctx->num_sms / kvx : 1;
    const int P = groups * kvx;
    const size_t entries = (size_t)kvx * (B + groups) * G;
    const size_t need = entries * (hd + 2) * sizeof(float);
    if (workspace != nullptr && workspace_bytes >= need) {
      part_o = reinterpret_cast<float*>(workspace);
      part_ml = part_o + entries * hd;
      NVC_CUDA_CHECK(ctx, nvc_launch(ctx, pdmma::paged_decode_flat_kernel, dim3(P), dim3(pdmma::kThreads), pdmma::kSmemBytes,
                                     st, q, ldq, k_cache, v_cache, block_tables, max_blocks, context_lens, B, nh, kv, group,
                                     G, block_size, scale_log2e, out, part_o, part_ml, ctx->pd_counters));
      NVC_CUDA_CHECK(ctx, cudaGetLastError());
      ctx->launches++;
      return NVC_OK;
    }
  }
  if (splits > 1) {
    const size_t need = (size_t)B * nh * splits * (hd + 2) * sizeof(float);
    NVC_REQUIRE(ctx, workspace != nullptr && workspace_bytes >= need, NVC_ENOMEM, "paged decode workspace too small");
    part_o = reinterpret_cast<float*>(workspace);
    part_ml = part_o + (size_t)B * nh * splits * hd;
  }
  dim3 grid(kv * chunks, B, splits);
  // splits merged by the last CTA to arrive (tickets in pd_counters) instead of the combine kernel
  // (opt-in, NVC_PD_FUSED_MERGE=1: measured 2 - 4 us SLOWER than the combine kernel on every shape of the sweep - the
  // last arriver walks heads x splits dependent L2 loads alone, the combine kernel spreads them over nh x B CTAs)
  const bool fused_merge = !use_simt && splits > 1 && ctx->pd_counters != nullptr && (int64_t)kv * chunks * B <= kFlatMaxPairs &&
                           getenv("NVC_PD_FUSED_MERGE") && atoi(getenv("NVC_PD_FUSED_MERGE")) != 0;
  auto kc = reinterpret_cast<const uint4*>(k_cache);
  auto vc = reinterpret_cast<const uint4*>(v_cache);
#define NVC_LAUNCH_PD(GV)                                                                                   \
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, paged_decode_kernel<GV>, grid, dim3(kThreads), 0, st, q, ldq, kc, vc,      \
                                 block_tables, max_blocks, context_lens, nh, kv, group, block_size,            \
                                 scale_log2e, out, part_o, part_ml, splits))
  if (!use_simt) {
    NVC_CUDA_CHECK(ctx, nvc_launch(ctx, pdmma::paged_decode_mma_kernel, grid, dim3(pdmma::kThreads), pdmma::kSmemBytes, st,
                                   q, ldq, k_cache, v_cache, block_tables, max_blocks, context_lens, nh, kv, group, G,
                                   block_size, scale_log2e, out, part_o, part_ml, splits, fused_merge ? ctx->pd_counters : nullptr));
  } else {
    switch (G) {
      case 4: NVC_LAUNCH_PD(4); break;
      case 5: NVC_LAUNCH_PD(5); break;
      case 2: NVC_LAUNCH_PD(2); break;
      default: NVC_LAUNCH_PD(1); break;
    }
  }
#undef NVC_LAUNCH_PD
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  if (splits > 1 && !fused_merge) {
    NVC_CUDA_CHECK(ctx, nvc_launch(ctx, paged_decode_combine_kernel, dim3(nh, B), dim3(128), 0, st, part_o, part_ml, out,
                                   nh, splits));
    NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  }
  return NVC_OK;
}

extern "C" int nvc_paged_decode_attn(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                                     const nvc_bf16* v_cache, const int32_t* block_tables, int max_blocks,
                                     const int32_t* context_lens, int B, int nh, int kv, int hd, int block_size,
                                     float scale, nvc_bf16* out, void* workspace, size_t workspace_bytes,
                                     void* stream) {
  return nvc_paged_decode_attn_ex(ctx, q, ldq, k_cache, v_cache, block_tables, max_blocks, max_blocks, context_lens, B,
                                  nh, kv, hd, block_size, scale, out, workspace, workspace_bytes, stream);
}
