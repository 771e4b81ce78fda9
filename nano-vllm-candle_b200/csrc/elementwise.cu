// HBM-bound kernels of the Qwen3 forward path: embedding gather, fused add+RMSNorm,
// q/k-norm + RoPE + KV-cache scatter, stand-alone slot-mapping scatter, SiLU*mul,
// last-token row gather, argmax.  All use 16-byte vector accesses, one pass over the
// data, fp32 math, warp-shuffle reductions.  Each entry point cites the reference code it
// replaces in include/nvc_b200.h.
#include "common.cuh"
#include "internal.h"
#include "ptx.cuh"

extern int g_nvc_debug[8];   // prefill_attn.cu: [6] == 1 switches the sampling kernel's screening off (tests)

#include <cmath>
#include <vector>

// ------------------------------------------------------------------------------------
// K0 embed gather: out[n,:] = table[ids[n],:]        (src/models/qwen3.rs:465-468)
// one warp per 512-byte slice of a row; 16 B per lane.
// ------------------------------------------------------------------------------------
__global__ void embed_gather_kernel(const int32_t* __restrict__ ids, const uint4* __restrict__ table,
                                    uint4* __restrict__ out, int N, int vec_per_row) {
  pdl_trigger();
  pdl_wait();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t total = (int64_t)N * vec_per_row;
  if (i >= total) return;
  int n = (int)(i / vec_per_row);
  int c = (int)(i - (int64_t)n * vec_per_row);
  int id = __ldg(ids + n);
  out[i] = __ldg(table + (int64_t)id * vec_per_row + c);
}

extern "C" int nvc_embed_gather(nvc_ctx* ctx, const int32_t* ids, const nvc_bf16* table, nvc_bf16* out,
                                int N, int H, int V, void* stream) {
  NVC_REQUIRE(ctx, ctx && ids && table && out, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, N >= 0 && H > 0 && V > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, H % 8 == 0, NVC_ESHAPE, "H must be a multiple of 8");
  if (N == 0) return NVC_OK;
  int vpr = H / 8;
  int64_t total = (int64_t)N * vpr;
  int threads = 256;
  int64_t blocks = (total + threads - 1) / threads;
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, embed_gather_kernel, dim3((unsigned)blocks), dim3(threads), 0, nvc_stream(stream),
                                 ids, reinterpret_cast<const uint4*>(table), reinterpret_cast<uint4*>(out), N, vpr));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// ------------------------------------------------------------------------------------
// K1 fused add + RMSNorm                              (src/layers/layernorm.rs:44-60)
// one CTA per row, one 8-element vector per thread (H <= 8192) kept in registers, or a
// strided loop for wider rows (second pass re-reads residual_out from L2).
// Algorithmic bytes per row: H * (2 [x] + 4 [res in] + 4 [res out] + 2 [out]).
// ------------------------------------------------------------------------------------
template <bool kLoop>
__global__ void __launch_bounds__(1024)
fused_add_rmsnorm_kernel(const uint4* __restrict__ x, const float4* __restrict__ res_in,
                         float4* __restrict__ res_out, const uint4* __restrict__ w, float eps,
                         uint4* __restrict__ out, int H) {
  __shared__ float red[32];
  pdl_trigger();
  const int row = blockIdx.x;
  const int nvec = H >> 3;
  // the norm weight is not produced by a kernel: fetch it before the dependency wait (single-pass variant)
  uint4 wv = make_uint4(0, 0, 0, 0);
  if (!kLoop && (int)threadIdx.x < nvec) wv = __ldg(w + threadIdx.x);
  pdl_wait();
  const int64_t vbase = (int64_t)row * nvec;
  float s[8];
  float ss = 0.f;
  if (!kLoop) {
    const int v = threadIdx.x;
    if (v < nvec) {
      unpack8(ldg_nc_u4(x + vbase + v), s);
      if (res_in) {
        float4 a = res_in[(vbase + v) * 2], b = res_in[(vbase + v) * 2 + 1];
        s[0] += a.x; s[1] += a.y; s[2] += a.z; s[3] += a.w;
        s[4] += b.x; s[5] += b.y; s[6] += b.z; s[7] += b.w;
      }
      if (res_out) {
        res_out[(vbase + v) * 2] = make_float4(s[0], s[1], s[2], s[3]);
        res_out[(vbase + v) * 2 + 1] = make_float4(s[4], s[5], s[6], s[7]);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) ss += s[i] * s[i];
    }
  } else {
    for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
      unpack8(ldg_nc_u4(x + vbase + v), s);
      if (res_in) {
        float4 a = res_in[(vbase + v) * 2], b = res_in[(vbase + v) * 2 + 1];
        s[0] += a.x; s[1] += a.y; s[2] += a.z; s[3] += a.w;
        s[4] += b.x; s[5] += b.y; s[6] += b.z; s[7] += b.w;
      }
      if (res_out) {
        res_out[(vbase + v) * 2] = make_float4(s[0], s[1], s[2], s[3]);
        res_out[(vbase + v) * 2 + 1] = make_float4(s[4], s[5], s[6], s[7]);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) ss += s[i] * s[i];
    }
  }
  ss = warp_sum(ss);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) red[wid] = ss;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float tot = (lane < nw) ? red[lane] : 0.f;
  tot = warp_sum(tot);
  // reference: x * (1 / sqrt(mean + eps))
  const float rstd = 1.0f / sqrtf(tot / (float)H + eps);
  if (!kLoop) {
    const int v = threadIdx.x;
    if (v < nvec) {
      float wf[8];
      unpack8(wv, wf);
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] = s[i] * rstd * wf[i];
      out[vbase + v] = pack8(s);
    }
  } else {
    for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
      if (res_out) {   // re-read the sum we just wrote (same thread wrote it)
        float4 a = res_out[(vbase + v) * 2], b = res_out[(vbase + v) * 2 + 1];
        s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
      } else {
        unpack8(ldg_nc_u4(x + vbase + v), s);
        if (res_in) {
          float4 a = res_in[(vbase + v) * 2], b = res_in[(vbase + v) * 2 + 1];
          s[0] += a.x; s[1] += a.y; s[2] += a.z; s[3] += a.w;
          s[4] += b.x; s[5] += b.y; s[6] += b.z; s[7] += b.w;
        }
      }
      float wf[8];
      unpack8(__ldg(w + v), wf);
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] = s[i] * rstd * wf[i];
      out[vbase + v] = pack8(s);
    }
  }
}

extern "C" int nvc_fused_add_rmsnorm(nvc_ctx* ctx, const nvc_bf16* x, const float* residual_in,
                                     float* residual_out, const nvc_bf16* w, float eps, nvc_bf16* out,
                                     int N, int H, void* stream) {
  NVC_REQUIRE(ctx, ctx && x && w && out, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, N >= 0 && H > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, H % 8 == 0 && H <= 65536, NVC_ESHAPE, "H must be a multiple of 8 and <= 65536");
  if (N == 0) return NVC_OK;
  const int nvec = H / 8;
  auto xs = reinterpret_cast<const uint4*>(x);
  auto ri = reinterpret_cast<const float4*>(residual_in);
  auto ro = reinterpret_cast<float4*>(residual_out);
  auto ws = reinterpret_cast<const uint4*>(w);
  auto os = reinterpret_cast<uint4*>(out);
  if (nvec <= 1024) {
    int threads = ((nvec + 31) / 32) * 32;
    NVC_CUDA_CHECK(ctx, nvc_launch(ctx, fused_add_rmsnorm_kernel<false>, dim3(N), dim3(threads), 0, nvc_stream(stream),
                                   xs, ri, ro, ws, eps, os, H));
  } else {
    NVC_CUDA_CHECK(ctx, nvc_launch(ctx, fused_add_rmsnorm_kernel<true>, dim3(N), dim3(1024), 0, nvc_stream(stream),
                                   xs, ri, ro, ws, eps, os, H));
  }
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// ------------------------------------------------------------------------------------
// rope table (host math = the reference's f32 powf/cos/sin)   (rotary_embedding.rs:56-80)
// ------------------------------------------------------------------------------------
extern "C" int nvc_rope_table_init_sync(nvc_ctx* ctx, float theta, int head_dim, int max_pos, float* out) {
  NVC_REQUIRE(ctx, ctx && out, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, head_dim > 0 && head_dim % 2 == 0 && max_pos > 0, NVC_EINVAL, "bad sizes");
  const int half = head_dim / 2;
  std::vector<float> inv_freq(half);
  for (int j = 0; j < half; ++j) {
    float exponent = (2.0f * (float)j) / (float)head_dim;
    inv_freq[j] = 1.0f / powf(theta, exponent);
  }
  std::vector<float> tab((size_t)max_pos * head_dim);
  for (int p = 0; p < max_pos; ++p) {
    float* row = tab.data() + (size_t)p * head_dim;
    for (int j = 0; j < half; ++j) {
      float ang = (float)p * inv_freq[j];
      row[j] = cosf(ang);
      row[half + j] = sinf(ang);
    }
  }
  NVC_CUDA_CHECK(ctx, cudaMemcpy(out, tab.data(), tab.size() * sizeof(float), cudaMemcpyHostToDevice));
  return NVC_OK;
}

// ------------------------------------------------------------------------------------
// K2 q/k RMSNorm + RoPE + KV-cache scatter, one pass over the fused qkv buffer.
// (src/models/qwen3.rs:206-234, src/layers/rotary_embedding.rs:82-107, N3)
// 8 lanes per (token, head) unit: lane l owns elements [8l, 8l+8) and [64+8l, 64+8l+8).
// Algorithmic bytes per token: (nh+2kv)*256 read + (nh+kv)*256 written in place
//                              + 2*kv*256 written to the cache.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
qk_norm_rope_store_kernel(uint4* __restrict__ qkv, const uint4* __restrict__ qw, const uint4* __restrict__ kw,
                          float eps, const float4* __restrict__ cos_sin, const int32_t* __restrict__ positions,
                          const int32_t* __restrict__ slot_mapping, uint4* __restrict__ k_cache,
                          uint4* __restrict__ v_cache, int N, int nh, int kv) {
  pdl_trigger();
  const int units_per_tok = nh + 2 * kv;
  const int64_t unit = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const int l = threadIdx.x & 7;
  const bool active = unit < (int64_t)N * units_per_tok;
  // all 8 lanes of a unit share `active`; inactive units still take part in shuffles below
  const int n = active ? (int)(unit / units_per_tok) : 0;
  const int u = active ? (int)(unit - (int64_t)n * units_per_tok) : 0;
  uint4* row = qkv + ((int64_t)n * units_per_tok + u) * 16;   // 16 uint4 per 128-element head
  const int slot = (slot_mapping != nullptr && active) ? __ldg(slot_mapping + n) : -1;
  // Everything that is not produced by a kernel of this step - slot, position, the rope row, the norm weight.
  // Small grids (decode: latency-bound, the kernel sits on the step's critical path): fetched BEFORE the
  // dependency wait, so the two dependent loads positions -> cos/sin overlap the predecessor.  Large grids
  // (prefill: throughput-bound): fetched after the row loads have been issued, so both chains are in flight
  // together (fetching them first serialises the two latencies per thread: 120 -> 162 us at 32768 tokens).
  const bool is_qk = u < nh + kv;
  const uint4* wp = u < nh ? qw : kw;
  uint4 w_lo = make_uint4(0, 0, 0, 0), w_hi = w_lo;
  float4 c0 = make_float4(0, 0, 0, 0), c1 = c0, s0 = c0, s1 = c0;
  auto load_static = [&]() {
    if (active && is_qk) {
      if (wp != nullptr) {
        w_lo = __ldg(wp + l);
        w_hi = __ldg(wp + l + 8);
      }
      const float4* cs = cos_sin + (int64_t)__ldg(positions + n) * 32;   // 128 floats per position: cos[64], sin[64]
      c0 = __ldg(cs + 2 * l); c1 = __ldg(cs + 2 * l + 1);
      s0 = __ldg(cs + 16 + 2 * l); s1 = __ldg(cs + 16 + 2 * l + 1);
    }
  };
  const bool early = gridDim.x <= 1024;     // block-uniform
  if (early) load_static();
  pdl_wait();

  if (u >= nh + kv) {   // V head: copy to the cache (bit-exact)
    if (active && slot >= 0) {
      const int h = u - nh - kv;
      uint4* dst = v_cache + ((int64_t)slot * kv + h) * 16;
      dst[l] = row[l];
      dst[l + 8] = row[l + 8];
    }
    return;
  }
  // NOTE: a warp never mixes the early-returned V units with q/k units only if the unit
  // kinds align to warp boundaries, which they do not in general; the shuffles below use
  // an 8-lane mask so they only involve the lanes of this unit.
  const unsigned mask = 0xffu << ((threadIdx.x & 31) & ~7);
  uint4 ra = make_uint4(0, 0, 0, 0), rb = ra;
  if (active) {
    ra = row[l];
    rb = row[l + 8];
    if (!early) load_static();
  }
  uint4 o1, o2;
  qk_norm_rope8(ra, rb, mask, wp != nullptr, w_lo, w_hi, eps, c0, c1, s0, s1, o1, o2);
  if (!active) return;
  const bool is_q = u < nh;
  row[l] = o1;
  row[l + 8] = o2;
  if (!is_q && slot >= 0) {
    const int h = u - nh;
    uint4* dst = k_cache + ((int64_t)slot * kv + h) * 16;
    dst[l] = o1;
    dst[l + 8] = o2;
  }
}

extern "C" int nvc_qk_norm_rope_store(nvc_ctx* ctx, nvc_bf16* qkv, const nvc_bf16* qw, const nvc_bf16* kw,
                                      float eps, const float* cos_sin, const int32_t* positions,
                                      const int32_t* slot_mapping, nvc_bf16* k_cache, nvc_bf16* v_cache,
                                      int N, int nh, int kv, int hd, void* stream) {
  NVC_REQUIRE(ctx, ctx && qkv && cos_sin && positions, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, (qw == nullptr) == (kw == nullptr), NVC_EINVAL, "q/k norm weights must both be given or both be NULL");
  NVC_REQUIRE(ctx, slot_mapping == nullptr || (k_cache && v_cache), NVC_EINVAL, "slot_mapping without caches");
  NVC_REQUIRE(ctx, N >= 0 && nh > 0 && kv > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, hd == 128, NVC_ESHAPE, "head_dim must be 128");
  if (N == 0) return NVC_OK;
  const int64_t units = (int64_t)N * (nh + 2 * kv);
  const int threads = 256;
  const int64_t blocks = (units * 8 + threads - 1) / threads;
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, qk_norm_rope_store_kernel, dim3((unsigned)blocks), dim3(threads), 0,
                                 nvc_stream(stream), reinterpret_cast<uint4*>(qkv), reinterpret_cast<const uint4*>(qw),
                                 reinterpret_cast<const uint4*>(kw), eps, reinterpret_cast<const float4*>(cos_sin),
                                 positions, slot_mapping, reinterpret_cast<uint4*>(k_cache),
                                 reinterpret_cast<uint4*>(v_cache), N, nh, kv));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// ------------------------------------------------------------------------------------
// K2' stand-alone slot-mapping scatter (N3).  16 B per thread; bit-exact copy.
// Algorithmic bytes per token: 2 * kv * hd * 2 read + the same written.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
store_kvcache_kernel(const uint4* __restrict__ k, const uint4* __restrict__ v, int ld_vec,
                     uint4* __restrict__ k_cache, uint4* __restrict__ v_cache,
                     const int32_t* __restrict__ slot_mapping, int N, int row_vec) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)N * row_vec) return;
  const int n = (int)(i / row_vec);
  const int c = (int)(i - (int64_t)n * row_vec);
  const int slot = __ldg(slot_mapping + n);
  if (slot < 0) return;
  k_cache[(int64_t)slot * row_vec + c] = ldg_nc_u4(k + (int64_t)n * ld_vec + c);
  v_cache[(int64_t)slot * row_vec + c] = ldg_nc_u4(v + (int64_t)n * ld_vec + c);
}

extern "C" int nvc_store_kvcache(nvc_ctx* ctx, const nvc_bf16* k, const nvc_bf16* v, int ld,
                                 nvc_bf16* k_cache, nvc_bf16* v_cache, const int32_t* slot_mapping,
                                 int N, int kv, int hd, void* stream) {
  NVC_REQUIRE(ctx, ctx && k && v && k_cache && v_cache && slot_mapping, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, N >= 0 && kv > 0 && hd > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, (kv * hd) % 8 == 0 && ld % 8 == 0 && ld >= kv * hd, NVC_ESHAPE, "row must be 16-byte granular");
  if (N == 0) return NVC_OK;
  const int row_vec = kv * hd / 8;
  const int64_t total = (int64_t)N * row_vec;
  const int threads = 256;
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, store_kvcache_kernel, dim3((unsigned)((total + threads - 1) / threads)),
                                 dim3(threads), 0, nvc_stream(stream), reinterpret_cast<const uint4*>(k),
                                 reinterpret_cast<const uint4*>(v), ld / 8, reinterpret_cast<uint4*>(k_cache),
                                 reinterpret_cast<uint4*>(v_cache), slot_mapping, N, row_vec));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// ------------------------------------------------------------------------------------
// K6 SiLU(x[:, :I]) * x[:, I:]                          (src/layers/activation.rs:13-18)
// Algorithmic bytes per token: I * (4 read + 2 written).
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
silu_and_mul_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, int N, int ivec) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)N * ivec) return;
  const int n = (int)(i / ivec);
  const int c = (int)(i - (int64_t)n * ivec);
  float g[8], u[8];
  unpack8(ldg_nc_u4(x + (int64_t)n * 2 * ivec + c), g);
  unpack8(ldg_nc_u4(x + (int64_t)n * 2 * ivec + ivec + c), u);
#pragma unroll
  for (int j = 0; j < 8; ++j) g[j] = (g[j] / (1.0f + __expf(-g[j]))) * u[j];
  out[i] = pack8(g);
}

extern "C" int nvc_silu_and_mul(nvc_ctx* ctx, const nvc_bf16* x, nvc_bf16* out, int N, int I, void* stream) {
  NVC_REQUIRE(ctx, ctx && x && out, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, N >= 0 && I > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, I % 8 == 0, NVC_ESHAPE, "I must be a multiple of 8");
  if (N == 0) return NVC_OK;
  const int ivec = I / 8;
  const int64_t total = (int64_t)N * ivec;
  const int threads = 256;
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, silu_and_mul_kernel, dim3((unsigned)((total + threads - 1) / threads)),
                                 dim3(threads), 0, nvc_stream(stream), reinterpret_cast<const uint4*>(x),
                                 reinterpret_cast<uint4*>(out), N, ivec));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// ------------------------------------------------------------------------------------
// K7 last-token row gather (hidden bf16 + residual f32)
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gather_rows_kernel(const uint4* __restrict__ x, const float4* __restrict__ res, const int32_t* __restrict__ rows,
                   uint4* __restrict__ x_out, float4* __restrict__ res_out, int B, int hvec) {
  pdl_trigger();
  pdl_wait();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)B * hvec) return;
  const int b = (int)(i / hvec);
  const int c = (int)(i - (int64_t)b * hvec);
  const int64_t r = __ldg(rows + b);
  x_out[i] = x[r * hvec + c];
  if (res != nullptr) {
    res_out[i * 2] = res[(r * hvec + c) * 2];
    res_out[i * 2 + 1] = res[(r * hvec + c) * 2 + 1];
  }
}

extern "C" int nvc_gather_rows(nvc_ctx* ctx, const nvc_bf16* x, const float* residual, const int32_t* rows,
                               nvc_bf16* x_out, float* residual_out, int B, int H, void* stream) {
  NVC_REQUIRE(ctx, ctx && x && rows && x_out, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, (residual == nullptr) == (residual_out == nullptr), NVC_EINVAL, "residual in/out mismatch");
  NVC_REQUIRE(ctx, B >= 0 && H > 0 && H % 8 == 0, NVC_ESHAPE, "H must be a multiple of 8");
  if (B == 0) return NVC_OK;
  const int hvec = H / 8;
  const int64_t total = (int64_t)B * hvec;
  const int threads = 256;
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, gather_rows_kernel, dim3((unsigned)((total + threads - 1) / threads)),
                                 dim3(threads), 0, nvc_stream(stream), reinterpret_cast<const uint4*>(x),
                                 reinterpret_cast<const float4*>(residual), rows, reinterpret_cast<uint4*>(x_out),
                                 reinterpret_cast<float4*>(residual_out), B, hvec));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// ------------------------------------------------------------------------------------
// K8 argmax with the reference's tie rule (LAST maximal index)  (llm_engine.rs:135-142)
// float4 loads; (value, index) reduced with "greater, or equal and later".
// ------------------------------------------------------------------------------------
__device__ __forceinline__ void argmax_combine(float& v, int& i, float ov, int oi) {
  if (ov > v || (ov == v && oi > i)) { v = ov; i = oi; }
}

// Block reduce of the threads' (value, index) candidates, then - when the row is shared by the `csz` CTAs of a cluster -
// every CTA stores its pair into CTA 0's shared memory over DSMEM and CTA 0 merges after one cluster barrier.
__device__ __forceinline__ void argmax_finish_row(float bv, int bi, int r, int rank, int csz, int32_t* __restrict__ idx,
                                                  float* __restrict__ val, int index_offset) {
  __shared__ float sv[32];
  __shared__ int si[32];
  __shared__ float cv[4];
  __shared__ int ci[4];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    argmax_combine(bv, bi, ov, oi);
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { sv[wid] = bv; si[wid] = bi; }
  __syncthreads();
  if (wid == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    bv = lane < nw ? sv[lane] : -INFINITY;
    bi = lane < nw ? si[lane] : -1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      argmax_combine(bv, bi, ov, oi);
    }
    if (lane == 0 && csz > 1) {
      asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(ptx::mapa_u32(smem_u32(&cv[rank]), 0)), "f"(bv) : "memory");
      asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(ptx::mapa_u32(smem_u32(&ci[rank]), 0)), "r"(bi) : "memory");
    }
  }
  if (csz > 1) {
    ptx::cluster_sync_all();
    if (rank != 0) return;
    if (threadIdx.x == 0) {
      bv = cv[0];
      bi = ci[0];
      for (int c = 1; c < csz; ++c) argmax_combine(bv, bi, cv[c], ci[c]);
    }
  }
  if (threadIdx.x == 0) {
    idx[r] = bi + index_offset;
    if (val) val[r] = bv;
  }
}

// widest cluster (4 / 2 / 1 CTAs per row) that keeps the grid within one resident wave of two 1024-thread CTAs per SM
static int argmax_cluster_width(const nvc_ctx* ctx, int B, int V) {
  return (V >= 16384 && 4 * B <= 2 * ctx->num_sms) ? 4 : (V >= 16384 && 2 * B <= 2 * ctx->num_sms) ? 2 : 1;
}

// A row is shared by the `csz` CTAs of a cluster (csz = 1, 2 or 4: at batch 64 one CTA per row would leave 84 SMs idle):
// CTA r scans the r-th slice of the row, stores its (value, index) into CTA 0's shared memory over DSMEM, and CTA 0
// merges after one cluster barrier.  The "last maximal index" rule is order-independent, so the split does not show.
__global__ void __launch_bounds__(1024)
argmax_last_kernel(const float* __restrict__ logits, int ld, int32_t* __restrict__ idx, float* __restrict__ val,
                   int V, int index_offset, int csz) {
  pdl_trigger();
  pdl_wait();
  const int r = blockIdx.x / csz, rank = blockIdx.x - r * csz;
  const float* row = logits + (int64_t)r * ld;
  float bv = -INFINITY;
  int bi = -1;
  const int v4 = V >> 2;
  const int per = (v4 + csz - 1) / csz;
  const int j1 = min(v4, (rank + 1) * per);
  const float4* row4 = reinterpret_cast<const float4*>(row);
#pragma unroll 4
  for (int j = rank * per + threadIdx.x; j < j1; j += blockDim.x) {
    const float4 f = __ldg(row4 + j);
    const int base = j << 2;
    argmax_combine(bv, bi, f.x, base);
    argmax_combine(bv, bi, f.y, base + 1);
    argmax_combine(bv, bi, f.z, base + 2);
    argmax_combine(bv, bi, f.w, base + 3);
  }
  if (rank == csz - 1)
    for (int j = (v4 << 2) + threadIdx.x; j < V; j += blockDim.x) argmax_combine(bv, bi, __ldg(row + j), j);
  argmax_finish_row(bv, bi, r, rank, csz, idx, val, index_offset);
}

extern "C" int nvc_argmax_last(nvc_ctx* ctx, const float* logits, int ld, int32_t* idx, float* val,
                               int B, int V, int index_offset, void* stream) {
  NVC_REQUIRE(ctx, ctx && logits && idx, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, B >= 0 && V > 0 && ld >= V, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, ld % 4 == 0, NVC_ESHAPE, "ld must be a multiple of 4");
  if (B == 0) return NVC_OK;
  const int csz = argmax_cluster_width(ctx, B, V);
  NVC_CUDA_CHECK(ctx, nvc_launch_cluster(ctx, argmax_last_kernel, dim3(B * csz), dim3(1024), 0, nvc_stream(stream), csz,
                                         logits, ld, idx, val, V, index_offset, csz));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// ------------------------------------------------------------------------------------
// K8s temperature sampling on device.  Replaces Qwen3ModelRunner::sample_token
// (llm_engine.rs:97-133: softmax((l - max) / max(T, 1e-6)) then WeightedIndex with an UNSEEDED
// rand::rng(), i.e. not reproducible) by the Gumbel-max form of the same categorical draw:
//   token = argmax_i ( l_i / T + g_i ),  g_i = -log(-log(u_i)),  u_i uniform in (0, 1)
// with counter-based noise: Philox4x32-10, key = seed, counter = (i / 4, stream_lo, stream_hi, 0),
// word i % 4, where i is the GLOBAL vocab index and stream identifies (sequence, step).  The draw
// therefore does not depend on batch composition or on the tensor-parallel vocab split, and the
// (value, index) pair per shard merges exactly like the greedy argmax.
// ------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

__device__ __forceinline__ float gumbel_from_bits(uint32_t x) {
  // u = (2 * (x >> 9) + 1) * 2^-24: an odd 24-bit integer over 2^24, exact in fp32, in [2^-24, 1 - 2^-24]
  // (a 24-bit draw + half a step would round to 1.0f at the top and make g = +inf)
  const float u = (float)(((x >> 9) << 1) | 1u) * 5.9604644775390625e-08f;
  return -logf(-logf(u));
}

// Screening: almost every element loses by a wide margin, and the exact score costs two logf + one IEEE division.  After
// its first (exact) float4 every thread knows the CTA's best score so far (`s_thr`, kept current with shared-memory
// atomics) and evaluates an UPPER BOUND of each later score with the fast intrinsics: x * (1/T) is within 2 ulp of x / T
// (margin |x| * 1e-6), and for u <= 1 - 2^-10 the two MUFU.LG2 give the Gumbel term to 2e-4 (margin 2e-3); draws nearer to
// one than that (1 in 1024) always take the exact path.  Only a float4 with an element whose bound reaches the threshold
// is evaluated exactly, so the result is the one the unscreened kernel gives, bit for bit (GPU test).
__device__ __forceinline__ int float_order_key(float x) {
  const int b = __float_as_int(x);
  return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float float_from_order_key(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

__device__ __forceinline__ bool gumbel_may_reach(float l, float rT, uint32_t bits, float thr) {
  const float u = (float)(((bits >> 9) << 1) | 1u) * 5.9604644775390625e-08f;
  const float inner = -0.69314718f * __log2f(u);
  const float g = -0.69314718f * __log2f(inner);
  const float x = l * rT;
  const float ub = x + g + 2e-3f + fabsf(x) * 1e-6f;
  return (bits >> 22) == 0x3ffu || !(ub < thr);           // NaN / inf in the bound: not "below", so exact
}

__global__ void __launch_bounds__(1024)
sample_gumbel_kernel(const float* __restrict__ logits, int ld, const float* __restrict__ temperatures,
                     const uint64_t* __restrict__ streams, uint64_t seed, int32_t* __restrict__ idx,
                     float* __restrict__ val, int V, int index_offset, int csz, int screen) {
  pdl_trigger();
  pdl_wait();
  __shared__ int s_thr;
  const int b = blockIdx.x / csz, rank = blockIdx.x - b * csz;     // a row is shared by a cluster, as in the argmax
  const float* row = logits + (int64_t)b * ld;
  const float T = fmaxf(__ldg(temperatures + b), 1e-6f);          // llm_engine.rs:104
  const float rT = 1.0f / T;
  const uint64_t stream = __ldg(streams + b);
  const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  const uint32_t s_lo = (uint32_t)stream, s_hi = (uint32_t)(stream >> 32);
  float bv = -INFINITY;
  int bi = -1;
  if (threadIdx.x == 0) s_thr = float_order_key(-INFINITY);
  __syncthreads();
  // index_offset % 4 == 0 and V % 4 == 0: one Philox block covers the float4 at local index 4j
  const int v4 = V >> 2;
  const int per = (v4 + csz - 1) / csz;
  const int j1 = min(v4, (rank + 1) * per);
  const int jbeg = rank * per;
  const float4* row4 = reinterpret_cast<const float4*>(row);
  for (int j0 = jbeg; j0 < j1; j0 += blockDim.x) {                 // block-uniform trip count (the barrier below)
    const int j = j0 + threadIdx.x;
    const bool first = j0 == jbeg;
    if (j < j1) {
      const float4 f = __ldg(row4 + j);
      const uint4 n = philox4x32_10(make_uint4((uint32_t)((index_offset >> 2) + j), s_lo, s_hi, 0u), key);
      bool exact = first || !screen;
      if (!exact) {
        const float thr = float_from_order_key(*reinterpret_cast<volatile int*>(&s_thr));
        exact = gumbel_may_reach(f.x, rT, n.x, thr) | gumbel_may_reach(f.y, rT, n.y, thr) |
                gumbel_may_reach(f.z, rT, n.z, thr) | gumbel_may_reach(f.w, rT, n.w, thr);
      }
      if (exact) {
        const int base = j << 2;
        const float before = bv;
        argmax_combine(bv, bi, __fdiv_rn(f.x, T) + gumbel_from_bits(n.x), base);
        argmax_combine(bv, bi, __fdiv_rn(f.y, T) + gumbel_from_bits(n.y), base + 1);
        argmax_combine(bv, bi, __fdiv_rn(f.z, T) + gumbel_from_bits(n.z), base + 2);
        argmax_combine(bv, bi, __fdiv_rn(f.w, T) + gumbel_from_bits(n.w), base + 3);
        if (!first && bv > before) atomicMax(&s_thr, float_order_key(bv));
      }
    }
    if (first && screen) {           // the CTA's best over everybody's first float4 becomes the first threshold
      const float wm = warp_max(bv);
      if ((threadIdx.x & 31) == 0) atomicMax(&s_thr, float_order_key(wm));
      __syncthreads();
    }
  }
  argmax_finish_row(bv, bi, b, rank, csz, idx, val, index_offset);
}

extern "C" int nvc_sample_gumbel(nvc_ctx* ctx, const float* logits, int ld, const float* temperatures,
                                 const uint64_t* streams, uint64_t seed, int32_t* idx, float* val, int B, int V,
                                 int index_offset, void* stream) {
  NVC_REQUIRE(ctx, ctx && logits && temperatures && streams && idx, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, B >= 0 && V > 0 && ld >= V, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, ld % 4 == 0 && V % 4 == 0 && index_offset % 4 == 0 && index_offset >= 0, NVC_ESHAPE,
              "ld, V and index_offset must be multiples of 4");
  if (B == 0) return NVC_OK;
  const int csz = argmax_cluster_width(ctx, B, V);
  NVC_CUDA_CHECK(ctx, nvc_launch_cluster(ctx, sample_gumbel_kernel, dim3(B * csz), dim3(1024), 0, nvc_stream(stream), csz,
                                         logits, ld, temperatures, streams, seed, idx, val, V, index_offset, csz,
                                         g_nvc_debug[6] == 1 ? 0 : 1));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// (max, idx) of a vocab shard as one 8-byte record per sequence (TP argmax exchange)
__global__ void pack_argmax_pairs_kernel(const float* __restrict__ val, const int32_t* __restrict__ idx,
                                         float2* __restrict__ pairs, int B) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) pairs[b] = make_float2(val[b], __int_as_float(idx[b]));
}

extern "C" int nvc_pack_argmax_pairs(nvc_ctx* ctx, const float* val, const int32_t* idx, float* pairs, int B,
                                     void* stream) {
  NVC_REQUIRE(ctx, ctx && val && idx && pairs && B >= 0, NVC_EINVAL, "bad arguments");
  if (B == 0) return NVC_OK;
  NVC_CUDA_CHECK(ctx, nvc_launch(ctx, pack_argmax_pairs_kernel, dim3((B + 127) / 128), dim3(128), 0, nvc_stream(stream),
                                 val, idx, reinterpret_cast<float2*>(pairs), B));
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}
