// K5: paged decode attention over block_tables with GQA head grouping (N2 + the decode half
// of src/models/qwen3.rs:236-277; the reference recomputes dense attention over the whole
// sequence each step and has no KV cache, src/models/qwen3.rs:402).
//
// HBM-bound: every K and V row of a (sequence, kv head) is read exactly once and shared by
// the G = nh/kv query heads of the group.  Algorithmic bytes per (sequence, layer):
//   context_len * 2 * kv * 128 * 2  (K+V)  +  2 * nh * 128 * 2  (q in, out).
//
// The kernel is paged_decode_mma.cuh (mma.sync over the GQA group, cp.async rings); this file holds
// the split-KV combine kernel, the grid / split / warps-per-CTA choice and the C ABI.  With KV
// splits > 1 the partial (m, l, o) states go to a workspace and the combine kernel merges them in
// split order.
#include "common.cuh"
#include "internal.h"
#include "paged_decode_mma.cuh"

#include <cstdlib>

extern int g_nvc_debug[8];

// Scratch of nvc_paged_decode_attn_static (the entry point that allocates nothing): the KV-split partials (fp32 O rows +
// (max, sum) pairs) live in this file-scope array instead of a caller workspace.  4 MiB per GPU that loads the library,
// held until exit.  In HOST code `g_pd_static_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress (nvc_paged_decode_init, once per device, that device current) and is what the kernels receive.
#define NVC_PD_STATIC_SCRATCH_BYTES (4u << 20)
static __device__ __align__(256) unsigned char g_pd_static_scratch[NVC_PD_STATIC_SCRATCH_BYTES];
static void* g_pd_static_scratch_addr[64] = {nullptr};   // device address of the array, per device ordinal

namespace {

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
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

int max_splits_for_batch(int B, int num_sms) {
  int s = (4 * num_sms + B - 1) / B;
  return s < 1 ? 1 : (s > 64 ? 64 : s);
}

template <int kWarps>
cudaError_t launch_pd(nvc_ctx* ctx, dim3 grid, cudaStream_t st, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                      const nvc_bf16* v_cache, const int32_t* block_tables, int max_blocks, const int32_t* context_lens,
                      int nh, int kv, int group, int G, int block_size, float scale_log2e, nvc_bf16* out, float* part_o,
                      float* part_ml, int splits) {
  return nvc_launch(ctx, pdmma::paged_decode_mma_kernel<kWarps>, grid, dim3(kWarps * 32),
                    (size_t)pdmma::smem_bytes<kWarps>(), st, q, ldq, k_cache, v_cache, block_tables, max_blocks,
                    context_lens, nh, kv, group, G, block_size, scale_log2e, out, part_o, part_ml, splits);
}

int paged_decode_impl(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache, const nvc_bf16* v_cache,
                      const int32_t* block_tables, int max_blocks, int max_blocks_hint, const int32_t* context_lens, int B,
                      int nh, int kv, int hd, int block_size, float scale, nvc_bf16* out, void* workspace,
                      size_t workspace_bytes, void* stream) {
  NVC_REQUIRE(ctx, ctx && q && k_cache && v_cache && block_tables && context_lens && out, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, B >= 0 && nh > 0 && kv > 0 && max_blocks > 0 && max_blocks_hint > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, hd == 128, NVC_ESHAPE, "head_dim must be 128");
  NVC_REQUIRE(ctx, nh % kv == 0, NVC_ESHAPE, "nh must be a multiple of kv");
  NVC_REQUIRE(ctx, block_size > 0 && block_size % 16 == 0, NVC_ESHAPE, "block_size must be a multiple of 16");
  NVC_REQUIRE(ctx, ldq % 8 == 0 && ldq >= nh * hd, NVC_ESHAPE, "ldq must be >= nh*hd and a multiple of 8");
  if (B == 0) return NVC_OK;
  const int group = nh / kv;
  const int G = group < 16 ? group : 16;
  const int chunks = (group + G - 1) / G;
  const int base = B * kv * chunks;                    // CTAs without KV splits
  // Warps per CTA.  4 warps / 96 KB: 2 CTAs per SM; 2 warps / 48 KB: 4 CTAs per SM.  A grid beyond the 2-per-SM
  // residency runs as 2-warp CTAs: Qwen3-0.6B at batch 64 (512 CTAs on 148 SMs) is ONE resident wave instead of
  // 1.7 waves of 4-warp CTAs (28.5 -> 26.6 us), 384 CTAs 26.5 -> 21.1 us, 1024 CTAs 54.7 -> 49.8 us, 2048 CTAs equal;
  // below it the 4-warp CTAs (with KV splits) win (profiles/r02_pd_warps_split_sweep.txt).  NVC_PD_WARPS / nvc_debug_set(5, w) force it (microbench, tests).
  int warps = base > 2 * ctx->num_sms ? 2 : 4;
  if (const char* e = getenv("NVC_PD_WARPS")) {
    const int f = atoi(e);
    if (f == 2 || f == 4) warps = f;
  }
  if (g_nvc_debug[5] == 2 || g_nvc_debug[5] == 4) warps = g_nvc_debug[5];
  const int cta_tok = warps * pdmma::kTileTok;
  const int max_tiles = (int)(((int64_t)max_blocks_hint * block_size + cta_tok - 1) / cta_tok);
  // KV splits: only when the (sequence, kv head) grid leaves most of the resident slots empty, and never
  // beyond them (tools/decode_microbench.py --splits, profiles/r02_pd_split_sweep.txt: 64 CTAs -> 4 splits 10.3 ->
  // 7.6 us, 128 -> 2, 256 and up -> 1; rounding the split count UP to cover the slots cost 3 - 5 us on the
  // TP-sharded shapes through the second wave it created)
  int splits = 2 * ctx->num_sms / base;
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
  if (splits > 1) {
    const size_t need = (size_t)B * nh * splits * (hd + 2) * sizeof(float);
    NVC_REQUIRE(ctx, workspace != nullptr && workspace_bytes >= need, NVC_ENOMEM, "paged decode workspace too small");
    part_o = reinterpret_cast<float*>(workspace);
    part_ml = part_o + (size_t)B * nh * splits * hd;
  }
  dim3 grid(kv * chunks, B, splits);
#define NVC_PD_ARGS ctx, grid, st, q, ldq, k_cache, v_cache, block_tables, max_blocks, context_lens, nh, kv, group, G, \
                    block_size, scale_log2e, out, part_o, part_ml, splits
  cudaError_t ce;
  ce = warps == 2 ? launch_pd<2>(NVC_PD_ARGS) : launch_pd<4>(NVC_PD_ARGS);
#undef NVC_PD_ARGS
  NVC_CUDA_CHECK(ctx, ce);
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  if (splits > 1) {
    NVC_CUDA_CHECK(ctx, nvc_launch(ctx, paged_decode_combine_kernel, dim3(nh, B), dim3(128), 0, st, part_o, part_ml, out,
                                   nh, splits));
    NVC_CUDA_CHECK(ctx, cudaGetLastError());
    ctx->launches++;
  }
  return NVC_OK;
}

}  // namespace

extern "C" int nvc_paged_decode_init(nvc_ctx* ctx) {
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(pdmma::paged_decode_mma_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           pdmma::smem_bytes<4>()));
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(pdmma::paged_decode_mma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           pdmma::smem_bytes<2>()));
  // device address of the static scratch (ctx creation runs with ctx->device current)
  if (ctx->device >= 0 && ctx->device < 64 && g_pd_static_scratch_addr[ctx->device] == nullptr) {
    void* p = nullptr;
    NVC_CUDA_CHECK(ctx, cudaGetSymbolAddress(&p, g_pd_static_scratch));
    g_pd_static_scratch_addr[ctx->device] = p;
  }
  return NVC_OK;
}

extern "C" size_t nvc_paged_decode_workspace_bytes(int B, int nh, int hd, int num_sms) {
  if (B <= 0 || nh <= 0 || hd <= 0) return 0;
  const int sms = num_sms > 0 ? num_sms : 148;
  const size_t S = (size_t)max_splits_for_batch(B, sms);
  return (size_t)B * nh * S * (hd + 2) * sizeof(float);
}

extern "C" int nvc_paged_decode_attn_ex(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                                        const nvc_bf16* v_cache, const int32_t* block_tables, int max_blocks,
                                        int max_blocks_hint, const int32_t* context_lens, int B, int nh, int kv,
                                        int hd, int block_size, float scale, nvc_bf16* out, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  return paged_decode_impl(ctx, q, ldq, k_cache, v_cache, block_tables, max_blocks, max_blocks_hint, context_lens, B, nh, kv,
                           hd, block_size, scale, out, workspace, workspace_bytes, stream);
}

extern "C" int nvc_paged_decode_attn(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                                     const nvc_bf16* v_cache, const int32_t* block_tables, int max_blocks,
                                     const int32_t* context_lens, int B, int nh, int kv, int hd, int block_size,
                                     float scale, nvc_bf16* out, void* workspace, size_t workspace_bytes,
                                     void* stream) {
  return nvc_paged_decode_attn_ex(ctx, q, ldq, k_cache, v_cache, block_tables, max_blocks, max_blocks, context_lens, B,
                                  nh, kv, hd, block_size, scale, out, workspace, workspace_bytes, stream);
}

// K5 without any allocation (no counterpart in the reference): the same
// kernels as nvc_paged_decode_attn, the KV-split partials in the library's static device array (4 MiB) instead of a caller
// workspace.  Calls no cudaMalloc / cudaFree; every call on a device shares the one array, so calls that overlap on the GPU
// (two streams) race - serialise them.  Nothing has to start at zero: every (sequence, head, split) partial the combine
// kernel reads is written by the attention kernel of the same call.
extern "C" size_t nvc_paged_decode_static_scratch_bytes(void) { return NVC_PD_STATIC_SCRATCH_BYTES; }

extern "C" int nvc_paged_decode_attn_static(nvc_ctx* ctx, const nvc_bf16* q, int ldq, const nvc_bf16* k_cache,
                                            const nvc_bf16* v_cache, const int32_t* block_tables, int max_blocks,
                                            const int32_t* context_lens, int B, int nh, int kv, int hd, int block_size,
                                            float scale, nvc_bf16* out, void* stream) {
  NVC_REQUIRE(ctx, ctx != nullptr, NVC_EINVAL, "null context");
  void* scratch = (ctx->device >= 0 && ctx->device < 64) ? g_pd_static_scratch_addr[ctx->device] : nullptr;
  NVC_REQUIRE(ctx, scratch != nullptr, NVC_EDEVICE, "static scratch address was not resolved for this device");
  const int rc = paged_decode_impl(ctx, q, ldq, k_cache, v_cache, block_tables, max_blocks, max_blocks, context_lens, B, nh,
                                   kv, hd, block_size, scale, out, scratch, NVC_PD_STATIC_SCRATCH_BYTES, stream);
  if (rc == NVC_ENOMEM)      // refused before any launch: the KV-split partials of this shape exceed the static array
    nvc_set_error(ctx, "nvc_paged_decode_attn_static: the KV-split partials of B=" + std::to_string(B) + ", nh=" +
                           std::to_string(nh) + " exceed the " + std::to_string(NVC_PD_STATIC_SCRATCH_BYTES) +
                           "-byte static scratch; use nvc_paged_decode_attn with a caller workspace");
  return rc;
}
