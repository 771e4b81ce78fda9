// K3: C[M,N] = A[M,K] * W[N,K]^T, bf16 x bf16 -> fp32 accumulate in tensor memory.
// Replaces candle_nn::Linear::forward under every *ParallelLinear::forward
// (src/layers/linear.rs:33-38,72-77,99-104,171-175,184-198) and compute_logits
// (src/models/qwen3.rs:542-550).
//
// Structure (one 128 x BN output tile per CTA, optional split-K over gridDim.z):
//   warp 0   : TMA producer  - cp.async.bulk.tensor 2D, 128B-swizzled [128 x 64] A tile and
//              [BN x 64] W tile per stage, mbarrier expect_tx ring of kStages stages
//   warp 1   : allocates TMEM, one elected lane issues tcgen05.mma (M=128, N=BN, K=16) x4 per
//              stage, tcgen05.commit frees the stage / publishes the accumulator
//   warps 2-5: epilogue - tcgen05.ld 32 lanes x 16 columns, convert, 16-byte stores
// Split-K: each split writes its fp32 partial tile to a workspace; the CTA that takes the last
// ticket of a tile sums the partials in split order (deterministic) and writes the output.
// Rows >= M and columns >= N of edge tiles are zero-filled by TMA and masked at the store.
#include "common.cuh"
#include "ptx.cuh"
#include "internal.h"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kThreads = 192;
constexpr int kEpilogueThreads = 128;

template <int BN>
struct Cfg {
  static constexpr int kBytesA = BM * BK * 2;
  static constexpr int kBytesB = BN * BK * 2;
  static constexpr int kStageBytes = kBytesA + kBytesB;
  static constexpr int kStages = (BN >= 256) ? 4 : (BN >= 128) ? 6 : 8;
  static constexpr int kTmemCols = BN < 32 ? 32 : BN;   // BN in {32,64,128,256}: already a power of two
  // tiles + (2*stages + 1) mbarriers + tmem slot + ticket, plus 1024 for manual alignment
  static constexpr int kSmemBytes = kStages * kStageBytes + (2 * kStages + 1) * 8 + 16 + 1024;
};

template <int BN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         void* __restrict__ C, int ldc, int out_f32, int M, int N, int k_blocks_total,
                         int k_blocks_per_split, float* __restrict__ ws, int* __restrict__ counters) {
  using cfg = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t bar_base = smem_base + cfg::kStages * cfg::kStageBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (cfg::kStages + s); };
  const uint32_t tmem_full_bar = bar_base + 8u * (2 * cfg::kStages);
  const uint32_t tmem_slot = tmem_full_bar + 8u;
  volatile uint32_t* tmem_slot_gen =
      reinterpret_cast<volatile uint32_t*>(smem_gen + cfg::kStages * cfg::kStageBytes + (2 * cfg::kStages + 1) * 8);
  volatile int* ticket_gen = reinterpret_cast<volatile int*>(const_cast<uint32_t*>(tmem_slot_gen) + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x, m_tile = blockIdx.y, split = blockIdx.z;
  const int num_splits = gridDim.z;
  const int kb_begin = split * k_blocks_per_split;
  const int kb_end = min(k_blocks_total, kb_begin + k_blocks_per_split);
  const int num_kb = kb_end - kb_begin;

  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tma_a);
    ptx::tma_prefetch_desc(&tma_b);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < cfg::kStages; ++s) {
        ptx::mbar_init(full_bar(s), 1);
        ptx::mbar_init(empty_bar(s), 1);
      }
      ptx::mbar_init(tmem_full_bar, 1);
      ptx::fence_barrier_init();
    }
    __syncwarp();
    ptx::tmem_alloc(tmem_slot, cfg::kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % cfg::kStages;
        const uint32_t ph = (kb / cfg::kStages) & 1;
        ptx::mbar_wait(empty_bar(s), ph ^ 1u);
        ptx::mbar_arrive_expect_tx(full_bar(s), cfg::kStageBytes);
        const uint32_t sa = smem_base + s * cfg::kStageBytes;
        const uint32_t sb = sa + cfg::kBytesA;
        const int kcoord = (kb_begin + kb) * BK;
        ptx::tma_load_2d(sa, &tma_a, full_bar(s), kcoord, m_tile * BM);
        ptx::tma_load_2d(sb, &tma_b, full_bar(s), kcoord, n_tile * BN);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(BM, BN, 0, 0);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % cfg::kStages;
        const uint32_t ph = (kb / cfg::kStages) & 1;
        ptx::mbar_wait(full_bar(s), ph);
        ptx::tcgen05_fence_after();
        const uint32_t sa = smem_base + s * cfg::kStageBytes;
        const uint32_t sb = sa + cfg::kBytesA;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          // advance 16 bf16 = 32 bytes along K inside the 128-byte swizzle span
          const uint64_t da = ptx::umma_desc_sw128(sa + k * 32, 16, 1024);
          const uint64_t db = ptx::umma_desc_sw128(sb + k * 32, 16, 1024);
          ptx::umma_bf16_ss(tmem_base, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        ptx::umma_commit(empty_bar(s));   // stage reusable once these MMAs have read it
      }
      ptx::umma_commit(tmem_full_bar);    // accumulator complete
    }
  } else {
    // ------------------------------------------------------------ epilogue (warps 2..5)
    const int q = warp & 3;                 // TMEM lane quadrant this warp may access
    const int r = q * 32 + lane;            // row inside the tile
    const int grow = m_tile * BM + r;
    const bool row_ok = grow < M;
    ptx::mbar_wait(tmem_full_bar, 0);
    ptx::tcgen05_fence_after();
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const int tile_id = m_tile * gridDim.x + n_tile;

    if (num_splits == 1) {
#pragma unroll 1
      for (int c = 0; c < BN / 16; ++c) {
        uint32_t v[16];
        ptx::tmem_ld_32x32b_x16(taddr + c * 16, v);
        ptx::tmem_ld_wait();
        const int n0 = n_tile * BN + c * 16;
        if (row_ok && n0 < N) {
          if (out_f32) {
            float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(C) + (int64_t)grow * ldc + n0);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              dst[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                   __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
          } else {
            uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<nvc_bf16*>(C) + (int64_t)grow * ldc + n0);
            uint4 o0, o1;
            o0.x = pack_bf16(__uint_as_float(v[0]), __uint_as_float(v[1]));
            o0.y = pack_bf16(__uint_as_float(v[2]), __uint_as_float(v[3]));
            o0.z = pack_bf16(__uint_as_float(v[4]), __uint_as_float(v[5]));
            o0.w = pack_bf16(__uint_as_float(v[6]), __uint_as_float(v[7]));
            o1.x = pack_bf16(__uint_as_float(v[8]), __uint_as_float(v[9]));
            o1.y = pack_bf16(__uint_as_float(v[10]), __uint_as_float(v[11]));
            o1.z = pack_bf16(__uint_as_float(v[12]), __uint_as_float(v[13]));
            o1.w = pack_bf16(__uint_as_float(v[14]), __uint_as_float(v[15]));
            dst[0] = o0;
            dst[1] = o1;
          }
        }
      }
    } else {
      // ---- split-K: publish the partial, last arriver reduces in split order
      float* my = ws + ((int64_t)tile_id * num_splits + split) * (BM * BN) + (int64_t)r * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 16; ++c) {
        uint32_t v[16];
        ptx::tmem_ld_32x32b_x16(taddr + c * 16, v);
        ptx::tmem_ld_wait();
        if (row_ok) {
          float4* dst = reinterpret_cast<float4*>(my + c * 16);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            __stcg(dst + j, make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                        __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3])));
        }
      }
      __threadfence();
      ptx::named_bar_sync(1, kEpilogueThreads);
      if (threadIdx.x == 64) *ticket_gen = atomicAdd(counters + tile_id, 1);
      ptx::named_bar_sync(1, kEpilogueThreads);
      if (*ticket_gen == num_splits - 1) {
        __threadfence();
        const float* base = ws + (int64_t)tile_id * num_splits * (BM * BN) + (int64_t)r * BN;
        if (row_ok) {
#pragma unroll 1
          for (int c = 0; c < BN / 16; ++c) {
            const int n0 = n_tile * BN + c * 16;
            if (n0 >= N) break;
            float acc[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] = 0.f;
            for (int z = 0; z < num_splits; ++z) {
              const float4* src = reinterpret_cast<const float4*>(base + (int64_t)z * (BM * BN) + c * 16);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float4 f = __ldcg(src + j);
                acc[4 * j] += f.x; acc[4 * j + 1] += f.y; acc[4 * j + 2] += f.z; acc[4 * j + 3] += f.w;
              }
            }
            if (out_f32) {
              float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(C) + (int64_t)grow * ldc + n0);
#pragma unroll
              for (int j = 0; j < 4; ++j) dst[j] = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
            } else {
              uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<nvc_bf16*>(C) + (int64_t)grow * ldc + n0);
              dst[0] = pack8(acc);
              dst[1] = pack8(acc + 8);
            }
          }
        }
        if (threadIdx.x == 64) counters[tile_id] = 0;   // re-arm for the next launch
      }
    }
    ptx::tcgen05_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    ptx::tcgen05_fence_after();
    ptx::tmem_dealloc(tmem_base, cfg::kTmemCols);
  }
}

int make_tma_2d(nvc_ctx* ctx, CUtensorMap* map, const void* base, uint64_t inner, uint64_t rows,
                uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_rows) {
  cuuint64_t gdim[2] = {inner, rows};
  cuuint64_t gstr[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_inner, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = ctx->encode_tiled(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box,
                                 estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    nvc_set_error(ctx, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    return NVC_EINVAL;
  }
  return NVC_OK;
}

template <int BN>
int launch(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc, int out_f32, int M,
           int N, int K, int splits, cudaStream_t stream) {
  using cfg = Cfg<BN>;
  CUtensorMap ta, tb;
  int rc = make_tma_2d(ctx, &ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, BK, BM);
  if (rc) return rc;
  rc = make_tma_2d(ctx, &tb, W, (uint64_t)K, (uint64_t)N, (uint64_t)K * 2, BK, BN);
  if (rc) return rc;
  const int kb_total = K / BK;
  const int kb_per = (kb_total + splits - 1) / splits;
  const int eff_splits = (kb_total + kb_per - 1) / kb_per;   // no empty splits
  dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM, eff_splits);
  gemm_bf16_tcgen05_kernel<BN><<<grid, kThreads, cfg::kSmemBytes, stream>>>(
      ta, tb, C, ldc, out_f32, M, N, kb_total, kb_per, ctx->gemm_ws, ctx->gemm_counters);
  NVC_CUDA_CHECK(ctx, cudaGetLastError());
  ctx->launches++;
  return NVC_OK;
}

// Tile / split-K choice: minimise waves * per-CTA time under a crude cycle model
// (MMA floor 2*BN cycles per 64-wide k-block, ~48 B/clk/SM of L2->SM bandwidth when the
// chip is full, a fixed prologue/epilogue cost, and the split-K round trip through L2).
void choose_config(const nvc_ctx* ctx, int M, int N, int K, int* bn_out, int* splits_out) {
  const int kb = K / BK;
  const int m_tiles = (M + BM - 1) / BM;
  const int rows_in_tile = M < BM ? M : BM;
  const int bns[4] = {256, 128, 64, 32};
  double best = 1e30;
  int best_bn = 128, best_s = 1;
  for (int bi = 0; bi < 4; ++bi) {
    const int bn = bns[bi];
    if (bn > 32 && bn / 2 >= N) continue;            // tile mostly empty
    const int tiles = m_tiles * ((N + bn - 1) / bn);
    for (int s = 1; s <= 16 && s <= kb; ++s) {
      const int kb_per = (kb + s - 1) / s;
      const int eff_s = (kb + kb_per - 1) / kb_per;
      if (eff_s != s) continue;
      if (s > 1 && kb_per < 2) continue;
      if (s > 1 && (size_t)tiles * s * BM * bn * 4 > ctx->gemm_ws_bytes) continue;
      if (s > 1 && tiles > ctx->gemm_num_counters) continue;
      const double ctas = (double)tiles * s;
      const double waves = ceil(ctas / ctx->num_sms);
      const double mma = 2.0 * bn;
      const double load = (rows_in_tile * 128.0 + bn * 128.0) / 48.0;
      double per = kb_per * (mma > load ? mma : load) + 2500.0 + bn * 6.0;
      if (s > 1) per += 1500.0 + (double)rows_in_tile * bn * 4.0 * (1.0 + 1.0) / 48.0;
      const double t = waves * per;
      if (t < best) { best = t; best_bn = bn; best_s = s; }
    }
  }
  *bn_out = best_bn;
  *splits_out = best_s;
}

}  // namespace

extern "C" int nvc_gemm_init(nvc_ctx* ctx) {
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<256>::kSmemBytes));
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<128>::kSmemBytes));
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<64>::kSmemBytes));
  NVC_CUDA_CHECK(ctx, cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<32>::kSmemBytes));
  return NVC_OK;
}

extern "C" int nvc_gemm_bf16_ex(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc,
                                int out_dtype, int M, int Nout, int K, int force_bn, int force_splits,
                                void* stream) {
  NVC_REQUIRE(ctx, ctx && A && W && C, NVC_EINVAL, "null pointer");
  NVC_REQUIRE(ctx, M >= 0 && Nout > 0 && K > 0, NVC_EINVAL, "bad sizes");
  NVC_REQUIRE(ctx, K % 64 == 0, NVC_ESHAPE, "K must be a multiple of 64");
  NVC_REQUIRE(ctx, Nout % 16 == 0, NVC_ESHAPE, "Nout must be a multiple of 16");
  NVC_REQUIRE(ctx, lda % 8 == 0 && lda >= K, NVC_ESHAPE, "lda must be >= K and a multiple of 8");
  NVC_REQUIRE(ctx, ldc >= Nout && ldc % 8 == 0, NVC_ESHAPE, "ldc must be >= Nout and a multiple of 8");
  NVC_REQUIRE(ctx, out_dtype == NVC_DT_BF16 || out_dtype == NVC_DT_F32, NVC_EINVAL, "bad out_dtype");
  NVC_REQUIRE(ctx, ((uintptr_t)A % 16) == 0 && ((uintptr_t)W % 16) == 0 && ((uintptr_t)C % 16) == 0, NVC_EINVAL,
              "pointers must be 16-byte aligned");
  NVC_REQUIRE(ctx, ctx->encode_tiled != nullptr, NVC_EDEVICE, "cuTensorMapEncodeTiled unavailable");
  if (M == 0) return NVC_OK;
  int bn, splits;
  choose_config(ctx, M, Nout, K, &bn, &splits);
  if (force_bn) bn = force_bn;
  if (force_splits) splits = force_splits;
  if (splits > K / BK) splits = K / BK;
  const int tiles = ((M + BM - 1) / BM) * ((Nout + bn - 1) / bn);
  if (splits > 1) {
    NVC_REQUIRE(ctx, (size_t)tiles * splits * BM * bn * 4 <= ctx->gemm_ws_bytes && tiles <= ctx->gemm_num_counters,
                NVC_ENOMEM, "split-K workspace too small");
  }
  cudaStream_t st = nvc_stream(stream);
  const int f32 = out_dtype == NVC_DT_F32;
  switch (bn) {
    case 256: return launch<256>(ctx, A, lda, W, C, ldc, f32, M, Nout, K, splits, st);
    case 128: return launch<128>(ctx, A, lda, W, C, ldc, f32, M, Nout, K, splits, st);
    case 64: return launch<64>(ctx, A, lda, W, C, ldc, f32, M, Nout, K, splits, st);
    case 32: return launch<32>(ctx, A, lda, W, C, ldc, f32, M, Nout, K, splits, st);
    default: NVC_REQUIRE(ctx, false, NVC_EINVAL, "unsupported tile width");
  }
  return NVC_OK;
}

extern "C" int nvc_gemm_bf16(nvc_ctx* ctx, const nvc_bf16* A, int lda, const nvc_bf16* W, void* C, int ldc,
                             int out_dtype, int M, int Nout, int K, void* stream) {
  return nvc_gemm_bf16_ex(ctx, A, lda, W, C, ldc, out_dtype, M, Nout, K, 0, 0, stream);
}
