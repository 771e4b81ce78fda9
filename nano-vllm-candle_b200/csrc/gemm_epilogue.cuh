// Epilogue helpers shared by the 1-CTA and the CTA-pair GEMM: one thread owns one accumulator row
// and emits 16 or 32 consecutive columns of it.
#pragma once
#include "common.cuh"

// NVC_EPI_QK_NORM_ROPE (pair kernel, persistent schedule, 256-wide tiles): the GEMM is the fused QKV projection and its
// epilogue is K2 - per-head RMSNorm of q / k, neox RoPE at the row's position, q / k / v rows written to C and the k / v rows
// scattered to the paged cache at slot_mapping[row] - so the separate pass over the QKV buffer (120-160 us for 600 MB at
// 32 768 tokens, 10-13 % of a 0.6B prefill step) disappears.  A 256-column tile is two heads of one row; the warp parks a head
// (32 rows x 256 B) in its staging buffer and works on it with the lane mapping and arithmetic of qk_norm_rope_store_kernel.
// NVC_EPI_TRANSPOSE (pair kernel, persistent schedule): the operands are swapped by the host - the kernel's "A" rows are
// weight rows, its "W" rows are tokens - and the epilogue stores its accumulator row r as COLUMN r of C: lm_head at decode
// batch sizes, where 64 tokens would fill a quarter of a 256-row MMA and only one CTA of the pair would hold valid rows.
enum { NVC_EPI_NONE = 0, NVC_EPI_SILU_MUL = 1, NVC_EPI_TRANSPOSE = 2, NVC_EPI_QK_NORM_ROPE = 3 };

struct QkEpi {
  const uint4* qw = nullptr;         // q_norm weight [128] bf16
  const uint4* kw = nullptr;         // k_norm weight
  const float4* cos_sin = nullptr;   // [max_pos][128] f32: cos[64], sin[64]
  const int32_t* positions = nullptr;
  const int32_t* slot_mapping = nullptr;   // may be NULL: no cache scatter
  uint4* k_cache = nullptr;
  uint4* v_cache = nullptr;
  float eps = 0.f;
  int nh = 0, kv = 0;                // local q / kv heads: C columns [0, nh) q heads, [nh, nh + kv) k heads, then v heads
};

// 16 accumulator columns of one row -> C (bf16 or f32), 16-byte stores
__device__ __forceinline__ void store16(void* C, int out_f32, int64_t off, const float* f) {
  if (out_f32) {
    float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(C) + off);
#pragma unroll
    for (int j = 0; j < 4; ++j) dst[j] = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
  } else {
    uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<nvc_bf16*>(C) + off);
    dst[0] = pack8(f);
    dst[1] = pack8(f + 8);
  }
}

// 32 consecutive accumulator columns [n0, n0 + 32) of one row.
//  NVC_EPI_NONE:     C[row, n0 .. n0+32) = f                                   (columns >= N masked per 16)
//  NVC_EPI_SILU_MUL: the weight rows are interleaved in groups of 16 (gate rows [16g,16g+16) then up
//                    rows [16g,16g+16)), so f[0..16) = gate and f[16..32) = up of features n0/2 .. n0/2+16:
//                    C[row, n0/2 .. n0/2+16) = silu(gate) * up   (src/layers/activation.rs:13-18,
//                    silu(x) = x / (1 + exp(-x)) in f32), bf16.  row_off = row * ldc in elements of C.
__device__ __forceinline__ void emit32(void* C, int out_f32, int epi, int64_t row_off, int n0, int N, const float* f) {
  if (epi == NVC_EPI_SILU_MUL) {
    if (n0 < N) {
      float o[16];
#pragma unroll
      // fast reciprocal (MUFU.RCP, 2 ulp): an IEEE division per element makes the epilogue of a K = 1024
      // prefill tile (128 of them per thread) as long as its main loop
      for (int j = 0; j < 16; ++j) o[j] = __fdividef(f[j], 1.0f + __expf(-f[j])) * f[16 + j];
      store16(C, 0, row_off + (n0 >> 1), o);
    }
  } else {
    if (n0 < N) store16(C, out_f32, row_off + n0, f);
    if (n0 + 16 < N) store16(C, out_f32, row_off + n0 + 16, f + 16);
  }
}

// ---------------------------------------------------------------------------------------------
// Coalesced epilogue.  One thread owns one accumulator row, so storing straight from registers
// makes every warp-level store touch 32 different 128-byte lines with 16 bytes each (measured:
// ~1.3k cycles per 32 accumulator columns, an 11k-cycle exposed epilogue per 256-wide tile).
// Instead each warp parks its 32 rows x 128 output bytes in a private 4 KB shared-memory buffer
// (16-byte chunks XOR-swizzled by row, conflict-free both ways) and writes them back with 8 lanes
// per row: every store instruction covers 4 rows x 128 contiguous bytes.
// ---------------------------------------------------------------------------------------------
struct RowStage {
  uint8_t* buf;            // this warp's 4 KB (generic address)
  uint8_t* g_row0;         // global address of (first row of this warp, first output column of the tile)
  int64_t row_stride;      // bytes between output rows
  int rows_valid;          // rows of this warp inside M (0..32)
  int tile_bytes_valid;    // output bytes of the tile inside N (per row)
  int filled = 0;          // bytes per row parked since the last flush
  int flushed = 0;         // bytes per row already written

  __device__ __forceinline__ void put(int lane, uint4 v) {
    *reinterpret_cast<uint4*>(buf + lane * 128 + (((filled >> 4) ^ (lane & 7)) << 4)) = v;
    filled += 16;
  }
  __device__ __forceinline__ void flush(int lane) {
    __syncwarp();
    const int c = lane & 7, sub = lane >> 3;
    const int nbytes = min(filled, tile_bytes_valid - flushed);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int row = i * 4 + sub;
      if (row < rows_valid && (c << 4) < nbytes) {
        const uint4 v = *reinterpret_cast<const uint4*>(buf + row * 128 + ((c ^ (row & 7)) << 4));
        *reinterpret_cast<uint4*>(g_row0 + row * row_stride + flushed + (c << 4)) = v;
      }
    }
    __syncwarp();
    flushed += filled;
    filled = 0;
  }
  // 32 consecutive accumulator columns of this thread's row (same semantics as emit32); flushes every 128 bytes
  __device__ __forceinline__ void emit32(int lane, int out_f32, int epi, const float* f) {
    if (epi == NVC_EPI_SILU_MUL) {
      float o[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) o[j] = __fdividef(f[j], 1.0f + __expf(-f[j])) * f[16 + j];
      put(lane, pack8(o));
      put(lane, pack8(o + 8));
    } else if (out_f32) {
#pragma unroll
      for (int j = 0; j < 8; ++j) put(lane, make_uint4(__float_as_uint(f[4 * j]), __float_as_uint(f[4 * j + 1]),
                                                       __float_as_uint(f[4 * j + 2]), __float_as_uint(f[4 * j + 3])));
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) put(lane, pack8(f + 8 * j));
    }
    if (filled == 128) flush(lane);
  }
  __device__ __forceinline__ void finish(int lane) {
    if (filled) flush(lane);
  }
};

// Output geometry of a tile for RowStage: bytes per accumulator column and the tile's byte window.
__device__ __forceinline__ RowStage make_row_stage(uint8_t* buf, void* C, int out_f32, int epi, int64_t ldc, int row0, int M,
                                                   int n_base, int n_tile_cols, int N) {
  // SiLU-mul: 32 accumulator columns -> 16 bf16; plain: bf16 or f32 per column
  const int esz = out_f32 ? 4 : 2;
  const int col0 = epi == NVC_EPI_SILU_MUL ? (n_base >> 1) : n_base;
  const int n_out = epi == NVC_EPI_SILU_MUL ? (N >> 1) : N;
  const int tile_out = epi == NVC_EPI_SILU_MUL ? (n_tile_cols >> 1) : n_tile_cols;
  RowStage s;
  s.buf = buf;
  s.g_row0 = reinterpret_cast<uint8_t*>(C) + ((int64_t)row0 * ldc + col0) * esz;
  s.row_stride = ldc * esz;
  s.rows_valid = max(0, min(32, M - row0));
  s.tile_bytes_valid = max(0, min(tile_out, n_out - col0)) * esz;
  return s;
}
