// Epilogue helpers shared by the 1-CTA and the CTA-pair GEMM: one thread owns one accumulator row
// and emits 16 or 32 consecutive columns of it.
#pragma once
#include "common.cuh"

// NVC_EPI_LL_ALLREDUCE (tensor parallel, decode): the GEMM is a row-parallel projection and its epilogue IS the all-reduce.
// Every epilogue thread multimem.st's its 16-byte pieces of this rank's partial rows into range `rank` of the current
// low-latency buffer of the NVLS multicast object (one store reaches every rank), then waits for the same pieces of the
// other ranks to land in its OWN copy of the buffer (spinning on the data: the buffers are kept filled with a sentinel no
// partial can contain), sums the tp pieces in rank order (bit-identical on every rank) and stores the bf16 sums to C.
// The NVLink transfer therefore overlaps the GEMM's own tail and no separate all-reduce kernel runs (a kernel boundary
// would wait for the remote stores to drain: pushing from the epilogue but reducing in a second kernel measured no gain,
// profiles/r02_nvls_gemm_push_tp2.jsonl).  Protocol state as in nvls_allreduce.cu (same buffers, same epochs).
enum { NVC_EPI_NONE = 0, NVC_EPI_SILU_MUL = 1, NVC_EPI_LL_ALLREDUCE = 2 };

struct LLReduce {
  uint8_t* mc_ll = nullptr;          // multicast address of LL buffer 0
  uint8_t* uc_ll = nullptr;          // this rank's copy of LL buffer 0
  size_t ll_buf_bytes = 0;           // bytes per LL buffer (three rotate by epoch)
  uint32_t* state = nullptr;         // [0] epoch of the last completed LL call, [1] CTAs finished, [2..4] vectors in use per buffer
  uint32_t epoch_mod = 0;
  int rank = 0, tp = 1;
};

constexpr uint32_t kLLSentinel = 0xFFFFFFFFu;
constexpr unsigned kLLSpinLimit = 1u << 24;
// words of 0xFFFF / 0xFFFFFFFF are the LL protocol's "not arrived yet": a bf16 NaN with that payload is sent as 0x7FFF
__device__ __forceinline__ uint32_t ll_scrub(uint32_t w) {
  if ((w & 0xFFFFu) == 0xFFFFu) w &= ~0x8000u;
  if ((w >> 16) == 0xFFFFu) w &= ~0x80000000u;
  return w;
}
__device__ __forceinline__ uint4 ll_ld_volatile(const void* p) {     // written by peers through NVLink: never from L1
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
// vector index of 8 columns [n0, n0 + 8) of row `row` inside one rank's [M, N] message
__device__ __forceinline__ size_t ll_vec(int N, int row, int n0) { return (size_t)row * ((uint32_t)N >> 3) + (n0 >> 3); }
// push 8 columns (16 bytes) of this rank's partial row to every rank
__device__ __forceinline__ uint4 ll_push8(const LLReduce& ll, uint32_t epoch, size_t T, size_t vec, const float* f) {
  uint4 p = pack8(f);
  p.x = ll_scrub(p.x); p.y = ll_scrub(p.y); p.z = ll_scrub(p.z); p.w = ll_scrub(p.w);
  uint8_t* dst = ll.mc_ll + (size_t)(epoch % 3) * ll.ll_buf_bytes + ((size_t)ll.rank * T + vec) * 16;
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(dst), "r"(p.x), "r"(p.y), "r"(p.z),
               "r"(p.w) : "memory");
  return p;
}
// Wait for the tp pieces of V vectors (16 bytes each, V * tp loads in flight at once: one NVLink / L2 latency, not V * tp)
// and sum them in rank order (fp32).  mine[v] = what this thread pushed for vector v: taken from registers instead of
// waiting for the multicast store to come back to this rank's own copy.  false: timed out.
template <int V>
__device__ __forceinline__ bool ll_wait_sum(const LLReduce& ll, uint32_t epoch, size_t T, const size_t* vec, const uint4* mine,
                                            int nvalid, float (*s)[8]) {
  constexpr int kMaxTp = 8;
  const uint8_t* cur = ll.uc_ll + (size_t)(epoch % 3) * ll.ll_buf_bytes;
  uint4 part[V][kMaxTp];
  unsigned pending = 0;                                // bit v * 8 + p: piece not there yet
#pragma unroll
  for (int v = 0; v < V; ++v) {
#pragma unroll
    for (int p = 0; p < kMaxTp; ++p) {
      if (v < nvalid && p < ll.tp) {
        if (p == ll.rank) {
          part[v][p] = mine[v];
        } else {
          part[v][p] = ll_ld_volatile(cur + ((size_t)p * T + vec[v]) * 16);
          pending |= 1u << (v * kMaxTp + p);
        }
      }
    }
  }
  unsigned spins = 0;
  while (pending) {
#pragma unroll
    for (int v = 0; v < V; ++v) {
#pragma unroll
      for (int p = 0; p < kMaxTp; ++p) {
        const unsigned bit = 1u << (v * kMaxTp + p);
        if (pending & bit) {
          const uint4 x = part[v][p];
          if (x.x != kLLSentinel && x.y != kLLSentinel && x.z != kLLSentinel && x.w != kLLSentinel) pending &= ~bit;
          else part[v][p] = ll_ld_volatile(cur + ((size_t)p * T + vec[v]) * 16);
        }
      }
    }
    if (++spins > kLLSpinLimit) return false;
  }
#pragma unroll
  for (int v = 0; v < V; ++v) {
#pragma unroll
    for (int k = 0; k < 8; ++k) s[v][k] = 0.f;
#pragma unroll
    for (int p = 0; p < kMaxTp; ++p) {
      if (v < nvalid && p < ll.tp) {
        float f[8];
        unpack8(part[v][p], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) s[v][k] += f[k];
      }
    }
  }
  return true;
}

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
