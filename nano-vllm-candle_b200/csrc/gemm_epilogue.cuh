// Epilogue helpers shared by the 1-CTA and the CTA-pair GEMM: one thread owns one accumulator row
// and emits 16 or 32 consecutive columns of it.
#pragma once
#include "common.cuh"

enum { NVC_EPI_NONE = 0, NVC_EPI_SILU_MUL = 1 };

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
