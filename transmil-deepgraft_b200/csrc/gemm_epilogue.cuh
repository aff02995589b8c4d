// Fused GEMM epilogues of the path, shared by the CUDA-core (gemm_simt.cu) and the
// tcgen05 (gemm_tc.cu) engines: each caller hands over 4 consecutive accumulator
// columns of one logical row.
#pragma once
#include "common.cuh"
#include <cuda_fp16.h>

#include "tc_common.cuh"

namespace tmil {

// logical row r -> C row, or -1 when the row is not stored (front-pad rows of to_out)
__device__ __forceinline__ long long map_c_row(const GemmParams& p, int r, int* i_out) {
  if (p.rows_per_bag <= 0) { *i_out = r; return r; }
  const int bag = r / p.rows_per_bag;
  const int i = r - bag * p.rows_per_bag;
  *i_out = i;
  if (i < p.c_skip) return -1;
  return (long long)bag * p.c_bag_rows + p.c_row_off + (i - p.c_skip);
}

__device__ __forceinline__ void epilogue_store4(const GemmParams& p, int r, int c, float4 acc, const float* __restrict__ Lb,
                                                float* __restrict__ Cb) {
  switch (p.epi) {
    case EPI_AXPBY: {
      float4 o = make_float4(p.alpha * acc.x, p.alpha * acc.y, p.alpha * acc.z, p.alpha * acc.w);
      if (Lb != nullptr) {
        float4 l = *reinterpret_cast<const float4*>(Lb + (long long)r * p.ldl + c);
        o.x = fmaf(p.beta, l.x, o.x); o.y = fmaf(p.beta, l.y, o.y);
        o.z = fmaf(p.beta, l.z, o.z); o.w = fmaf(p.beta, l.w, o.w);
      }
      if (p.round_out) { o.x = round_tf32(o.x); o.y = round_tf32(o.y); o.z = round_tf32(o.z); o.w = round_tf32(o.w); }
      *reinterpret_cast<float4*>(Cb + (long long)r * p.ldc + c) = o;
      break;
    }
    case EPI_AXPBY_T: {   // the value was finished (scaled, second term, rounding) in the accumulator layout; plain store
      if (Cb != nullptr) *reinterpret_cast<float4*>(Cb + (long long)r * p.ldc + c) = acc;
      break;
    }
    case EPI_BIAS_GELU: {
      float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + c));
      float4 o = make_float4(gelu_erf(acc.x + b.x), gelu_erf(acc.y + b.y), gelu_erf(acc.z + b.z), gelu_erf(acc.w + b.w));
      int i;
      const long long cr = map_c_row(p, r, &i);
      *reinterpret_cast<float4*>(Cb + cr * p.ldc + c) = o;
      if (i < p.dup) *reinterpret_cast<float4*>(Cb + (cr + p.rows_per_bag) * p.ldc + c) = o;  // duplicate pad (TransMIL.py:180)
      break;
    }
    case EPI_QKV: {
      const int which = c / TM_DIM;
      const int cc = c - which * TM_DIM;
      const int hh = cc / TM_DHEAD, d = cc % TM_DHEAD;
      const int bag = r / p.Sp, t = r - bag * p.Sp;
      float* base = which == 0 ? p.q : (which == 1 ? p.k : p.v);
      const float s = which == 0 ? 0.125f : 1.0f;  // q * dim_head^-0.5
      float4 o = make_float4(acc.x * s, acc.y * s, acc.z * s, acc.w * s);
      if (p.round_out) { o.x = round_tf32(o.x); o.y = round_tf32(o.y); o.z = round_tf32(o.z); o.w = round_tf32(o.w); }
      *reinterpret_cast<float4*>(base + (((long long)bag * TM_HEADS + hh) * p.Sp + t) * TM_DHEAD + d) = o;
      break;
    }
    case EPI_QKV16: {
      const int which = c / TM_DIM;
      const int cc = c - which * TM_DIM;
      const int hh = cc / TM_DHEAD, d = cc % TM_DHEAD;
      const int bag = r / p.Sp, t = r - bag * p.Sp;
      __half* base = reinterpret_cast<__half*>(which == 0 ? p.q : (which == 1 ? p.k : p.v));
      const float s = which == 0 ? 0.125f : 1.0f;  // q * dim_head^-0.5 (exact in fp16: power of two)
      uint2 o;
      o.x = pack_f16x2(acc.x * s, acc.y * s);
      o.y = pack_f16x2(acc.z * s, acc.w * s);
      *reinterpret_cast<uint2*>(base + (((long long)bag * TM_HEADS + hh) * p.Sp + t) * TM_DHEAD + d) = o;
      break;
    }
    case EPI_BIAS_RESID: {
      int i;
      const long long cr = map_c_row(p, r, &i);
      if (cr < 0) break;
      float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + c));
      float4 l = *reinterpret_cast<const float4*>(p.L + cr * p.ldl + c);
      float4 o = make_float4(l.x + (acc.x + b.x), l.y + (acc.y + b.y), l.z + (acc.z + b.z), l.w + (acc.w + b.w));
      *reinterpret_cast<float4*>(Cb + cr * p.ldc + c) = o;
      break;
    }
  }
}

}  // namespace tmil
