// K9+K10+K11 on the 5th-gen tensor cores (fp16 operands, fp32 accumulate):
//   merged[b,t,h*64:(h+1)*64] = softmax_j(q_t . k_lm_j) @ M_h + res_conv(v)[t]        (M = Z @ W3)
// One CTA owns a run of 128-token tiles of one (bag, head).  Per tile:
//   MMA1  S[128,256]  = Q[128,64] . K_lm[256,64]^T          smem x smem -> TMEM (256 fp32 columns: a full softmax row)
//   softmax over the 256 landmarks straight out of TMEM (one thread per row), P written back IN PLACE as packed fp16
//   MMA2  O[128,64]   = P[128,256] . M^T[64,256]^T          TMEM x smem -> TMEM  (the [Sp,256] attn1 matrix never leaves the SM)
//   MMA3  O2[128,64]  = T[128,160] . Vhalo[160,64]          the 33-tap depthwise conv as a banded Toeplitz GEMM:
//                                                           T[i,j] = w_h[j-i]; Vhalo = rows t0-16 .. t0+143 of v (TMA zero-fills
//                                                           rows outside [0,Sp): exactly the conv's zero padding)
//   epilogue  merged = O / rowsum + O2, head-merged 'b h n d -> b n (h d)'
// Warp roles: warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 / 6-9 two softmax+epilogue groups that ping-pong on two
// TMEM slots (2 x 256 columns) so the tensor pipe, the MUFU pipe and the TMA engine overlap across consecutive tiles.
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

constexpr int NT_TOK = 128;
constexpr int NT_HALO_ROWS = 160;
constexpr int NT_KLM_BYTES = 256 * 128;       // [256 landmarks][64 d] fp16
constexpr int NT_MT_BYTES = 4 * 64 * 128;     // 4 k-blocks of [64 d][64 landmarks] fp16
constexpr int NT_T_BYTES = 3 * 128 * 128;     // 3 k-blocks of [128 rows][64 halo cols] fp16
constexpr int NT_Q_BYTES = 128 * 128;         // [128 tokens][64 d]
constexpr int NT_V_BYTES = NT_HALO_ROWS * 128;
constexpr int NT_OFF_KLM = 0;
constexpr int NT_OFF_MT = NT_OFF_KLM + NT_KLM_BYTES;
constexpr int NT_OFF_T = NT_OFF_MT + NT_MT_BYTES;
constexpr int NT_OFF_Q = NT_OFF_T + NT_T_BYTES;                 // 2 slots of (Q, Vhalo)
constexpr int NT_SLOT_BYTES = NT_Q_BYTES + NT_V_BYTES;
constexpr int NT_OFF_BAR = NT_OFF_Q + 2 * NT_SLOT_BYTES;
constexpr size_t NT_SMEM = 1024 + NT_OFF_BAR + 256;
constexpr int NT_THREADS = 320;
constexpr float NT_LOG2E = 1.4426950408889634f;

template <bool ROUND>
__global__ void __launch_bounds__(NT_THREADS, 1)
nystrom_out_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmV,
                      const __grid_constant__ CUtensorMap tmKlm, const __grid_constant__ CUtensorMap tmMT,
                      const float* __restrict__ conv_w, int Sp, int tiles_per_chunk, int tiles_per_bh, float* __restrict__ merged) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sKlm = smem + NT_OFF_KLM;
  uint8_t* sMT = smem + NT_OFF_MT;
  uint8_t* sT = smem + NT_OFF_T;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NT_OFF_BAR);
  uint64_t* q_full = bars;          // [2] TMA -> MMA      (Q tile + V halo landed)
  uint64_t* q_empty = bars + 2;     // [2] MMA -> TMA      (MMA1 and MMA3 of the tile retired)
  uint64_t* s_full = bars + 4;      // [2] MMA -> softmax  (S in TMEM)
  uint64_t* p_full = bars + 6;      // [2] softmax -> MMA  (P in TMEM), 4 arrivals
  uint64_t* o_full = bars + 8;      // [2] MMA -> epilogue (O and O2 in TMEM)
  uint64_t* t_empty = bars + 10;    // [2] epilogue -> MMA (TMEM slot drained), 4 arrivals
  uint64_t* c_full = bars + 12;     // K_lm and M^T landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bh = blockIdx.y;
  const int h = bh % TM_HEADS, b = bh / TM_HEADS;
  const int tile0 = blockIdx.x * tiles_per_chunk;
  const int ntile = min(tiles_per_chunk, tiles_per_bh - tile0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmKlm); tma_prefetch_desc(&tmMT);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], 4); mbar_init(&o_full[i], 1); mbar_init(&t_empty[i], 4);
    }
    mbar_init(c_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  if (warp >= 2) {
    // banded Toeplitz operand of the conv: T[i][j] = w_h[j - i] for 0 <= j - i <= 32 (halo row j <-> token t0 - 16 + j),
    // K-major SWIZZLE_128B blocks of [128 rows][64 cols]
    const float* w = conv_w + h * TM_CONV_TAPS;
    for (int ci = threadIdx.x - 64; ci < 128 * 24; ci += NT_THREADS - 64) {
      const int i = ci / 24, cc = ci - i * 24;
      const int kb = cc >> 3, c = cc & 7;
      const int j0 = kb * 64 + c * 8;
      uint32_t pk[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int d0 = j0 + 2 * e - i, d1 = d0 + 1;
        const float w0 = (d0 >= 0 && d0 < TM_CONV_TAPS) ? __ldg(w + d0) : 0.f;
        const float w1 = (d1 >= 0 && d1 < TM_CONV_TAPS) ? __ldg(w + d1) : 0.f;
        pk[e] = pack_f16x2(w0, w1);
      }
      *reinterpret_cast<uint4*>(sT + kb * 16384 + (i >> 3) * 1024 + (i & 7) * 128 + ((c ^ (i & 7)) << 4)) =
          make_uint4(pk[0], pk[1], pk[2], pk[3]);
    }
    fence_proxy_async();   // generic-proxy smem writes -> visible to the tensor core's async-proxy reads
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(c_full, NT_KLM_BYTES + NT_MT_BYTES);
      tma_load_3d(sKlm, &tmKlm, c_full, 0, 0, bh);
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) tma_load_3d(sMT + kb * 8192, &tmMT, c_full, kb * 64, 0, bh);
      for (int i = 0; i < ntile; ++i) {
        const int s = i & 1;
        const uint32_t par = (i >> 1) & 1;
        mbar_wait(&q_empty[s], par ^ 1);
        uint8_t* sQ = smem + NT_OFF_Q + s * NT_SLOT_BYTES;
        mbar_arrive_expect_tx(&q_full[s], NT_SLOT_BYTES);
        tma_load_3d(sQ, &tmQ, &q_full[s], 0, (tile0 + i) * NT_TOK, bh);
        tma_load_3d(sQ + NT_Q_BYTES, &tmV, &q_full[s], 0, (tile0 + i) * NT_TOK - 16, bh);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_f16(128, 256);
      constexpr uint32_t idesc_o = umma_idesc_f16(128, 64);
      constexpr uint32_t idesc_c = umma_idesc_f16(128, 64, 0, 1);   // B = V halo, MN-major (d contiguous)
      const uint64_t dK = umma_desc_k_sw128(smem_u32(sKlm));
      mbar_wait(c_full, 0);
      auto front = [&](int i) {
        const int s = i & 1;
        const uint32_t par = (i >> 1) & 1;
        mbar_wait(&q_full[s], par);
        mbar_wait(&t_empty[s], par ^ 1);
        tc_fence_after();
        const uint64_t dQ = umma_desc_k_sw128(smem_u32(smem + NT_OFF_Q + s * NT_SLOT_BYTES));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_base + s * 256, dQ + 2 * k, dK + 2 * k, idesc_s, k > 0 ? 1u : 0u);
        umma_commit(&s_full[s]);
      };
      if (ntile > 0) front(0);
      for (int i = 0; i < ntile; ++i) {
        if (i + 1 < ntile) front(i + 1);
        const int s = i & 1;
        const uint32_t par = (i >> 1) & 1;
        mbar_wait(&p_full[s], par);
        tc_fence_after();
        const uint32_t slot = tmem_base + s * 256;
#pragma unroll
        for (int ks = 0; ks < 16; ++ks) {
          const uint64_t dM = umma_desc_k_sw128(smem_u32(sMT + (ks >> 2) * 8192)) + 2 * (ks & 3);
          umma_f16_ts(slot + 128, slot + 8 * ks, dM, idesc_o, ks > 0 ? 1u : 0u);
        }
        const uint32_t sV = smem_u32(smem + NT_OFF_Q + s * NT_SLOT_BYTES + NT_Q_BYTES);
#pragma unroll
        for (int ks = 0; ks < 10; ++ks) {
          const uint64_t dT = umma_desc_k_sw128(smem_u32(sT + (ks >> 2) * 16384)) + 2 * (ks & 3);
          const uint64_t dV = umma_desc_mn_sw128(sV + ks * 2048, NT_V_BYTES, 1024);
          umma_f16_ss(slot + 192, dT, dV, idesc_c, ks > 0 ? 1u : 0u);
        }
        umma_commit(&q_empty[s]);
        umma_commit(&o_full[s]);
      }
    }
  } else {
    const int g = (warp - 2) >> 2;      // softmax group = TMEM slot
    const int qd = warp & 3;            // TMEM lane quarter this warp may access
    const uint32_t slot = tmem_base + ((uint32_t)(qd * 32) << 16) + g * 256;
    float* out_row0 = merged + ((long long)b * Sp + qd * 32 + lane) * TM_DIM + h * TM_DHEAD;
    for (int i = g; i < ntile; i += 2) {
      const uint32_t par = (i >> 1) & 1;
      mbar_wait(&s_full[g], par);
      tc_fence_after();
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 8; c += 2) {
        uint32_t r0[32], r1[32];
        tmem_ld_32x32_nowait(slot + c * 32, r0);
        tmem_ld_32x32_nowait(slot + c * 32 + 32, r1);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) mx = fmaxf(mx, fmaxf(__uint_as_float(r0[e]), __uint_as_float(r1[e])));
      }
      const float ml = mx * NT_LOG2E;
      float sum = 0.f;
#pragma unroll 1
      for (int c = 0; c < 8; ++c) {
        uint32_t r[32], pk[16];
        tmem_ld_32x32_nowait(slot + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float p0 = ex2_approx(fmaf(__uint_as_float(r[2 * e]), NT_LOG2E, -ml));
          const float p1 = ex2_approx(fmaf(__uint_as_float(r[2 * e + 1]), NT_LOG2E, -ml));
          sum += p0 + p1;
          pk[e] = pack_f16x2(p0, p1);
        }
        tmem_st_32x16(slot + c * 16, pk);   // P overwrites S columns that were already consumed
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[g]);

      mbar_wait(&o_full[g], par);
      tc_fence_after();
      const float inv = 1.0f / sum;
      float* orow = out_row0 + (long long)(tile0 + i) * NT_TOK * TM_DIM;
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        uint32_t o[32], c2[32];
        tmem_ld_32x32_nowait(slot + 128 + hf * 32, o);
        tmem_ld_32x32_nowait(slot + 192 + hf * 32, c2);
        tmem_ld_wait();
        if (hf == 1) {   // both halves are in registers: hand the TMEM slot back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&t_empty[g]);
        }
#pragma unroll
        for (int e = 0; e < 32; e += 4) {
          float4 r4;
          r4.x = fmaf(__uint_as_float(o[e]), inv, __uint_as_float(c2[e]));
          r4.y = fmaf(__uint_as_float(o[e + 1]), inv, __uint_as_float(c2[e + 1]));
          r4.z = fmaf(__uint_as_float(o[e + 2]), inv, __uint_as_float(c2[e + 2]));
          r4.w = fmaf(__uint_as_float(o[e + 3]), inv, __uint_as_float(c2[e + 3]));
          if (ROUND) { r4.x = round_tf32(r4.x); r4.y = round_tf32(r4.y); r4.z = round_tf32(r4.z); r4.w = round_tf32(r4.w); }
          *reinterpret_cast<float4*>(orow + hf * 32 + e) = r4;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// M^T in fp16 for MMA2's K-major B operand: mt16[bh][d][j] = half(M[bh][j][d])
__global__ void __launch_bounds__(256) mt16_kernel(const float* __restrict__ M, __half* __restrict__ mt16) {
  __shared__ float t[64][65];
  const int bh = blockIdx.y, j0 = blockIdx.x * 64;
  const float* src = M + ((long long)bh * TM_LANDMARKS + j0) * TM_DHEAD;
  for (int i = threadIdx.x; i < 64 * 64; i += 256) t[i >> 6][i & 63] = src[i];
  __syncthreads();
  __half* dst = mt16 + (long long)bh * TM_DHEAD * TM_LANDMARKS + j0;
  for (int i = threadIdx.x; i < 64 * 64; i += 256) {
    const int d = i >> 6, j = i & 63;
    dst[(long long)d * TM_LANDMARKS + j] = __float2half_rn(t[j][d]);
  }
}

// M = Z @ W3 in exact fp32 (256x256 @ 256x64 per head) with both consumers' layouts written at once:
// M [bh][j][d] fp32 (stage buffer) and mt16 [bh][d][j] fp16 (K-major B operand of attn1 @ M).
// grid (16, B*8): 16 rows per CTA; thread = (row, 4 columns); Z rows staged in smem, W3 streamed through L1/L2.
__global__ void __launch_bounds__(256) zw16_kernel(const float* __restrict__ Z, const float* __restrict__ W3, float* __restrict__ M,
                                                   __half* __restrict__ mt16) {
  __shared__ __align__(16) float zs[16][TM_LANDMARKS];
  const int bh = blockIdx.y, r0 = blockIdx.x * 16, tid = threadIdx.x;
  const float4* zg = reinterpret_cast<const float4*>(Z + ((long long)bh * TM_LANDMARKS + r0) * TM_LANDMARKS);
  for (int i = tid; i < 16 * 64; i += 256) reinterpret_cast<float4*>(&zs[0][0])[i] = __ldg(zg + i);
  __syncthreads();
  const int ry = tid >> 4, cx = tid & 15;
  const float4* wg = reinterpret_cast<const float4*>(W3 + (long long)bh * TM_LANDMARKS * TM_DHEAD) + cx;
  float4 acc0 = make_float4(0.f, 0.f, 0.f, 0.f), acc1 = acc0;
#pragma unroll 4
  for (int k = 0; k < TM_LANDMARKS; k += 2) {
    const float4 w0 = __ldg(wg + (long long)k * 16), w1 = __ldg(wg + (long long)(k + 1) * 16);
    const float z0 = zs[ry][k], z1 = zs[ry][k + 1];
    acc0.x = fmaf(z0, w0.x, acc0.x); acc0.y = fmaf(z0, w0.y, acc0.y); acc0.z = fmaf(z0, w0.z, acc0.z); acc0.w = fmaf(z0, w0.w, acc0.w);
    acc1.x = fmaf(z1, w1.x, acc1.x); acc1.y = fmaf(z1, w1.y, acc1.y); acc1.z = fmaf(z1, w1.z, acc1.z); acc1.w = fmaf(z1, w1.w, acc1.w);
  }
  const float4 o = make_float4(acc0.x + acc1.x, acc0.y + acc1.y, acc0.z + acc1.z, acc0.w + acc1.w);
  const int r = r0 + ry;
  reinterpret_cast<float4*>(M + ((long long)bh * TM_LANDMARKS + r) * TM_DHEAD)[cx] = o;
  __half* mt = mt16 + ((long long)bh * TM_DHEAD + cx * 4) * TM_LANDMARKS + r;
  mt[0] = __float2half_rn(o.x);
  mt[TM_LANDMARKS] = __float2half_rn(o.y);
  mt[2 * TM_LANDMARKS] = __float2half_rn(o.z);
  mt[3 * TM_LANDMARKS] = __float2half_rn(o.w);
}

int attn_chunks(int64_t nbh, int64_t tiles_per_bh) {
  int64_t c = (148 + nbh / 2) / nbh;   // about one CTA per SM in total
  if (c < 1) c = 1;
  if (c > tiles_per_bh) c = tiles_per_bh;
  return (int)c;
}

}  // namespace tmil

extern "C" int tm_mt16_fwd(const float* M, void* mt16, int64_t B, void* stream) {
  using namespace tmil;
  TM_CHECK_ARG(M && mt16 && B > 0, TM_ERR_BAD_ARGUMENT, "tm_mt16_fwd: bad arguments");
  mt16_kernel<<<dim3(4, (unsigned)(B * TM_HEADS)), 256, 0, (cudaStream_t)stream>>>(M, (__half*)mt16);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

extern "C" int tm_zw16_fwd(const float* Z, const float* W3, int64_t B, float* M, void* mt16, void* stream) {
  using namespace tmil;
  TM_CHECK_ARG(Z && W3 && M && mt16 && B > 0, TM_ERR_BAD_ARGUMENT, "tm_zw16_fwd: bad arguments");
  zw16_kernel<<<dim3(16, (unsigned)(B * TM_HEADS)), 256, 0, (cudaStream_t)stream>>>(Z, W3, M, (__half*)mt16);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

extern "C" int tm_nystrom_out_tc_fwd(const void* q16, const void* klm16, const void* mt16, const void* v16, const float* conv_w,
                                     int64_t B, int64_t Sp, float* merged, uint32_t flags, void* stream) {
  using namespace tmil;
  TM_CHECK_ARG(B > 0 && Sp > 0 && Sp % TM_LANDMARKS == 0, TM_ERR_BAD_SHAPE, "tm_nystrom_out_tc_fwd: Sp must be a multiple of 256 (Sp=%lld)", (long long)Sp);
  TM_CHECK_ARG(q16 && klm16 && mt16 && v16 && conv_w && merged, TM_ERR_BAD_ARGUMENT, "tm_nystrom_out_tc_fwd: NULL pointer");
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "tm_nystrom_out_tc_fwd needs an sm_100 device");
  const int64_t nbh = B * TM_HEADS;
  CUtensorMap tmQ, tmV, tmK, tmM;
  TM_PROPAGATE(make_tmap_3d_f16(&tmQ, q16, nbh, Sp, TM_DHEAD, Sp * TM_DHEAD, TM_DHEAD, NT_TOK, TM_DHEAD));
  TM_PROPAGATE(make_tmap_3d_f16(&tmV, v16, nbh, Sp, TM_DHEAD, Sp * TM_DHEAD, TM_DHEAD, NT_HALO_ROWS, TM_DHEAD));
  TM_PROPAGATE(make_tmap_3d_f16(&tmK, klm16, nbh, TM_LANDMARKS, TM_DHEAD, TM_LANDMARKS * TM_DHEAD, TM_DHEAD, TM_LANDMARKS, TM_DHEAD));
  TM_PROPAGATE(make_tmap_3d_f16(&tmM, mt16, nbh, TM_DHEAD, TM_LANDMARKS, TM_DHEAD * TM_LANDMARKS, TM_LANDMARKS, TM_DHEAD, 64));
  static bool attr_done = false;
  if (!attr_done) {
    TM_CUDA(cudaFuncSetAttribute(nystrom_out_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NT_SMEM));
    TM_CUDA(cudaFuncSetAttribute(nystrom_out_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NT_SMEM));
    attr_done = true;
  }
  const int tiles_per_bh = (int)(Sp / NT_TOK);
  const int chunks = attn_chunks(nbh, tiles_per_bh);
  const int tpc = (int)ceil_div(tiles_per_bh, chunks);
  const int used = (int)ceil_div(tiles_per_bh, tpc);
  dim3 grid((unsigned)used, (unsigned)nbh);
  if (flags & TM_FLAG_TF32_OPERANDS)
    nystrom_out_tc_kernel<true><<<grid, NT_THREADS, NT_SMEM, (cudaStream_t)stream>>>(tmQ, tmV, tmK, tmM, conv_w, (int)Sp, tpc, tiles_per_bh, merged);
  else
    nystrom_out_tc_kernel<false><<<grid, NT_THREADS, NT_SMEM, (cudaStream_t)stream>>>(tmQ, tmV, tmK, tmM, conv_w, (int)Sp, tpc, tiles_per_bh, merged);
  TM_LAUNCH_CHECK();
  return TM_OK;
}
