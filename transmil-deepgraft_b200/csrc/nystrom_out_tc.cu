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
#ifndef TM_POLY_NO
#define TM_POLY_NO (e % 3 == 2)   // which exponential pairs (e = 0..15 of a 32-column chunk) use ex2_poly_pair
#endif

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
constexpr int NT_STAGES = 3;                                     // smem stages of (Q, Vhalo): a stage is held for the whole
                                                                 // chain of its tile, so 3 keep the HBM latency off the chain
constexpr int NT_OFF_Q = NT_OFF_T + NT_T_BYTES;
constexpr int NT_SLOT_BYTES = NT_Q_BYTES + NT_V_BYTES;
constexpr int NT_OFF_BAR = NT_OFF_Q + NT_STAGES * NT_SLOT_BYTES;
constexpr size_t NT_SMEM = 1024 + NT_OFF_BAR + 256;
constexpr int NT_THREADS = 320;
constexpr float NT_LOG2E = 1.4426950408889634f;

template <int OUT>   // 0: fp32, 1: fp32 rounded to tf32, 2: fp16 (operand of the fp16 to_out GEMM)
__global__ void __launch_bounds__(NT_THREADS, 1)
nystrom_out_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmV,
                      const __grid_constant__ CUtensorMap tmKlm, const __grid_constant__ CUtensorMap tmMT,
                      const float* __restrict__ conv_w, int Sp, int tiles_per_chunk, int tiles_per_bh, void* __restrict__ merged_v) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sKlm = smem + NT_OFF_KLM;
  uint8_t* sMT = smem + NT_OFF_MT;
  uint8_t* sT = smem + NT_OFF_T;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NT_OFF_BAR);
  uint64_t* q_full = bars;          // [3] TMA -> MMA      (Q tile + V halo landed)
  uint64_t* q_empty = bars + 3;     // [3] MMA -> TMA      (MMA1 and MMA3 of the tile retired)
  uint64_t* s_full = bars + 6;      // [2] MMA -> softmax  (S in TMEM)
  uint64_t* p_full = bars + 8;      // [2] softmax -> MMA  (P in TMEM), 4 arrivals
  uint64_t* o_full = bars + 10;     // [2] MMA -> epilogue (O and O2 in TMEM)
  uint64_t* t_empty = bars + 12;    // [2] epilogue -> MMA (TMEM slot drained), 4 arrivals
  uint64_t* c_full = bars + 14;     // K_lm and M^T landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bh = blockIdx.y;
  const int h = bh % TM_HEADS, b = bh / TM_HEADS;
  const int tile0 = blockIdx.x * tiles_per_chunk;
  const int ntile = min(tiles_per_chunk, tiles_per_bh - tile0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmKlm); tma_prefetch_desc(&tmMT);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < NT_STAGES; ++i) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 4); mbar_init(&o_full[i], 1); mbar_init(&t_empty[i], 4);
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
  pdl_sync();   // conv_w above is a module parameter (no kernel of the forward writes it); q, v, k_lm, M are read from here on
  const uint32_t tmem_base = *tmem_slot;
#ifdef TM_TIMELINE
  __shared__ long long tl[2][12][6];
#define TL(role, j, e) do { if ((j) < 12) tl[role][j][e] = clock64(); } while (0)
#else
#define TL(role, j, e) do { } while (0)
#endif

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(c_full, NT_KLM_BYTES + NT_MT_BYTES);
      tma_load_3d(sKlm, &tmKlm, c_full, 0, 0, bh);
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) tma_load_3d(sMT + kb * 8192, &tmMT, c_full, kb * 64, 0, bh);
      for (int i = 0; i < ntile; ++i) {
        const int st = i % NT_STAGES;
        const uint32_t spar = (i / NT_STAGES) & 1;
        mbar_wait(&q_empty[st], spar ^ 1);
        uint8_t* sQ = smem + NT_OFF_Q + st * NT_SLOT_BYTES;
        mbar_arrive_expect_tx(&q_full[st], NT_SLOT_BYTES);
        tma_load_3d(sQ, &tmQ, &q_full[st], 0, (tile0 + i) * NT_TOK, bh);
        tma_load_3d(sQ + NT_Q_BYTES, &tmV, &q_full[st], 0, (tile0 + i) * NT_TOK - 16, bh);
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
        const int st = i % NT_STAGES;
        mbar_wait(&q_full[st], (i / NT_STAGES) & 1);
        TL(1, i, 0);
        mbar_wait(&t_empty[s], par ^ 1);
        tc_fence_after();
        const uint64_t dQ = umma_desc_k_sw128(smem_u32(smem + NT_OFF_Q + st * NT_SLOT_BYTES));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_base + s * 256, dQ + 2 * k, dK + 2 * k, idesc_s, k > 0 ? 1u : 0u);
        umma_commit(&s_full[s]);
        TL(1, i, 1);
      };
      if (ntile > 0) front(0);
      for (int i = 0; i < ntile; ++i) {
        if (i + 1 < ntile) front(i + 1);
        const int s = i & 1;
        const uint32_t par = (i >> 1) & 1;
        mbar_wait(&p_full[s], par);
        TL(1, i, 2);
        tc_fence_after();
        const uint32_t slot = tmem_base + s * 256;
#pragma unroll
        for (int ks = 0; ks < 16; ++ks) {
          const uint64_t dM = umma_desc_k_sw128(smem_u32(sMT + (ks >> 2) * 8192)) + 2 * (ks & 3);
          umma_f16_ts(slot + 128, slot + 8 * ks, dM, idesc_o, ks > 0 ? 1u : 0u);
        }
        const uint32_t sV = smem_u32(smem + NT_OFF_Q + (i % NT_STAGES) * NT_SLOT_BYTES + NT_Q_BYTES);
#pragma unroll
        for (int ks = 0; ks < 10; ++ks) {
          const uint64_t dT = umma_desc_k_sw128(smem_u32(sT + (ks >> 2) * 16384)) + 2 * (ks & 3);
          const uint64_t dV = umma_desc_mn_sw128(sV + ks * 2048, NT_V_BYTES, 1024);
          umma_f16_ss(slot + 192, dT, dV, idesc_c, ks > 0 ? 1u : 0u);
        }
        umma_commit(&q_empty[i % NT_STAGES]);
        umma_commit(&o_full[s]);
        TL(1, i, 3);
      }
    }
  } else {
    const int g = (warp - 2) >> 2;      // softmax group = TMEM slot
    const int qd = warp & 3;            // TMEM lane quarter this warp may access
    const uint32_t slot = tmem_base + ((uint32_t)(qd * 32) << 16) + g * 256;
    const long long out_row0 = ((long long)b * Sp + qd * 32 + lane) * TM_DIM + h * TM_DHEAD;   // element offset into merged
    for (int i = g; i < ntile; i += 2) {
      const uint32_t par = (i >> 1) & 1;
      mbar_wait(&s_full[g], par);
      if (qd == 2 && lane == 0) TL(0, i, 0);
      tc_fence_after();
      // pass 1: row max over the 256 landmarks.  128 columns per TMEM round trip, four independent max chains.
      float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll 1
      for (int c = 0; c < 8; c += 4) {
        uint32_t r0[32], r1[32], r2[32], r3[32];
        tmem_ld_32x32_nowait(slot + c * 32, r0);
        tmem_ld_32x32_nowait(slot + c * 32 + 32, r1);
        tmem_ld_32x32_nowait(slot + c * 32 + 64, r2);
        tmem_ld_32x32_nowait(slot + c * 32 + 96, r3);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          m4[0] = fmaxf(m4[0], __uint_as_float(r0[e])); m4[1] = fmaxf(m4[1], __uint_as_float(r1[e]));
          m4[2] = fmaxf(m4[2], __uint_as_float(r2[e])); m4[3] = fmaxf(m4[3], __uint_as_float(r3[e]));
        }
      }
      const float mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      if (qd == 2 && lane == 0) TL(0, i, 1);
      const float ml = mx * NT_LOG2E;
      // pass 2: P = exp(S - max) as packed fp16, written over the consumed S columns; the next chunk's TMEM load is in
      // flight while the current one is exponentiated; four independent row-sum chains.
      float s4[4] = {0.f, 0.f, 0.f, 0.f};
      uint32_t ra[32], rb[32];
      tmem_ld_32x32_nowait(slot, ra);
      tmem_ld_wait();
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        uint32_t (&cur)[32] = (c & 1) ? rb : ra;
        uint32_t (&nxt)[32] = (c & 1) ? ra : rb;
        if (c + 1 < 8) tmem_ld_32x32_nowait(slot + (c + 1) * 32, nxt);
        uint32_t pk[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          float p0, p1;
          if (TM_POLY_NO) {    // some of the 16 pairs on the FMA pipe, the rest on the MUFU pipe
            ex2_poly_pair(fmaf(__uint_as_float(cur[2 * e]), NT_LOG2E, -ml), fmaf(__uint_as_float(cur[2 * e + 1]), NT_LOG2E, -ml), p0, p1);
          } else {
            p0 = ex2_approx(fmaf(__uint_as_float(cur[2 * e]), NT_LOG2E, -ml));
            p1 = ex2_approx(fmaf(__uint_as_float(cur[2 * e + 1]), NT_LOG2E, -ml));
          }
          s4[e & 3] += p0 + p1;
          pk[e] = pack_f16x2(p0, p1);
        }
        if (c + 1 < 8) tmem_ld_wait();      // chunk c+1 is in registers before chunk c's columns... 
        tmem_st_32x16(slot + c * 16, pk);   // ...P overwrites S columns [16c,16c+16), all read by now (16c+16 <= 32(c+1))
      }
      const float sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[g]);
      if (qd == 2 && lane == 0) TL(0, i, 2);

      mbar_wait(&o_full[g], par);
      if (qd == 2 && lane == 0) TL(0, i, 3);
      tc_fence_after();
      const float inv = 1.0f / sum;
      const long long orow = out_row0 + (long long)(tile0 + i) * NT_TOK * TM_DIM;
      {
        uint32_t oa[32], ob[32], ca[32], cb[32];   // O and the conv accumulator, one TMEM round trip
        tmem_ld_32x32_nowait(slot + 128, oa);
        tmem_ld_32x32_nowait(slot + 160, ob);
        tmem_ld_32x32_nowait(slot + 192, ca);
        tmem_ld_32x32_nowait(slot + 224, cb);
        tmem_ld_wait();
        tc_fence_before();   // everything is in registers: hand the TMEM slot back to the MMA warp
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[g]);
        if (qd == 2 && lane == 0) TL(0, i, 4);
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          const uint32_t (&o)[32] = hf ? ob : oa;
          const uint32_t (&c2)[32] = hf ? cb : ca;
          if (OUT == 2) {
            __half* op = reinterpret_cast<__half*>(merged_v) + orow + hf * 32;
#pragma unroll
            for (int e = 0; e < 32; e += 8) {
              uint4 u;
              u.x = pack_f16x2(fmaf(__uint_as_float(o[e]), inv, __uint_as_float(c2[e])), fmaf(__uint_as_float(o[e + 1]), inv, __uint_as_float(c2[e + 1])));
              u.y = pack_f16x2(fmaf(__uint_as_float(o[e + 2]), inv, __uint_as_float(c2[e + 2])), fmaf(__uint_as_float(o[e + 3]), inv, __uint_as_float(c2[e + 3])));
              u.z = pack_f16x2(fmaf(__uint_as_float(o[e + 4]), inv, __uint_as_float(c2[e + 4])), fmaf(__uint_as_float(o[e + 5]), inv, __uint_as_float(c2[e + 5])));
              u.w = pack_f16x2(fmaf(__uint_as_float(o[e + 6]), inv, __uint_as_float(c2[e + 6])), fmaf(__uint_as_float(o[e + 7]), inv, __uint_as_float(c2[e + 7])));
              *reinterpret_cast<uint4*>(op + e) = u;
            }
          } else {
            float* op = reinterpret_cast<float*>(merged_v) + orow + hf * 32;
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              float4 r4;
              r4.x = fmaf(__uint_as_float(o[e]), inv, __uint_as_float(c2[e]));
              r4.y = fmaf(__uint_as_float(o[e + 1]), inv, __uint_as_float(c2[e + 1]));
              r4.z = fmaf(__uint_as_float(o[e + 2]), inv, __uint_as_float(c2[e + 2]));
              r4.w = fmaf(__uint_as_float(o[e + 3]), inv, __uint_as_float(c2[e + 3]));
              if (OUT == 1) { r4.x = round_tf32(r4.x); r4.y = round_tf32(r4.y); r4.z = round_tf32(r4.z); r4.w = round_tf32(r4.w); }
              *reinterpret_cast<float4*>(op + e) = r4;
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
#ifdef TM_TIMELINE
  if (threadIdx.x == 64 && blockIdx.x == 1 && blockIdx.y == 0) {
    const long long t0 = tl[0][0][0];
    for (int j = 0; j < 12 && j < ntile; ++j)
      printf("tile %2d grp%d: s_full %6lld max %6lld p_arrive %6lld o_full %6lld done %6lld | mma: q_full %6lld front %6lld p_full %6lld back %6lld\n", j, j & 1,
             tl[0][j][0] - t0, tl[0][j][1] - t0, tl[0][j][2] - t0, tl[0][j][3] - t0, tl[0][j][4] - t0, tl[1][j][0] - t0, tl[1][j][1] - t0,
             tl[1][j][2] - t0, tl[1][j][3] - t0);
  }
#endif
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
// grid (16, B*8): 16 rows per CTA; thread = (row, 4 columns); the CTA's 16 Z rows and the head's whole W3 (64 KB) are
// staged in shared memory with every load in flight at once -- the kernel sits on the critical path between the pinv /
// attn3.v join and nystrom_out, so its latency, not its 67 MFLOP, is what counts.
__global__ void __launch_bounds__(256) zw16_kernel(const float* __restrict__ Z, const float* __restrict__ W3, float* __restrict__ M,
                                                   __half* __restrict__ mt16) {
  extern __shared__ __align__(16) float zw_smem[];
  pdl_sync();
  float* ws = zw_smem;                                   // [256][64]
  float (*zs)[TM_LANDMARKS] = reinterpret_cast<float (*)[TM_LANDMARKS]>(zw_smem + TM_LANDMARKS * TM_DHEAD);   // [16][256]
  const int bh = blockIdx.y, r0 = blockIdx.x * 16, tid = threadIdx.x;
  const float4* zg = reinterpret_cast<const float4*>(Z + ((long long)bh * TM_LANDMARKS + r0) * TM_LANDMARKS);
  const float4* wg = reinterpret_cast<const float4*>(W3 + (long long)bh * TM_LANDMARKS * TM_DHEAD);
#pragma unroll
  for (int i = 0; i < 4; ++i) reinterpret_cast<float4*>(&zs[0][0])[tid + 256 * i] = __ldg(zg + tid + 256 * i);
#pragma unroll
  for (int i = 0; i < 16; ++i) reinterpret_cast<float4*>(ws)[tid + 256 * i] = __ldg(wg + tid + 256 * i);
  __syncthreads();
  const int ry = tid >> 4, cx = tid & 15;
  const float4* wsv = reinterpret_cast<const float4*>(ws) + cx;
  float4 acc0 = make_float4(0.f, 0.f, 0.f, 0.f), acc1 = acc0;
#pragma unroll 8
  for (int k = 0; k < TM_LANDMARKS; k += 2) {
    const float4 w0 = wsv[k * 16], w1 = wsv[(k + 1) * 16];
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
  // one CTA per SM (512 TMEM columns each): never more CTAs than SMs unless the heads alone exceed them -- 152 CTAs on
  // 148 SMs is a second wave that doubles the kernel time
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
      sms = 148;
  }
  int64_t c = sms / nbh;
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
  const size_t smem = (size_t)(TM_LANDMARKS * TM_DHEAD + 16 * TM_LANDMARKS) * sizeof(float);
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done)) {
    TM_CUDA(cudaFuncSetAttribute(zw16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  launch_pdl(zw16_kernel, dim3(16, (unsigned)(B * TM_HEADS)), dim3(256), smem, (cudaStream_t)stream, Z, W3, M, (__half*)mt16);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

namespace tmil {
int nystrom_out_tc_tiles(const void* q16, const void* klm16, const void* mt16, const void* v16, const float* conv_w, int64_t B, int64_t Sp,
                         int64_t row_tiles, void* merged, uint32_t flags, cudaStream_t stream);
}

extern "C" int tm_nystrom_out_tc_fwd(const void* q16, const void* klm16, const void* mt16, const void* v16, const float* conv_w,
                                     int64_t B, int64_t Sp, void* merged, uint32_t flags, void* stream) {
  return tmil::nystrom_out_tc_tiles(q16, klm16, mt16, v16, conv_w, B, Sp, 0, merged, flags, (cudaStream_t)stream);
}

// row_tiles > 0: only the first row_tiles tiles of 128 padded rows of every bag are produced (the rest of `merged` is left
// untouched); every tile is computed exactly as in the full call (same operands, same order), so its rows have the same bits.
int tmil::nystrom_out_tc_tiles(const void* q16, const void* klm16, const void* mt16, const void* v16, const float* conv_w, int64_t B,
                               int64_t Sp, int64_t row_tiles, void* merged, uint32_t flags, cudaStream_t stream) {
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
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done)) {
    TM_CUDA(cudaFuncSetAttribute(nystrom_out_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NT_SMEM));
    TM_CUDA(cudaFuncSetAttribute(nystrom_out_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NT_SMEM));
    TM_CUDA(cudaFuncSetAttribute(nystrom_out_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)NT_SMEM));
  }
  int tiles_per_bh = (int)(Sp / NT_TOK);
  if (row_tiles > 0 && row_tiles < tiles_per_bh) tiles_per_bh = (int)row_tiles;
  const int chunks = attn_chunks(nbh, tiles_per_bh);
  const int tpc = (int)ceil_div(tiles_per_bh, chunks);
  const int used = (int)ceil_div(tiles_per_bh, tpc);
  dim3 grid((unsigned)used, (unsigned)nbh);
  if (flags & TM_FLAG_F16_ACT)
    launch_pdl(nystrom_out_tc_kernel<2>, grid, dim3(NT_THREADS), NT_SMEM, (cudaStream_t)stream, tmQ, tmV, tmK, tmM, conv_w, (int)Sp, (int)tpc, (int)tiles_per_bh, merged);
  else if (flags & TM_FLAG_TF32_OPERANDS)
    launch_pdl(nystrom_out_tc_kernel<1>, grid, dim3(NT_THREADS), NT_SMEM, (cudaStream_t)stream, tmQ, tmV, tmK, tmM, conv_w, (int)Sp, (int)tpc, (int)tiles_per_bh, merged);
  else
    launch_pdl(nystrom_out_tc_kernel<0>, grid, dim3(NT_THREADS), NT_SMEM, (cudaStream_t)stream, tmQ, tmV, tmK, tmM, conv_w, (int)Sp, (int)tpc, (int)tiles_per_bh, merged);
  TM_LAUNCH_CHECK();
  return TM_OK;
}
