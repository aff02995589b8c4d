// K8 on the 5th-gen tensor cores (fp16 operands, fp32 accumulate), flash style:
//   W3 = softmax_t(q_lm k^T) v   over ALL Sp padded tokens (zero-score front-pad rows included, like the reference)
// The [256, Sp] score/probability matrix never reaches HBM.  One CTA owns one (bag, head) and a contiguous chunk of
// 128-token tiles; it keeps the 256 landmark queries in smem and streams K/V tiles through a 4-stage TMA ring.  Per tile
// and per landmark half (rows 0-127 / 128-255):
//   MMA  S[128 lm,128 tok] = Q_lm[128,64] . K_tile[128,64]^T      smem x smem -> TMEM
//   online softmax along the token axis out of TMEM (one thread per landmark row): running max with LAZY rescaling
//   (the accumulator is only rescaled when the max grows by more than TAU; P <= e^TAU stays well inside fp16), P written
//   back in place as packed fp16
//   MMA  O[128,64] += P[128,128] . V_tile[128,64]                  TMEM x smem (V is the MN-major B operand) -> TMEM
// The chunk emits (max, sum, numerator) partials; attn3v_combine_kernel (attn3v.cu) merges them in a fixed order.  The same
// triple is what token-sharded ranks exchange.
// Warp roles: warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 softmax of landmark half a, warps 6-9 of half b.
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_common.cuh"
#ifndef TM_POLY_A3
#define TM_POLY_A3 (false)   // which exponential pairs (e = 0..15 of a 32-column chunk) use ex2_poly_pair: none here -- measured on B200, any share made this kernel slower (64.5 us -> 66.9 with 5 of 16, 67.3 with 4 of 16); nystrom_out gains (68.0 -> 59.6)
#endif

namespace tmil {

int attn_chunks(int64_t nbh, int64_t tiles_per_bh);   // nystrom_out_tc.cu

constexpr int A3_TOK = 128;
constexpr int A3_STAGES = 4;
constexpr int A3_QLM_BYTES = 256 * 128;
constexpr int A3_K_BYTES = 128 * 128;
constexpr int A3_STAGE_BYTES = 2 * A3_K_BYTES;   // K tile + V tile
constexpr int A3_OFF_STAGE = A3_QLM_BYTES;
constexpr int A3_OFF_BAR = A3_OFF_STAGE + A3_STAGES * A3_STAGE_BYTES;
constexpr size_t A3_SMEM = 1024 + A3_OFF_BAR + 256;
constexpr int A3_THREADS = 320;
constexpr float A3_LOG2E = 1.4426950408889634f;
constexpr float A3_TAU = 5.0f;   // lazy-rescale threshold (natural-log units): P <= e^5 ~ 148

__device__ __forceinline__ void tmem_st_32x32_u(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
      "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
      "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]),
      "r"(v[31])
      : "memory");
}

__global__ void __launch_bounds__(A3_THREADS, 1)
attn3v_tc_kernel(const __grid_constant__ CUtensorMap tmQlm, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, int tiles_per_chunk, int tiles_per_bh, int nchunks,
                 float* __restrict__ pm, float* __restrict__ ps, float* __restrict__ pO) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + A3_OFF_BAR);
  uint64_t* kv_full = bars;                    // [4] TMA -> MMA
  uint64_t* kv_empty = bars + A3_STAGES;       // [4] MMA -> TMA
  uint64_t* s_full = bars + 2 * A3_STAGES;     // [2] MMA -> softmax half
  uint64_t* p_full = s_full + 2;               // [2] softmax half -> MMA, 4 arrivals
  uint64_t* o_full = p_full + 2;               // [2] last PV of the half retired
  uint64_t* c_full = o_full + 2;               // landmark queries landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(c_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bh = blockIdx.y, chunk = blockIdx.x;
  const int tile0 = chunk * tiles_per_chunk;
  const int ntile = min(tiles_per_chunk, tiles_per_bh - tile0);

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmQlm); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < A3_STAGES; ++i) { mbar_init(&kv_full[i], 1); mbar_init(&kv_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 4); mbar_init(&o_full[i], 1); }
    mbar_init(c_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_sync();
  const uint32_t tmem_base = *tmem_slot;
#ifdef TM_TIMELINE
  __shared__ long long tl[2][12][6];
#define TL(role, j, e) do { if ((j) < 12) tl[role][j][e] = clock64(); } while (0)
#else
#define TL(role, j, e) do { } while (0)
#endif

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(c_full, A3_QLM_BYTES);
      tma_load_3d(sQ, &tmQlm, c_full, 0, 0, bh);
      for (int j = 0; j < ntile; ++j) {
        const int st = j % A3_STAGES;
        const uint32_t par = (j / A3_STAGES) & 1;
        mbar_wait(&kv_empty[st], par ^ 1);
        uint8_t* sK = smem + A3_OFF_STAGE + st * A3_STAGE_BYTES;
        mbar_arrive_expect_tx(&kv_full[st], A3_STAGE_BYTES);
        tma_load_3d(sK, &tmK, &kv_full[st], 0, (tile0 + j) * A3_TOK, bh);
        tma_load_3d(sK + A3_K_BYTES, &tmV, &kv_full[st], 0, (tile0 + j) * A3_TOK, bh);
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && ntile > 0) {
      constexpr uint32_t idesc_s = umma_idesc_f16(128, 128);
      constexpr uint32_t idesc_o = umma_idesc_f16(128, 64, 0, 1);   // B = V tile, MN-major (d contiguous)
      mbar_wait(c_full, 0);
      auto issue_s = [&](int j, int hf) {
        const int st = j % A3_STAGES;
        const uint64_t dQ = umma_desc_k_sw128(smem_u32(sQ + hf * (128 * 128)));
        const uint64_t dK = umma_desc_k_sw128(smem_u32(smem + A3_OFF_STAGE + st * A3_STAGE_BYTES));
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_base + hf * 128, dQ + 2 * k, dK + 2 * k, idesc_s, k > 0 ? 1u : 0u);
        umma_commit(&s_full[hf]);
      };
      mbar_wait(&kv_full[0], 0);
      tc_fence_after();
      issue_s(0, 0);
      issue_s(0, 1);
      for (int j = 0; j < ntile; ++j) {
        const int st = j % A3_STAGES;
        const uint32_t sV = smem_u32(smem + A3_OFF_STAGE + st * A3_STAGE_BYTES + A3_K_BYTES);
        for (int hf = 0; hf < 2; ++hf) {
          mbar_wait(&p_full[hf], j & 1);
          tc_fence_after();
          if (hf == 0) TL(1, j, 0);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t dV = umma_desc_mn_sw128(sV + ks * 2048, A3_K_BYTES, 1024);
            umma_f16_ts(tmem_base + 256 + hf * 64, tmem_base + hf * 128 + 8 * ks, dV, idesc_o, (j > 0 || ks > 0) ? 1u : 0u);
          }
          if (hf == 1) umma_commit(&kv_empty[st]);   // K_j and V_j have no reader left
          if (j + 1 < ntile) {
            if (hf == 0) {
              mbar_wait(&kv_full[(j + 1) % A3_STAGES], ((j + 1) / A3_STAGES) & 1);
              tc_fence_after();
            }
            issue_s(j + 1, hf);
            if (hf == 0) TL(1, j, 1);
          } else {
            umma_commit(&o_full[hf]);
          }
        }
      }
    }
  } else {
    const int hf = (warp - 2) >> 2;     // landmark half
    const int qd = warp & 3;            // TMEM lane quarter
    const uint32_t lane_base = (uint32_t)(qd * 32) << 16;
    const uint32_t tS = tmem_base + lane_base + hf * 128;
    const uint32_t tO = tmem_base + lane_base + 256 + hf * 64;
    float m_used = -INFINITY, sum = 0.f;
    for (int j = 0; j < ntile; ++j) {
      mbar_wait(&s_full[hf], j & 1);
      tc_fence_after();
      if (warp == 2 && lane == 0) TL(0, j, 0);
      // the 128 scores of this landmark row come to registers in ONE TMEM round trip and are used for both the max and
      // the exponentials (the TMEM round trip, ~100 cycles per wait, is what the serial chain of a tile is made of)
      uint32_t r0[32], r1[32], r2[32], r3[32];
      tmem_ld_32x32_nowait(tS, r0);
      tmem_ld_32x32_nowait(tS + 32, r1);
      tmem_ld_32x32_nowait(tS + 64, r2);
      tmem_ld_32x32_nowait(tS + 96, r3);
      tmem_ld_wait();
      if (warp == 2 && lane == 0) TL(0, j, 1);
      float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        m4[0] = fmaxf(m4[0], __uint_as_float(r0[e])); m4[1] = fmaxf(m4[1], __uint_as_float(r1[e]));
        m4[2] = fmaxf(m4[2], __uint_as_float(r2[e])); m4[3] = fmaxf(m4[3], __uint_as_float(r3[e]));
      }
      const float tmax = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      if (j == 0) {
        m_used = tmax;
      } else {
        const bool need = tmax > m_used + A3_TAU;
        if (__any_sync(0xffffffffu, need)) {
          // the accumulator is quiescent here: s_full(j) was committed after PV(j-1)
          const float f = need ? ex2_approx((m_used - tmax) * A3_LOG2E) : 1.0f;
          if (need) { m_used = tmax; sum *= f; }
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            uint32_t o[32];
            tmem_ld_32x32_nowait(tO + c * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * f);
            tmem_st_32x32_u(tO + c * 32, o);
          }
        }
      }
      const float ml = m_used * A3_LOG2E;
      if (warp == 2 && lane == 0) TL(0, j, 2);
      float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const uint32_t (&r)[32] = c == 0 ? r0 : (c == 1 ? r1 : (c == 2 ? r2 : r3));
        uint32_t pk[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          float p0, p1;
          if (TM_POLY_A3) {    // some of the 16 pairs on the FMA pipe, the rest on the MUFU pipe
            ex2_poly_pair(fmaf(__uint_as_float(r[2 * e]), A3_LOG2E, -ml), fmaf(__uint_as_float(r[2 * e + 1]), A3_LOG2E, -ml), p0, p1);
          } else {
            p0 = ex2_approx(fmaf(__uint_as_float(r[2 * e]), A3_LOG2E, -ml));
            p1 = ex2_approx(fmaf(__uint_as_float(r[2 * e + 1]), A3_LOG2E, -ml));
          }
          s4[e & 3] += p0 + p1;
          pk[e] = pack_f16x2(p0, p1);
        }
        tmem_st_32x16(tS + c * 16, pk);
      }
      sum += (s4[0] + s4[1]) + (s4[2] + s4[3]);
      if (warp == 2 && lane == 0) TL(0, j, 3);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[hf]);
      if (warp == 2 && lane == 0) TL(0, j, 4);
    }
    if (ntile > 0) {
      mbar_wait(&o_full[hf], 0);
      tc_fence_after();
      const int row = hf * 128 + qd * 32 + lane;
      const long long pi = ((long long)bh * nchunks + chunk) * TM_LANDMARKS + row;
      pm[pi] = m_used;
      ps[pi] = sum;
      float4* od = reinterpret_cast<float4*>(pO + pi * TM_DHEAD);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t o[32];
        tmem_ld_32x32_nowait(tO + c * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; e += 4)
          od[c * 8 + (e >> 2)] = make_float4(__uint_as_float(o[e]), __uint_as_float(o[e + 1]), __uint_as_float(o[e + 2]), __uint_as_float(o[e + 3]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
#ifdef TM_TIMELINE
  if (threadIdx.x == 0 && blockIdx.x == 1 && blockIdx.y == 0) {
    const long long t0 = tl[0][0][0];
    for (int j = 0; j < 12 && j < ntile; ++j)
      printf("tile %2d softmax: s_full %6lld ld %6lld max %6lld exp %6lld arrive %6lld | mma: p_full %6lld issued %6lld\n", j,
             tl[0][j][0] - t0, tl[0][j][1] - t0, tl[0][j][2] - t0, tl[0][j][3] - t0, tl[0][j][4] - t0, tl[1][j][0] - t0, tl[1][j][1] - t0);
  }
#endif
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

__global__ void attn3v_combine_kernel(const float* __restrict__ pm, const float* __restrict__ ps, const float* __restrict__ pO,
                                      int nchunks, long long cs_ms, long long bs_ms, long long cs_o, long long bs_o,
                                      float* __restrict__ W3, float* __restrict__ out_m, float* __restrict__ out_s);   // attn3v.cu

}  // namespace tmil

extern "C" int64_t tm_attn3v_tc_scratch_floats(int64_t B, int64_t Sp) {
  const int64_t nbh = B * TM_HEADS;
  return nbh * tmil::attn_chunks(nbh, Sp / tmil::A3_TOK) * TM_LANDMARKS * (TM_DHEAD + 2);
}

// partial == NULL: W3 = normalised result.  partial != NULL: the rank-level (max [nbh,256], sum [nbh,256], numerator
// [nbh,256,64]) triple is written there, unnormalised, and W3 is not touched.
static int attn3v_tc_run(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t Sp, float* W3, float* partial,
                         float* scratch, void* stream, int reserved_sms = 0) {
  using namespace tmil;
  TM_CHECK_ARG(B > 0 && Sp > 0 && Sp % TM_LANDMARKS == 0, TM_ERR_BAD_SHAPE, "tm_attn3v_tc_fwd: Sp must be a multiple of 256 (Sp=%lld)", (long long)Sp);
  TM_CHECK_ARG(qlm16 && k16 && v16 && (W3 || partial) && scratch, TM_ERR_BAD_ARGUMENT, "tm_attn3v_tc_fwd: NULL pointer");
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "tm_attn3v_tc_fwd needs an sm_100 device");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t nbh = B * TM_HEADS;
  CUtensorMap tmQ, tmK, tmV;
  TM_PROPAGATE(make_tmap_3d_f16(&tmQ, qlm16, nbh, TM_LANDMARKS, TM_DHEAD, TM_LANDMARKS * TM_DHEAD, TM_DHEAD, TM_LANDMARKS, TM_DHEAD));
  TM_PROPAGATE(make_tmap_3d_f16(&tmK, k16, nbh, Sp, TM_DHEAD, Sp * TM_DHEAD, TM_DHEAD, A3_TOK, TM_DHEAD));
  TM_PROPAGATE(make_tmap_3d_f16(&tmV, v16, nbh, Sp, TM_DHEAD, Sp * TM_DHEAD, TM_DHEAD, A3_TOK, TM_DHEAD));
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done)) {
    TM_CUDA(cudaFuncSetAttribute(attn3v_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)A3_SMEM));
  }
  const int tiles_per_bh = (int)(Sp / A3_TOK);
  int chunks = attn_chunks(nbh, tiles_per_bh);
  if (reserved_sms > 0) {   // another kernel of the same forward holds `reserved_sms` SMs: stay a single wave on the rest
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int c = (int)((sms - reserved_sms) / nbh);
    if (c >= 1 && c < chunks) chunks = c;
  }
  const int tpc = (int)ceil_div(tiles_per_bh, chunks);
  const int used = (int)ceil_div(tiles_per_bh, tpc);
  float* pm = scratch;
  float* ps = pm + nbh * used * TM_LANDMARKS;
  float* pO = ps + nbh * used * TM_LANDMARKS;
  launch_pdl(attn3v_tc_kernel, dim3((unsigned)used, (unsigned)nbh), dim3(A3_THREADS), A3_SMEM, st, tmQ, tmK, tmV, tpc, (int)tiles_per_bh, used, pm, ps, pO);
  TM_LAUNCH_CHECK();
  float* out_m = partial;
  float* out_s = partial ? partial + nbh * TM_LANDMARKS : nullptr;
  float* out_n = partial ? partial + 2 * nbh * TM_LANDMARKS : W3;
  launch_pdl(attn3v_combine_kernel, dim3(16, (unsigned)nbh), dim3(256), 0, st, (const float*)pm, (const float*)ps, (const float*)pO, (int)used,
             (long long)TM_LANDMARKS, (long long)used * TM_LANDMARKS, (long long)(TM_LANDMARKS * TM_DHEAD),
             (long long)used * TM_LANDMARKS * TM_DHEAD, out_n, out_m, out_s);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

extern "C" int tm_attn3v_tc_fwd(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t Sp, float* W3,
                                float* scratch, uint32_t flags, void* stream) {
  (void)flags;
  TM_CHECK_ARG(W3 != nullptr, TM_ERR_BAD_ARGUMENT, "tm_attn3v_tc_fwd: NULL pointer");
  return attn3v_tc_run(qlm16, k16, v16, B, Sp, W3, nullptr, scratch, stream);
}

namespace tmil {
// tm_attn3v_tc_fwd sized to leave `reserved_sms` SMs to a concurrently running kernel (forward.cu: the pinv cluster kernel).
int attn3v_tc_fwd_budget(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t Sp, float* W3, float* scratch,
                         int reserved_sms, cudaStream_t st) {
  return attn3v_tc_run(qlm16, k16, v16, B, Sp, W3, nullptr, scratch, st, reserved_sms);
}
// tm_attn3v_tc_partial_fwd with the same SM budget (sharded_fwd.cu: the replicated pinv chain runs beside it)
int attn3v_tc_partial_budget(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t rows, float* partial,
                             float* scratch, int reserved_sms, cudaStream_t st) {
  return attn3v_tc_run(qlm16, k16, v16, B, rows, nullptr, partial, scratch, st, reserved_sms);
}
}  // namespace tmil

extern "C" int64_t tm_attn3v_partial_floats(int64_t B) { return B * TM_HEADS * TM_LANDMARKS * (TM_DHEAD + 2); }

// Token-sharded K8, local half: this rank's rows [0, rows) of k16 / v16 (rows % 256 == 0) against the REPLICATED landmark
// queries -> partial = (max, sum, numerator), tm_attn3v_partial_floats(B) floats.
extern "C" int tm_attn3v_tc_partial_fwd(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t rows,
                                        float* partial, float* scratch, void* stream) {
  TM_CHECK_ARG(partial != nullptr, TM_ERR_BAD_ARGUMENT, "tm_attn3v_tc_partial_fwd: NULL pointer");
  return attn3v_tc_run(qlm16, k16, v16, B, rows, nullptr, partial, scratch, stream);
}

// Token-sharded K8, merge: `gathered` holds the partials of R ranks back to back (rank order); W3 = the softmax-weighted
// combination, visited in rank order on every rank (bit-identical results on all ranks).
extern "C" int tm_attn3v_merge_fwd(const float* gathered, int64_t R, int64_t B, float* W3, void* stream) {
  using namespace tmil;
  TM_CHECK_ARG(gathered && W3 && R > 0 && B > 0, TM_ERR_BAD_ARGUMENT, "tm_attn3v_merge_fwd: bad arguments");
  const long long nbh = B * TM_HEADS, per_rank = nbh * TM_LANDMARKS * (TM_DHEAD + 2);
  attn3v_combine_kernel<<<dim3(16, (unsigned)nbh), 256, 0, (cudaStream_t)stream>>>(
      gathered, gathered + nbh * TM_LANDMARKS, gathered + 2 * nbh * TM_LANDMARKS, (int)R, per_rank, TM_LANDMARKS, per_rank,
      TM_LANDMARKS * TM_DHEAD, W3, nullptr, nullptr);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

// ---- tm_attn3v_tc_fwd_static: the same launch pair with the chunk partials in library-owned static device memory ------
// For callers that may not allocate (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc): no cudaMalloc,
// cudaFree or allocating library routine on this path.  The array below exists once per GPU that loads the library and is
// held until the process exits (10 002 432 bytes).  In HOST code its name is only a stand-in, not a device address: the
// address each device gave it is looked up once per device, with that device current, by cudaGetSymbolAddress.
#define TM_ATTN3V_STATIC_SLOTS 148   /* (bag, head, chunk) partial slots: one per SM of a B200 */
#define TM_ATTN3V_STATIC_FLOATS (TM_ATTN3V_STATIC_SLOTS * TM_LANDMARKS * (TM_DHEAD + 2))
static __device__ __align__(256) float g_attn3v_static_scratch[TM_ATTN3V_STATIC_FLOATS];

extern "C" int tm_attn3v_tc_fwd_static(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t Sp, float* W3,
                                       uint32_t flags, void* stream) {
  (void)flags;
  TM_CHECK_ARG(W3 != nullptr, TM_ERR_BAD_ARGUMENT, "tm_attn3v_tc_fwd_static: NULL pointer");
  TM_CHECK_ARG(B > 0 && Sp > 0 && Sp % TM_LANDMARKS == 0, TM_ERR_BAD_SHAPE,
               "tm_attn3v_tc_fwd_static: Sp must be a multiple of 256 (Sp=%lld)", (long long)Sp);
  const int64_t need = tm_attn3v_tc_scratch_floats(B, Sp);
  TM_CHECK_ARG(need <= (int64_t)TM_ATTN3V_STATIC_FLOATS, TM_ERR_BAD_SHAPE,
               "tm_attn3v_tc_fwd_static: B=%lld, Sp=%lld needs %lld scratch floats, the static array holds %lld "
               "(148 partial slots: B <= 18 on a 148-SM device); use tm_attn3v_tc_fwd with caller-owned scratch",
               (long long)B, (long long)Sp, (long long)need, (long long)TM_ATTN3V_STATIC_FLOATS);
  static float* dev_ptr[64] = {nullptr};   // per device: the array's address on that device
  int dev = 0;
  TM_CUDA(cudaGetDevice(&dev));
  TM_CHECK_ARG(dev >= 0 && dev < 64, TM_ERR_BAD_ARGUMENT, "tm_attn3v_tc_fwd_static: device ordinal %d out of range", dev);
  if (dev_ptr[dev] == nullptr) {
    float* p = nullptr;
    TM_CUDA(cudaGetSymbolAddress((void**)&p, g_attn3v_static_scratch));
    TM_CHECK_ARG(p != nullptr, TM_ERR_CUDA, "tm_attn3v_tc_fwd_static: cudaGetSymbolAddress returned NULL");
    dev_ptr[dev] = p;
  }
  // every slot the combine kernel reads is written in full by the partial kernel of the same call: nothing has to start at zero
  return attn3v_tc_run(qlm16, k16, v16, B, Sp, W3, nullptr, dev_ptr[dev], stream);
}
