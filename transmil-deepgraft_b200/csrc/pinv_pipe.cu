// K7, third generation: the 24 dependent 256^3 products of the Moore-Penrose iteration inside a 4-CTA cluster, exchanged through
// distributed shared memory (like the second generation, r1) and PIPELINED BY COLUMN HALF.
//
//   XZ = A Z;  T1 = 7 XZ - XZ XZ;  T2 = 15 XZ - XZ T1;  Z' = 3.25 Z - 0.25 Z T2      (moore_penrose_iter_pinv, 0.0.11)
//
// CTA (mi, ni) owns the 128x128 block (rows 128 mi, columns 128 ni) of every product.  The block is two 64-column halves
// gg = 0, 1, each with its own TMEM accumulator, its own four epilogue warps and its own barriers.  The B operand of step
// s+1 is the product of step s, and the MMA of output half gg reads only column group gg of B -- i.e. only what the gg
// epilogues (this CTA's and its B peer's) produced.  So the step is two nearly independent chains
//      MMA(s, gg) -> epilogue(s, gg) -> 4 x 4 KB pushes into the B peer -> MMA(s+1, gg)
// that interleave on the one MMA issuer and the one DSMEM port: while half 0 is being packed and pushed, the tensor core
// works on half 1.  The second generation ran these phases back to back (MMA, then epilogue, then one 32 KB push issued by
// one thread: 5 800 - 7 800 cycles per step, profiles/r1_pinv_bulk_push_timeline.txt).  The chains meet once per iteration,
// at XZ . XZ, whose A operand needs all of XZ.
// Other changes against the second generation:
//   * every epilogue warp pushes its own 4 KB piece (32 rows x 128 B, contiguous in the operand layout) as soon as it has
//     packed it, and arms the receiver's transaction barrier remotely (mbarrier.arrive.expect_tx.shared::cluster), so the
//     receiver never has to re-arm a barrier in time;
//   * the constant A (attn2 rows) stays resident in its own 64 KB buffer instead of being re-loaded every iteration; XZ / Z
//     use a second A buffer, so producing XZ never waits for the readers of A;
//   * the TMA store of Z' is waited for one step later, off the critical path.
#include <cuda_fp16.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

namespace {
constexpr int PQ_THREADS = 320;                          // warp 0 producer, warp 1 MMA, warps 2-5 epilogue half 0, 6-9 half 1
constexpr int PQ_OP = 65536;
constexpr int PQ_STG = 32768;
constexpr size_t PQ_SMEM = 1024 + 3 * PQ_OP + PQ_STG + 256;
constexpr int PQ_STEPS = 4 * TM_PINV_ITERS;

__device__ __forceinline__ void pq_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t pq_mapa(uint32_t cta_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_addr), "r"(rank));
  return r;
}
// one arrival + `bytes` expected on a PEER's barrier, then the copy that completes them
__device__ __forceinline__ void pq_push(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile("mbarrier.arrive.expect_tx.relaxed.cluster.shared::cluster.b64 _, [%0], %1;" ::"r"(bar_cluster), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster),
               "r"(src_cta), "r"(bytes), "r"(bar_cluster)
               : "memory");
}
__device__ __forceinline__ void pq_commit_mc(uint64_t* bar, uint16_t cta_mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"(cta_mask)
               : "memory");
}
}  // namespace

struct PinvPipeMaps {
  CUtensorMap A, Za, Zb;   // [nbh][256][256] fp16, box [1][64][64]
};

__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(PQ_THREADS, 1)
pinv_pipe_kernel(const __grid_constant__ PinvPipeMaps maps, const __half* __restrict__ Z0, float* __restrict__ Zout) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA0 = smem;                  // attn2 rows 128 mi (K-major SWIZZLE_128B, 4 k-blocks of [128][64]); resident
  uint8_t* sA1 = smem + PQ_OP;          // XZ rows 128 mi (steps g = 1, 2), then Z rows 128 mi (g = 3)
  uint8_t* sB = smem + 2 * PQ_OP;       // [gg][256 k rows][128 B]  MN-major SWIZZLE_128B
  uint8_t* sStage = smem + 3 * PQ_OP;   // [gg][128 rows][128 B]   K-major staging of the own Z' block for its TMA store
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 3 * PQ_OP + PQ_STG);
  uint64_t* tmaA0 = bars;               // 1
  uint64_t* tmaA1 = bars + 1;           // [4]  Z k-blocks by TMA
  uint64_t* pushA = bars + 5;           // [2]  the A peer's XZ k-block of its half gg: 4 warp pieces
  uint64_t* fullB = bars + 7;           // [4][2]  k-chunk kc of column group gg: 2 pieces (TMA at step 0, the B peer afterwards)
  uint64_t* ready = bars + 15;          // [2]  accumulator gg complete here AND at the B peer (whose B rows we overwrite)
  uint64_t* own_ready = bars + 17;      // [2]  epilogue gg -> MMA: own rows written, accumulator drained
  uint64_t* afree = bars + 19;          // the A peer is done reading its XZ / Z buffer (it may be overwritten with the next XZ)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);

#ifdef TM_PINV_TIMELINE
  __shared__ long long tl[12][2][6];   // per half: mma start, mma committed | epi ready, packed, pushed, done
#define TLQ(s, gg, i) if ((s) < 12) tl[s][gg][i] = clock64()
#else
#define TLQ(s, gg, i)
#endif
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = blockIdx.x;          // cluster spans gridDim.x = 4: rank == %cluster_ctarank
  const int mi = rank >> 1, ni = rank & 1;
  const int rankB = (1 - mi) * 2 + ni, rankA = mi * 2 + (1 - ni);
  const int bh = blockIdx.y;
  const long long mat = (long long)bh * TM_LANDMARKS * TM_LANDMARKS;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.A); tma_prefetch_desc(&maps.Za); tma_prefetch_desc(&maps.Zb);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(tmaA0, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&tmaA1[i], 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&pushA[i], 4); mbar_init(&ready[i], 2); mbar_init(&own_ready[i], 1); }
    for (int i = 0; i < 8; ++i) mbar_init(&fullB[i], 2);
    mbar_init(afree, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  pq_cluster_sync();               // every CTA's barriers exist before any remote arrive / push can reach them
  tc_fence_after();
  pdl_sync();                      // A16 / Z0 of the prep kernel are read from here on
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---- producer: the operands that come from global memory ------------------------------------------------------
    if (lane == 0) {
      mbar_arrive_expect_tx(tmaA0, PQ_OP);
      for (int kb = 0; kb < 4; ++kb) {
        tma_load_3d(sA0 + kb * 16384, &maps.A, tmaA0, kb * 64, 128 * mi, bh);
        tma_load_3d(sA0 + kb * 16384 + 8192, &maps.A, tmaA0, kb * 64, 128 * mi + 64, bh);
      }
      for (int kc = 0; kc < 4; ++kc)       // Z0 (prep kernel) is the only B operand that comes from global memory
        for (int gg = 0; gg < 2; ++gg) {
          uint64_t* b = &fullB[kc * 2 + gg];
          mbar_arrive(b);
          mbar_arrive_expect_tx(b, 8192);
          tma_load_3d(sB + gg * 32768 + kc * 8192, &maps.Za, b, 128 * ni + 64 * gg, kc * 64, bh);
        }
      for (int s = 0; s < PQ_STEPS; ++s) {
        // A operand of step g = 3: the current iterate Z, rows 128 mi (global, L2 resident), into the XZ buffer once both
        // halves of step g = 2 have been multiplied (ready[1] of that step: the MMA issuer commits half 1 last).  Every
        // phase of the barrier is waited for in turn: a parity wait must never fall two phases behind.
        mbar_wait(&ready[1], s & 1);
        if ((s & 3) != 2) continue;
        const int it = s >> 2;
        const CUtensorMap* m = (it & 1) ? &maps.Zb : &maps.Za;
        for (int kb = 0; kb < 4; ++kb) {
          mbar_arrive_expect_tx(&tmaA1[kb], 16384);
          tma_load_3d(sA1 + kb * 16384, m, &tmaA1[kb], kb * 64, 128 * mi, bh);
          tma_load_3d(sA1 + kb * 16384 + 8192, m, &tmaA1[kb], kb * 64, 128 * mi + 64, bh);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ---- MMA issuer ----------------------------------------------------------------------------------------------
    constexpr uint32_t idesc = umma_idesc_f16(128, 64, 0, 1);   // A K-major, B MN-major
    const uint16_t maskB = (uint16_t)((1u << rank) | (1u << rankB));
    const uint16_t maskA = (uint16_t)(1u << rankA);
    if (lane == 0) {
      mbar_wait(tmaA0, 0);
      for (int s = 0; s < PQ_STEPS; ++s) {
        const int it = s >> 2, g = s & 3;
        const uint32_t sAc = smem_u32(g == 0 ? sA0 : sA1);
        for (int gg = 0; gg < 2; ++gg) {
          if (s > 0) mbar_wait(&own_ready[gg], (s - 1) & 1);
          TLQ(s, gg, 0);
          tc_fence_after();
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            // own rows of B first (already in shared memory), the peer's while they are still arriving
            const int kc = (s == 0) ? i : ((2 * mi + i) & 3);
            if (s == 0 || (kc >> 1) != mi) mbar_wait(&fullB[kc * 2 + gg], s & 1);
            if (g == 1) {                       // A = XZ: k-block kc is column group kc & 1 of the own or of the A peer's block
              if ((kc >> 1) == ni) mbar_wait(&own_ready[kc & 1], (s - 1) & 1);
              else mbar_wait(&pushA[kc & 1], it & 1);
            } else if (g == 3) {
              mbar_wait(&tmaA1[kc], it & 1);
            }
            tc_fence_after();
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
              const uint64_t dA = umma_desc_k_sw128(sAc + kc * 16384) + 2 * k4;
              const uint64_t dB = umma_desc_mn_sw128(smem_u32(sB + gg * 32768 + (kc * 4 + k4) * 2048), 32768, 1024);
              umma_f16_ss(tmem_base + gg * 64, dA, dB, idesc, (i > 0 || k4 > 0) ? 1u : 0u);
            }
          }
          pq_commit_mc(&ready[gg], maskB);
          if (gg == 1 && g == 3 && s + 1 < PQ_STEPS) pq_commit_mc(afree, maskA);   // my XZ / Z buffer is free for the A peer's next XZ
          TLQ(s, gg, 1);
        }
      }
    }
    __syncwarp();
  } else {
    // ---- epilogue: one thread per row of the 128x64 half -----------------------------------------------------------------
    const int qd = warp & 3;                  // TMEM lane quarter this warp may read
    const int gg = (warp - 2) >> 2;
    const int r = qd * 32 + lane;
    const uint32_t t_row = tmem_base + ((uint32_t)(qd * 32) << 16) + gg * 64;
    const long long row_off = mat + (long long)(128 * mi + r) * TM_LANDMARKS + 128 * ni + 64 * gg;
    const int kk = 128 * mi + r;              // k row of the own block inside a B operand
    uint32_t xz[32], zs[32];                  // own XZ / own Z half row, packed fp16 (the second terms of the updates)
    {
      const uint4* z0 = reinterpret_cast<const uint4*>(Z0 + row_off);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint4 u = __ldcg(z0 + i);
        zs[4 * i] = u.x; zs[4 * i + 1] = u.y; zs[4 * i + 2] = u.z; zs[4 * i + 3] = u.w;
      }
    }
    const bool leader = qd == 0 && lane == 0;      // one thread per half: own_ready, the TMA store of Z'
    bool store_pending = false;
    // this warp's 4 KB piece: rows kk0 .. +32 of column group gg in a B operand / rows r0 .. +32 of a K-major k-block
    const uint32_t pieceB = smem_u32(sB) + gg * 32768 + (128 * mi + qd * 32) * 128;
    const uint32_t pieceA = smem_u32(sA1) + (2 * ni + gg) * 16384 + qd * 32 * 128;
    const uint32_t barB = smem_u32(&fullB[(2 * mi + (qd >> 1)) * 2 + gg]);
    const uint32_t barA = smem_u32(&pushA[gg]);
    for (int s = 0; s < PQ_STEPS; ++s) {
      const int it = s >> 2, g = s & 3;
      const bool last = s == PQ_STEPS - 1;
      mbar_wait(&ready[gg], s & 1);   // accumulator complete AND the B peer's MMA done reading the rows we will overwrite there
      if (leader) { TLQ(s, gg, 2); }
      tc_fence_after();
      uint32_t raw[2][32];
      tmem_ld_32x32_nowait(t_row, raw[0]);
      tmem_ld_32x32_nowait(t_row + 32, raw[1]);
      tmem_ld_wait();
      uint32_t pk[32];
      if (g == 0) {
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          pk[e] = pack_f16x2(__uint_as_float(raw[e >> 4][(2 * e) & 31]), __uint_as_float(raw[e >> 4][(2 * e + 1) & 31]));
          xz[e] = pk[e];
        }
      } else if (g == 3) {
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float2 z = __half22float2(*reinterpret_cast<const __half2*>(&zs[e]));
          const float a = fmaf(3.25f, z.x, -0.25f * __uint_as_float(raw[e >> 4][(2 * e) & 31]));
          const float b = fmaf(3.25f, z.y, -0.25f * __uint_as_float(raw[e >> 4][(2 * e + 1) & 31]));
          raw[e >> 4][(2 * e) & 31] = __float_as_uint(a);
          raw[e >> 4][(2 * e + 1) & 31] = __float_as_uint(b);
          pk[e] = pack_f16x2(a, b);
          zs[e] = pk[e];
        }
      } else {
        const float c = g == 1 ? 7.0f : 15.0f;
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&xz[e]));
          pk[e] = pack_f16x2(fmaf(c, x.x, -__uint_as_float(raw[e >> 4][(2 * e) & 31])), fmaf(c, x.y, -__uint_as_float(raw[e >> 4][(2 * e + 1) & 31])));
        }
      }
      if (last) {   // final iterate: fp32, unrounded, straight into the caller's Z
        float4* zo = reinterpret_cast<float4*>(Zout + row_off);
#pragma unroll
        for (int e = 0; e < 16; ++e)
          zo[e] = make_float4(__uint_as_float(raw[e >> 3][(4 * e) & 31]), __uint_as_float(raw[e >> 3][(4 * e + 1) & 31]),
                              __uint_as_float(raw[e >> 3][(4 * e + 2) & 31]), __uint_as_float(raw[e >> 3][(4 * e + 3) & 31]));
        break;
      }
      // own rows of the next B operand (MN-major); for XZ also the own k-block of the next A operand (K-major); for Z' a
      // K-major staging tile that a TMA store publishes to global memory (A operand of the NEXT iteration's last step)
      uint8_t* bdst = sB + gg * 32768 + kk * 128;
      uint8_t* kdst = g == 0 ? sA1 + (2 * ni + gg) * 16384 + r * 128 : (g == 3 ? sStage + gg * 16384 + r * 128 : nullptr);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const uint4 u = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
        *reinterpret_cast<uint4*>(bdst + ((c ^ (kk & 7)) << 4)) = u;
        if (kdst != nullptr) *reinterpret_cast<uint4*>(kdst + ((c ^ (r & 7)) << 4)) = u;
      }
      fence_proxy_async();         // generic-proxy smem writes -> visible to the bulk-copy engine, the TMA store and the MMA
      __syncwarp();
      if (leader) { TLQ(s, gg, 3); }
      if (lane == 0) {
        if (leader && store_pending) {   // the Z' block stored at the end of the previous step is in global memory (and its
          tma_store_wait_all<0>();       // staging tile read) before this thread publishes anything else: the peers' loads of
          store_pending = false;         // Z, and the next writes of the staging tile, are ordered behind this push
        }
        pq_push(pq_mapa(pieceB, rankB), pieceB, 4096, pq_mapa(barB, rankB));
        if (g == 0) {
          if (it > 0) mbar_wait(afree, (it - 1) & 1);   // the A peer's MMAs of the previous iteration have released its XZ / Z buffer
          pq_push(pq_mapa(pieceA, rankA), pieceA, 4096, pq_mapa(barA, rankA));
        }
      }
      if (leader) { TLQ(s, gg, 4); }
      tc_fence_before();             // TMEM reads retire before the next step's MMA overwrites the accumulator
      named_bar_sync(1 + gg, 128);
      if (leader) {
        mbar_arrive(&own_ready[gg]);   // this CTA's MMA warp may start on the own rows of half gg
        if (g == 3) {                  // Z' -> global (the other buffer of the Za / Zb pair)
          const CUtensorMap* md = (it & 1) ? &maps.Za : &maps.Zb;
          tma_store_3d(md, sStage + gg * 16384, 128 * ni + 64 * gg, 128 * mi, bh);
          tma_store_3d(md, sStage + gg * 16384 + 8192, 128 * ni + 64 * gg, 128 * mi + 64, bh);
          tma_store_commit();
          store_pending = true;
        }
        TLQ(s, gg, 5);
      }
    }
  }
  pq_cluster_sync();               // no CTA leaves while a peer could still signal or push into it
#ifdef TM_PINV_TIMELINE
  if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) {
    const long long t0 = tl[0][0][0];
    for (int s = 0; s < 12; ++s)
      for (int h = 0; h < 2; ++h)
        printf("step %2d g%d half %d | mma: start %6lld committed %6lld | epi: ready %6lld packed %6lld pushed %6lld done %6lld\n", s, s & 3, h,
               tl[s][h][0] - t0, tl[s][h][1] - t0, tl[s][h][2] - t0, tl[s][h][3] - t0, tl[s][h][4] - t0, tl[s][h][5] - t0);
  }
#endif
  tc_fence_after();
  if (warp == 2) tmem_dealloc(tmem_base, 128);
}

// A16, Za (= Z0, written by the prep kernel), Zb: [nbh][256][256] fp16.  Z: [nbh][256][256] fp32 result.
int pinv_pipe_launch(__half* A16, __half* Za, __half* Zb, int nbh, float* Z, cudaStream_t st) {
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "pinv_pipe_launch needs an sm_100 device");
  const long long mat = (long long)TM_LANDMARKS * TM_LANDMARKS;
  PinvPipeMaps maps;
  __half* ptrs[3] = {A16, Za, Zb};
  CUtensorMap* ms[3] = {&maps.A, &maps.Za, &maps.Zb};
  for (int i = 0; i < 3; ++i)
    TM_PROPAGATE(make_tmap_3d_f16(ms[i], ptrs[i], nbh, TM_LANDMARKS, TM_LANDMARKS, mat, TM_LANDMARKS, 64, 64));
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done)) {
    TM_CUDA(cudaFuncSetAttribute(pinv_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PQ_SMEM));
  }
  launch_pdl(pinv_pipe_kernel, dim3(4, nbh), dim3(PQ_THREADS), PQ_SMEM, st, maps, (const __half*)Za, Z);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

}  // namespace tmil
