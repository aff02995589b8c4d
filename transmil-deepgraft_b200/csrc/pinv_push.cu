// K7, second generation: the 24 dependent 256^3 products of the Moore-Penrose iteration inside a 4-CTA cluster, with the
// operand exchange done entirely through DISTRIBUTED SHARED MEMORY -- no global-memory round trip and no cluster-wide
// barrier on the critical path.
//
//   XZ = A Z;  T1 = 7 XZ - XZ XZ;  T2 = 15 XZ - XZ T1;  Z' = 3.25 Z - 0.25 Z T2      (moore_penrose_iter_pinv, 0.0.11)
//
// CTA (mi, ni) owns the 128x128 block (rows 128 mi, columns 128 ni) of every product X @ Y:
//   A operand = X[128 mi .. +128, :]  K-major  SWIZZLE_128B, 4 k-blocks of [128 rows][64 k]      (sA, 64 KB)
//   B operand = Y[:, 128 ni .. +128]  MN-major SWIZZLE_128B, 2 column groups of [256 k][64 n]    (sB, 64 KB)
// The B operand of step s+1 is always the product of step s; its k rows 128 mi .. are this CTA's own block, the other
// half is the block of the "B peer" (1-mi, ni).  The A operand is fresh only for XZ (steps g = 1, 2): k-blocks 2 ni, 2 ni+1
// are the own block, the rest is the block of the "A peer" (mi, 1-ni).
// Epilogue of step s (8 warps): TMEM -> registers -> polynomial term -> fp16, written in operand layout into the CTA's OWN
// sB (and sA for XZ); then one thread pushes those same bytes into the peers' shared memory with
// cp.async.bulk.shared::cluster (the layouts are position independent, so source and destination offsets coincide),
// completing the peers' fullB / fullA transaction barriers -- the peers' MMA warps wait on them exactly as they would on a
// TMA load.  The reverse dependency (a peer may only be overwritten once ITS MMA has finished reading) is carried by the
// MMA completion itself: tcgen05.commit is multicast to the `ready` barrier of the CTA and of both its peers, and an
// epilogue starts when all three arrivals are in.
// Only Z (A operand of the last product of an iteration, needed two steps after it is produced) and the constant A go
// through global memory, by TMA, off the critical path.
#include <cuda_fp16.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

namespace {
constexpr int PP_THREADS = 320;                          // warp 0 producer, warp 1 MMA, warps 2-9 epilogue
constexpr int PP_OP_BYTES = 65536;
constexpr int PP_STG_BYTES = 32768;                      // K-major staging of the own Z' block for its TMA store
constexpr size_t PP_SMEM_BYTES = 1024 + 2 * PP_OP_BYTES + PP_STG_BYTES + 256;
constexpr int PP_STEPS = 4 * TM_PINV_ITERS;

__device__ __forceinline__ void cluster_sync_all_() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t cta_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_addr), "r"(rank));
  return r;
}
// own shared memory -> a peer's shared memory (same offsets), completing `bytes` on the peer's transaction barrier
__device__ __forceinline__ void dsmem_push(uint32_t dst_cluster_addr, uint32_t src_cta_addr, uint32_t bytes, uint32_t mbar_cluster_addr) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster_addr),
               "r"(src_cta_addr), "r"(bytes), "r"(mbar_cluster_addr)
               : "memory");
}
// remote (peer CTA) 16-byte store and remote barrier arrive, for the REMOTE_ST variant
__device__ __forceinline__ void st_cluster_v4(uint32_t cluster_addr, const uint4& u) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(cluster_addr), "r"(u.x), "r"(u.y), "r"(u.z), "r"(u.w) : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t mbar_cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(mbar_cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "W_%=:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra D_%=;\n\t"
      "bra W_%=;\n\t"
      "D_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async_full() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void umma_commit_multicast(uint64_t* bar, uint16_t cta_mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"(cta_mask)
               : "memory");
}
}  // namespace

struct PinvPushMaps {
  CUtensorMap A, Za, Zb;   // [nbh][256][256] fp16, box [1][64][64]
};

// REMOTE_ST = false (default): the own halves are pushed with cp.async.bulk.shared::cluster after the epilogue has packed them.
// REMOTE_ST = true (TM_PINV_KERNEL=remote): the epilogue threads write the peers' halves themselves with st.shared::cluster
// while they pack, and one remote barrier arrive per peer replaces the transaction count.  Measured on B200 (r1): the
// thread-issued remote stores stall the packing loop (3300-6300 cycles per step instead of 1100-1900) and the kernel is
// 25 % SLOWER than the bulk-copy form (0.142 vs 0.113 ms for the whole pinv); kept as the recorded experiment.
template <bool REMOTE_ST>
__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(PP_THREADS, 1)
pinv_push_kernel(const __grid_constant__ PinvPushMaps maps, const __half* __restrict__ Z0, float* __restrict__ Zout) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = smem + PP_OP_BYTES;
  uint8_t* sStage = smem + 2 * PP_OP_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * PP_OP_BYTES + PP_STG_BYTES);
  uint64_t* fullA = bars;          // [4] per k-block of the A operand: TMA bytes or a peer's push
  uint64_t* fullB = bars + 4;      // [4] per 64-row k-chunk of the B operand
  uint64_t* ready = bars + 8;      // 3 arrivals per step: MMA of this CTA, of its B peer and of its A peer complete
  uint64_t* own_ready = bars + 9;  // epilogue -> MMA warp: own halves are in shared memory, TMEM is drained
  uint64_t* pushB = bars + 10;     // REMOTE_ST: the B peer has written its half of the next B operand into sB
  uint64_t* pushA = bars + 11;     // REMOTE_ST: the A peer has written its half of XZ into sA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

#ifdef TM_PINV_TIMELINE
  __shared__ long long tl[PP_STEPS][8];   // mma: own_ready seen, operands complete, committed | epi: ready seen, packed, synced, pushed
#define TLP(s, i) tl[s][i] = clock64()
#else
#define TLP(s, i)
#endif
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = blockIdx.x;          // cluster spans gridDim.x = 4: rank == %cluster_ctarank
  const int mi = rank >> 1, ni = rank & 1;
  const int rankB = (1 - mi) * 2 + ni, rankA = mi * 2 + (1 - ni);
  const int bh = blockIdx.y;
  const long long mat = (long long)bh * TM_LANDMARKS * TM_LANDMARKS;
  const int pB = 2 * (1 - mi), pA = 2 * (1 - ni);   // first k-chunk of the peer's half of B / of A

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.A); tma_prefetch_desc(&maps.Za); tma_prefetch_desc(&maps.Zb);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 4; ++i) { mbar_init(&fullA[i], 1); mbar_init(&fullB[i], 1); }
    mbar_init(ready, 3);
    mbar_init(own_ready, 1);
    mbar_init(pushB, 1);
    mbar_init(pushA, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all_();             // every CTA's barriers exist before any remote arrive / push can reach them
  tc_fence_after();
  pdl_sync();                      // A16 / Z0 of the prep kernel are read from here on
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---- producer: TMA loads of the operands that come from global memory, and the expect-tx side of the peers' pushes ----
    auto loadA = [&](const CUtensorMap* m) {   // rows 128 mi .. +128 of a row-major matrix, K-major k-blocks
      for (int kb = 0; kb < 4; ++kb) {
        mbar_arrive_expect_tx(&fullA[kb], 16384);
        tma_load_3d(sA + kb * 16384, m, &fullA[kb], kb * 64, 128 * mi, bh);
        tma_load_3d(sA + kb * 16384 + 8192, m, &fullA[kb], kb * 64, 128 * mi + 64, bh);
      }
    };
    if (lane == 0) {
      loadA(&maps.A);
      for (int kc = 0; kc < 4; ++kc) {   // Z0 (prep kernel) is the only B operand that comes from global memory
        mbar_arrive_expect_tx(&fullB[kc], 16384);
        tma_load_3d(sB + kc * 8192, &maps.Za, &fullB[kc], 128 * ni, kc * 64, bh);
        tma_load_3d(sB + 32768 + kc * 8192, &maps.Za, &fullB[kc], 128 * ni + 64, kc * 64, bh);
      }
      for (int s = 1; s < PP_STEPS; ++s) {
        const int it = s >> 2, g = s & 3;
        // the peers push the halves of this step's fresh operands during their epilogue of step s-1, i.e. after `ready`
        // of step s-1; by then this CTA's MMA of step s-1 has consumed the previous phase of every barrier re-armed here
        mbar_wait(ready, (s - 1) & 1);   // also keeps this loop in lockstep with the phases of `ready` (parity waits must not run ahead)
        if (!REMOTE_ST) {
          mbar_arrive_expect_tx(&fullB[pB], 16384);
          mbar_arrive_expect_tx(&fullB[pB + 1], 16384);
          if (g == 1) {
            mbar_arrive_expect_tx(&fullA[pA], 16384);
            mbar_arrive_expect_tx(&fullA[pA + 1], 16384);
          }
        }
        if (g == 2) {                        // A operand of step g = 3: the current iterate Z, rows 128 mi (global, L2 resident)
          mbar_wait(ready, s & 1);           // sA (XZ) is free once this step's MMA is done
          loadA((it & 1) ? &maps.Zb : &maps.Za);
        } else if (g == 3 && s + 1 < PP_STEPS) {
          mbar_wait(ready, s & 1);
          loadA(&maps.A);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ---- MMA issuer ----------------------------------------------------------------------------------------------
    constexpr uint32_t idesc = umma_idesc_f16(128, 64, 0, 1);   // A K-major, B MN-major
    const uint16_t mask = (uint16_t)((1u << rank) | (1u << rankB) | (1u << rankA));
    if (lane == 0) {
      uint32_t phA = 0, phB = 0;   // bit kc = parity of the next phase of fullA[kc] / fullB[kc]
      for (int s = 0; s < PP_STEPS; ++s) {
        const int g = s & 3;
        if (s > 0) mbar_wait(own_ready, (s - 1) & 1);
        TLP(s, 0);
        tc_fence_after();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          // own half first (already in shared memory), the peer's half while it is still arriving
          const int kc = (s == 0) ? i : ((2 * mi + i) & 3);
          const bool peerA = g == 1 && (kc >> 1) == (pA >> 1), peerB = s > 0 && (kc >> 1) == (pB >> 1);
          const bool waitA = (g == 0 || g == 3) || (peerA && !REMOTE_ST);
          const bool waitB = (s == 0) || (peerB && !REMOTE_ST);
          if (waitA) { mbar_wait(&fullA[kc], (phA >> kc) & 1); phA ^= 1u << kc; }
          if (waitB) { mbar_wait(&fullB[kc], (phB >> kc) & 1); phB ^= 1u << kc; }
          if (REMOTE_ST && s > 0) {
            const bool needB = i == 2;                // i = 2, 3 are the peer's k-chunks of B (own half first)
            const bool needA = peerA && !(i & 1);     // first of the A peer's two k-blocks of XZ (they are adjacent in this order)
            if (needB) mbar_wait_cluster(pushB, (s - 1) & 1);
            if (needA) mbar_wait_cluster(pushA, (s >> 2) & 1);
            if (needB || needA) fence_proxy_async_full();   // peers' generic-proxy remote stores -> the MMA's async-proxy reads
          }
          if (i == 3) { TLP(s, 1); }
          tc_fence_after();
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) {
            const int ks = kc * 4 + k4;
            const uint64_t dA = umma_desc_k_sw128(smem_u32(sA + kc * 16384)) + 2 * k4;
#pragma unroll
            for (int gg = 0; gg < 2; ++gg) {
              const uint64_t dB = umma_desc_mn_sw128(smem_u32(sB + gg * 32768 + ks * 2048), 32768, 1024);
              umma_f16_ss(tmem_base + gg * 64, dA, dB, idesc, (i > 0 || k4 > 0) ? 1u : 0u);
            }
          }
        }
        umma_commit_multicast(ready, mask);
        TLP(s, 2);
      }
    }
    __syncwarp();
  } else {
    // ---- epilogue: one thread per row and 64-column half of the 128x128 block -----------------------------------------
    const int qd = warp & 3;
    const int r = qd * 32 + lane;
    const int ch0 = ((warp - 2) >> 2) * 2;   // this warp's two 32-column chunks: ch0, ch0 + 1
    const uint32_t t_row = tmem_base + ((uint32_t)(qd * 32) << 16);
    const long long row_off = mat + (long long)(128 * mi + r) * TM_LANDMARKS + 128 * ni;
    const int kk = 128 * mi + r;             // k row of the own block inside a B operand
    uint32_t xz[32], zs[32];                 // own XZ / own Z half row, packed fp16 (the second terms of the updates)
    {
      const uint4* z0 = reinterpret_cast<const uint4*>(Z0 + row_off + ch0 * 32);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint4 u = __ldcg(z0 + i);
        zs[4 * i] = u.x; zs[4 * i + 1] = u.y; zs[4 * i + 2] = u.z; zs[4 * i + 3] = u.w;
      }
    }
    const bool leader = threadIdx.x == 64;
    const uint32_t sB_u = smem_u32(sB), sA_u = smem_u32(sA);
    for (int s = 0; s < PP_STEPS; ++s) {
      const int it = s >> 2, g = s & 3;
      const bool last = s == PP_STEPS - 1;
      mbar_wait(ready, s & 1);       // accumulator complete AND both peers' MMAs done reading what will be overwritten
      if (leader) { TLP(s, 3); }
      tc_fence_after();
      uint32_t raw[2][32];
      tmem_ld_32x32_nowait(t_row + ch0 * 32, raw[0]);
      tmem_ld_32x32_nowait(t_row + ch0 * 32 + 32, raw[1]);
      tmem_ld_wait();
#pragma unroll
      for (int cl = 0; cl < 2; ++cl) {
        const int ch = ch0 + cl;
        float v[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(raw[cl][e]);
        uint32_t pk[16];
        if (g == 0) {
#pragma unroll
          for (int e = 0; e < 16; ++e) { pk[e] = pack_f16x2(v[2 * e], v[2 * e + 1]); xz[cl * 16 + e] = pk[e]; }
        } else if (g == 3) {
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const float2 z = __half22float2(*reinterpret_cast<const __half2*>(&zs[cl * 16 + e]));
            v[2 * e] = fmaf(3.25f, z.x, -0.25f * v[2 * e]);
            v[2 * e + 1] = fmaf(3.25f, z.y, -0.25f * v[2 * e + 1]);
            pk[e] = pack_f16x2(v[2 * e], v[2 * e + 1]);
            zs[cl * 16 + e] = pk[e];
          }
        } else {
          const float c = g == 1 ? 7.0f : 15.0f;
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&xz[cl * 16 + e]));
            pk[e] = pack_f16x2(fmaf(c, x.x, -v[2 * e]), fmaf(c, x.y, -v[2 * e + 1]));
          }
        }
        if (last) {   // final iterate: fp32, unrounded, straight into the caller's Z
          float4* zo = reinterpret_cast<float4*>(Zout + row_off + ch * 32);
#pragma unroll
          for (int e = 0; e < 8; ++e) zo[e] = make_float4(v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]);
        } else {
          // own half of the next B operand (MN-major); for XZ also the own k-blocks of the next A operand (K-major); for Z'
          // a K-major staging tile that a TMA store publishes to global memory (A operand of the NEXT iteration's last step)
          uint8_t* bdst = sB + (ch >> 1) * 32768 + kk * 128;
          uint8_t* kdst = g == 0 ? sA + (2 * ni + (ch >> 1)) * 16384 + r * 128 : (g == 3 ? sStage + (ch >> 1) * 16384 + r * 128 : nullptr);
          const uint32_t rb = REMOTE_ST ? mapa_u32(smem_u32(bdst), rankB) : 0u;
          const uint32_t ra = (REMOTE_ST && g == 0) ? mapa_u32(smem_u32(kdst), rankA) : 0u;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint4 u = make_uint4(pk[4 * e], pk[4 * e + 1], pk[4 * e + 2], pk[4 * e + 3]);
            const int c = (ch & 1) * 4 + e;
            *reinterpret_cast<uint4*>(bdst + ((c ^ (kk & 7)) << 4)) = u;
            if (kdst != nullptr) *reinterpret_cast<uint4*>(kdst + ((c ^ (r & 7)) << 4)) = u;
            if (REMOTE_ST) {
              st_cluster_v4(rb + ((c ^ (kk & 7)) << 4), u);
              if (g == 0) st_cluster_v4(ra + ((c ^ (r & 7)) << 4), u);
            }
          }
        }
      }
      tc_fence_before();             // TMEM reads retire before the next step's MMA overwrites the accumulator
      if (leader) { TLP(s, 4); }
      if (!last) {
        fence_proxy_async();         // generic-proxy smem writes -> visible to the bulk-copy engine, the TMA store and the MMA
        named_bar_sync(1, 256);
        if (leader) {
          TLP(s, 5);
          mbar_arrive(own_ready);    // this CTA's MMA warp may start on the own halves
          if (REMOTE_ST) {
            // every epilogue thread's remote stores happen-before this point (bar.sync); the release is cumulative
            mbar_arrive_remote(mapa_u32(smem_u32(pushB), rankB));
            if (g == 0) mbar_arrive_remote(mapa_u32(smem_u32(pushA), rankA));
          } else {
            // B: k rows 128 mi .. +128 = k-chunks 2 mi, 2 mi + 1, both column groups -> B peer, same offsets
#pragma unroll
            for (int gg = 0; gg < 2; ++gg)
#pragma unroll
              for (int q2 = 0; q2 < 2; ++q2) {
                const int kc = 2 * mi + q2;
                const uint32_t off = gg * 32768 + kc * 8192;
                dsmem_push(mapa_u32(sB_u + off, rankB), sB_u + off, 8192, mapa_u32(smem_u32(&fullB[kc]), rankB));
              }
            if (g == 0) {              // XZ: own k-blocks 2 ni, 2 ni + 1 -> A peer
#pragma unroll
              for (int q2 = 0; q2 < 2; ++q2) {
                const int kb = 2 * ni + q2;
                dsmem_push(mapa_u32(sA_u + kb * 16384, rankA), sA_u + kb * 16384, 16384, mapa_u32(smem_u32(&fullA[kb]), rankA));
              }
            }
          }
          if (g == 3) {              // Z' -> global (the other buffer of the Za / Zb pair)
            const CUtensorMap* md = (it & 1) ? &maps.Za : &maps.Zb;
#pragma unroll
            for (int bx = 0; bx < 2; ++bx)
#pragma unroll
              for (int hf = 0; hf < 2; ++hf)
                tma_store_3d(md, sStage + bx * 16384 + hf * 8192, 128 * ni + 64 * bx, 128 * mi + 64 * hf, bh);
            tma_store_commit();
            tma_store_wait_all<0>();
          }
          TLP(s, 6);
        }
      }
    }
  }
  cluster_sync_all_();               // no CTA leaves while a peer could still signal or push into it
#ifdef TM_PINV_TIMELINE
  if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) {
    const long long t0 = tl[0][0];
    for (int s = 0; s < 12; ++s)
      printf("step %2d g%d | mma: start %6lld operands %6lld committed %6lld | epi: ready %6lld packed %6lld synced %6lld pushed %6lld\n", s, s & 3,
             tl[s][0] - t0, tl[s][1] - t0, tl[s][2] - t0, tl[s][3] - t0, tl[s][4] - t0, tl[s][5] - t0, tl[s][6] - t0);
  }
#endif
  tc_fence_after();
  if (warp == 2) tmem_dealloc(tmem_base, 128);
}

// A16, Za (= Z0, written by the prep kernel), Zb: [nbh][256][256] fp16.  Z: [nbh][256][256] fp32 result.
int pinv_push_launch(__half* A16, __half* Za, __half* Zb, int nbh, float* Z, cudaStream_t st) {
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "pinv_push_launch needs an sm_100 device");
  const long long mat = (long long)TM_LANDMARKS * TM_LANDMARKS;
  PinvPushMaps maps;
  __half* ptrs[3] = {A16, Za, Zb};
  CUtensorMap* ms[3] = {&maps.A, &maps.Za, &maps.Zb};
  for (int i = 0; i < 3; ++i)
    TM_PROPAGATE(make_tmap_3d_f16(ms[i], ptrs[i], nbh, TM_LANDMARKS, TM_LANDMARKS, mat, TM_LANDMARKS, 64, 64));
  // default: cp.async.bulk pushes; TM_PINV_KERNEL=remote: st.shared::cluster
  static const int remote_st = [] { const char* e = getenv("TM_PINV_KERNEL"); return (e != nullptr && strcmp(e, "remote") == 0) ? 1 : 0; }();
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done)) {
    TM_CUDA(cudaFuncSetAttribute(pinv_push_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PP_SMEM_BYTES));
    TM_CUDA(cudaFuncSetAttribute(pinv_push_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PP_SMEM_BYTES));
  }
  if (remote_st) launch_pdl(pinv_push_kernel<true>, dim3(4, nbh), dim3(PP_THREADS), PP_SMEM_BYTES, st, maps, (const __half*)Za, Z);
  else launch_pdl(pinv_push_kernel<false>, dim3(4, nbh), dim3(PP_THREADS), PP_SMEM_BYTES, st, maps, (const __half*)Za, Z);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

}  // namespace tmil
