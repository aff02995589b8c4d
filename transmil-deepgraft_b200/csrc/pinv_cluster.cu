// K7 as ONE kernel: the 6 Moore-Penrose iterations (24 dependent 256^3 products) of every (bag, head) run inside a
// 4-CTA thread-block cluster, fp16 operands / fp32 accumulation on the tensor cores.
//
//   Z <- 1/4 Z (13 I - XZ (15 I - XZ (7 I - XZ))),  XZ = A Z        (nystrom-attention 0.0.11 moore_penrose_iter_pinv)
//   evaluated as  XZ = A Z;  T1 = 7 XZ - XZ XZ;  T2 = 15 XZ - XZ T1;  Z' = 3.25 Z - 0.25 Z T2   (same as pinv.cu)
//
// CTA (mi, ni) of the cluster owns the 128x128 output block (rows 128 mi.., columns 128 ni..) of every product X @ Y:
//   A operand = X[128 mi .. +128, :]   K-major   (4 k-blocks of [128 rows][64 k])
//   B operand = Y[:, 128 ni .. +128]   MN-major  (2 column groups of [256 k rows][64 n]) -- the row-major iterate serves both
//   roles, so every iterate is stored once, row-major fp16, in global memory (L2 resident: 128 KB each).
// A product needs blocks written by the other CTAs of the cluster.  The epilogue of a step therefore (1) writes the CTA's own
// block straight into its shared memory in the operand layouts of the next step (its half of B, and of A when the product
// is XZ), and (2) publishes the block to the peers with a TMA store from a swizzled staging tile; after ONE hardware
// cluster barrier per product the peers' TMA loads fetch only the half they do not own (32 KB instead of 64 KB per
// operand), and the MMA starts on the local half while that load is in flight.  There is no host round trip and no relaunch.  The second term of each update (XZ, Z) is the CTA's own previous output block: it stays in registers as
// packed fp16.  Operand loads that do not depend on the step in flight (A rows, Z rows) are prefetched behind the barrier.
// Warp roles: warp 0 TMA, warp 1 MMA issue, warps 2-9 epilogue (one thread per output row and 64-column half).
#include <cuda_fp16.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

constexpr int PC_THREADS = 320;   // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue (two warps per TMEM lane quarter: column halves)
constexpr int PC_OP_BYTES = 65536;                       // one operand tile: 128 x 256 (A) or 256 x 128 (B) fp16
constexpr int PC_STAGE_BYTES = 32768;                    // the CTA's own 128 x 128 output block, K-major, on its way to global
constexpr size_t PC_SMEM = 1024 + 2 * PC_OP_BYTES + PC_STAGE_BYTES + 256;
constexpr int PC_STEPS = 4 * TM_PINV_ITERS;

__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

struct PinvMaps {
  CUtensorMap A, Za, Zb, XZ, T1, T2;   // [nbh][256][256] fp16, box [1][64][64]
};

__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(PC_THREADS, 1)
pinv_cluster_kernel(const __grid_constant__ PinvMaps maps, __half* __restrict__ Za, __half* __restrict__ Zb,
                    __half* __restrict__ XZ, __half* __restrict__ T1, __half* __restrict__ T2, float* __restrict__ Zout) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = smem + PC_OP_BYTES;
  uint8_t* sStage = smem + 2 * PC_OP_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * PC_OP_BYTES + PC_STAGE_BYTES);
  uint64_t* fullA = bars;         // [4] per k-block of the A operand
  uint64_t* fullB = bars + 4;     // [4] per 64-row k-chunk of the B operand
  uint64_t* mma_done = bars + 8;  // accumulator of the step complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

#ifdef TM_PINV_TIMELINE
  __shared__ long long tl[PC_STEPS][10];   // per step: producer barrier exit, B issued | mma: barrier exit, B[3] full, issued | epi: barrier exit, mma_done, stores done
#define TL(s, i) tl[s][i] = clock64()
#else
#define TL(s, i)
#endif
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = blockIdx.x;          // cluster spans gridDim.x = 4
  const int mi = rank >> 1, ni = rank & 1;
  const int bh = blockIdx.y;
  const long long mat = (long long)bh * TM_LANDMARKS * TM_LANDMARKS;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.A); tma_prefetch_desc(&maps.Za); tma_prefetch_desc(&maps.Zb);
    tma_prefetch_desc(&maps.XZ); tma_prefetch_desc(&maps.T1); tma_prefetch_desc(&maps.T2);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 4; ++i) { mbar_init(&fullA[i], 1); mbar_init(&fullB[i], 1); }
    mbar_init(mma_done, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---- TMA producer ------------------------------------------------------------------------------------------
    // Only the PEER's half of a fresh operand travels through global memory: the CTA's own 128x128 block of the previous
    // product was written straight into sB (and, for XZ, into sA) by its epilogue.
    auto loadA = [&](const CUtensorMap* m, int kb0, int kb1) {   // rows 128 mi .. +128 of a row-major iterate, K-major k-blocks
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_arrive_expect_tx(&fullA[kb], 16384);
        tma_load_3d(sA + kb * 16384, m, &fullA[kb], kb * 64, 128 * mi, bh);
        tma_load_3d(sA + kb * 16384 + 8192, m, &fullA[kb], kb * 64, 128 * mi + 64, bh);
      }
    };
    auto loadB = [&](const CUtensorMap* m, int kc0, int kc1) {   // columns 128 ni .. +128, MN-major: 2 groups of [256 k rows][64 n]
      for (int kc = kc0; kc < kc1; ++kc) {
        mbar_arrive_expect_tx(&fullB[kc], 16384);
        tma_load_3d(sB + kc * 8192, m, &fullB[kc], 128 * ni, kc * 64, bh);
        tma_load_3d(sB + 32768 + kc * 8192, m, &fullB[kc], 128 * ni + 64, kc * 64, bh);
      }
    };
    const int pB = 2 * (1 - mi), pA = 2 * (1 - ni);   // first k-chunk of the peer's half of B / of A
    if (lane == 0) loadA(&maps.A, 0, 4);
    __syncwarp();
    for (int s = 0; s < PC_STEPS; ++s) {
      cluster_sync_all();
      if (lane == 0) {
        TL(s, 0);
        // the peers published their blocks with TMA stores (async proxy, completed before they arrived at the barrier) and
        // this TMA load is async proxy too: no generic <-> async proxy fence is needed on this side
        const int it = s >> 2, g = s & 3;
        const CUtensorMap* zc = (it & 1) ? &maps.Zb : &maps.Za;
        const CUtensorMap* mb = g == 0 ? zc : (g == 1 ? &maps.XZ : (g == 2 ? &maps.T1 : &maps.T2));
        if (s == 0) loadB(mb, 0, 4);        // Z0 comes from the prep kernel: nothing is in shared memory yet
        else loadB(mb, pB, pB + 2);
        if (g == 1) loadA(&maps.XZ, pA, pA + 2);
        TL(s, 1);
        if (g == 2) {                       // next A operand = Z rows: independent of the step in flight
          mbar_wait(mma_done, s & 1);
          loadA(zc, 0, 4);
        } else if (g == 3 && s + 1 < PC_STEPS) {
          mbar_wait(mma_done, s & 1);
          loadA(&maps.A, 0, 4);
        }
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ---- MMA issuer --------------------------------------------------------------------------------------------
    constexpr uint32_t idesc = umma_idesc_f16(128, 64, 0, 1);   // A K-major, B MN-major
    uint32_t phA = 0, phB = 0;   // bit kc = parity of the next TMA phase of fullA[kc] / fullB[kc]
    const int pB = 2 * (1 - mi), pA = 2 * (1 - ni);
    for (int s = 0; s < PC_STEPS; ++s) {
      cluster_sync_all();
      if (lane == 0) {
        TL(s, 2);
        tc_fence_after();
        const int g = s & 3;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          // the CTA's own half of B (k rows 128 mi ..) is already in shared memory: start there, the peer's half arrives by TMA
          const int kc = (s == 0) ? i : ((2 * mi + i) & 3);
          const bool waitA = (g == 0 || g == 3) || (g == 1 && (kc >> 1) == (pA >> 1));
          const bool waitB = (s == 0) || (kc >> 1) == (pB >> 1);
          if (waitA) { mbar_wait(&fullA[kc], (phA >> kc) & 1); phA ^= 1u << kc; }
          if (waitB) { mbar_wait(&fullB[kc], (phB >> kc) & 1); phB ^= 1u << kc; }
          if (i == 0) { TL(s, 6); }
          if (i == 3) { TL(s, 3); }
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
        umma_commit(mma_done);
        TL(s, 4);
      }
      __syncwarp();
    }
  } else {
    // ---- epilogue: one thread per row of the 128x128 block ---------------------------------------------------------
    const int qd = warp & 3;
    const int r = qd * 32 + lane;
    const int ch0 = ((warp - 2) >> 2) * 2;   // this warp's two 32-column chunks of the 128-column block: ch0, ch0 + 1
    const uint32_t t_row = tmem_base + ((uint32_t)(qd * 32) << 16);
    const long long row_off = mat + (long long)(128 * mi + r) * TM_LANDMARKS + 128 * ni;
    uint32_t xz[32], zs[32];   // own XZ / own Z half row, packed fp16 (the second terms of the updates)
    {
      const uint4* z0 = reinterpret_cast<const uint4*>(Za + row_off + ch0 * 32);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint4 u = __ldcg(z0 + i);
        zs[4 * i] = u.x; zs[4 * i + 1] = u.y; zs[4 * i + 2] = u.z; zs[4 * i + 3] = u.w;
      }
    }
    for (int s = 0; s < PC_STEPS; ++s) {
      cluster_sync_all();
      const int it = s >> 2, g = s & 3;
      const bool last = s == PC_STEPS - 1;
      mbar_wait(mma_done, s & 1);
      if (threadIdx.x == 64) { TL(s, 5); }
      tc_fence_after();
      uint32_t raw[2][32];         // both 32-column chunks in flight before the single wait
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
          // (1) own half of the NEXT step's B operand, MN-major SWIZZLE_128B: k row kk = 128 mi + r, n group = ch / 2
          // (2) the block in K-major SWIZZLE_128B form [128 rows][64 cols] x 2: source of the TMA store that publishes it
          //     to the peers; for XZ (g == 0) that form, placed at k-blocks 2 ni, 2 ni + 1 of sA, IS the own half of the
          //     next step's A operand
          const int kk = 128 * mi + r;
          uint8_t* bdst = sB + (ch >> 1) * 32768 + kk * 128;
          uint8_t* kdst = (g == 0 ? sA + (2 * ni + (ch >> 1)) * 16384 : sStage + (ch >> 1) * 16384) + r * 128;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint4 u = make_uint4(pk[4 * e], pk[4 * e + 1], pk[4 * e + 2], pk[4 * e + 3]);
            const int c = (ch & 1) * 4 + e;
            *reinterpret_cast<uint4*>(bdst + ((c ^ (kk & 7)) << 4)) = u;
            *reinterpret_cast<uint4*>(kdst + ((c ^ (r & 7)) << 4)) = u;
          }
        }
      }
      tc_fence_before();           // TMEM reads retire before the next step's MMA overwrites the accumulator
      if (threadIdx.x == 64) { TL(s, 8); }
      if (!last) {
        fence_proxy_async();       // generic-proxy smem writes -> visible to the TMA store and to the next step's MMA
        named_bar_sync(1, 256);
        if (threadIdx.x == 64) {
          TL(s, 9);
          const CUtensorMap* md = g == 0 ? &maps.XZ : (g == 1 ? &maps.T1 : (g == 2 ? &maps.T2 : ((it & 1) ? &maps.Za : &maps.Zb)));
          const uint8_t* src = g == 0 ? sA + (2 * ni) * 16384 : sStage;
#pragma unroll
          for (int bx = 0; bx < 2; ++bx)
#pragma unroll
            for (int hf = 0; hf < 2; ++hf)
              tma_store_3d(md, src + bx * 16384 + hf * 8192, 128 * ni + 64 * bx, 128 * mi + 64 * hf, bh);
          tma_store_commit();
          tma_store_wait_all<0>();   // the block is in global memory before this thread arrives at the cluster barrier
        }
      }
      if (threadIdx.x == 64) { TL(s, 7); }
    }
  }
  cluster_sync_all();
#ifdef TM_PINV_TIMELINE
  if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) {
    const long long t0 = tl[0][0];
    for (int s = 0; s < 12; ++s)
      printf("step %2d g%d | tma: bar_exit %6lld issued %6lld | mma: bar_exit %6lld B0_full %6lld B3_full %6lld committed %6lld | epi: mma_done %6lld packed %6lld synced %6lld stored %6lld\n",
             s, s & 3, tl[s][0] - t0, tl[s][1] - t0, tl[s][2] - t0, tl[s][6] - t0, tl[s][3] - t0, tl[s][4] - t0, tl[s][5] - t0, tl[s][8] - t0, tl[s][9] - t0, tl[s][7] - t0);
  }
#endif
  tc_fence_after();
  if (warp == 2) tmem_dealloc(tmem_base, 128);
}

// A16 = half(attn2);  Z0 = half(attn2^T / scale)      grid (8, 8, B*8), block (32, 8)
// scale = max_i rowsum * max_j colsum over ALL bags and heads of the call (moore_penrose_iter_pinv's global scalar): every
// CTA reduces the 2 x nbh x 256 sums itself (max is order independent, so all CTAs get the same bits) instead of waiting
// for a one-block kernel in between.
// With many (bag, head) pairs that per-CTA reduction would be quadratic work: the caller then passes `scale` computed by
// pinv_scale_kernel (scale != NULL) and the loop below is skipped.
__global__ void __launch_bounds__(256) pinv_prep16_kernel(const float* __restrict__ A, const float* __restrict__ rowsum,
                                                          const float* __restrict__ colpart, int nbh, int nparts,
                                                          const float* __restrict__ scale, __half* __restrict__ A16,
                                                          __half* __restrict__ Z16) {
  __shared__ float t[32][33];
  __shared__ float red[2][8];
  pdl_sync();
  const int tid = threadIdx.y * 32 + threadIdx.x;
  float mr = 0.f, mc = 0.f;
  if (scale == nullptr) {
    // global init scalar: every CTA needs it, so the scan is kept short -- four columns per thread, the 16 row-tile partials
    // of a head as 16 independent 16-byte loads (fixed summation order p = 0 .. 15: same bits in every CTA and every run)
    const int j4 = (tid & 63) * 4;
    for (int bh = tid >> 6; bh < nbh; bh += 4) {
      const float4 rs = __ldg(reinterpret_cast<const float4*>(rowsum + bh * TM_LANDMARKS + j4));
      mr = fmaxf(mr, fmaxf(fmaxf(rs.x, rs.y), fmaxf(rs.z, rs.w)));
      float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
      if (nparts == 16) {
        float4 v[16];
#pragma unroll
        for (int p = 0; p < 16; ++p) v[p] = __ldg(reinterpret_cast<const float4*>(colpart + ((long long)bh * 16 + p) * TM_LANDMARKS + j4));
#pragma unroll
        for (int p = 0; p < 16; ++p) { c.x += v[p].x; c.y += v[p].y; c.z += v[p].z; c.w += v[p].w; }
      } else {
        for (int p = 0; p < nparts; ++p) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(colpart + ((long long)bh * nparts + p) * TM_LANDMARKS + j4));
          c.x += v.x; c.y += v.y; c.z += v.z; c.w += v.w;
        }
      }
      mc = fmaxf(mc, fmaxf(fmaxf(c.x, c.y), fmaxf(c.z, c.w)));
    }
  }
  mr = warp_max(mr); mc = warp_max(mc);
  if (threadIdx.x == 0) { red[0][threadIdx.y] = mr; red[1][threadIdx.y] = mc; }
  const long long base = (long long)blockIdx.z * TM_LANDMARKS * TM_LANDMARKS;
  const int x0 = blockIdx.x * 32, y0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const long long o = base + (long long)(y0 + i) * TM_LANDMARKS + x0 + threadIdx.x;
    const float a = A[o];
    t[i][threadIdx.x] = a;
    A16[o] = __float2half_rn(a);
  }
  __syncthreads();
  mr = red[0][0]; mc = red[1][0];
#pragma unroll
  for (int w = 1; w < 8; ++w) { mr = fmaxf(mr, red[0][w]); mc = fmaxf(mc, red[1][w]); }
  const float sc = scale != nullptr ? scale[0] : mr * mc;
  for (int i = threadIdx.y; i < 32; i += 8)
    Z16[base + (long long)(x0 + i) * TM_LANDMARKS + y0 + threadIdx.x] = __float2half_rn(t[threadIdx.x][i] / sc);
}

int pinv_push_launch(__half* A16, __half* Za, __half* Zb, int nbh, float* Z, cudaStream_t st);   // pinv_push.cu
int pinv_pipe_launch(__half* A16, __half* Za, __half* Zb, int nbh, float* Z, cudaStream_t st);   // pinv_pipe.cu

// scratch16: 6 fp16 matrices per (bag, head): A16, Za, Zb, XZ, T1, T2
int pinv_cluster_run(const float* attn2, const float* rowsum, const float* colpart, int nparts, const float* scale, int nbh,
                     void* scratch16, float* Z, cudaStream_t st) {
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "pinv_cluster_run needs an sm_100 device");
  const long long mat = (long long)TM_LANDMARKS * TM_LANDMARKS;
  __half* A16 = (__half*)scratch16;
  __half* Za = A16 + nbh * mat;
  __half* Zb = Za + nbh * mat;
  __half* XZ = Zb + nbh * mat;
  __half* T1 = XZ + nbh * mat;
  __half* T2 = T1 + nbh * mat;
  launch_pdl(pinv_prep16_kernel, dim3(8, 8, nbh), dim3(32, 8), 0, st, attn2, rowsum, colpart, nbh, nparts, scale, A16, Za);
  TM_LAUNCH_CHECK();
  // default: the distributed-shared-memory kernel (pinv_push.cu); TM_PINV_KERNEL=tma selects the first-generation kernel
  // of this file (peers exchange their blocks through global memory with TMA stores / loads), kept as a cross-check
  static int use_push = -1;
  if (use_push < 0) {
    const char* e = getenv("TM_PINV_KERNEL");
    use_push = (e != nullptr && strcmp(e, "tma") == 0) ? 0 : ((e != nullptr && (strcmp(e, "push") == 0 || strcmp(e, "remote") == 0)) ? 1 : 2);
  }
  if (use_push == 2) return pinv_pipe_launch(A16, Za, Zb, nbh, Z, st);
  if (use_push) return pinv_push_launch(A16, Za, Zb, nbh, Z, st);
  PinvMaps maps;
  __half* ptrs[6] = {A16, Za, Zb, XZ, T1, T2};
  CUtensorMap* ms[6] = {&maps.A, &maps.Za, &maps.Zb, &maps.XZ, &maps.T1, &maps.T2};
  for (int i = 0; i < 6; ++i)
    TM_PROPAGATE(make_tmap_3d_f16(ms[i], ptrs[i], nbh, TM_LANDMARKS, TM_LANDMARKS, mat, TM_LANDMARKS, 64, 64));
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done)) {
    TM_CUDA(cudaFuncSetAttribute(pinv_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PC_SMEM));
  }
  pinv_cluster_kernel<<<dim3(4, nbh), PC_THREADS, PC_SMEM, st>>>(maps, Za, Zb, XZ, T1, T2, Z);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

}  // namespace tmil
