// tcgen05 strided batched GEMM of the BACKWARD:  C[z] = alpha * op(A[z]) op(B[z]) + beta * L[z] + diag * I,
// fp32 tensors in HBM, tf32 tensor-core contraction, fp32 accumulation in TMEM.
//
// The backward's contractions come in all four transposition forms (dX = dY W is NN, dW = dY^T X is TN, the attention
// score gradients are NT, the pinv chain has TT).  tcgen05 takes tf32 operands K-major (k contiguous) in the plain
// SWIZZLE_128B shared-memory layout; MN-major tf32 would need the 32-byte-atom swizzle variant.  So the operand tiles are
// always laid out K-major in shared memory and the transposition is done by the LOADER on the way in:
//   * source k-contiguous:  16-byte cp.async per (row, 4 k) chunk straight to its swizzled place;
//   * source m-contiguous:  4-byte cp.async per element, lanes arranged 8 rows x 4 k so that the global side reads whole
//     32-byte sectors and the shared side hits 32 different banks (chunk index XOR row, as the TMA unit would write it).
// Out-of-range rows / k are zero-filled by the copy (src-size 0), so ragged M, N, K need no separate path.
//
// One persistent CTA per SM, 13 warps:
//   warp 0      MMA issuer (one lane): 4 x tcgen05.mma.kind::tf32 (M128 x Nt x K8) per 32-wide k block, accumulators double
//               buffered in TMEM (2 x Nt columns) so the epilogue of tile i runs under the main loop of tile i+1
//   warps 1-4   epilogue: tcgen05.ld -> per-warp smem transpose -> coalesced 128-bit stores (alpha, beta*L, diag applied)
//   warps 5-12  loaders: 4-stage ring, 3 cp.async groups in flight per thread; a stage is published with
//               cp.async.wait_group -> fence.proxy.async -> mbarrier.arrive (256 arrivals per stage)
// Operands that satisfy the TMA rules (16-byte aligned base and strides) skip the loader warps: <TMA = true> feeds the same
// ring with cp.async.bulk.tensor from ONE thread.  A k-contiguous operand arrives as a [rows x 32 k] SWIZZLE_128B box (K-major
// descriptor); an m-contiguous operand arrives as [32 k x 32 m] boxes written with the 32-byte-atom 128B swizzle
// (CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B) and is multiplied MN-major (descriptor layout type 1, the only MN-major form tf32 has):
// no transposition anywhere.
// Split-K (weight gradients: M, N small, K = all tokens of the batch): the K range is cut into `splits` slices that become
// extra tiles writing fp32 partials; a second kernel sums them in a fixed order (deterministic, no atomics).
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace tmil {

typedef long long i64;

struct BGemmTc {
  const float* A; const float* B; const float* L; float* C;
  int M, N, K;
  i64 sam, sak, sbk, sbn, ldc, ldl;
  i64 strA, strB, strC, strL;
  float alpha, beta, diag;
  float* C2; float alpha2, diag2;   // optional second output C2 = alpha2 * acc + diag2 * I (ldc / strC of C; not with split-K)
  int a_km, b_km;          // operand is k-contiguous in memory
  int bn;                  // N tile (64 / 128 / 256)
  int m_tiles, n_tiles;
  int splits, kb_per_split, k_blocks;   // k blocks of 32
  int total;               // tiles = batch * splits * m_tiles * n_tiles
  uint32_t idesc;
  int c2_vec;
  int c_vec;               // C base and batch stride allow 16-byte stores
  int l_vec;               // L base and batch stride allow 16-byte loads
  int a_mn, b_mn;          // <TMA>: operand is multiplied MN-major (m-contiguous boxes); idesc carries the matching major bits
};
struct BGemmMaps { CUtensorMap a, b; };

constexpr int BG_BM = 128;
constexpr int BG_BK = 32;
constexpr int BG_STAGES = 4;
constexpr int BG_LOOKAHEAD = 3;                 // cp.async groups in flight per loader thread
constexpr int BG_A_BYTES = BG_BM * 128;         // 16 KB
constexpr int BG_B_BYTES = 256 * 128;           // 32 KB (sized for the widest N tile)
constexpr int BG_STG_LD = 36;
constexpr int BG_STG_BYTES = 32 * BG_STG_LD * 4;
constexpr int BG_LOADERS = 256;
constexpr int BG_THREADS = 32 + 128 + BG_LOADERS;
constexpr size_t BG_SMEM = 1024 + BG_STAGES * (BG_A_BYTES + BG_B_BYTES) + 4 * BG_STG_BYTES + 256;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// One operand tile [rows x 32 k] into the K-major SWIZZLE_128B layout at `sdst` (shared address):
//   element (r, k) -> r * 128 + (((k >> 2) ^ (r & 7)) << 4) + (k & 3) * 4
// src(r, k) = base + r * s_row + k * s_k (elements); rows >= row_lim and k >= k_lim read as zero.
// lt = loader thread index (0..255).
__device__ __forceinline__ void load_tile(uint32_t sdst, const float* base, int rows, int row_lim, int k_lim, i64 s_row, i64 s_k,
                                          bool k_major, int lt) {
  if (k_major) {
    // thread: chunk c = lt & 7 (k = 4c .. 4c+3), rows (lt >> 3) + 32 i
    const int c = lt & 7;
    const bool k_ok = 4 * c < k_lim;            // K % 4 == 0: a chunk is entirely inside or outside
    const float* src_c = base + 4 * c;
    for (int r = lt >> 3; r < rows; r += 32) {
      const bool ok = k_ok && r < row_lim;
      const float* src = ok ? src_c + (i64)r * s_row : base;
      cp_async16(sdst + r * 128 + ((c ^ (r & 7)) << 4), src, ok ? 16u : 0u);
    }
  } else {
    // warp w: k = 4 w + ki; lane: mi = lane & 7, ki = lane >> 3; row blocks of 8
    const int w = lt >> 5, lane = lt & 31;
    const int mi = lane & 7, ki = lane >> 3;
    const int k = 4 * w + ki;
    const bool k_ok = k < k_lim;
    const float* src_k = base + (i64)k * s_k;
    const uint32_t d0 = sdst + mi * 128 + ((w ^ mi) << 4) + ki * 4;
    for (int rb = 0; rb < rows; rb += 8) {
      const int r = rb + mi;
      const bool ok = k_ok && r < row_lim;
      const float* src = ok ? src_k + (i64)r * s_row : base;
      cp_async4(d0 + rb * 128, src, ok ? 4u : 0u);
    }
  }
}

// MN-major tf32 operand tile: [32 k rows][32 mn floats] boxes, 128 B per k row, 32-byte chunks XOR-ed with (row & 3); the canonical
// UMMA layout ((8,n),(4,k)):((1,LBO),(8,SBO)) in 16-byte units: LBO = distance between 32-float MN groups (one box = 4096 B),
// SBO = distance between 4-row k groups (512 B).  One MMA (K = 8) reads two k groups: start address + ks * 1024.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128_32b(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(lbo_bytes >> 4) << 16;
  d |= (uint64_t)(sbo_bytes >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)1 << 61;   // SWIZZLE_128B_BASE32B
  return d;
}

template <bool TMA>
__global__ void __launch_bounds__(BG_THREADS, 1) bgemm_tc_kernel(const BGemmTc p, const __grid_constant__ BGemmMaps maps) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = sA + BG_STAGES * BG_A_BYTES;
  float* stg = reinterpret_cast<float*>(sB + BG_STAGES * BG_B_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(stg) + 4 * BG_STG_BYTES);
  uint64_t* full = bars;                   // [STAGES] loaders -> MMA
  uint64_t* empty = bars + BG_STAGES;      // [STAGES] MMA -> loaders
  uint64_t* tfull = bars + 2 * BG_STAGES;  // [2] MMA -> epilogue
  uint64_t* tempty = tfull + 2;            // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bn = p.bn;
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < BG_STAGES; ++i) { mbar_init(&full[i], TMA ? 1 : BG_LOADERS); mbar_init(&empty[i], 1); }
    if (TMA) { tma_prefetch_desc(&maps.a); tma_prefetch_desc(&maps.b); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 4); }
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 2 * bn);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int tiles_mn = p.m_tiles * p.n_tiles;

  if (warp == 0) {
    // ---- MMA issuer ------------------------------------------------------------------------------------------------
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x, ++it) {
        const int zs = tile / tiles_mn;
        const int split = zs % p.splits;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(p.k_blocks, kb0 + p.kb_per_split);
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tempty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * bn;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * BG_A_BYTES), b_addr = smem_u32(sB + stage * BG_B_BYTES);
          // per K = 8 step: K-major advances 32 bytes inside the 128-byte row, MN-major two 4-row k groups (1024 bytes)
          const uint64_t a_desc = (TMA && p.a_mn) ? umma_desc_mn_sw128_32b(a_addr, 4096, 512) : umma_desc_k_sw128(a_addr);
          const uint64_t b_desc = (TMA && p.b_mn) ? umma_desc_mn_sw128_32b(b_addr, 4096, 512) : umma_desc_k_sw128(b_addr);
          const uint64_t a_step = (TMA && p.a_mn) ? 64 : 2, b_step = (TMA && p.b_mn) ? 64 : 2;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_tf32_ss(d_tmem, a_desc + a_step * k, b_desc + b_step * k, p.idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit(&empty[stage]);
          if (++stage == BG_STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull[as]);
      }
    }
    __syncwarp();
  } else if (warp >= 5) {
    // ---- loaders -----------------------------------------------------------------------------------------------------
    if (TMA) {
      // one thread feeds the ring: 3-D maps (inner, rows, batch); coordinates outside the extents are zero-filled
      if (warp == 5 && lane == 0) {
        int stage = 0; uint32_t phase = 0;
        const uint32_t tx = BG_A_BYTES + (uint32_t)bn * 128u;
        for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x) {
          const int zs = tile / tiles_mn, rem = tile - zs * tiles_mn;
          const int z = zs / p.splits, split = zs - z * p.splits;
          const int m_blk = rem / p.n_tiles, n_blk = rem - m_blk * p.n_tiles;
          const int kb0 = split * p.kb_per_split;
          const int kb1 = min(p.k_blocks, kb0 + p.kb_per_split);
          const int m0 = m_blk * BG_BM, n0 = n_blk * bn;
          for (int kb = kb0; kb < kb1; ++kb) {
            const int k0 = kb * BG_BK;
            mbar_wait(&empty[stage], phase ^ 1);
            mbar_arrive_expect_tx(&full[stage], tx);
            uint8_t* a_dst = sA + stage * BG_A_BYTES;
            uint8_t* b_dst = sB + stage * BG_B_BYTES;
            if (p.a_mn) {
#pragma unroll
              for (int g = 0; g < BG_BM / 32; ++g) tma_load_3d(a_dst + g * 4096, &maps.a, &full[stage], m0 + 32 * g, k0, z);
            } else {
              tma_load_3d(a_dst, &maps.a, &full[stage], k0, m0, z);
            }
            if (p.b_mn) {
              for (int g = 0; g < (bn >> 5); ++g) tma_load_3d(b_dst + g * 4096, &maps.b, &full[stage], n0 + 32 * g, k0, z);
            } else {
              tma_load_3d(b_dst, &maps.b, &full[stage], k0, n0, z);
            }
            if (++stage == BG_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    } else {
    const int lt = threadIdx.x - 160;
    int stage = 0; uint32_t phase = 0;
    int pub = 0;        // next stage to publish
    int inflight = 0;   // committed groups not yet published
    for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x) {
      const int zs = tile / tiles_mn, rem = tile - zs * tiles_mn;
      const int z = zs / p.splits, split = zs - z * p.splits;
      const int m_blk = rem / p.n_tiles, n_blk = rem - m_blk * p.n_tiles;
      const int kb0 = split * p.kb_per_split;
      const int kb1 = min(p.k_blocks, kb0 + p.kb_per_split);
      const int m0 = m_blk * BG_BM, n0 = n_blk * bn;
      const float* Ab = p.A + (i64)z * p.strA + (i64)m0 * p.sam;
      const float* Bb = p.B + (i64)z * p.strB + (i64)n0 * p.sbn;
      const int m_lim = p.M - m0, n_lim = p.N - n0;
      for (int kb = kb0; kb < kb1; ++kb) {
        const int k0 = kb * BG_BK;
        mbar_wait(&empty[stage], phase ^ 1);
        load_tile(smem_u32(sA + stage * BG_A_BYTES), Ab + (i64)k0 * p.sak, BG_BM, m_lim, p.K - k0, p.sam, p.sak, p.a_km != 0, lt);
        load_tile(smem_u32(sB + stage * BG_B_BYTES), Bb + (i64)k0 * p.sbk, bn, n_lim, p.K - k0, p.sbn, p.sbk, p.b_km != 0, lt);
        cp_async_commit();
        if (++stage == BG_STAGES) { stage = 0; phase ^= 1; }
        if (++inflight == BG_LOOKAHEAD) {
          cp_async_wait<BG_LOOKAHEAD - 1>();
          fence_proxy_async();
          mbar_arrive(&full[pub]);
          if (++pub == BG_STAGES) pub = 0;
          --inflight;
        }
      }
    }
    // drain
    if (inflight == 2) {
      cp_async_wait<1>();
      fence_proxy_async();
      mbar_arrive(&full[pub]);
      if (++pub == BG_STAGES) pub = 0;
      --inflight;
    }
    if (inflight == 1) {
      cp_async_wait<0>();
      fence_proxy_async();
      mbar_arrive(&full[pub]);
    }
    }
  } else {
    // ---- epilogue: TMEM -> registers -> per-warp smem transpose -> coalesced 128-bit global accesses ------------------------
    const int q = warp & 3;   // TMEM lane quarter this warp may access
    float* my_stg = stg + (warp - 1) * (32 * BG_STG_LD);
    const int sub = lane >> 3, cl = (lane & 7) * 4;   // store phase: rows sub, sub+4, ...; 4 columns at cl
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x, ++it) {
      const int zs = tile / tiles_mn, rem = tile - zs * tiles_mn;
      const int m_blk = rem / p.n_tiles, n_blk = rem - m_blk * p.n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      // with split-K the partials of (z, split) are stored back to back: zs indexes them directly
      const int zl = p.splits > 1 ? zs / p.splits : zs;
      float* Cz = p.C + (i64)zs * p.strC;
      float* C2z = p.C2 != nullptr ? p.C2 + (i64)zs * p.strC : nullptr;
      const float* Lz = p.L != nullptr ? p.L + (i64)zl * p.strL : nullptr;
      mbar_wait(&tfull[as], aphase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + as * bn;
      const int nchunks = bn >> 5;
#pragma unroll 1
      for (int chunk = 0; chunk < nchunks; ++chunk) {
        float v[32];
        tmem_ld_32x32(t_row + chunk * 32, v);
        if (chunk == nchunks - 1) {   // accumulator fully read: hand the TMEM buffer back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty[as]);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
          *reinterpret_cast<float4*>(my_stg + lane * BG_STG_LD + i * 4) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        __syncwarp();
        const int c = n_blk * bn + chunk * 32 + cl;
        if (c < p.N) {
          float4 acc[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] = *reinterpret_cast<const float4*>(my_stg + (i * 4 + sub) * BG_STG_LD + cl);
          const bool vec = (c + 3 < p.N) && ((p.ldc & 3) == 0) && p.c_vec;
          // second term: every load of the chunk in flight before its first store (L may alias C, so the compiler cannot
          // hoist them itself; a thread only ever reads the elements it is about to overwrite)
          float4 lv[8];
          if (Lz != nullptr) {
            const bool lvec = (c + 3 < p.N) && ((p.ldl & 3) == 0) && p.l_vec;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int r = m_blk * BG_BM + q * 32 + i * 4 + sub;
              lv[i] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (r >= p.M) continue;
              const float* lp = Lz + (i64)r * p.ldl + c;
              if (lvec) {
                lv[i] = *reinterpret_cast<const float4*>(lp);
              } else {
                lv[i].x = lp[0];
                if (c + 1 < p.N) lv[i].y = lp[1];
                if (c + 2 < p.N) lv[i].z = lp[2];
                if (c + 3 < p.N) lv[i].w = lp[3];
              }
            }
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int r = m_blk * BG_BM + q * 32 + i * 4 + sub;
            if (r >= p.M) continue;
            float o[4] = {p.alpha * acc[i].x, p.alpha * acc[i].y, p.alpha * acc[i].z, p.alpha * acc[i].w};
            if (Lz != nullptr) {
              o[0] = fmaf(p.beta, lv[i].x, o[0]); o[1] = fmaf(p.beta, lv[i].y, o[1]);
              o[2] = fmaf(p.beta, lv[i].z, o[2]); o[3] = fmaf(p.beta, lv[i].w, o[3]);
            }
            if (p.diag != 0.f) {
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (r == c + e) o[e] += p.diag;
            }
            float* dst = Cz + (i64)r * p.ldc + c;
            if (vec) {
              *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (c + e < p.N) dst[e] = o[e];
            }
            if (C2z != nullptr) {
              float o2[4] = {p.alpha2 * acc[i].x, p.alpha2 * acc[i].y, p.alpha2 * acc[i].z, p.alpha2 * acc[i].w};
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (r == c + e) o2[e] += p.diag2;
              float* dst2 = C2z + (i64)r * p.ldc + c;
              if (vec && p.c2_vec) {
                *reinterpret_cast<float4*>(dst2) = make_float4(o2[0], o2[1], o2[2], o2[3]);
              } else {
#pragma unroll
                for (int e = 0; e < 4; ++e)
                  if (c + e < p.N) dst2[e] = o2[e];
              }
            }
          }
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * bn);
  }
}

// out[z][m,n] = alpha * sum_s P[z][s][m,n] + beta * L[z][m*ldl + n] + (m == n ? diag : 0); fixed summation order
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ P, int splits, i64 mn, int N, i64 batch,
                                                            const float* L, i64 ldl, i64 strL, float* C, i64 ldc, i64 strC,
                                                            float alpha, float beta, float diag) {
  const i64 i = (i64)blockIdx.x * 256 + threadIdx.x;
  if (i >= batch * mn) return;
  const i64 z = i / mn, e = i - z * mn;
  const i64 m = e / N, n = e - m * N;
  const float* src = P + z * splits * mn + e;
  float a = 0.f;
  for (int s = 0; s < splits; ++s) a += src[(i64)s * mn];
  float o = alpha * a;
  if (L != nullptr) o = fmaf(beta, L[z * strL + m * ldl + n], o);
  if (m == n) o += diag;
  C[z * strC + m * ldc + n] = o;
}

int g_bgemm_force_cpasync = 0;   // tm_bgemm_engine(2): loader-warp path even for TMA-able operands

static int bg_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  return sms;
}

// Split-K plan of one call: number of K slices (1 = none) and k blocks per slice.
static void bg_split_plan(i64 M, i64 N, i64 K, i64 batch, int bn, int* splits, int* kb_per_split) {
  const int k_blocks = (int)ceil_div(K, BG_BK);
  const i64 tiles = batch * ceil_div(M, BG_BM) * ceil_div(N, bn);
  const int sms = bg_sms() > 0 ? bg_sms() : 148;
  int s = 1;
  if (tiles * 2 <= sms && k_blocks >= 16) {
    s = (int)(sms / tiles);
    const int max_s = k_blocks / 8;          // at least 8 k blocks (256 k) per slice
    if (s > max_s) s = max_s;
    if (s < 1) s = 1;
  }
  int per = (int)ceil_div(k_blocks, s);
  s = (int)ceil_div(k_blocks, per);           // no empty slice
  *splits = s;
  *kb_per_split = per;
}
static int bg_pick_bn(i64 N) { return N <= 64 ? 64 : (N <= 128 ? 128 : 256); }

i64 bgemm_tc_scratch_floats(i64 M, i64 N, i64 K, i64 batch) {
  int s, per;
  bg_split_plan(M, N, K, batch, bg_pick_bn(N), &s, &per);
  return s > 1 ? batch * s * M * N : 0;
}

// Same contract as the CUDA-core bgemm of backward.cu.  scratch: bgemm_tc_scratch_floats() floats (may be NULL: then no split-K).
int bgemm_tc(const float* A, const float* B, float* C, const float* L, i64 M, i64 N, i64 K, i64 sam, i64 sak, i64 sbk, i64 sbn, i64 ldc,
             i64 ldl, i64 batch, i64 strA, i64 strB, i64 strC, i64 strL, float alpha, float beta, float diag, float* C2, float alpha2,
             float diag2, float* scratch, i64 scratch_floats, cudaStream_t st) {
  TM_CHECK_ARG(M > 0 && N > 0 && K > 0 && batch > 0, TM_ERR_BAD_SHAPE, "bgemm_tc: empty problem");
  TM_CHECK_ARG(M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), TM_ERR_BAD_SHAPE, "bgemm_tc: dimension too large");
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "bgemm_tc needs an sm_100 device");
  static unsigned long long attr_done = 0;
  if (first_use_on_device(&attr_done)) {
    TM_CUDA(cudaFuncSetAttribute(bgemm_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BG_SMEM));
    TM_CUDA(cudaFuncSetAttribute(bgemm_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BG_SMEM));
  }
  BGemmTc p;
  memset(&p, 0, sizeof(p));
  p.A = A; p.B = B; p.L = L; p.C = C;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.sam = sam; p.sak = sak; p.sbk = sbk; p.sbn = sbn; p.ldc = ldc; p.ldl = ldl;
  p.strA = strA; p.strB = strB; p.strC = strC; p.strL = strL;
  p.alpha = alpha; p.beta = beta; p.diag = diag;
  p.C2 = C2; p.alpha2 = alpha2; p.diag2 = diag2;
  p.c2_vec = (C2 != nullptr && aligned16(C2)) ? 1 : 0;
  if (C2 != nullptr) { scratch = nullptr; scratch_floats = 0; }   // a second output rules out split-K
  // the 16-byte copy path needs k-contiguous, 16-byte aligned rows and K % 4 == 0; anything else takes the element path
  auto vec_ok = [&](const float* ptr, i64 s_row, i64 s_k, i64 str) {
    return s_k == 1 && (K & 3) == 0 && (s_row & 3) == 0 && (str & 3) == 0 && aligned16(ptr);
  };
  p.a_km = vec_ok(A, sam, sak, strA) ? 1 : 0;
  p.b_km = vec_ok(B, sbn, sbk, strB) ? 1 : 0;
  p.l_vec = (L != nullptr && aligned16(L) && (strL & 3) == 0) ? 1 : 0;
  p.bn = bg_pick_bn(N);
  p.idesc = umma_idesc_tf32(BG_BM, p.bn);
  p.m_tiles = (int)ceil_div(M, BG_BM);
  p.n_tiles = (int)ceil_div(N, p.bn);
  p.k_blocks = (int)ceil_div(K, BG_BK);
  bg_split_plan(M, N, K, batch, p.bn, &p.splits, &p.kb_per_split);
  if (p.splits > 1 && (scratch == nullptr || scratch_floats < batch * p.splits * M * N)) {
    p.splits = 1;
    p.kb_per_split = p.k_blocks;
  }
  p.c_vec = (p.splits > 1) ? ((((i64)M * N) & 3) == 0 && aligned16(scratch)) : (aligned16(C) && (strC & 3) == 0);
  const i64 total = batch * p.splits * p.m_tiles * p.n_tiles;
  TM_CHECK_ARG(total < (1ll << 31), TM_ERR_BAD_SHAPE, "bgemm_tc: too many tiles");
  p.total = (int)total;
  if (p.splits > 1) {   // partials: [batch][splits][M][N], plain stores
    p.C = scratch; p.ldc = N; p.strC = M * N; p.L = nullptr; p.alpha = 1.f; p.beta = 0.f; p.diag = 0.f;
  }
  const int sms = bg_sms();
  const int grid = total < sms ? (int)total : sms;
  // TMA feed when both operands have a unit stride and obey the tensor-map rules (16-byte aligned base and pitches);
  // TM_BGEMM_LOADER=cpasync keeps everything on the loader-warp path (the cross-check of the TMA path)
  static const bool tma_off = [] { const char* e = getenv("TM_BGEMM_LOADER"); return e && strcmp(e, "cpasync") == 0; }();
  auto tma_ok = [&](const float* ptr, i64 s_row, i64 s_k, i64 str, int* mn) {
    if (!aligned16(ptr) || (batch > 1 && (str & 3) != 0)) return false;
    if (s_k == 1 && (s_row & 3) == 0) { *mn = 0; return true; }
    if (s_row == 1 && (s_k & 3) == 0) { *mn = 1; return true; }
    return false;
  };
  BGemmMaps maps;
  memset(&maps, 0, sizeof(maps));
  bool use_tma = !tma_off && !g_bgemm_force_cpasync && tma_ok(A, sam, sak, strA, &p.a_mn) && tma_ok(B, sbn, sbk, strB, &p.b_mn);
  if (use_tma) {
    // single problem: the batch extent is 1, its pitch only has to be a legal number (rows x row pitch)
    const uint64_t bsA = batch > 1 ? (uint64_t)strA : (p.a_mn ? (uint64_t)K * sak : (uint64_t)M * sam);
    const uint64_t bsB = batch > 1 ? (uint64_t)strB : (p.b_mn ? (uint64_t)K * sbk : (uint64_t)N * sbn);
    int rc;
    if (p.a_mn) rc = make_tmap_3d_mn32(&maps.a, A, (uint64_t)batch, (uint64_t)K, (uint64_t)M, bsA, (uint64_t)sak, BG_BK);
    else rc = make_tmap_3d(&maps.a, A, (uint64_t)batch, (uint64_t)M, (uint64_t)K, bsA, (uint64_t)sam, BG_BM, BG_BK);
    if (rc == TM_OK) {
      if (p.b_mn) rc = make_tmap_3d_mn32(&maps.b, B, (uint64_t)batch, (uint64_t)K, (uint64_t)N, bsB, (uint64_t)sbk, BG_BK);
      else rc = make_tmap_3d(&maps.b, B, (uint64_t)batch, (uint64_t)N, (uint64_t)K, bsB, (uint64_t)sbn, (uint32_t)p.bn, BG_BK);
    }
    if (rc != TM_OK) use_tma = false;   // an extent the encoder refuses: the loader-warp path takes anything
  }
  if (use_tma) {
    p.idesc = umma_idesc_tf32(BG_BM, p.bn, p.a_mn, p.b_mn);
    bgemm_tc_kernel<true><<<grid, BG_THREADS, BG_SMEM, st>>>(p, maps);
  } else {
    p.a_mn = p.b_mn = 0;
    bgemm_tc_kernel<false><<<grid, BG_THREADS, BG_SMEM, st>>>(p, maps);
  }
  TM_LAUNCH_CHECK();
  if (p.splits > 1) {
    const i64 n_out = batch * M * N;
    splitk_reduce_kernel<<<(unsigned)ceil_div(n_out, 256), 256, 0, st>>>(scratch, p.splits, M * N, (int)N, batch, L, ldl, strL, C, ldc, strC,
                                                                        alpha, beta, diag);
    TM_LAUNCH_CHECK();
  }
  return TM_OK;
}

}  // namespace tmil
