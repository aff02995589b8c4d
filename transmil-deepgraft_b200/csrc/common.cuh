// Shared host/device helpers of libtransmil_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "transmil_b200.h"

namespace tmil {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

enum Stage : int { ST_FC1_GEMM0 = 0, ST_FC1_LN, ST_FC1_GEMM1, ST_FC1_TAIL, ST_LN_PAD, ST_QKV, ST_LANDMARKS, ST_PINV, ST_ATTN3V,
                   ST_ZW, ST_NYSTROM_OUT, ST_TO_OUT, ST_PPEG, ST_HEAD,
                   // token-sharded forward only: operand preparation and the exchanges (publish + wait + pull)
                   ST_PREP, ST_X_LANDMARKS, ST_X_ATTN3V, ST_X_CONV_HALO, ST_X_PPEG_HALO, ST_X_LOGITS, ST_END, ST_COUNT };
void prof_mark(int stage, cudaStream_t st);
bool prof_enabled();
// Library-owned high-priority side stream of the current device plus a fork/join event pair (created on first use): lets
// two independent stages of one forward run concurrently on disjoint SMs.  Fork/join through events, so the pattern is also
// legal inside a stream capture.
struct SideStream { cudaStream_t stream; cudaEvent_t fork, join; };
int side_stream(SideStream* out);

#define TM_CHECK_ARG(cond, code, ...)  \
  do {                                 \
    if (!(cond)) {                     \
      tmil::set_error(__VA_ARGS__);      \
      return (code);                   \
    }                                  \
  } while (0)

#define TM_CUDA(expr)                                                              \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) {                                                       \
      tmil::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return TM_ERR_CUDA;                                                          \
    }                                                                              \
  } while (0)

#define TM_LAUNCH_CHECK()                                                          \
  do {                                                                             \
    tmil::count_launch();                                                            \
    cudaError_t _e = cudaGetLastError();                                           \
    if (_e != cudaSuccess) {                                                       \
      tmil::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return TM_ERR_CUDA;                                                          \
    }                                                                              \
  } while (0)

#define TM_PROPAGATE(expr)     \
  do {                         \
    int _s = (expr);           \
    if (_s != TM_OK) return _s; \
  } while (0)

// Programmatic dependent launch.  A kernel launched through launch_pdl() may be scheduled while the previous kernel of the
// stream is still running: its CTAs take SM slots as the predecessor's CTAs exit, run their prologue (barrier init, TMEM
// allocation, tensor-map prefetch) and block in pdl_wait() until the predecessor has completed and flushed its memory.
// Rules kept by every kernel that is launched this way: nothing produced by another kernel is read, and nothing is written
// to global memory, before pdl_wait(); pdl_trigger() comes after the kernel's own pdl_wait(), so at most one generation of
// dependents is ever resident early and "my wait returned" implies "every earlier kernel of the stream finished".
// TM_PDL=0 in the environment launches everything the ordinary way (the device-side calls are then no-ops).
bool pdl_enabled();
// The next launch_pdl() of this thread is an ordinary launch (used for the first kernel after a cross-stream event wait).
void pdl_plain_next();
void pdl_suspend(bool on);   // every launch_pdl() of this thread is ordinary until switched back
bool pdl_take_plain();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() { pdl_wait(); pdl_trigger(); }
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl_cluster(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster_x,
                                             Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  int n = 0;
  if (cluster_x > 1) {
    at[n].id = cudaLaunchAttributeClusterDimension;
    at[n].val.clusterDim.x = (unsigned)cluster_x;
    at[n].val.clusterDim.y = 1;
    at[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl_enabled() && !pdl_take_plain()) {
    at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = at;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  return launch_pdl_cluster(kern, grid, block, smem, st, 1, static_cast<Args&&>(args)...);
}
#endif

// cudaFuncSetAttribute applies to the current device only: `done` is a per-call-site bit mask of the devices that already
// have the attribute (one process per GPU is the normal deployment, but a process may also drive several devices).
static inline bool first_use_on_device(unsigned long long* done) {
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (*done & bit) return false;
  *done |= bit;
  return true;
}

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

__device__ __forceinline__ float gelu_erf(float x) {
  // nn.GELU() default (approximate='none'): 0.5 x (1 + erf(x / sqrt 2))
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- generic fp32 CUDA-core GEMM (gemm_simt.cu) ---------------------------
enum GemmEpi : int {
  EPI_AXPBY = 0,      // C = alpha*acc + beta*L[r,c]           (pinv steps, Z@W3)
  EPI_BIAS_GELU = 1,  // C[map(r)] = gelu(acc + bias[c]) (+ duplicate-pad rows)   fc1
  EPI_QKV = 2,        // head-major q(*0.125),k,v split
  EPI_BIAS_RESID = 3, // C = acc + bias[c] + R[r,c]            to_out + residual
  EPI_AXPBY_T = 5,    // o = alpha*acc + beta*LT[c,r]; stores C[r,c] (if C) and CT[c,r] (if CT): the pinv chain keeps every
                      // iterate in both orientations so each product is an NT GEMM (B operand = transposed copy)
  EPI_QKV16 = 4,      // head-major q(*0.125),k,v split stored as fp16 (p.q/k/v point at __half arrays)
};

struct GemmParams {
  const float* A; const float* Bm; float* C;
  int M, N, K;
  int lda, ldb, ldc;
  int b_is_kn;                 // 0: B is [N,K] (weights, K contiguous); 1: B is [K,N]
  int batch;                   // grid.z
  long long strideA, strideB, strideC, strideL;
  int epi;
  float alpha, beta;
  const float* L; int ldl;     // EPI_AXPBY second term / EPI_BIAS_RESID residual / EPI_AXPBY_T transposed second term
  float* CT;                   // EPI_AXPBY_T transposed output (same ldc / strideC as C)
  const float* bias;
  // C row map: logical row r in [0,M): bag = r / rows_per_bag, i = r % rows_per_bag;
  // rows with i < c_skip are not stored; C row = bag*c_bag_rows + c_row_off + (i - c_skip)
  int rows_per_bag;            // 0 = identity
  long long c_bag_rows; int c_row_off; int c_skip;
  int dup;                     // EPI_BIAS_GELU: rows i < dup are also written rows_per_bag rows further down
  int round_out;               // round stored values to tf32 (consumer is a tensor-core contraction)
  int f16_in;                  // A and Bm point at fp16 operands (tensor-core engine only)
  int bn128;                   // tensor-core engine: use the 128-wide N tile (small batched problems)
  // EPI_QKV
  float* q; float* k; float* v; int Sp;
  int col_off;                 // EPI_QKV16: logical column of C column 0 (a GEMM over a column slice of to_qkv: Bm points at row col_off)
  int max_ctas;                // tensor-core engine: cap on the persistent grid (0 = one CTA per SM); leaves SMs to a concurrent stream
  // EPI_QKV16, TMA-store epilogue only: per-tile column sums of the stored fp16 q / k boxes, split at the landmark-group
  // boundary inside the tile (needs lm_l >= 128 so that a 128-row tile meets at most two groups):
  // lm_part[((bag * Sp/128 + tile) * 4 + row quarter) * 2 + (0: first group of the tile, 1: next group)][1024 q|k columns]
  float* lm_part; int lm_l;
  int dbg;                     // measurement only (TM_GEMM_DBG): 1 = TMA-store epilogues release the accumulator and do nothing
                               // else, 2 = they do the TMEM loads and the math but no smem writes / barriers / stores
};
int gemm_simt(const GemmParams& p, cudaStream_t st);

// ---- tcgen05 / TMA GEMM family (gemm_tc.cu) -------------------------------
bool tc_available();
int gemm_tc(const GemmParams& p, cudaStream_t st);   // same contract as gemm_simt for NT weight GEMMs

}  // namespace tmil
