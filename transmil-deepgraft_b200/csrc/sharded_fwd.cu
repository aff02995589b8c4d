// Token-sharded forward of ONE giant bag behind the C ABI (SURVEY 8e/E3, 8b last row): the whole per-rank schedule -- kernels,
// the pinv / attn3.v fork, and every exchange with the other ranks -- is issued from here on the caller's stream, so a host in any
// language runs a sharded bag with one call per rank, and the call can be captured into a CUDA graph.
//
// Exchanges go over NVLink PEER MEMORY, not through a collective library: every rank owns a small communication region
// (cudaMalloc, exported with cudaIpcGetMemHandle; the host passes the 64-byte handles around once, any way it likes) that the
// other ranks map into their address space.  A producer kernel writes its contribution into the rank's OWN region; a
// one-warp kernel then publishes it by storing the forward's sequence number into a flag slot inside every CONSUMER's region
// (fence + st.release.sys over NVLink: a posted write, the consumer polls its local memory).  A consumer is a pull kernel that
// spins on its local flags (ld.acquire.sys; bounded: a peer that never arrives raises the error word instead of hanging the
// GPU) and copies the payload out of the peers' regions with 16-byte uncached loads.  Per layer:
//   landmark partial sums   [2,8,256,64] fp32 (1 MB)  all ranks -> all ranks, summed in rank order (identical bits everywhere)
//   attn3.v partials        (max, sum, numerator) 540 KB  all -> all, log-sum-exp merge in rank order
//   res_conv halo           first / last 16 rows of v (2 x 16 KB fp16)  to the neighbours
// between the layers the PPEG halo (<= 4G token rows each way, neighbours); at the end the logits (rank 0 -> all).
// Buffers need no double buffering: every layer opens with an all-to-all exchange, so no rank can reach the point where it
// overwrites a buffer before every reader of the previous forward has moved past it (argument in DESIGN.md section 7).
#include <cuda_fp16.h>

#include <new>

#include "common.cuh"
#include "tc_common.cuh"

extern "C" int64_t tm_pinv_scratch_floats(int64_t B);

namespace tmil {

typedef long long i64;
constexpr int SH_MAX_WORLD = TM_SHARD_MAX_WORLD;
constexpr int SH_FLAG_STRIDE = 32;     // uint32 per flag slot (128 bytes: one line per flag)
enum ShFlag : int { FL_LM0 = 0, FL_LM1, FL_A30, FL_A31, FL_VE0, FL_VE1, FL_PPEG, FL_LOGITS, FL_COUNT };

int prep_operands(const float* const* srcs, void* const* dsts, const long long* counts, const int* kinds, int njobs,
                  const tm_weights* w, float* packed, cudaStream_t st);
int attn3v_tc_partial_budget(const void* qlm16, const void* k16, const void* v16, int64_t B, int64_t rows, float* partial,
                             float* scratch, int reserved_sms, cudaStream_t st);
int ppeg_rows_prepacked(const tm_weights* w, const float* in_rows, int64_t y_in0, int64_t y_in1, int64_t G, int64_t tok0, int64_t tok1,
                        float* out, float* packed, cudaStream_t st);

struct CommLayout {
  size_t flags, epoch, err, lm[2], a3[2], ve[2], ppeg_lo, ppeg_hi, logits, total;   // byte offsets into a rank's region
};
static size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }
static size_t a3_partial_bytes() { return (size_t)TM_HEADS * TM_LANDMARKS * (TM_DHEAD + 2) * 4; }
static CommLayout comm_layout(i64 halo_rows_max) {
  CommLayout L;
  size_t cur = 0;
  auto take = [&](size_t bytes) { size_t at = cur; cur += up256(bytes); return at; };
  L.flags = take((size_t)FL_COUNT * SH_MAX_WORLD * SH_FLAG_STRIDE * 4);   // [kind][source rank]
  L.epoch = take(256);
  L.err = take(256);
  for (int l = 0; l < 2; ++l) {
    L.lm[l] = take((size_t)2 * TM_HEADS * TM_LANDMARKS * TM_DHEAD * 4);
    L.a3[l] = take(a3_partial_bytes());
    L.ve[l] = take((size_t)2 * TM_HEADS * 16 * TM_DHEAD * 2);             // [first 16 rows | last 16 rows] of v, fp16
  }
  L.ppeg_lo = take((size_t)halo_rows_max * TM_DIM * 4);                   // my first tokens, for rank - 1
  L.ppeg_hi = take((size_t)halo_rows_max * TM_DIM * 4);                   // my last tokens, for rank + 1
  L.logits = take(4096);
  L.total = cur;
  return L;
}

}  // namespace tmil

struct tm_shard_comm {
  int rank, world, attached, device;
  long long max_n, halo_rows_max;
  tmil::CommLayout L;
  uint8_t* base[tmil::SH_MAX_WORLD];   // [rank] = own region, the others = mappings of the peers' regions
  bool ipc_mapped[tmil::SH_MAX_WORLD];
};

namespace tmil {

struct PeerPtrs { uint8_t* p[SH_MAX_WORLD]; };

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ uint4 ld_peer16(const uint4* p) {   // never served from this SM's L1
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

__global__ void epoch_bump_kernel(uint32_t* epoch) { *epoch += 1; }

// Publishing a contribution = storing this rank's epoch into flag [kind][my_rank] of every consumer's region.  Everything the
// rank's stream wrote before the launch is ordered before the flag (kernel boundary, then fence + release at system scope).
// There is no separate publish kernel: the store is the first thing the next kernel of the schedule does.
constexpr unsigned long long SH_TIMEOUT_NS = 4000000000ull;   // 4 s: a dead peer becomes an error code, not a hung GPU
__device__ __forceinline__ void wait_flag(const uint32_t* flags, int kind, int src, uint32_t want, uint32_t* err) {
  const uint32_t* f = flags + (kind * SH_MAX_WORLD + src) * SH_FLAG_STRIDE;
  if ((int)(ld_acquire_sys(f) - want) >= 0) return;
  const unsigned long long t0 = globaltimer_ns();
  while ((int)(ld_acquire_sys(f) - want) < 0) {
    if (globaltimer_ns() - t0 > SH_TIMEOUT_NS) { atomicExch(err, 1u + (uint32_t)kind); return; }
  }
}

// Pull: for every source s in `mask` (grid.y = world), wait for its flag, then copy n16 16-byte words from its region at
// src_off[s] to dst + dst_off16[s].
// sig_mask != 0: block (0, 0) first publishes this rank's own contribution of the same exchange (see above), so
// publish + wait + copy is one launch.  Every block of the grid is resident at once (<= 32 x 16 blocks), so the publishing
// block always runs even while the others spin.
struct PullJob { size_t src_off[SH_MAX_WORLD]; i64 dst_off16[SH_MAX_WORLD]; i64 n16[SH_MAX_WORLD]; };
__global__ void __launch_bounds__(256) pull_kernel(PeerPtrs regions, const uint32_t* flags, uint32_t mask, int kind, const uint32_t* epoch,
                                                   uint4* __restrict__ dst, PullJob job, uint32_t* err, size_t flags_off, uint32_t sig_mask,
                                                   int my_rank) {
  if (sig_mask != 0 && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x < SH_MAX_WORLD && ((sig_mask >> threadIdx.x) & 1)) {
    __threadfence_system();
    uint32_t* f = reinterpret_cast<uint32_t*>(regions.p[threadIdx.x] + flags_off) + (kind * SH_MAX_WORLD + my_rank) * SH_FLAG_STRIDE;
    st_release_sys(f, *epoch);
  }
  const int s = blockIdx.y;
  if (!((mask >> s) & 1)) return;
  if (threadIdx.x == 0) wait_flag(flags, kind, s, *epoch, err);
  __syncthreads();
  const uint4* from = reinterpret_cast<const uint4*>(regions.p[s] + job.src_off[s]);
  uint4* to = dst + job.dst_off16[s];
  const i64 n16 = job.n16[s];
  // four loads in flight per thread: a 3.6 MB PPEG halo over NVLink is latency x concurrency bound, not link bound
  const i64 stride = (i64)gridDim.x * 256;
  i64 i = (i64)blockIdx.x * 256 + threadIdx.x;
  for (; i + 3 * stride < n16; i += 4 * stride) {
    const uint4 a = ld_peer16(from + i), b = ld_peer16(from + i + stride), c = ld_peer16(from + i + 2 * stride), d = ld_peer16(from + i + 3 * stride);
    to[i] = a; to[i + stride] = b; to[i + 2 * stride] = c; to[i + 3 * stride] = d;
  }
  for (; i < n16; i += stride) to[i] = ld_peer16(from + i);
}

// Landmark means straight from the peers' regions: group g of (tensor, head) receives the partial sums of the ranks whose
// row range meets the group (one or two of them), added in rank order -> the same bits on every rank.
struct GroupRanges { int lo[SH_MAX_WORLD], hi[SH_MAX_WORLD]; };
__global__ void __launch_bounds__(256) landmark_gather_finish_kernel(PeerPtrs regions, size_t lm_off, const uint32_t* flags, int kind,
                                                                    int world, const uint32_t* epoch, uint32_t* err, GroupRanges gr,
                                                                    float fl, float* __restrict__ q_lm, float* __restrict__ k_lm,
                                                                    __half* __restrict__ qlm16, __half* __restrict__ klm16) {
  constexpr i64 per_tensor = (i64)TM_HEADS * TM_LANDMARKS * TM_DHEAD;
  if ((int)threadIdx.x < world) wait_flag(flags, kind, threadIdx.x, *epoch, err);
  __syncthreads();
  const i64 i = ((i64)blockIdx.x * 256 + threadIdx.x) * 2;
  if (i >= 2 * per_tensor) return;
  const int g = (int)((i / TM_DHEAD) % TM_LANDMARKS);
  float2 s = make_float2(0.f, 0.f);
  for (int r = 0; r < world; ++r) {
    if (g < gr.lo[r] || g >= gr.hi[r]) continue;
    const float* src = reinterpret_cast<const float*>(regions.p[r] + lm_off) + i;
    float2 v;
    asm volatile("ld.volatile.global.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(src));
    s.x += v.x; s.y += v.y;
  }
  const bool is_k = i >= per_tensor;
  const i64 o = is_k ? i - per_tensor : i;
  const float a = s.x / fl, b = s.y / fl;
  *reinterpret_cast<float2*>((is_k ? k_lm : q_lm) + o) = make_float2(a, b);
  *reinterpret_cast<uint32_t*>((is_k ? klm16 : qlm16) + o) = pack_f16x2(a, b);
}
// One block, launched behind landmark_sums: publishes the landmark partial sums (flag kind_lm, to lm_mask) at once, then packs
// the first / last 16 rows of v [8, rows, 64] fp16 into edges [2][8][16][64] of the own region and publishes those (kind_ve,
// to the neighbours in ve_mask).
__global__ void __launch_bounds__(1024) publish_layer_kernel(PeerPtrs regions, size_t flags_off, int my_rank, const uint32_t* epoch,
                                                             uint32_t lm_mask, int kind_lm, uint32_t ve_mask, int kind_ve,
                                                             const __half* __restrict__ v, i64 rows, __half* __restrict__ edges) {
  const int t = threadIdx.x;
  const uint32_t ep = *epoch;
  if (t < SH_MAX_WORLD && ((lm_mask >> t) & 1)) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(regions.p[t] + flags_off) + (kind_lm * SH_MAX_WORLD + my_rank) * SH_FLAG_STRIDE, ep);
  }
  if (ve_mask == 0) return;
  for (int i = t; i < 2 * TM_HEADS * 16 * 8; i += 1024) {      // uint4 index
    const int c = i & 7, tt = (i >> 3) & 15, h = (i >> 7) & 7, e = i >> 10;
    const i64 row = e == 0 ? tt : rows - 16 + tt;
    reinterpret_cast<uint4*>(edges)[i] = reinterpret_cast<const uint4*>(v + ((i64)h * rows + row) * TM_DHEAD)[c];
  }
  __syncthreads();
  if (t < SH_MAX_WORLD && ((ve_mask >> t) & 1)) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(regions.p[t] + flags_off) + (kind_ve * SH_MAX_WORLD + my_rank) * SH_FLAG_STRIDE, ep);
  }
}
// up to two plain device copies in one launch (grid.y selects the job)
__global__ void __launch_bounds__(256) copy2_kernel(const uint4* __restrict__ s0, uint4* __restrict__ d0, i64 n0,
                                                    const uint4* __restrict__ s1, uint4* __restrict__ d1, i64 n1) {
  const uint4* s = blockIdx.y ? s1 : s0;
  uint4* d = blockIdx.y ? d1 : d0;
  const i64 n = blockIdx.y ? n1 : n0;
  for (i64 i = (i64)blockIdx.x * 256 + threadIdx.x; i < n; i += (i64)gridDim.x * 256) d[i] = s[i];
}

}  // namespace tmil

using namespace tmil;

// ---- plan -------------------------------------------------------------------------------------------------------------
// Same integers as ShardPlan / RankShard of transmil-deepgraft_b200/sharded.py (tests/test_sharded.py compares them).
static void rank_shard(const tm_plan& p, i64 world, i64 r, tm_shard_plan* s) {
  memset(s, 0, sizeof(*s));
  s->bag = p;
  s->world = world; s->rank = r;
  const i64 row0 = 256 * ((p.l * r) / world), row1 = 256 * ((p.l * (r + 1)) / world);
  s->row0 = row0; s->rows = row1 - row0;
  s->P_local = r == 0 ? p.P : 0;
  const i64 s0 = row0 - p.P > 0 ? row0 - p.P : 0, s1 = row1 - p.P;
  s->S_local = s1 - s0;
  s->has_cls = r == 0;
  const i64 tok0 = s0 - 1 > 0 ? s0 - 1 : 0, tok1 = s1 - 1;
  s->tok0 = tok0; s->ntok = tok1 - tok0;
  i64 nr = 0;
  if (tok0 < p.n) { s->x_range[nr][0] = tok0; s->x_range[nr][1] = tok1 < p.n ? tok1 : p.n; ++nr; }
  if (tok1 > p.n) { s->x_range[nr][0] = (tok0 > p.n ? tok0 : p.n) - p.n; s->x_range[nr][1] = tok1 - p.n; ++nr; }
  s->n_x_ranges = nr;
  const i64 y0 = tok0 / p.G - 3, y1 = (tok1 - 1) / p.G + 4;
  s->y_in0 = y0 > 0 ? y0 : 0;
  s->y_in1 = y1 < p.G ? y1 : p.G;
  s->halo_lo = tok0 - s->y_in0 * p.G;
  s->halo_hi = s->y_in1 * p.G - tok1;
}

extern "C" int tm_make_shard_plan(int64_t n, int64_t world, int64_t rank, tm_shard_plan* out) {
  TM_CHECK_ARG(out != nullptr, TM_ERR_BAD_ARGUMENT, "tm_make_shard_plan: out is NULL");
  TM_CHECK_ARG(world >= 1 && world <= SH_MAX_WORLD && rank >= 0 && rank < world, TM_ERR_BAD_ARGUMENT,
               "tm_make_shard_plan: world must be 1..%d and 0 <= rank < world (world=%lld rank=%lld)", SH_MAX_WORLD, (long long)world, (long long)rank);
  tm_plan p;
  TM_PROPAGATE(tm_make_plan(n, &p));
  TM_CHECK_ARG(p.l >= world, TM_ERR_BAD_SHAPE, "a bag of %lld patches has only %lld blocks of 256 padded rows: cannot shard over %lld ranks",
               (long long)n, (long long)p.l, (long long)world);
  tm_shard_plan prev, cur, next;
  for (i64 r = 0; r < world; ++r) {
    rank_shard(p, world, r, &cur);
    if (r > 0) rank_shard(p, world, r - 1, &prev);
    if (r + 1 < world) rank_shard(p, world, r + 1, &next);
    const bool lo_ok = cur.halo_lo == 0 || (r > 0 && cur.halo_lo <= prev.ntok);
    const bool hi_ok = cur.halo_hi == 0 || (r + 1 < world && cur.halo_hi <= next.ntok);
    TM_CHECK_ARG(lo_ok && hi_ok && cur.rows >= 32, TM_ERR_BAD_SHAPE,
                 "a bag of %lld patches is too small to shard over %lld ranks (the PPEG halo of rank %lld would span more than one neighbour)",
                 (long long)n, (long long)world, (long long)r);
    if (r == rank) {
      *out = cur;
      out->n_up = r + 1 < world ? next.halo_lo : 0;   // my last tokens the next rank needs
      out->n_dn = r > 0 ? prev.halo_hi : 0;           // my first tokens the previous rank needs
    }
  }
  return TM_OK;
}

// ---- communication regions ----------------------------------------------------------------------------------------------
static i64 halo_rows_for(i64 max_n) {
  tm_plan p;
  if (tm_make_plan(max_n, &p) != TM_OK) return 0;
  return 4 * p.G;     // 3 image rows + the partial row on either side of a cut
}

extern "C" int tm_shard_comm_create(int rank, int world, int64_t max_n, tm_shard_comm** out) {
  TM_CHECK_ARG(out != nullptr, TM_ERR_BAD_ARGUMENT, "tm_shard_comm_create: out is NULL");
  TM_CHECK_ARG(world >= 1 && world <= SH_MAX_WORLD && rank >= 0 && rank < world && max_n > 0, TM_ERR_BAD_ARGUMENT,
               "tm_shard_comm_create: world 1..%d, 0 <= rank < world, max_n > 0", SH_MAX_WORLD);
  tm_shard_comm* c = new (std::nothrow) tm_shard_comm();
  TM_CHECK_ARG(c != nullptr, TM_ERR_BAD_ARGUMENT, "tm_shard_comm_create: out of host memory");
  memset(c, 0, sizeof(*c));
  c->rank = rank; c->world = world; c->max_n = max_n;
  c->halo_rows_max = halo_rows_for(max_n);
  c->L = comm_layout(c->halo_rows_max);
  cudaError_t e = cudaGetDevice(&c->device);
  void* base = nullptr;
  if (e == cudaSuccess) e = cudaMalloc(&base, c->L.total);
  if (e == cudaSuccess) e = cudaMemset(base, 0, c->L.total);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("tm_shard_comm_create: %s", cudaGetErrorString(e));
    if (base) cudaFree(base);
    delete c;
    return TM_ERR_CUDA;
  }
  c->base[rank] = (uint8_t*)base;
  c->attached = world == 1;
  *out = c;
  return TM_OK;
}

extern "C" int64_t tm_shard_comm_bytes(const tm_shard_comm* c) { return c ? (int64_t)c->L.total : 0; }

extern "C" int tm_shard_comm_handle(const tm_shard_comm* c, void* handle64) {
  TM_CHECK_ARG(c && handle64, TM_ERR_BAD_ARGUMENT, "tm_shard_comm_handle: NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == TM_SHARD_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t h;
  TM_CUDA(cudaIpcGetMemHandle(&h, c->base[c->rank]));
  memcpy(handle64, &h, sizeof(h));
  return TM_OK;
}

// handles: world x 64 bytes in rank order (the own entry is ignored).  One process per GPU.
extern "C" int tm_shard_comm_attach(tm_shard_comm* c, const void* handles) {
  TM_CHECK_ARG(c && handles, TM_ERR_BAD_ARGUMENT, "tm_shard_comm_attach: NULL argument");
  TM_CHECK_ARG(!c->attached || c->world == 1, TM_ERR_BAD_ARGUMENT, "tm_shard_comm_attach: already attached");
  for (int r = 0; r < c->world; ++r) {
    if (r == c->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const uint8_t*)handles + (size_t)r * TM_SHARD_HANDLE_BYTES, sizeof(h));
    void* p = nullptr;
    TM_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c->base[r] = (uint8_t*)p;
    c->ipc_mapped[r] = true;
  }
  c->attached = 1;
  return TM_OK;
}

// Single-process hosts (one thread per GPU): the peers' regions by device pointer (tm_shard_comm_base of their comm
// objects); peer access between the devices is enabled here.
extern "C" void* tm_shard_comm_base(const tm_shard_comm* c) { return c ? c->base[c->rank] : nullptr; }
extern "C" int tm_shard_comm_attach_ptrs(tm_shard_comm* c, void* const* bases, const int* devices) {
  TM_CHECK_ARG(c && bases && devices, TM_ERR_BAD_ARGUMENT, "tm_shard_comm_attach_ptrs: NULL argument");
  for (int r = 0; r < c->world; ++r) {
    if (r == c->rank) continue;
    TM_CHECK_ARG(bases[r] != nullptr, TM_ERR_BAD_ARGUMENT, "tm_shard_comm_attach_ptrs: base of rank %d is NULL", r);
    if (devices[r] != c->device) {
      cudaError_t e = cudaDeviceEnablePeerAccess(devices[r], 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
      TM_CUDA(e);
    }
    c->base[r] = (uint8_t*)bases[r];
  }
  c->attached = 1;
  return TM_OK;
}

// Synchronises the device.  *code = 0, or 1 + the exchange (0/1 landmarks, 2/3 attn3.v, 4/5 conv halo, 6 PPEG halo, 7 logits)
// whose producer did not arrive within the time-out in some forward since the last call; the word is cleared.
extern "C" int tm_shard_comm_status(tm_shard_comm* c, int32_t* code) {
  TM_CHECK_ARG(c && code, TM_ERR_BAD_ARGUMENT, "tm_shard_comm_status: NULL argument");
  uint32_t v = 0;
  TM_CUDA(cudaDeviceSynchronize());
  TM_CUDA(cudaMemcpy(&v, c->base[c->rank] + c->L.err, 4, cudaMemcpyDeviceToHost));
  if (v) TM_CUDA(cudaMemset(c->base[c->rank] + c->L.err, 0, 4));
  *code = (int32_t)v;
  return TM_OK;
}

extern "C" int tm_shard_comm_destroy(tm_shard_comm* c) {
  if (!c) return TM_OK;
  for (int r = 0; r < c->world; ++r)
    if (r != c->rank && c->ipc_mapped[r] && c->base[r]) cudaIpcCloseMemHandle(c->base[r]);
  if (c->base[c->rank]) cudaFree(c->base[c->rank]);
  delete c;
  return TM_OK;
}

// ---- the per-rank schedule ------------------------------------------------------------------------------------------------
namespace tmil {

struct ShLayout {
  size_t total;   // floats
  size_t h0, ext, h2, fc1s, xn, q, k, v, merged, q_lm, k_lm, attn2, Z, W3, M, qlm16, klm16, mt16, a3_gather, a3_s, v_prev, v_next,
      pinv_s, packed, wr_fc1_0, wr_fc1_1, wr_qkv[2], wr_out[2], logits;
};
static size_t shbump(size_t& cur, size_t n) {
  size_t at = cur;
  cur += (n + 63) & ~(size_t)63;
  return at;
}
static int sh_fc1_width(const tm_weights* w) { return w->fc1_kind == TM_FC1_2048 ? 1024 : w->fc1_kind == TM_FC1_768 ? 768 : 0; }
static void sh_layout(const tm_weights* w, const tm_shard_plan& s, ShLayout* L) {
  size_t cur = 0;
  const int f0 = sh_fc1_width(w);
  const size_t hs = (size_t)s.S_local * TM_DIM, half_rows = (size_t)s.rows * TM_DIM / 2;
  const size_t lm = (size_t)TM_HEADS * TM_LANDMARKS * TM_DHEAD, mm = (size_t)TM_HEADS * 65536;
  L->h0 = shbump(cur, hs);
  L->ext = shbump(cur, (size_t)(s.halo_lo + s.S_local + s.halo_hi) * TM_DIM);
  L->h2 = shbump(cur, hs);
  L->fc1s = shbump(cur, (size_t)s.ntok * f0);
  L->xn = shbump(cur, half_rows);
  L->q = shbump(cur, half_rows);
  L->k = shbump(cur, half_rows);
  L->v = shbump(cur, half_rows);
  L->merged = shbump(cur, half_rows);
  L->q_lm = shbump(cur, lm); L->k_lm = shbump(cur, lm);
  L->attn2 = shbump(cur, mm); L->Z = shbump(cur, mm);
  L->W3 = shbump(cur, lm); L->M = shbump(cur, lm);
  L->qlm16 = shbump(cur, lm / 2); L->klm16 = shbump(cur, lm / 2); L->mt16 = shbump(cur, lm / 2);
  L->a3_gather = shbump(cur, (size_t)s.world * (a3_partial_bytes() / 4));
  L->a3_s = shbump(cur, (size_t)tm_attn3v_tc_scratch_floats(1, s.rows));
  L->v_prev = shbump(cur, (size_t)TM_HEADS * 16 * TM_DHEAD / 2);
  L->v_next = shbump(cur, (size_t)TM_HEADS * 16 * TM_DHEAD / 2);
  L->pinv_s = shbump(cur, (size_t)tm_pinv_scratch_floats(1));
  L->packed = shbump(cur, 50 * TM_DIM);
  L->wr_fc1_0 = shbump(cur, (size_t)(f0 ? f0 : TM_DIM) * w->in_features);
  L->wr_fc1_1 = shbump(cur, (size_t)TM_DIM * f0);
  for (int l = 0; l < 2; ++l) {
    L->wr_qkv[l] = shbump(cur, (size_t)3 * TM_DIM * TM_DIM);
    L->wr_out[l] = shbump(cur, (size_t)TM_DIM * TM_DIM);
  }
  L->logits = shbump(cur, 1024);
  L->total = cur;
}

struct ShCtx {
  tm_shard_comm* c;
  PeerPtrs regions;
  uint8_t* own;
  const uint32_t* flags;
  uint32_t* epoch;
  uint32_t* err;
  uint32_t all_mask, nb_mask;
  cudaStream_t st;
};
static int sh_pull(const ShCtx& x, int kind, uint32_t mask, void* dst, const PullJob& job, uint32_t sig_mask = 0) {
  if (!mask && !sig_mask) return TM_OK;
  i64 most = 0;
  for (int s = 0; s < x.c->world; ++s)
    if (((mask >> s) & 1) && job.n16[s] > most) most = job.n16[s];
  int blocks = (int)ceil_div(most, 256 * 8);
  if (blocks < 1) blocks = 1;
  const int cap = 592 / x.c->world > 16 ? 592 / x.c->world : 16;     // the whole grid stays resident (4 CTAs of 256 threads per SM)
  if (blocks > cap) blocks = cap;
  pull_kernel<<<dim3((unsigned)blocks, (unsigned)x.c->world), 256, 0, x.st>>>(x.regions, x.flags, mask, kind, x.epoch, (uint4*)dst, job, x.err,
                                                                              x.c->L.flags, sig_mask, x.c->rank);
  TM_LAUNCH_CHECK();
  return TM_OK;
}

// One TransLayer (TransMIL.py:45-57) on this rank's rows.
static int sh_layer(const ShCtx& x, const tm_layer_weights& lw, const tm_shard_plan& s, int li, const float* h_in, float* h_out, float* ws,
                    const ShLayout& L, uint32_t flags) {
  const tm_shard_comm* c = x.c;
  cudaStream_t st = x.st;
  const int rank = c->rank, world = c->world;
  void* xn = ws + L.xn; void* q16 = ws + L.q; void* k16 = ws + L.k; void* v16 = ws + L.v; void* merged = ws + L.merged;
  float* q_lm = ws + L.q_lm; float* k_lm = ws + L.k_lm; void* qlm16 = ws + L.qlm16; void* klm16 = ws + L.klm16; void* mt16 = ws + L.mt16;
  prof_mark(ST_LN_PAD, st);
  TM_PROPAGATE(tm_ln_pad_fwd(h_in, lw.norm_w, lw.norm_b, 1, s.S_local, s.P_local, (float*)xn, flags, st));
  prof_mark(ST_QKV, st);
  TM_PROPAGATE(tm_qkv16_fwd(xn, lw.qkv_w, 1, s.rows, q16, k16, v16, flags, st));
  // landmark partial sums -> own region -> everyone
  prof_mark(ST_LANDMARKS, st);
  TM_PROPAGATE(tm_landmark_sums16_fwd(q16, k16, 1, s.rows, s.row0, s.bag.l, (float*)(x.own + c->L.lm[li]), st));
  // publish them; the same launch packs the res_conv halo rows of v for the neighbours (consumed late in the layer)
  prof_mark(ST_X_LANDMARKS, st);
  publish_layer_kernel<<<1, 1024, 0, st>>>(x.regions, c->L.flags, rank, x.epoch, x.all_mask, FL_LM0 + li, x.nb_mask, FL_VE0 + li,
                                           (const __half*)v16, s.rows, (__half*)(x.own + c->L.ve[li]));
  TM_LAUNCH_CHECK();
  {
    GroupRanges gr;
    memset(&gr, 0, sizeof(gr));
    for (int r = 0; r < world; ++r) {
      const i64 r0 = 256 * ((s.bag.l * r) / world), r1 = 256 * ((s.bag.l * (r + 1)) / world);
      gr.lo[r] = (int)(r0 / s.bag.l);
      gr.hi[r] = (int)((r1 - 1) / s.bag.l) + 1;
    }
    landmark_gather_finish_kernel<<<(unsigned)(TM_HEADS * TM_LANDMARKS * TM_DHEAD / 256), 256, 0, st>>>(
        x.regions, c->L.lm[li], x.flags, FL_LM0 + li, world, x.epoch, x.err, gr, (float)s.bag.l, q_lm, k_lm, (__half*)qlm16, (__half*)klm16);
    TM_LAUNCH_CHECK();
  }
  // replicated sim2 + pseudo-inverse on the side stream, beside this rank's share of attn3.v and its exchange
  const bool overlap = !(flags & TM_FLAG_NO_OVERLAP) && !prof_enabled();
  prof_mark(ST_PINV, st);
  SideStream ss;
  if (overlap) {
    TM_PROPAGATE(side_stream(&ss));
    TM_CUDA(cudaEventRecord(ss.fork, st));
    TM_CUDA(cudaStreamWaitEvent(ss.stream, ss.fork, 0));
    pdl_plain_next();
    TM_PROPAGATE(tm_pinv_fwd(q_lm, k_lm, 1, ws + L.attn2, ws + L.Z, ws + L.pinv_s, flags, ss.stream));
    TM_CUDA(cudaEventRecord(ss.join, ss.stream));
    pdl_plain_next();
  } else {
    TM_PROPAGATE(tm_pinv_fwd(q_lm, k_lm, 1, ws + L.attn2, ws + L.Z, ws + L.pinv_s, flags, st));
  }
  prof_mark(ST_ATTN3V, st);
  TM_PROPAGATE(attn3v_tc_partial_budget(qlm16, k16, v16, 1, s.rows, (float*)(x.own + c->L.a3[li]), ws + L.a3_s, overlap ? 4 * TM_HEADS : 0, st));
  prof_mark(ST_X_ATTN3V, st);
  {
    PullJob job;
    memset(&job, 0, sizeof(job));
    const i64 n16 = (i64)(a3_partial_bytes() / 16);
    for (int r = 0; r < world; ++r) { job.src_off[r] = c->L.a3[li]; job.dst_off16[r] = r * n16; job.n16[r] = n16; }
    TM_PROPAGATE(sh_pull(x, FL_A30 + li, x.all_mask, ws + L.a3_gather, job, x.all_mask));   // publish own partial + gather all
  }
  prof_mark(ST_ATTN3V, st);
  TM_PROPAGATE(tm_attn3v_merge_fwd(ws + L.a3_gather, world, 1, ws + L.W3, st));
  if (overlap) {
    TM_CUDA(cudaStreamWaitEvent(st, ss.join, 0));
    pdl_plain_next();
  }
  prof_mark(ST_ZW, st);
  TM_PROPAGATE(tm_zw16_fwd(ws + L.Z, ws + L.W3, 1, ws + L.M, mt16, st));
  prof_mark(ST_NYSTROM_OUT, st);
  TM_PROPAGATE(tm_nystrom_out_tc_fwd(q16, klm16, mt16, v16, lw.conv_w, 1, s.rows, merged, flags, st));
  if (x.nb_mask) {
    PullJob job;
    memset(&job, 0, sizeof(job));
    prof_mark(ST_X_CONV_HALO, st);
    const i64 n16 = (i64)TM_HEADS * 16 * TM_DHEAD * 2 / 16;    // one edge: [8,16,64] fp16
    if (rank > 0) { job.src_off[rank - 1] = c->L.ve[li] + (size_t)n16 * 16; job.dst_off16[rank - 1] = 0; job.n16[rank - 1] = n16; }
    if (rank + 1 < world) {
      job.src_off[rank + 1] = c->L.ve[li];
      job.dst_off16[rank + 1] = (i64)((L.v_next - L.v_prev) * 4 / 16);
      job.n16[rank + 1] = n16;
    }
    TM_PROPAGATE(sh_pull(x, FL_VE0 + li, x.nb_mask, ws + L.v_prev, job));
    TM_PROPAGATE(tm_conv_halo_fix_fwd(rank > 0 ? (void*)(ws + L.v_prev) : nullptr, rank + 1 < world ? (void*)(ws + L.v_next) : nullptr,
                                      lw.conv_w, 1, s.rows, merged, st));
  }
  prof_mark(ST_TO_OUT, st);
  TM_PROPAGATE(tm_to_out_residual_fwd((const float*)merged, lw.out_w, lw.out_b, h_in, 1, s.S_local, s.P_local, h_out, flags, st));
  return TM_OK;
}

}  // namespace tmil

extern "C" int tm_sharded_workspace_bytes(const tm_weights* w, const tm_shard_plan* plan, size_t* bytes) {
  TM_CHECK_ARG(w && plan && bytes, TM_ERR_BAD_ARGUMENT, "tm_sharded_workspace_bytes: NULL argument");
  ShLayout L;
  sh_layout(w, *plan, &L);
  *bytes = L.total * sizeof(float);
  return TM_OK;
}

// TransMIL.forward (TransMIL.py:167-211) of ONE bag, this rank's share.  x_rows: the rows of the bag this rank projects
// (plan->x_range, concatenated: [ntok, D] fp32 on the device).  Every rank of the group calls it once per bag with the plan
// of the same (n, world); logits [n_classes] arrive on every rank.  Asynchronous on `stream`; capturable into a CUDA graph.
extern "C" int tm_transmil_fwd_sharded(const tm_weights* w, const float* x_rows, const tm_shard_plan* plan, tm_shard_comm* comm,
                                       float* logits, void* workspace, size_t workspace_bytes, uint32_t flags, void* stream) {
  TM_CHECK_ARG(w && x_rows && plan && comm && logits && workspace, TM_ERR_BAD_ARGUMENT, "tm_transmil_fwd_sharded: NULL argument");
  TM_CHECK_ARG(tc_available(), TM_ERR_UNSUPPORTED_ARCH, "tm_transmil_fwd_sharded needs an sm_100 device");
  TM_CHECK_ARG(!(flags & (TM_FLAG_SIMT_ONLY | TM_FLAG_X_F16 | TM_FLAG_KEEP_STAGES | TM_FLAG_PINV_GEMM_CHAIN)), TM_ERR_BAD_ARGUMENT,
               "tm_transmil_fwd_sharded: tensor-core engine, fp32 bag only (flags: 0 or TM_FLAG_NO_OVERLAP)");
  TM_CHECK_ARG(w->fc1_kind != TM_FC1_TOKENS, TM_ERR_BAD_ARGUMENT, "tm_transmil_fwd_sharded: TM_FC1_TOKENS (per-tile bags of CTMIL) is not sharded");
  const tm_shard_plan& s = *plan;
  TM_CHECK_ARG(s.world == comm->world && s.rank == comm->rank, TM_ERR_BAD_ARGUMENT, "tm_transmil_fwd_sharded: plan is for rank %lld of %lld, comm for %d of %d",
               (long long)s.rank, (long long)s.world, comm->rank, comm->world);
  TM_CHECK_ARG(comm->attached, TM_ERR_BAD_ARGUMENT, "tm_transmil_fwd_sharded: the comm object has not been attached to its peers");
  TM_CHECK_ARG(s.halo_lo <= comm->halo_rows_max && s.halo_hi <= comm->halo_rows_max && s.n_up <= comm->halo_rows_max &&
                   s.n_dn <= comm->halo_rows_max, TM_ERR_BAD_SHAPE, "tm_transmil_fwd_sharded: bag of %lld patches exceeds the comm object's max_n=%lld",
               (long long)s.bag.n, comm->max_n);
  TM_CHECK_ARG(w->n_classes > 0 && w->n_classes <= 1024, TM_ERR_BAD_SHAPE, "tm_transmil_fwd_sharded: n_classes must be 1..1024");
  TM_CHECK_ARG(aligned16(x_rows) && ((uintptr_t)workspace & 255u) == 0, TM_ERR_BAD_ALIGNMENT, "tm_transmil_fwd_sharded: x_rows must be 16-byte, workspace 256-byte aligned");
  ShLayout L;
  sh_layout(w, s, &L);
  TM_CHECK_ARG(workspace_bytes >= L.total * sizeof(float), TM_ERR_WORKSPACE_TOO_SMALL, "tm_transmil_fwd_sharded: workspace too small (%zu < %zu bytes)",
               workspace_bytes, L.total * sizeof(float));
  cudaStream_t st = (cudaStream_t)stream;
  float* ws = (float*)workspace;
  const int rank = comm->rank, world = comm->world;
  flags |= TM_FLAG_TF32_OPERANDS | TM_FLAG_F16_ACT;

  ShCtx x;
  x.c = comm; x.st = st;
  for (int r = 0; r < SH_MAX_WORLD; ++r) x.regions.p[r] = r < world ? comm->base[r] : nullptr;
  x.own = comm->base[rank];
  x.flags = (const uint32_t*)(x.own + comm->L.flags);
  x.epoch = (uint32_t*)(x.own + comm->L.epoch);
  x.err = (uint32_t*)(x.own + comm->L.err);
  x.all_mask = (1u << world) - 1;
  x.nb_mask = (rank > 0 ? 1u << (rank - 1) : 0u) | (rank + 1 < world ? 1u << (rank + 1) : 0u);

  prof_mark(ST_PREP, st);
  epoch_bump_kernel<<<1, 1, 0, st>>>(x.epoch);
  TM_LAUNCH_CHECK();
  // operand copies of the parameters (tf32-rounded fc1, fp16 to_qkv / to_out, packed PPEG filter) in one launch
  tm_weights wl = *w;
  {
    const int f0 = sh_fc1_width(w);
    const float* srcs[6]; void* dsts[6]; long long cnt[6]; int kinds[6];
    int nj = 0;
    srcs[nj] = w->fc1_w0; dsts[nj] = ws + L.wr_fc1_0; cnt[nj] = (long long)(f0 ? f0 : TM_DIM) * w->in_features; kinds[nj++] = 0;
    wl.fc1_w0 = ws + L.wr_fc1_0;
    if (f0) {
      srcs[nj] = w->fc1_w1; dsts[nj] = ws + L.wr_fc1_1; cnt[nj] = (long long)TM_DIM * f0; kinds[nj++] = 0;
      wl.fc1_w1 = ws + L.wr_fc1_1;
    }
    tm_layer_weights* lws[2] = {&wl.layer1, &wl.layer2};
    for (int l = 0; l < 2; ++l) {
      srcs[nj] = lws[l]->qkv_w; dsts[nj] = ws + L.wr_qkv[l]; cnt[nj] = (long long)3 * TM_DIM * TM_DIM; kinds[nj++] = 1;
      srcs[nj] = lws[l]->out_w; dsts[nj] = ws + L.wr_out[l]; cnt[nj] = (long long)TM_DIM * TM_DIM; kinds[nj++] = 1;
      lws[l]->qkv_w = ws + L.wr_qkv[l];
      lws[l]->out_w = ws + L.wr_out[l];
    }
    pdl_plain_next();
    TM_PROPAGATE(prep_operands(srcs, dsts, cnt, kinds, nj, w, ws + L.packed, st));
  }
  const i64 off = s.has_cls ? 1 : 0;
  float* h0 = ws + L.h0;
  float* ext = ws + L.ext;
  float* h1 = ext + s.halo_lo * TM_DIM;
  float* h2 = ws + L.h2;
  float* h3 = h0;                      // h0 is dead once layer 1 has added its residual
  // fc1 on this rank's rows; the duplicate pad is just more rows of x (plan->x_range), the cls row lives on rank 0
  TM_PROPAGATE(tm_fc1_rows_fwd(&wl, x_rows, s.ntok, h0 + off * TM_DIM, ws + L.fc1s, flags, st));
  if (s.has_cls) {
    TM_CUDA(cudaMemcpyAsync(h0, w->cls_token, TM_DIM * sizeof(float), cudaMemcpyDeviceToDevice, st));
    pdl_plain_next();
  }
  TM_PROPAGATE(sh_layer(x, wl.layer1, s, 0, h0, h1, ws, L, flags));
  // PPEG halo (TransMIL.py:60-75 on the G x G token image): my first n_dn tokens are the previous rank's upper halo, my last
  // n_up tokens the next rank's lower halo
  if (x.nb_mask) {
    prof_mark(ST_X_PPEG_HALO, st);
    const float* toks = h1 + off * TM_DIM;
    const i64 n_dn16 = s.n_dn * (TM_DIM / 4), n_up16 = s.n_up * (TM_DIM / 4);
    if (n_dn16 || n_up16) {
      copy2_kernel<<<dim3(148, 2), 256, 0, st>>>((const uint4*)toks, (uint4*)(x.own + comm->L.ppeg_lo), n_dn16,
                                               (const uint4*)(toks + (s.ntok - s.n_up) * TM_DIM), (uint4*)(x.own + comm->L.ppeg_hi), n_up16);
      TM_LAUNCH_CHECK();
    }
    PullJob job;
    memset(&job, 0, sizeof(job));
    if (rank > 0) { job.src_off[rank - 1] = comm->L.ppeg_hi; job.dst_off16[rank - 1] = 0; job.n16[rank - 1] = s.halo_lo * (TM_DIM / 4); }
    if (rank + 1 < world) {
      job.src_off[rank + 1] = comm->L.ppeg_lo;
      job.dst_off16[rank + 1] = (s.halo_lo + s.S_local) * (TM_DIM / 4);
      job.n16[rank + 1] = s.halo_hi * (TM_DIM / 4);
    }
    TM_PROPAGATE(sh_pull(x, FL_PPEG, x.nb_mask, ext, job, x.nb_mask));
  }
  prof_mark(ST_PPEG, st);
  TM_PROPAGATE(ppeg_rows_prepacked(&wl, ext + off * TM_DIM, s.y_in0, s.y_in1, s.bag.G, s.tok0, s.tok0 + s.ntok, h2 + off * TM_DIM, ws + L.packed, st));
  if (s.has_cls) {
    TM_CUDA(cudaMemcpyAsync(h2, h1, TM_DIM * sizeof(float), cudaMemcpyDeviceToDevice, st));
    pdl_plain_next();
  }
  TM_PROPAGATE(sh_layer(x, wl.layer2, s, 1, h2, h3, ws, L, flags));
  // final norm + head on the cls row (TransMIL.py:202-204): rank 0, into its region; everyone pulls the logits
  prof_mark(ST_HEAD, st);
  if (s.has_cls) TM_PROPAGATE(tm_head_fwd(&wl, h3, 1, s.S_local, (float*)(x.own + comm->L.logits), st));
  {
    PullJob job;
    memset(&job, 0, sizeof(job));
    prof_mark(ST_X_LOGITS, st);
    job.src_off[0] = comm->L.logits; job.dst_off16[0] = 0; job.n16[0] = ceil_div((i64)w->n_classes * 4, 16);
    TM_PROPAGATE(sh_pull(x, FL_LOGITS, 1u, ws + L.logits, job, s.has_cls ? x.all_mask : 0u));
  }
  TM_CUDA(cudaMemcpyAsync(logits, ws + L.logits, (size_t)w->n_classes * sizeof(float), cudaMemcpyDeviceToDevice, st));
  prof_mark(ST_END, st);
  return TM_OK;
}
