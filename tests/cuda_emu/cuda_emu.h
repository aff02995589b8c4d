// TEST INFRASTRUCTURE ONLY -- a minimal host-side execution harness for CUDA kernel SOURCE.
//
// The build container has nvcc but no GPU.  The backward kernels of csrc/backward.cu are written in plain SIMT CUDA
// (threadIdx / blockIdx, __shared__ arrays, __syncthreads, no warp intrinsics, no atomics), so the very same source
// compiles with g++ against this header (-DTM_HOST_EMU) into tests/cuda_emu/_build/libtransmil_b200_emu.so, where every
// launch runs its grid on the host: blocks are spread over a few OS threads, the threads of one block are
// fibers (a small x86-64 stack switch) that hand over at __syncthreads().  The CPU tests use it to check the kernels' indexing and arithmetic and the
// host orchestration against the backward oracle before a GPU is available.  It is never loaded by the product path
// (transmil-deepgraft_b200/capi.py only ever opens libtransmil_b200.so, which is built by nvcc for sm_100a).
#pragma once
#include <math.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <functional>
#include <thread>
#include <vector>

#include "transmil_b200.h"

struct emu_uint3 { unsigned x, y, z; };
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
typedef void* cudaStream_t;

inline thread_local emu_uint3 threadIdx, blockIdx;
inline thread_local dim3 blockDim, gridDim;

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__ static thread_local

// Minimal x86-64 SysV context switch (callee-saved registers + stack pointer); ucontext's swapcontext makes two
// sigprocmask system calls per switch, which dominated the run time of barrier-heavy kernels.
extern "C" void tm_emu_switch(void** save_sp, void* load_sp);
asm(R"(
.text
.globl tm_emu_switch
.hidden tm_emu_switch
.type tm_emu_switch,@function
tm_emu_switch:
  pushq %rbp
  pushq %rbx
  pushq %r12
  pushq %r13
  pushq %r14
  pushq %r15
  movq %rsp, (%rdi)
  movq %rsi, %rsp
  popq %r15
  popq %r14
  popq %r13
  popq %r12
  popq %rbx
  popq %rbp
  ret
.size tm_emu_switch,.-tm_emu_switch
)");

namespace emu {

struct Fiber {
  void* sp;
  char* stack;
  bool done;
};
inline thread_local void* g_sched_sp = nullptr;
inline thread_local Fiber* g_cur = nullptr;
inline thread_local const std::function<void()>* g_body = nullptr;

inline void trampoline() {
  (*g_body)();
  g_cur->done = true;
  tm_emu_switch(&g_cur->sp, g_sched_sp);
  abort();   // a finished fiber is never resumed
}

inline void sync() { tm_emu_switch(&g_cur->sp, g_sched_sp); }

constexpr size_t kStack = 96 * 1024;

// Runs one block: every live fiber is resumed in thread order until its next __syncthreads() (or its end).
inline void run_block(std::vector<Fiber>& fibers, dim3 block, const std::function<void()>& body) {
  const unsigned n = block.x * block.y * block.z;
  g_body = &body;
  for (unsigned t = 0; t < n; ++t) {
    Fiber& f = fibers[t];
    uintptr_t top = ((uintptr_t)f.stack + kStack) & ~(uintptr_t)15;
    void** sp = (void**)top;
    *--sp = nullptr;                    // fake return address of trampoline (keeps rsp = 8 mod 16 at its entry)
    *--sp = (void*)&trampoline;         // popped by tm_emu_switch's ret
    for (int i = 0; i < 6; ++i) *--sp = nullptr;   // rbp rbx r12-r15
    f.sp = sp;
    f.done = false;
  }
  unsigned live = n;
  while (live > 0) {
    for (unsigned t = 0; t < n; ++t) {
      Fiber& f = fibers[t];
      if (f.done) continue;
      threadIdx.x = t % block.x;
      threadIdx.y = (t / block.x) % block.y;
      threadIdx.z = t / (block.x * block.y);
      g_cur = &f;
      tm_emu_switch(&g_sched_sp, f.sp);
      if (f.done) --live;
    }
  }
}

template <class F>
inline void launch(dim3 grid, dim3 block, F fn) {
  const long long total = (long long)grid.x * grid.y * grid.z;
  const unsigned nthr = block.x * block.y * block.z;
  std::atomic<long long> next{0};
  const std::function<void()> body = fn;
  auto worker = [&]() {
    std::vector<Fiber> fibers(nthr);
    for (auto& f : fibers) f.stack = (char*)malloc(kStack);
    blockDim = block;
    gridDim = grid;
    for (;;) {
      const long long b = next.fetch_add(1);
      if (b >= total) break;
      blockIdx.x = (unsigned)(b % grid.x);
      blockIdx.y = (unsigned)((b / grid.x) % grid.y);
      blockIdx.z = (unsigned)(b / ((long long)grid.x * grid.y));
      run_block(fibers, block, body);
    }
    for (auto& f : fibers) free(f.stack);
  };
  unsigned hw = std::thread::hardware_concurrency();
  long long nw = hw ? hw : 4;
  if (nw > total) nw = total;
  if (nw <= 1) { worker(); return; }
  std::vector<std::thread> pool;
  for (long long i = 0; i < nw; ++i) pool.emplace_back(worker);
  for (auto& t : pool) t.join();
}

}  // namespace emu

#define __syncthreads() emu::sync()
inline float rsqrtf(float x) { return 1.0f / sqrtf(x); }

namespace tmil {
inline thread_local char g_err[512] = "";
inline void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
}  // namespace tmil

#define TM_CHECK_ARG(cond, code, ...) \
  do {                                \
    if (!(cond)) {                    \
      tmil::set_error(__VA_ARGS__);   \
      return (code);                  \
    }                                 \
  } while (0)
#define TM_PROPAGATE(expr)      \
  do {                          \
    int _s = (expr);            \
    if (_s != TM_OK) return _s; \
  } while (0)
