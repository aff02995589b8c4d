// Micro-benchmarks behind DESIGN.md's kernel models (run on the B200 box):
//   tcgen05.ld / tcgen05.st register<->TMEM throughput per SM, MUFU.EX2 throughput per SM.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/microbench tools/microbench.cu && gpurun_out/microbench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int MODE>   // 0: ld x32 + wait each, 1: 4 ld in flight then wait, 2: st x32, 3: ex2, 4: ld x32 + 32 fmax, 5: cvt pack
__global__ void bench(int iters, long long* out, float* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t r[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) r[i] = threadIdx.x * 7 + i;
  float acc = 0.f, x = threadIdx.x * 1e-3f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0 || MODE == 4) {
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
            "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
            "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
            "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
          : "r"(base + (it & 7) * 32)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (MODE == 4) {
#pragma unroll
        for (int i = 0; i < 32; ++i) acc = fmaxf(acc, __uint_as_float(r[i]));
      } else acc += __uint_as_float(r[it & 31]);
    } else if (MODE == 1) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
              "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
              "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
              "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
            : "r"(base + k * 32)
            : "memory");
      }
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      acc += __uint_as_float(r[it & 31]);
    } else if (MODE == 2) {
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
          "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
          "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(base + (it & 7) * 32),
          "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
          "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]),
          "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]),
          "r"(r[31])
          : "memory");
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    } else if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        float y;
        asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x + i * 0.01f));
        acc += y;
      }
    } else if (MODE == 5) {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        uint32_t y;
        asm volatile("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(y) : "f"(x + i), "f"(acc + i));
        r[i] ^= y;
      }
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) s ^= r[i];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc + __uint_as_float(s);
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512) : "memory");
}

template <int MODE>
void run(const char* name, int warps, double units_per_iter_per_warp, const char* unit) {
  long long* d;
  float* sink;
  cudaMalloc(&d, 8 * 148);
  cudaMalloc(&sink, 4 * 148 * 1024);
  const int iters = 4096;
  bench<MODE><<<148, warps * 32>>>(iters, d, sink);
  bench<MODE><<<148, warps * 32>>>(iters, d, sink);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  double clk = (double)h[0];
  printf("%-28s warps/SM=%2d  %8.1f clk/iter  -> %8.1f %s per clk per SM   (%s)\n", name, warps, clk / iters,
         units_per_iter_per_warp * warps * iters / clk, unit, cudaGetErrorString(e));
  cudaFree(d);
  cudaFree(sink);
}

int main() {
  for (int w : {4, 8, 16}) {
    run<0>("tmem ld x32 (wait each)", w, 4096, "B");
    run<1>("tmem ld x32 (4 in flight)", w, 4 * 4096, "B");
    run<4>("tmem ld x32 + 32 fmax", w, 4096, "B");
    run<2>("tmem st x32 (wait each)", w, 4096, "B");
    run<3>("ex2.approx x32", w, 32 * 32, "ex2");
    run<5>("cvt f16x2 x32", w, 32 * 32, "cvt");
  }
  return 0;
}
