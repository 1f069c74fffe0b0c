// Micro-benchmark: does bar.sync wait for a warp's outstanding global stores?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bar_stg bar_stg.cu && ./bar_stg
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* out, long long* cyc, int mode, int iters, int work) {
  const int tid = threadIdx.x, warp = tid >> 5;
  float acc = tid;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    if (mode == 1 && warp == 0) out[(size_t)blockIdx.x * 1048576 + (i & 1023) * 32 + (tid & 31)] = acc;          // store right before the barrier
    if (mode == 2 && warp == 0) out[(size_t)blockIdx.x * 1048576 + (i & 1023) * 32 + (tid & 31)] = acc;          // store, then work, then barrier
    if (mode == 3) out[(size_t)blockIdx.x * 1048576 + (i & 1023) * 1024 + tid] = acc;                            // every warp stores
    if (mode == 2 || mode == 4)
      for (int j = 0; j < work; ++j) acc = acc * 1.0001f + 0.5f;
    asm volatile("bar.sync 0;" ::: "memory");
  }
  long long t1 = clock64();
  if (tid == 0) cyc[blockIdx.x] = t1 - t0;
  if (acc == 12345.f) out[0] = acc;
}
int main() {
  float* out; long long* cyc;
  cudaMalloc(&out, (size_t)64 * 1048576 * 4 + 4096 * 1024);
  cudaMalloc(&cyc, 64 * 8);
  const int iters = 2000;
  for (int threads : {256, 736})
    for (int mode = 0; mode <= 4; ++mode) {
      k<<<64, threads>>>(out, cyc, mode, iters, 100);
      cudaDeviceSynchronize();
      k<<<64, threads>>>(out, cyc, mode, iters, 100);
      long long h[64];
      cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
      printf("threads %d mode %d: %.1f cycles/iter  (%s)\n", threads, mode, (double)h[0] / iters, cudaGetErrorString(cudaGetLastError()));
    }
  return 0;
}
