// Micro-benchmark: cost of one CTA-wide barrier per loop iteration vs CTA size (1 CTA per SM).
#include <cstdio>
#include <cuda_runtime.h>
template <int KIND>
__global__ void k(long long* cyc, int iters, float* sink) {
  long long t0 = clock64();
  float acc = threadIdx.x;
  for (int i = 0; i < iters; ++i) {
    if (KIND == 0) asm volatile("bar.sync 0;" ::: "memory");
    if (KIND == 1) __syncthreads();
    if (KIND == 2) asm volatile("bar.sync 0;");           // no memory clobber
    if (KIND == 3) { acc = acc * 1.0001f + 0.5f; }        // no barrier: loop overhead only
    if (KIND == 4) { asm volatile("bar.sync 0;" ::: "memory"); asm volatile("bar.sync 0;" ::: "memory"); }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  if (acc == 12345.f) sink[0] = acc;
}
int main() {
  long long* cyc; float* sink;
  cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 4);
  const int iters = 4000;
  for (int threads : {32, 64, 128, 256, 512, 736, 1024}) {
    long long h[148]; double r[5];
    for (int kind = 0; kind < 5; ++kind) {
      for (int rep = 0; rep < 2; ++rep) {
        if (kind == 0) k<0><<<64, threads>>>(cyc, iters, sink);
        if (kind == 1) k<1><<<64, threads>>>(cyc, iters, sink);
        if (kind == 2) k<2><<<64, threads>>>(cyc, iters, sink);
        if (kind == 3) k<3><<<64, threads>>>(cyc, iters, sink);
        if (kind == 4) k<4><<<64, threads>>>(cyc, iters, sink);
        cudaDeviceSynchronize();
      }
      cudaMemcpy(h, cyc, sizeof(long long) * 64, cudaMemcpyDeviceToHost);
      r[kind] = (double)h[0] / iters;
    }
    printf("threads %4d: bar.sync %.1f  __syncthreads %.1f  bar(no clobber) %.1f  loop only %.1f  2x bar %.1f cycles/iter\n",
           threads, r[0], r[1], r[2], r[3], r[4]);
  }
  return 0;
}
