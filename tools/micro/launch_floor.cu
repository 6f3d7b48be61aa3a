// launch_floor.cu — what an event-bracketed launch of 296 x 512 threads costs before any work is done,
// with and without 78 KB of dynamic shared memory: the part of a kernel's duration that no kernel code controls.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o launch_floor launch_floor.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ unsigned long long g_t[2048];
__global__ void k_empty(int stamp) {
  extern __shared__ unsigned char sm[];
  if (stamp && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    g_t[blockIdx.x] = t;
  }
}
int main() {
  cudaStream_t st;
  cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaFuncSetAttribute(k_empty, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(k_empty, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  const int grids[] = {1, 148, 296, 296}, smems[] = {0, 0, 0, 78 * 1024};
  for (int v = 0; v < 4; ++v) {
    float best = 1e9f, sum = 0;
    for (int it = 0; it < 50; ++it) {
      cudaEventRecord(a, st);
      k_empty<<<grids[v], 512, smems[v], st>>>(1);
      cudaEventRecord(b, st);
      cudaStreamSynchronize(st);
      float ms; cudaEventElapsedTime(&ms, a, b);
      if (it >= 10) { sum += ms; if (ms < best) best = ms; }
    }
    printf("grid %3d smem %6d: events around one launch: best %.2f us, mean %.2f us\n", grids[v], smems[v], best * 1e3f, sum / 40 * 1e3f);
  }
  // two back-to-back event records with nothing in between
  float sum = 0;
  for (int it = 0; it < 50; ++it) {
    cudaEventRecord(a, st); cudaEventRecord(b, st); cudaStreamSynchronize(st);
    float ms; cudaEventElapsedTime(&ms, a, b);
    if (it >= 10) sum += ms;
  }
  printf("two event records, nothing between: mean %.2f us\n", sum / 40 * 1e3f);
  return 0;
}
