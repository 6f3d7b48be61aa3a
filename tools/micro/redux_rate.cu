// REDUX.OR throughput per SM (results go to uniform registers).  Measured on B200: 0.5 warp instructions per
// cycle per SM from 16 warps on, ~12 cycles latency.  nvcc -gencode arch=compute_100a,code=sm_100a -o redux_rate redux_rate.cu
#include <cstdio>
#include <cstdint>
__global__ void k(uint32_t* out, int iters, long long* cyc) {
  uint32_t x = threadIdx.x * 2654435761u, acc = 0;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      acc += __reduce_or_sync(0xffffffffu, x);
      x = x * 1664525u + 1013904223u + u;
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  uint32_t* o; long long* c; cudaMalloc(&o, 148 * 1024 * 4); cudaMallocManaged(&c, 8);
  for (int threads : {32, 128, 256, 512, 1024}) {
    k<<<148, threads>>>(o, 2000, c); cudaDeviceSynchronize();
    k<<<148, threads>>>(o, 2000, c); cudaDeviceSynchronize();
    double per_sm = (double)(threads / 32) * 2000 * 8 / (double)*c;
    printf("threads %4d: %lld cycles, %.3f REDUX warp-instr / cycle / SM\n", threads, *c, per_sm);
  }
}
