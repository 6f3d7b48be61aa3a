// Micro-benchmark: read-only streaming of a 400 MB buffer with 296 CTAs x 512 threads, 16 B per
// thread per step, two partitionings: (0) one contiguous share per CTA, (1) 8 KB chunks grid-strided,
// (2) 64 KB super-chunks grid-strided.  Build: nvcc -O3 -arch=sm_100a stream_read.cu -o stream_read
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(512, 2) k_read(const uint4* __restrict__ p, int64_t nvec, int mode, int depth, unsigned* out) {
  const int64_t chunk = 512;  // vectors per chunk (8 KB)
  const int64_t nch = nvec / chunk;
  unsigned acc = 0;
  if (mode == 0) {
    int64_t c0 = nch * blockIdx.x / gridDim.x, c1 = nch * (blockIdx.x + 1) / gridDim.x;
    for (int64_t c = c0; c < c1; c += 4) {
      uint4 v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = (c + q < c1) ? __ldg(p + (c + q) * chunk + threadIdx.x) : make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int q = 0; q < 4; ++q) acc += v[q].x ^ v[q].y ^ v[q].z ^ v[q].w;
    }
  } else {
    const int64_t sc = (mode == 1) ? 1 : 8;  // chunks per super-chunk
    const int64_t nsc = nch / sc;
    for (int64_t s = blockIdx.x; s < nsc; s += gridDim.x) {
      for (int64_t c = s * sc; c < (s + 1) * sc; c += 4) {
        uint4 v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) v[q] = (c + q < (s + 1) * sc) ? __ldg(p + (c + q) * chunk + threadIdx.x) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int q = 0; q < 4; ++q) acc += v[q].x ^ v[q].y ^ v[q].z ^ v[q].w;
      }
    }
  }
  if (acc == 0x12345678u) out[0] = acc;
}
int main() {
  const int64_t bytes = 400ll << 20;
  uint4* p; unsigned* out;
  cudaMalloc(&p, bytes); cudaMalloc(&out, 4); cudaMemset(p, 1, bytes);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int grid : {148, 296, 592, 1184})
    for (int mode = 0; mode < 3; ++mode) {
      float best = 1e9;
      for (int it = 0; it < 6; ++it) {
        cudaEventRecord(e0);
        k_read<<<grid, 512>>>(p, bytes / 16, mode, 4, out);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
      }
      printf("grid %4d mode %d: %.1f us  %.0f GB/s\n", grid, mode, best * 1e3, bytes / best / 1e6);
    }
  return 0;
}
