// malloc_cost.cu — what a cudaMalloc / cudaFree pair costs on this box as a function of size and of
// whether the memory was written before it was freed (tools/micro; not part of the library).
//   nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o malloc_cost malloc_cost.cu && ./malloc_cost
#include <chrono>
#include <cstdio>
#include <cuda_runtime.h>

static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main() {
  cudaFree(0);
  const size_t sizes[] = {4ull << 20, 64ull << 20, 360ull << 20, 600ull << 20, 2048ull << 20};
  for (size_t sz : sizes) {
    for (int rep = 0; rep < 4; ++rep) {
      void* p = nullptr;
      double t0 = now_ms();
      cudaMalloc(&p, sz);
      double t1 = now_ms();
      if (rep >= 1) cudaMemset(p, 1, sz);  // rep 0: freed untouched
      cudaDeviceSynchronize();
      double t2 = now_ms();
      cudaFree(p);
      double t3 = now_ms();
      std::printf("%5zu MB rep %d: malloc %7.3f ms  memset %7.3f ms  free %7.3f ms\n", sz >> 20, rep, t1 - t0, t2 - t1, t3 - t2);
    }
  }
  // stream-ordered pool with a high release threshold
  cudaMemPool_t pool;
  cudaDeviceGetDefaultMemPool(&pool, 0);
  unsigned long long thr = ~0ull;
  cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
  cudaStream_t st;
  cudaStreamCreate(&st);
  for (int rep = 0; rep < 4; ++rep) {
    void* p = nullptr;
    double t0 = now_ms();
    cudaMallocAsync(&p, 600ull << 20, st);
    cudaStreamSynchronize(st);
    double t1 = now_ms();
    cudaMemsetAsync(p, 1, 600ull << 20, st);
    cudaStreamSynchronize(st);
    double t2 = now_ms();
    cudaFreeAsync(p, st);
    cudaStreamSynchronize(st);
    double t3 = now_ms();
    std::printf("pool 600 MB rep %d: malloc %7.3f ms  memset %7.3f ms  free %7.3f ms\n", rep, t1 - t0, t2 - t1, t3 - t2);
  }
  return 0;
}
