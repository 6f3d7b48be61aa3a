// Throughput of the integer ops the G-profile kernel leans on (one SM's worth per CTA, 148 CTAs):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o alu_rate alu_rate.cu && ./alu_rate
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void k(uint32_t* out, uint32_t seed, int iters) {
  uint32_t a[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a[j] = seed + threadIdx.x * 7 + j * 13;
  uint32_t b = seed ^ 0x9E3779B9u, c = seed * 3 + 1;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (OP == 0) a[j] = __sad((int)a[j], (int)b, 0u);                 // VABSDIFF
      if (OP == 1) a[j] = a[j] + b + c;                                // IADD3
      if (OP == 2) a[j] = min(a[j], b) + 0;                            // VIMNMX
      if (OP == 3) a[j] = __umulhi(a[j], b);                           // IMAD.HI
      if (OP == 4) a[j] = a[j] * b + c;                                // IMAD
      if (OP == 5) a[j] = (a[j] >> 3) ^ b;                             // SHF + LOP3
      if (OP == 6) a[j] = __usad(a[j], b, c);                          // VABSDIFF.U32 with accumulate
      if (OP == 7) a[j] = (uint32_t)abs((int)a[j] - (int)b);           // IADD + IABS
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) s ^= a[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
void run(const char* name, uint32_t* d) {
  const int iters = 20000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<OP><<<148, 1024>>>(d, 12345, 100);
  cudaEventRecord(e0);
  k<OP><<<148, 1024>>>(d, 12345, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double warp_instr = 148.0 * 32 * iters * 8;   // per op kind (ops with two SASS instructions count once)
  printf("%-28s %8.3f ms  %6.2f warp-ops/cycle/SM (at 1.965 GHz)\n", name, ms, warp_instr / 148 / (ms * 1e-3 * 1.965e9));
}

int main() {
  uint32_t* d; cudaMalloc(&d, 148 * 1024 * 4);
  run<0>("VABSDIFF", d); run<6>("VABSDIFF.U32 + acc", d); run<1>("IADD3", d); run<2>("VIMNMX", d);
  run<3>("IMAD.HI", d); run<4>("IMAD", d); run<5>("SHF + LOP3 (2 instr)", d); run<7>("IADD + IABS (2 instr)", d);
  return 0;
}
