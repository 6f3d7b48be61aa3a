// device_utils.cuh — PTX wrappers and small helpers shared by the scoring kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"

namespace lxy {

// ---- PTX helpers (mbarrier + TMA bulk copy, sm_90+) ---------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst_smem, const void* src, uint32_t bytes,
                                             uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst_smem)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
#ifndef LXY_LDG_STREAM
#define LXY_LDG_STREAM "ld.global.nc.L1::no_allocate.v4.u32"
#endif
__device__ __forceinline__ uint4 ldg_stream_v4(const void* p) {
  uint4 r;
  asm volatile(LXY_LDG_STREAM " {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// One 16-byte packet for the whole warp through the uniform datapath: lanes 0..3 load one word each
// (one 16-byte request; a warp-uniform 128-bit load hands 512 bytes to the register file, four
// wavefronts of the L1 data pipe), four REDUX.OR spread the words (results live in uniform registers).
__device__ __forceinline__ uint32_t ldg_stream_word(const void* p, int lane) {
  uint32_t r = 0;
  if (lane < 4) asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"((const uint32_t*)p + lane));
  return r;
}
__device__ __forceinline__ uint4 warp_packet(uint32_t mine, int lane) {
  uint4 v;
  v.x = __reduce_or_sync(0xffffffffu, lane == 0 ? mine : 0u);
  v.y = __reduce_or_sync(0xffffffffu, lane == 1 ? mine : 0u);
  v.z = __reduce_or_sync(0xffffffffu, lane == 2 ? mine : 0u);
  v.w = __reduce_or_sync(0xffffffffu, lane == 3 ? mine : 0u);
  return v;
}

__host__ __device__ inline int64_t tile_row_start(int64_t I, int64_t nblk) {
  return I * nblk - I * (I - 1) / 2;
}

template <int WM> struct Acc;          // per-lane accumulator type
template <> struct Acc<W_UNIT> { typedef int64_t type; };
template <> struct Acc<W_INT32> { typedef int64_t type; };
template <> struct Acc<W_F64> { typedef double type; };


inline int grid_for(int64_t n, int threads, int sm) {
  int64_t g = (n + threads - 1) / threads;
  int64_t cap = (int64_t)sm * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace lxy
