// scaff_kernels.cu — scaffolding fitness on sm_100a (R-profile = /root/reference/src/lxy/scaff/
// scaff.go:246-310 in its link-centric form, derivation in DESIGN.md §2).
//
// Pipeline of one lxy_scaff_score_batch call:
//   k_scaff_layout   ordering -> placement tables (inverse permutation; offsets = ordering index)
//   k_scaff_score    the link-streaming kernel: tables for kCand candidates x 2 contig blocks are
//                    staged in shared memory by TMA bulk copies, 2-byte link records are streamed
//                    with 128-bit loads, one lane owns CPL candidates, so every record byte is
//                    reused by kCand candidates and the table gathers are bank-conflict free.
//   k_scaff_head     the reference's `i-k > 0` quirk (scaff.go:256): pairs whose earlier member is
//                    at ordering index 0 count once, everything else twice.  Instead of testing it
//                    per evaluation the main kernel sums every record once and this kernel looks up
//                    the six pairs (g[0], g[k]) in the sorted tiles.
//   k_scaff_finish   fixed-order reduction of the per-CTA partials, fitness = -(2S - head)/100.
#include <cstring>

#include "common.cuh"
#include "device_utils.cuh"

namespace lxy {
namespace {

#ifndef LXY_SCORE_THREADS
#define LXY_SCORE_THREADS 768
#endif
constexpr int kThreads = LXY_SCORE_THREADS;
constexpr int kWarps = kThreads / 32;
constexpr int kPacket = 8;  // records per 128-bit load

// T*100 by v = d + kBias + 1; 0 at the far entries 0, 42.  Initialised with the module (no copy at
// run time: a lazily queued cudaMemcpyToSymbolAsync on one context's stream could be overtaken by a
// second context of the same device).
__constant__ uint8_t c_lut100[kLutLen] = {
    /* far */ 0, /* d=-20 */ 5, 0, 0, 0, 0, 0, 0, 0, 0, /* -11 */ 10, 0, 0, 0, 0, 0, /* -5 */ 20, 0,
    /* -3 */ 33, /* -2 */ 50, /* -1 */ 100, /* 0 */ 0, 100, 50, 33, 0, 20, 0, 0, 0, 0, 0,
    /* 11 */ 10, 0, 0, 0, 0, 0, 0, 0, 0, /* 20 */ 5, /* far */ 0, 0, 0, 0, 0, 0};
__constant__ int c_steps[6] = {1, 2, 3, 5, 11, 20};
__constant__ int c_w100[6] = {100, 50, 33, 20, 10, 5};

// ---- k_scaff_layout -----------------------------------------------------------------------------
// Tables are stored group-major: tab[g][contig][cand] with kCand = 32*CPL candidates per group, so
// the rows of one contig block are one contiguous TMA bulk copy.  tabA holds pos + kBias + 1 and
// tabB holds -pos (16-bit, read modulo 2^16, while n <= 65515, else 32-bit): one VIADDMNMX per
// candidate (pair) turns a row of each into decay-table indices (eval_record).
template <int WIDE>
__global__ void k_scaff_layout(const int32_t* __restrict__ perm, int32_t P, int32_t Ppad, int32_t n,
                               int32_t cand, void* __restrict__ tabA, void* __restrict__ tabB,
                               int* __restrict__ flag) {
  int64_t total = (int64_t)Ppad * n;
  int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; idx < total; idx += stride) {
    int32_t p = (int32_t)(idx / n), i = (int32_t)(idx - (int64_t)p * n);
    int32_t cid = (p < P) ? perm[idx] : i;  // pad candidates = identity ordering (discarded)
    if ((uint32_t)cid >= (uint32_t)n) {
      atomicOr(flag, 1);
      continue;
    }
    int32_t g = p / cand, pc = p - g * cand;
    size_t at = ((size_t)g * n + cid) * cand + pc;
    if (WIDE == 1) {
      ((uint32_t*)tabA)[at] = (uint32_t)(i + kBias + 1);
      ((uint32_t*)tabB)[at] = (uint32_t)(-i);
    } else {
      ((uint16_t*)tabA)[at] = (uint16_t)(i + kBias + 1);
      ((uint16_t*)tabB)[at] = (uint16_t)(-i);
    }
  }
}

// pos[perm[i]] == i for every i  <=>  the row is a permutation (given the range check above).
template <int WIDE>
__global__ void k_scaff_check(const int32_t* __restrict__ perm, int32_t P, int32_t n, int32_t cand,
                              const void* __restrict__ tabB, int* __restrict__ flag) {
  int64_t total = (int64_t)P * n;
  int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; idx < total; idx += stride) {
    int32_t p = (int32_t)(idx / n), i = (int32_t)(idx - (int64_t)p * n);
    int32_t cid = perm[idx];
    if ((uint32_t)cid >= (uint32_t)n) continue;
    int32_t g = p / cand, pc = p - g * cand;
    size_t at = ((size_t)g * n + cid) * cand + pc;
    const uint32_t got = WIDE == 1 ? ((const uint32_t*)tabB)[at] : ((const uint16_t*)tabB)[at];
    const uint32_t want = WIDE == 1 ? (uint32_t)(-i) : (uint32_t)(uint16_t)(-i);
    if (got != want) atomicOr(flag, 2);
  }
}

// ---- placement tables in two coalesced passes (16-bit tables) ---------------------------------------
// k_scaff_layout scatters 2-byte stores over the whole table (one 32-byte sector per store: 2.2 ms of a
// 41 ms step at 50,000 contigs x 1,024 candidates, and it does not shrink when the records are sharded).
// Here the inverse permutation of ONE candidate is built in shared memory (the scatter stays on chip),
// checked there (pos[perm[i]] == i for all i <=> permutation; one table per CTA, so the loser of two
// stores to a duplicated id always sees a mismatch), and written out as a contiguous row; a second
// kernel transposes [candidate][contig] into the group-major [group][contig][candidate] tables.
__global__ void __launch_bounds__(512) k_scaff_inv16(const int32_t* __restrict__ perm, int32_t n,
                                                     uint16_t* __restrict__ inv, int validate, int* __restrict__ flag) {
  extern __shared__ uint16_t s_inv[];
  const int p = blockIdx.x;
  const int32_t* row = perm + (size_t)p * n;
  int bad = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int32_t cid = row[i];
    if ((uint32_t)cid >= (uint32_t)n) bad |= 1;
    else s_inv[cid] = (uint16_t)i;
  }
  __syncthreads();
  if (validate) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int32_t cid = row[i];
      if ((uint32_t)cid < (uint32_t)n && (uint32_t)s_inv[cid] != (uint32_t)(i & 0xFFFF)) bad |= 2;
    }
  }
  if (bad) atomicOr(flag, bad);
  uint16_t* out = inv + (size_t)p * n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = s_inv[i];
}

// One CTA per (64 contigs, candidate group): inv[p][c] -> tabA[g][c][pc] = pos + kBias + 1, tabB = -pos.
template <int CAND>
__global__ void __launch_bounds__(256) k_scaff_tables16(const uint16_t* __restrict__ inv, int32_t P, int32_t n,
                                                        uint16_t* __restrict__ tabA, uint16_t* __restrict__ tabB) {
  constexpr int kC = 64;
  __shared__ uint16_t tile[CAND][kC + 2];
  const int c0 = blockIdx.x * kC, g = blockIdx.y;
  for (int k = threadIdx.x; k < CAND * kC; k += 256) {
    const int pc = k / kC, c = k - pc * kC, p = g * CAND + pc;
    uint16_t v = (uint16_t)(c0 + c);                       // pad candidates: identity ordering (discarded)
    if (p < P && c0 + c < n) v = inv[(size_t)p * n + c0 + c];
    tile[pc][c] = v;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < CAND * kC; k += 256) {
    const int c = k / CAND, pc = k - c * CAND;
    if (c0 + c < n) {
      const uint32_t pos = tile[pc][c];
      const size_t at = ((size_t)g * n + c0 + c) * CAND + pc;
      tabA[at] = (uint16_t)(pos + kBias + 1);
      tabB[at] = (uint16_t)(0u - pos);
    }
  }
}

// ---- k_scaff_score ------------------------------------------------------------------------------
struct ScoreParams {
  const uint16_t* rec;        // sorted records
  const void* w;              // sorted weights (int32 / double) or null
  const int64_t* tile_start;  // [ntiles + 1]
  const void* tabA;           // [groups][n][cand]
  const void* tabB;
  void* partial;              // [units][cand] int64 / double
  int64_t L;
  int32_t n, nblk;
  int32_t groups, splits;     // units = groups * splits
};

// Per-lane state of the streaming loop: CPL candidates, their running sums.
template <int CPL, int WM>
struct Lane {
  int32_t a32[CPL];   // W_UNIT: 100*T summed per record (flushed into a64 before it can overflow)
  int64_t a64[CPL];
  double ad[CPL];
  uint32_t p32[(CPL + 1) / 2];  // pair look-ups (W_UNIT): two 16-bit sums per word, flushed every <= 655 records
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int j = 0; j < CPL; ++j) { a32[j] = 0; a64[j] = 0; ad[j] = 0.0; }
#pragma unroll
    for (int q = 0; q < (CPL + 1) / 2; ++q) p32[q] = 0;
  }
  __device__ __forceinline__ void flush16() {
    if (WM == W_UNIT) {
#pragma unroll
      for (int q = 0; q < CPL / 2; ++q) {
        a32[2 * q] += (int32_t)(p32[q] & 0xFFFFu);
        a32[2 * q + 1] += (int32_t)(p32[q] >> 16);
        p32[q] = 0;
      }
    }
  }
  __device__ __forceinline__ void flush() {
    if (WM == W_UNIT) {
#pragma unroll
      for (int j = 0; j < CPL; ++j) { a64[j] += a32[j]; a32[j] = 0; }
    }
  }
};

// This lane's slice of one table row (CPL candidates) as up to four 32-bit words.
template <int CPL, int WIDE>
__device__ __forceinline__ void load_row(const uint8_t* row, uint32_t (&X)[4]) {
  constexpr int BYTES = CPL * (WIDE == 1 ? 4 : 2);
  if (BYTES == 16) {
    const uint4 a = *(const uint4*)row;
    X[0] = a.x; X[1] = a.y; X[2] = a.z; X[3] = a.w;
  } else if (BYTES == 8) {
    const uint2 a = *(const uint2*)row;
    X[0] = a.x; X[1] = a.y;
  } else if (BYTES == 4) {
    X[0] = *(const uint32_t*)row;
  } else {
    X[0] = *(const uint16_t*)row;
  }
}

// Row A of a record / packet.
template <int CPL, int WIDE>
__device__ __forceinline__ void load_row_a(const uint8_t* row, uint32_t (&A)[4]) {
  load_row<CPL, WIDE>(row, A);
}

// Pair decay table (integer modes, even CPL).  VIADDMNMX.U16x2 leaves the indices of two candidates in
// one register, v0 | v1 << 16; one IDP.2A turns that into the shared address of
// lut2[v1][v0] = lut[v0] | lut[v1] << 16, so ONE 32-bit look-up serves two candidates and the two
// 16-bit halves are summed in one IADD: 6.25 -> 4.25 shared-memory wavefronts per record-warp.  Rows
// are kLut2Stride words, an odd number, so that the entries a warp touches at once (mostly (43, 43),
// some (v, 43) and (43, v)) spread over all 32 banks.
constexpr int kLut2Stride = 45;
constexpr int kLut2Words = kLutLen * kLut2Stride;
template <int CPL, int WIDE, int WM>
struct PairLut { static constexpr bool on = (WM != W_F64) && (CPL % 2 == 0); };

__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

// One record against the CPL candidates of this lane, row A already in registers.
//   16-bit tables: A = pos + kBias + 1, B = -pos (mod 2^16); VIADDMNMX.U16x2 gives, per half-word,
//   v = min((pos a - pos b + kBias + 1) mod 2^16, kMiss + 2): one instruction for two candidates.
//   v in [1, 41] <=> |d| <= 20; 0 and 42 are d = -21 / +21, every other difference lands on 43
//   (exact while n <= 65515: |d| <= n - 1 cannot alias into [0, 42] modulo 2^16).
//   32-bit tables: VIADDMNMX.RELU on full words, v = clamp(d + 21, 0, 42); 0 and 42 are "far".
// v indexes the decay table (zero outside the six weighted steps).
template <int CPL, int WIDE, int WM>
__device__ __forceinline__ void eval_record(const uint32_t (&A)[4], const uint8_t* rowB,
                                            const uint8_t* lut, const double* lutd, uint32_t lut2,
                                            int32_t wi, double wd, Lane<CPL, WM>& L) {
  constexpr uint32_t kHi = kMiss + 1;  // 42
  uint32_t idx[CPL], B[4];
  load_row<CPL, WIDE>(rowB, B);
  if (PairLut<CPL, WIDE, WM>::on) {
#pragma unroll
    for (int q = 0; q < CPL / 2; ++q) {
      uint32_t addr;
      if (WIDE == 1) {  // 32-bit tables: one word per candidate, two IMADs for the address
        const uint32_t v0 = (uint32_t)__viaddmin_s32_relu((int)A[2 * q], (int)B[2 * q], (int)kHi);
        const uint32_t v1 = (uint32_t)__viaddmin_s32_relu((int)A[2 * q + 1], (int)B[2 * q + 1], (int)kHi);
        addr = v1 * (4u * kLut2Stride) + (v0 * 4u + lut2);
      } else {  // 16-bit tables, sums modulo 2^16: every miss is index 43
        const uint32_t m = __viaddmin_u16x2(A[q], B[q], (kHi + 1) | ((kHi + 1) << 16));
        addr = __dp2a_lo(m, 4u | ((4u * kLut2Stride) << 8), lut2);
      }
      const uint32_t e = lds_u32(addr);
      if (WM == W_UNIT) {
        L.p32[q] += e;
      } else {
        L.a64[2 * q] += (int64_t)(int32_t)(e & 0xFFFFu) * (int64_t)wi;
        L.a64[2 * q + 1] += (int64_t)(int32_t)(e >> 16) * (int64_t)wi;
      }
    }
    return;
  }
  if (WIDE == 1) {
#pragma unroll
    for (int j = 0; j < CPL; ++j) idx[j] = (uint32_t)__viaddmin_s32_relu((int)A[j], (int)B[j], (int)kHi);
  } else {
#pragma unroll
    for (int q = 0; q < (CPL + 1) / 2; ++q) {
      const uint32_t m = __viaddmin_u16x2(A[q], B[q], (kHi + 1) | ((kHi + 1) << 16));
      idx[2 * q] = m & 0xFFFFu;
      if (2 * q + 1 < CPL) idx[2 * q + 1] = m >> 16;
    }
  }
#pragma unroll
  for (int j = 0; j < CPL; ++j) {
    if (WM == W_UNIT) L.a32[j] += lut[idx[j]];
    else if (WM == W_INT32) L.a64[j] += (int64_t)((int32_t)lut[idx[j]]) * (int64_t)wi;
    else L.ad[j] += lutd[idx[j]] * wd;
  }
}

// Records k in [k0, k1) of one 8-record packet (k0 = 0, k1 = 8 compiles to straight-line code).
// SAME_A: every record of the packet has the same a (records are sorted by (a, b) inside a tile,
// so this is the common case): row A is gathered once for the whole packet.
template <int CPL, int WIDE, int WM, bool SAME_A>
__device__ __forceinline__ void eval_packet(const uint4 v, const uint8_t* laneA, const uint8_t* laneB,
                                            uint32_t lane_off, const uint8_t* lut, const double* lutd,
                                            uint32_t lut2, const int32_t* wi, const double* wd, int k0, int k1,
                                            Lane<CPL, WM>& L) {
  constexpr int ROWB = 32 * CPL * (WIDE == 1 ? 4 : 2);
  const uint32_t words[4] = {v.x, v.y, v.z, v.w};
  uint32_t A[4] = {0, 0, 0, 0};
  uint32_t a_prev = 0xFFFFFFFFu;
  if (SAME_A) {
    if (ROWB == 256) load_row_a<CPL, WIDE>(laneA - lane_off + __byte_perm(v.x, lane_off, 0x7614), A);
    else load_row_a<CPL, WIDE>(laneA + ((v.x >> 8) & 0xFFu) * ROWB, A);
  }
#pragma unroll
  for (int k = 0; k < kPacket; ++k) {
    if (k < k0 || k >= k1) continue;
    const uint32_t wv = words[k >> 1];
    const uint8_t *rowA, *rowB;
    if (ROWB == 256) {
      // one PRMT builds (a_loc << 8 | lane_off): lane_off < 256 occupies byte 0 only
      const uint32_t oa = __byte_perm(wv, lane_off, (k & 1) ? 0x7634 : 0x7614);
      const uint32_t ob = __byte_perm(wv, lane_off, (k & 1) ? 0x7624 : 0x7604);
      rowA = laneA - lane_off + oa;
      rowB = laneB - lane_off + ob;
    } else {
      const uint32_t r = (k & 1) ? (wv >> 16) : (wv & 0xFFFFu);
      rowA = laneA + (r >> 8) * ROWB;
      rowB = laneB + (r & 0xFFu) * ROWB;
    }
    if (!SAME_A) {  // a packet that crosses from one a to the next: row A again only where a changes
      const uint32_t a_k = (k & 1) ? (wv >> 24) : ((wv >> 8) & 0xFFu);   // warp-uniform
      if (a_k != a_prev) load_row_a<CPL, WIDE>(rowA, A);
      a_prev = a_k;
    }
    eval_record<CPL, WIDE, WM>(A, rowB, lut, lutd, lut2, WM == W_INT32 ? wi[k] : 0,
                               WM == W_F64 ? wd[k] : 0.0, L);
  }
}

template <int WM>
__device__ __forceinline__ void load_weights(const void* w, int64_t base, int32_t (&wi)[kPacket],
                                             double (&wd)[kPacket]) {
  // the weight arrays are padded past L with zeros (link_store.cu), so whole packets are readable
  if (WM == W_INT32) {
    const uint4* p = reinterpret_cast<const uint4*>((const int32_t*)w + base);
    const uint4 x = __ldg(p), y = __ldg(p + 1);
    wi[0] = x.x; wi[1] = x.y; wi[2] = x.z; wi[3] = x.w; wi[4] = y.x; wi[5] = y.y; wi[6] = y.z; wi[7] = y.w;
  } else if (WM == W_F64) {
    const double2* p = reinterpret_cast<const double2*>((const double*)w + base);
#pragma unroll
    for (int q = 0; q < 4; ++q) { const double2 d = __ldg(p + q); wd[2 * q] = d.x; wd[2 * q + 1] = d.y; }
  }
}

constexpr int kRingPackets = 64;                  // packets per ring slot: the 16-bit pair sums are unpacked once per slot
constexpr int kRingBytes = kRingPackets * 16;
__device__ __forceinline__ uint32_t lds_u32s(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void mbar_wait_u32(uint32_t bar_smem, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(bar_smem),
      "r"(parity)
      : "memory");
}

template <int CPL, int WIDE, int WM>
__global__ void __launch_bounds__(kThreads, 1) k_scaff_score(ScoreParams prm) {
  constexpr int CAND = 32 * CPL;
  constexpr int ES = WIDE == 1 ? 4 : 2;
  constexpr int ROWB = CAND * ES;           // bytes per contig row
  constexpr int TABB = kBlk * ROWB;         // bytes per table
  typedef typename Acc<WM>::type acc_t;
  // dynamic shared memory: table A | table B | decay tables (u8 and f64) | mbarrier
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + TABB;
  uint8_t* s_lut = smem + 2 * TABB;                       // kLutLen bytes
  double* s_lutd = (double*)(smem + 2 * TABB + 64);       // kLutLen doubles
  uint64_t* bar = (uint64_t*)(smem + 2 * TABB + 64 + 8 * kLutLen);
  uint32_t* s_lut2 = (uint32_t*)(smem + 2 * TABB + 64 + 8 * kLutLen + 16);  // [kLutLen][kLut2Stride]
  const uint32_t lut2 = smem_u32(s_lut2);
  // per-warp record ring: two slots of kRingPackets packets and their mbarriers
  const uint32_t ring_w = smem_u32(s_lut2 + kLut2Words) + (uint32_t)(threadIdx.x >> 5) * (2 * kRingBytes);
  const uint32_t ring_bar = smem_u32(s_lut2 + kLut2Words) + (uint32_t)kWarps * (2 * kRingBytes) + (uint32_t)(threadIdx.x >> 5) * 16;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < kLutLen) {
    s_lut[tid] = c_lut100[tid];
    s_lutd[tid] = (double)c_lut100[tid];
  }
  if (tid == 0) mbar_init(bar, 1);
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ring_bar) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ring_bar + 8) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  uint32_t ring_ph = 0;   // bit s: phase of slot s's mbarrier
  __syncthreads();
  if (PairLut<CPL, WIDE, WM>::on) {
    for (int i = tid; i < kLut2Words; i += kThreads) {
      const int v1 = i / kLut2Stride, v0 = i - v1 * kLut2Stride;
      s_lut2[i] = (v0 < kLutLen ? (uint32_t)s_lut[v0] : 0u) | ((uint32_t)s_lut[v1] << 16);
    }
    __syncthreads();
  }
  uint32_t parity = 0;
  const uint32_t lane_off = lane * (CPL * ES);
  const uint8_t* laneA = sA + lane_off;
  const uint8_t* laneB = sB + lane_off;

  const int units = prm.groups * prm.splits;
  const int64_t npk_all = (prm.L + kPacket - 1) / kPacket;  // packets overall
  const int64_t ntiles = (int64_t)prm.nblk * (prm.nblk + 1) / 2;
  for (int u = blockIdx.x; u < units; u += gridDim.x) {
    const int g = u / prm.splits, s = u - g * prm.splits;
    // even split of the packets, so every unit boundary is 16-byte aligned
    const int64_t lo = (npk_all * s / prm.splits) * kPacket;
    const int64_t hi = min((npk_all * (s + 1) / prm.splits) * kPacket, prm.L);
    Lane<CPL, WM> acc;
    acc.clear();
    int32_t loadedI = -1;   // block of table A in shared memory (none: the reduction below reuses the area)

    if (lo < hi) {
      // tile containing record `lo`: largest t with tile_start[t] <= lo
      int64_t tl = 0, th = ntiles;  // invariant: tile_start[tl] <= lo < tile_start[th]
      while (th - tl > 1) {
        const int64_t mid = (tl + th) >> 1;
        if (prm.tile_start[mid] <= lo) tl = mid; else th = mid;
      }
      int64_t t = tl;
      int32_t I = 0;
      {
        int32_t a = 0, b = prm.nblk;  // tile_row_start(a) <= t < tile_row_start(b)
        while (b - a > 1) {
          const int32_t m = (a + b) >> 1;
          if (tile_row_start(m, prm.nblk) <= t) a = m; else b = m;
        }
        I = a;
      }
      int32_t J = I + (int32_t)(t - tile_row_start(I, prm.nblk));

      int64_t cur = lo;
      while (cur < hi) {
        const int64_t tend = prm.tile_start[t + 1];
        if (tend <= cur) {  // empty tile (or finished): next
          ++t; ++J;
          if (J == prm.nblk) { ++I; J = I; }
          continue;
        }
        const int64_t seg_end = min(hi, tend);
        // ---- stage the two contig-block tables of tile (I, J) for group g (TMA bulk copies) ----
        __syncthreads();  // everyone is done with the previous tables
        if (tid == 0) {
          // tiles are walked row by row: block I of table A stays while J moves along the row
          const bool needA = (I != loadedI);
          const int rowsA = needA ? min(kBlk, prm.n - I * kBlk) : 0, rowsB = min(kBlk, prm.n - J * kBlk);
          const uint8_t* gA = (const uint8_t*)prm.tabA + ((size_t)g * prm.n + (size_t)I * kBlk) * ROWB;
          const uint8_t* gB = (const uint8_t*)prm.tabB + ((size_t)g * prm.n + (size_t)J * kBlk) * ROWB;
          fence_proxy_async();
          mbar_expect_tx(bar, (uint32_t)(rowsA + rowsB) * ROWB);
          if (needA) tma_bulk_g2s(sA, gA, (uint32_t)rowsA * ROWB, bar);
          tma_bulk_g2s(sB, gB, (uint32_t)rowsB * ROWB, bar);
        }
        loadedI = I;
        mbar_wait(bar, parity);
        parity ^= 1;

        // ---- stream the records of [cur, seg_end): 8-record packets, one 128-bit load each ----
        const int64_t p_first = cur / kPacket, p_last = (seg_end - 1) / kPacket;
        const bool head_partial = (p_first * kPacket < cur);
        const bool tail_partial = ((p_last + 1) * kPacket > seg_end);
        int32_t wi[kPacket];
        double wd[kPacket];
        // boundary packets (shared with a neighbouring tile / unit): masked, one warp each
        if (head_partial && warp == 0) {
          const int64_t base = p_first * kPacket;
          const uint4 v = ldg_stream_v4(prm.rec + base);
          load_weights<WM>(prm.w, base, wi, wd);
          const int k0 = (int)(cur - base), k1 = (int)min((int64_t)kPacket, seg_end - base);
#pragma unroll 1
          for (int k = k0; k < k1; ++k)
            eval_packet<CPL, WIDE, WM, false>(v, laneA, laneB, lane_off, s_lut, s_lutd, lut2, wi, wd, k, k + 1, acc);
          acc.flush16();
        }
        if (tail_partial && warp == 1 && !(head_partial && p_last == p_first)) {
          const int64_t base = p_last * kPacket;
          const uint4 v = ldg_stream_v4(prm.rec + base);
          load_weights<WM>(prm.w, base, wi, wd);
          const int k1 = (int)(seg_end - base);
#pragma unroll 1
          for (int k = 0; k < k1; ++k)
            eval_packet<CPL, WIDE, WM, false>(v, laneA, laneB, lane_off, s_lut, s_lutd, lut2, wi, wd, k, k + 1, acc);
          acc.flush16();
        }
        // full packets: straight-line code, next packet's load in flight while this one is scored
        const int64_t f0 = p_first + (head_partial ? 1 : 0);
        const int64_t f1 = p_last - (tail_partial ? 1 : 0);  // inclusive
        const int64_t nfull = f1 - f0 + 1;
        if (nfull > 0) {
          const uint16_t* recp = prm.rec + f0 * kPacket;
          const int64_t wbase = f0 * kPacket;
          // Every warp takes a CONTIGUOUS run of the segment's packets, in slices of kRingPackets that
          // arrive in its own two-slot ring, each by one bulk copy (UBLKCP) counted on the slot's
          // mbarrier: a slice lasts the warp ~30 us, so the next one is in shared memory long before
          // it is needed, with no other candidate group walking the same stretch (one group per GPU at
          // 8 GPUs: 3.57 ms against 3.37 per group when four groups share the stream through L2).
          const int64_t w_lo = nfull * warp / kWarps, w_hi = nfull * (warp + 1) / kWarps;
          const int64_t nsl = (w_hi - w_lo + kRingPackets - 1) / kRingPackets;
          auto fill = [&](int64_t sl) {   // lane 0
            const int64_t p0 = w_lo + sl * kRingPackets;
            const uint32_t bytes = (uint32_t)min((int64_t)kRingPackets, w_hi - p0) * 16u;
            const uint32_t b = ring_bar + (uint32_t)(sl & 1) * 8, dst = ring_w + (uint32_t)(sl & 1) * kRingBytes;
            asm volatile(
                "mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %3;\n\t"
                "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%1], [%2], %3, [%0];" ::"r"(b),
                "r"(dst), "l"(recp + p0 * kPacket), "r"(bytes)
                : "memory");
          };
          if (lane == 0) {
            if (nsl > 0) fill(0);
            if (nsl > 1) fill(1);
          }
          int32_t unflushed = 0;   // slices since the 32-bit sums went to 64 bits: 1024 * 64 * 8 * 100 < 2^31
#pragma unroll 1
          for (int64_t sl = 0; sl < nsl; ++sl) {
            const uint32_t sb = (uint32_t)(sl & 1);
            mbar_wait_u32(ring_bar + sb * 8, (ring_ph >> sb) & 1u);
            ring_ph ^= 1u << sb;
            const int64_t p0 = w_lo + sl * kRingPackets;
            const int32_t cnt = (int32_t)min((int64_t)kRingPackets, w_hi - p0);
            const uint32_t slot = ring_w + sb * kRingBytes + (uint32_t)lane * 4;
            uint32_t mine_n = lane < 4 ? lds_u32s(slot) : 0u;
#pragma unroll 1
            for (int32_t k = 0; k < cnt; ++k) {
              const uint4 v = warp_packet(mine_n, lane);
              if (k + 1 < cnt && lane < 4) mine_n = lds_u32s(slot + (uint32_t)(k + 1) * 16);
              load_weights<WM>(prm.w, wbase + (p0 + k) * kPacket, wi, wd);
              if (((v.x >> 8) & 0xFFu) == (v.w >> 24))
                eval_packet<CPL, WIDE, WM, true>(v, laneA, laneB, lane_off, s_lut, s_lutd, lut2, wi, wd, 0, kPacket, acc);
              else
                eval_packet<CPL, WIDE, WM, false>(v, laneA, laneB, lane_off, s_lut, s_lutd, lut2, wi, wd, 0, kPacket, acc);
            }
            // the slot's packets are all in registers (REDUX consumed the last one): hand it back
            if (lane == 0 && sl + 2 < nsl) fill(sl + 2);
            acc.flush16();   // kRingPackets * 8 * 100 < 2^16
            if (++unflushed == 1024) { acc.flush(); unflushed = 0; }
          }
          acc.flush();
        }
        cur = seg_end;
      }
    }
    acc.flush();
    // ---- cross-warp reduction in the (now idle) table area, fixed order ----
    __syncthreads();
    acc_t* red = (acc_t*)smem;  // [kWarps][CAND]
#pragma unroll
    for (int j = 0; j < CPL; ++j) {
      acc_t v;
      if (WM == W_F64) v = (acc_t)acc.ad[j]; else v = (acc_t)acc.a64[j];
      red[warp * CAND + lane * CPL + j] = v;
    }
    __syncthreads();
    if (tid < CAND) {
      acc_t sum = 0;
      for (int wq = 0; wq < kWarps; ++wq) sum += red[wq * CAND + tid];
      ((acc_t*)prm.partial)[(size_t)u * CAND + tid] = sum;
    }
  }
}

// ---- k_scaff_head -------------------------------------------------------------------------------
// head[p] = sum_k 100*T[k] * W(g[0], g[k]) with W the summed weight of the records of that pair:
// exactly the evaluations the reference's `i-k > 0` test drops (scaff.go:256,264,...,299).
// One warp per candidate; lane k < 6 bounds the run of pair (g[0], g[step k]) in its sorted tile
// with two binary searches (the six dependent-load chains run side by side).  Unit weights need
// only the run length; otherwise the warp sums each run with coalesced loads (lane-strided, then
// a fixed-order shuffle tree: deterministic).
template <int WM>
__global__ void k_scaff_head(const int32_t* __restrict__ perm, int32_t P, int32_t n, int32_t nblk,
                             const uint16_t* __restrict__ rec, const void* __restrict__ w,
                             const int64_t* __restrict__ tile_start, void* __restrict__ head) {
  typedef typename Acc<WM>::type acc_t;
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= P) return;
  int64_t first = 0, last = 0;
  if (lane < 6 && c_steps[lane] < n) {
    const uint32_t z = (uint32_t)perm[(size_t)p * n];
    const uint32_t y = (uint32_t)perm[(size_t)p * n + c_steps[lane]];
    if (z < (uint32_t)n && y < (uint32_t)n && z != y) {  // invalid rows are flagged by k_scaff_check
      const uint32_t lo = min(z, y), hi = max(z, y);
      const int64_t I = lo >> kBlkShift, J = hi >> kBlkShift;
      const int64_t t = tile_row_start(I, nblk) + (J - I);
      const uint32_t key = ((lo & (kBlk - 1)) << 8) | (hi & (kBlk - 1));
      const int64_t t1 = tile_start[t + 1];
      int64_t a = tile_start[t], b = t1;
      while (a < b) {  // first record >= key
        const int64_t m = (a + b) >> 1;
        if (rec[m] < key) a = m + 1; else b = m;
      }
      first = a;
      b = t1;
      while (a < b) {  // first record > key
        const int64_t m = (a + b) >> 1;
        if (rec[m] <= key) a = m + 1; else b = m;
      }
      last = a;
    }
  }
  acc_t tot = 0;
#pragma unroll 1
  for (int k = 0; k < 6; ++k) {
    const int64_t f = __shfl_sync(0xffffffffu, (long long)first, k);
    const int64_t l = __shfl_sync(0xffffffffu, (long long)last, k);
    acc_t sum = 0;
    if (WM == W_UNIT) {
      sum = (acc_t)(l - f);
    } else {
      for (int64_t i = f + lane; i < l; i += 32) {
        if (WM == W_INT32) sum += (acc_t)((const int32_t*)w)[i];
        else sum += (acc_t)((const double*)w)[i];
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        if (WM == W_F64) sum += (acc_t)__shfl_xor_sync(0xffffffffu, (double)sum, o);
        else sum += (acc_t)__shfl_xor_sync(0xffffffffu, (long long)sum, o);
      }
    }
    tot += sum * (acc_t)c_w100[k];
  }
  if (lane == 0) ((acc_t*)head)[p] = tot;
}

// ---- k_scaff_finish -----------------------------------------------------------------------------
// One warp per candidate: lane-strided sum of the per-unit partials, then a fixed shuffle tree
// (the order does not depend on timing, so fp64 results are reproducible run to run).
template <int WM>
__global__ void k_scaff_finish(int32_t P, int32_t cand, int32_t splits,
                               const void* __restrict__ partial, const void* __restrict__ head,
                               double* __restrict__ out, int64_t* __restrict__ x100) {
  typedef typename Acc<WM>::type acc_t;
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= P) return;
  const int g = p / cand, pc = p - g * cand;
  acc_t S = 0;
  for (int s = lane; s < splits; s += 32) S += ((const acc_t*)partial)[((size_t)g * splits + s) * cand + pc];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    if (WM == W_F64) S += (acc_t)__shfl_xor_sync(0xffffffffu, (double)S, o);
    else S += (acc_t)__shfl_xor_sync(0xffffffffu, (long long)S, o);
  }
  if (lane != 0) return;
  const acc_t tot = S + S - ((const acc_t*)head)[p];  // every pair twice, head-of-ordering pairs once
  if (WM == W_F64) {
    out[p] = -((double)tot / 100.0);
    if (x100) x100[p] = INT64_MIN;
  } else {
    out[p] = -((double)(int64_t)tot / 100.0);
    if (x100) x100[p] = (int64_t)tot;
  }
}

#include "scaff_stream1.cuh"
#include "scaff_stream1p.cuh"

// ---- integer checks -----------------------------------------------------------------------------
__global__ void k_positions(const int32_t* __restrict__ perm, int32_t P, int32_t n,
                            int32_t* __restrict__ pos, int* __restrict__ flag) {
  int64_t total = (int64_t)P * n;
  int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; idx < total; idx += stride) {
    int32_t p = (int32_t)(idx / n), i = (int32_t)(idx - (int64_t)p * n);
    int32_t cid = perm[idx];
    if ((uint32_t)cid >= (uint32_t)n) { atomicOr(flag, 1); continue; }
    pos[(size_t)p * n + cid] = i;
  }
}
__global__ void k_link_dist(int64_t L, const uint32_t* __restrict__ a, const uint32_t* __restrict__ b,
                            const int32_t* __restrict__ pos, int32_t n, int64_t* __restrict__ dist,
                            int* __restrict__ flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < L; i += stride) {
    uint32_t x = a[i], y = b[i];
    if (x >= (uint32_t)n || y >= (uint32_t)n) { atomicOr(flag, 1); continue; }
    int64_t d = (int64_t)pos[x] - (int64_t)pos[y];
    dist[i] = d < 0 ? -d : d;
  }
}

template <int CPL, int WIDE, int WM>
int launch_score(lxy_ctx* c, const ScoreParams& prm, int grid) {
  constexpr int smem = 2 * kBlk * 32 * CPL * (WIDE == 1 ? 4 : 2) + 64 + 8 * kLutLen + 16 + 4 * kLut2Words +
                       kWarps * (2 * kRingBytes + 16);   // tables | decay tables | mbarrier | pair table | record rings + their mbarriers
  static bool configured[64] = {};
  std::lock_guard<std::mutex> cfg_lock(config_mutex());   // process-global flags: not covered by the context mutex
  if (!configured[c->cfg.device & 63]) {
    LXY_CUDA(cudaFuncSetAttribute(k_scaff_score<CPL, WIDE, WM>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    LXY_CUDA(cudaFuncSetAttribute(k_scaff_score<CPL, WIDE, WM>,
                                  cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    configured[c->cfg.device & 63] = true;
  }
  k_scaff_score<CPL, WIDE, WM><<<grid, kThreads, smem, c->stream>>>(prm);
  LXY_CUDA(cudaGetLastError());
  return LXY_OK;
}

template <int WIDE, int WM>
int launch_score_cpl(lxy_ctx* c, int cpl, const ScoreParams& prm, int grid) {
  switch (cpl) {
    case 1: return launch_score<1, WIDE, WM>(c, prm, grid);
    case 2: return launch_score<2, WIDE, WM>(c, prm, grid);
    default:
      if constexpr (WIDE == 1) return launch_score<2, WIDE, WM>(c, prm, grid);
      else return launch_score<4, WIDE, WM>(c, prm, grid);
  }
}

}  // namespace

// The streaming kernels run with most of the SM's L1 carved out as shared memory; the small
// kernels around them (layout, check, head, finish) use none.  Left to the default, the driver
// switches the carve-out between consecutive launches and serialises them (measured: +20 us per
// call around k_scaff_stream1).  Asking for the same carve-out everywhere removes the switch.
static int prefer_max_carveout(lxy_ctx* c) {
  static bool done[64] = {};
  std::lock_guard<std::mutex> cfg_lock(config_mutex());
  if (done[c->cfg.device & 63]) return LXY_OK;
  const void* fns[] = {
      (const void*)k_scaff_layout<0>, (const void*)k_scaff_layout<1>,
      (const void*)k_scaff_check<0>,  (const void*)k_scaff_check<1>,
      (const void*)k_scaff_head<W_UNIT>,  (const void*)k_scaff_head<W_INT32>, (const void*)k_scaff_head<W_F64>,
      (const void*)k_scaff_finish<W_UNIT>, (const void*)k_scaff_finish<W_INT32>, (const void*)k_scaff_finish<W_F64>};
  for (const void* f : fns)
    LXY_CUDA(cudaFuncSetAttribute(f, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  done[c->cfg.device & 63] = true;
  return LXY_OK;
}

// Streaming regime (scaff_stream1.cuh): head look-ups on the side stream, one pass over the records
// with lane = record, the same fixed-order finish as the batch path.
// `first` / `last`: position of this pass in a multi-pass call (scaff_score_enqueue): the error flags
// are cleared once, the timing events bracket all passes.
static int scaff_score_stream1(lxy_ctx* c, const int32_t* perm_dev, int32_t P, int nc, size_t smem,
                               double* out_dev, int64_t* x100_dev, bool first, bool last) {
  ScaffLinks& S = c->scaff;
  cudaStream_t st = c->stream;
  const int n = S.n;
  // two orderings per table word where that is exact (scaff_stream1p.cuh)
  static const bool no_pack = getenv("LXY_S1_NOPACK") != nullptr;  // experiments
  const bool pack = !no_pack && stream1p_applies(S.wmode, nc, n) && stream1p_smem(nc, n) <= (size_t)c->max_smem_optin;
  // The ring: two 2 KB stages per warp while two CTAs per SM fit (64 bulk copies in flight per SM; with
  // three stages = 96 in flight single SMs ran 1.7 x slower than the rest in some launches -
  // tools/s1_trace.py, 250 M and 300 M records - and nothing was gained elsewhere), else up to four
  int stages = 2;
  {
    const size_t cap = (size_t)c->max_smem_optin;
    auto need = [&](int sg) { return pack ? stream1p_smem(nc, n, sg) : stream1_smem(nc, n, sg); };
    if (2 * (need(2) + 1024) > cap)
      while (stages < 4 && need(stages + 1) <= cap) ++stages;
    static const char* force_stages = getenv("LXY_S1_STAGES");   // experiments
    if (force_stages) stages = atoi(force_stages);
    smem = need(stages);
  }
  LXY_TRY(c->d_flag.reserve(64));
  LXY_TRY(c->d_corr.reserve(8 * (size_t)P));
  LXY_TRY(prefer_max_carveout(c));
  if (first) {
    LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, st));
    LXY_CUDA(record_timing_event(c, 0));
  }
  {
    cudaStream_t s2 = c->stream2;
    LXY_CUDA(cudaEventRecord(c->ev_fork, st));
    LXY_CUDA(cudaStreamWaitEvent(s2, c->ev_fork, 0));
    const int hb = (P * 32 + 255) / 256;
    const uint16_t* rec = S.rec.as<uint16_t>();
    const int64_t* ts = S.tile_start.as<int64_t>();
    if (S.wmode == W_UNIT) k_scaff_head<W_UNIT><<<hb, 256, 0, s2>>>(perm_dev, P, n, S.nblk, rec, S.w.p, ts, c->d_corr.p);
    else if (S.wmode == W_INT32) k_scaff_head<W_INT32><<<hb, 256, 0, s2>>>(perm_dev, P, n, S.nblk, rec, S.w.p, ts, c->d_corr.p);
    else k_scaff_head<W_F64><<<hb, 256, 0, s2>>>(perm_dev, P, n, S.nblk, rec, S.w.p, ts, c->d_corr.p);
    LXY_CUDA(cudaGetLastError());
    LXY_CUDA(cudaEventRecord(c->ev_join, s2));
  }
  if (first) LXY_CUDA(record_timing_event(c, 1));
  // persistent CTAs: as many as fit per SM (<= 2 x 512 threads), each an equal share of the packets
  const int per_sm = (kS1MinBlocks == 2 && 2 * (smem + 1024) <= (size_t)c->max_smem_optin) ? 2 : 1;
  const int64_t npk = (S.L + kPacket - 1) / kPacket;
  int64_t grid = (int64_t)c->sm_count * per_sm;
  if (grid * 64 > npk) grid = npk / 64 > 0 ? npk / 64 : 1;  // at least 512 records per CTA
  LXY_TRY(c->d_partial.reserve(8 * (size_t)grid * nc));
  Stream1Params prm;
  prm.rec = S.rec.as<uint16_t>();
  prm.w = S.w.p;
  prm.tile_start = S.tile_start.as<int64_t>();
  prm.perm = perm_dev;
  prm.partial = c->d_partial.p;
  prm.flag = c->d_flag.as<int>();
  prm.L = S.L;
  prm.n = n;
  prm.nblk = S.nblk;
  prm.P = P;
  prm.validate = (c->cfg.flags & LXY_FLAG_NO_VALIDATE) ? 0 : 1;
  prm.stages = stages;
  LXY_TRY(pack ? launch_stream1p_nc(c, nc, prm, (int)grid, smem)
        : S.wmode == W_UNIT ? launch_stream1_nc<W_UNIT>(c, nc, prm, (int)grid, smem)
        : S.wmode == W_INT32 ? launch_stream1_nc<W_INT32>(c, nc, prm, (int)grid, smem)
                             : launch_stream1_nc<W_F64>(c, nc, prm, (int)grid, smem));
  if (last) LXY_CUDA(record_timing_event(c, 2));
  LXY_CUDA(cudaStreamWaitEvent(st, c->ev_join, 0));
  const int splits = (int)grid;
  if (S.wmode == W_UNIT) k_scaff_finish<W_UNIT><<<(P + 7) / 8, 256, 0, st>>>(P, nc, splits, c->d_partial.p, c->d_corr.p, out_dev, x100_dev);
  else if (S.wmode == W_INT32) k_scaff_finish<W_INT32><<<(P + 7) / 8, 256, 0, st>>>(P, nc, splits, c->d_partial.p, c->d_corr.p, out_dev, x100_dev);
  else k_scaff_finish<W_F64><<<(P + 7) / 8, 256, 0, st>>>(P, nc, splits, c->d_partial.p, c->d_corr.p, out_dev, x100_dev);
  LXY_CUDA(cudaGetLastError());
  if (last) LXY_CUDA(record_timing_event(c, 3));
  const int64_t link_bytes = S.L * (2 + (S.wmode == W_INT32 ? 4 : S.wmode == W_F64 ? 8 : 0));
  c->timing.launches = (first ? 0 : c->timing.launches) + 3;
  c->timing.grid = (int)grid;
  c->timing.link_bytes = (first ? 0 : c->timing.link_bytes) + link_bytes;
  c->timing.table_bytes = (first ? 0 : c->timing.table_bytes) + (int64_t)grid * nc * n * 4;
  return LXY_OK;
}

static int scaff_score_enqueue(lxy_ctx* c, const int32_t* perm_dev, int32_t P, double* out_dev,
                               int64_t* x100_dev);

// The scoring pass, replayed from a CUDA graph when the same call (same buffers, same population
// size, same link store) comes again: captured on the second identical call, when every buffer is
// sized and every kernel attribute set, so that nothing in the pass needs an API call that a capture
// forbids.  One graph launch replaces ~15 runtime calls; on the small configurations the host was the
// slower side of the pipeline.
int scaff_score(lxy_ctx* c, const int32_t* perm_dev, const uint8_t* ori_dev, int32_t P,
                double* out_dev, int64_t* x100_dev) {
  ScaffLinks& S = c->scaff;
  if (!S.ready) { set_error("scaff score: links not set"); return LXY_ERR_STATE; }
  if (S.gprofile) { set_error("scaff score: G-profile context"); return LXY_ERR_STATE; }
  (void)ori_dev;
  if (P <= 0) return LXY_OK;
  if (c->cfg.flags & LXY_FLAG_NO_GRAPH) return scaff_score_enqueue(c, perm_dev, P, out_dev, x100_dev);
  ScoreGraph* g = nullptr;
  ScoreGraph* oldest = &c->graphs[0];
  for (auto& e : c->graphs) {
    if (e.perm == perm_dev && e.out == out_dev && e.x100 == x100_dev && e.P == P && e.version == S.version &&
        e.seen > 0) g = &e;
    if (e.stamp < oldest->stamp) oldest = &e;
  }
  if (!g) {  // first call with these arguments: direct path (it sizes the buffers)
    if (oldest->exec) cudaGraphExecDestroy(oldest->exec);
    *oldest = ScoreGraph();
    oldest->perm = perm_dev; oldest->out = out_dev; oldest->x100 = x100_dev; oldest->P = P;
    oldest->version = S.version; oldest->seen = 1; oldest->stamp = ++c->graph_clock;
    return scaff_score_enqueue(c, perm_dev, P, out_dev, x100_dev);
  }
  g->stamp = ++c->graph_clock;
  ++g->seen;
  if (g->failed) return scaff_score_enqueue(c, perm_dev, P, out_dev, x100_dev);
  const void* scratch[6] = {c->d_posA.p, c->d_posB.p, c->d_partial.p, c->d_corr.p, c->d_flag.p, c->d_start.p};
  if (g->exec && std::memcmp(scratch, g->scratch, sizeof(scratch)) != 0) {  // a scratch buffer moved
    cudaGraphExecDestroy(g->exec);
    g->exec = nullptr;
  }
  if (!g->exec) {
    if (cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
      (void)cudaGetLastError();
      g->failed = true;
      return scaff_score_enqueue(c, perm_dev, P, out_dev, x100_dev);
    }
    c->capturing = true;
    const int rc = scaff_score_enqueue(c, perm_dev, P, out_dev, x100_dev);
    c->capturing = false;
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(c->stream, &graph);
    if (rc != LXY_OK || ce != cudaSuccess || !graph ||
        cudaGraphInstantiate(&g->exec, graph, 0) != cudaSuccess) {
      (void)cudaGetLastError();
      if (graph) cudaGraphDestroy(graph);
      g->exec = nullptr;
      g->failed = true;
      return scaff_score_enqueue(c, perm_dev, P, out_dev, x100_dev);
    }
    cudaGraphDestroy(graph);
    g->timing = c->timing;
    const void* now[6] = {c->d_posA.p, c->d_posB.p, c->d_partial.p, c->d_corr.p, c->d_flag.p, c->d_start.p};
    std::memcpy(g->scratch, now, sizeof(now));
  }
  LXY_CUDA(cudaGraphLaunch(g->exec, c->stream));
  c->timing.launches = g->timing.launches;
  c->timing.grid = g->timing.grid;
  c->timing.link_bytes = g->timing.link_bytes;
  c->timing.table_bytes = g->timing.table_bytes;
  return LXY_OK;
}

// Populations of up to 128 orderings as passes of <= 8 through the streaming-regime kernels.
static int scaff_enqueue_passes(lxy_ctx* c, const int32_t* perm_dev, int32_t P, double* out_dev,
                                int64_t* x100_dev, bool first, bool last) {
  const int n = c->scaff.n;
  constexpr int kS1Pass = 8;
  for (int32_t p0 = 0; p0 < P; p0 += kS1Pass) {
    const int32_t pp = P - p0 < kS1Pass ? P - p0 : kS1Pass;
    int ncp = 1;
    while (ncp < pp) ncp *= 2;
    LXY_TRY(scaff_score_stream1(c, perm_dev + (size_t)p0 * n, pp, ncp, stream1_smem(ncp, n, 2), out_dev + p0,
                                x100_dev ? x100_dev + p0 : nullptr, first && p0 == 0, last && p0 + kS1Pass >= P));
  }
  return LXY_OK;
}

static int scaff_enqueue_batch(lxy_ctx* c, const int32_t* perm_dev, int32_t P, double* out_dev,
                               int64_t* x100_dev, bool first, bool last);

// Measured on B200 (tools/p_sweep.py, n = 5,000, unit weights; 5 M, 50 M and 200 M records; L in
// millions of records), after both kernels moved to the bulk-copy rings: a pass of <= 8 orderings on the
// packed kernel (scaff_stream1p.cuh) costs 40 / 108 / 259 us = 33 + 1.5 L up to 50 M records and
// 108 + 1.0 (L - 50) beyond; one group of the batch kernel 55 + 11.1 L us up to 32 orderings, 58 + 10.5 L up
// to 64, 58 + 15.4 L up to 128; a further group of a launch that already has full groups 15.4 L.
// The model is only consulted inside that regime (unit weights, packed kernel applicable); weighted
// stores keep the single-pass rule (<= 8 orderings stream, everything else in groups).
static double pass_cost_us(int32_t P, double l_m) {
  return (double)((P + 7) / 8) * (l_m <= 50.0 ? 33.0 + 1.5 * l_m : 108.0 + 1.0 * (l_m - 50.0));
}
static double group_cost_us(int32_t P, double l_m) {   // one group of up to 128 orderings, its own launch
  return P <= 32 ? 55.0 + 11.1 * l_m : P <= 64 ? 58.0 + 10.5 * l_m : 58.0 + 15.4 * l_m;
}
constexpr double kNextGroupUsPerM = 15.4;             // one more group in a launch that has full groups

static int scaff_score_enqueue(lxy_ctx* c, const int32_t* perm_dev, int32_t P, double* out_dev,
                               int64_t* x100_dev) {
  ScaffLinks& S = c->scaff;
  const int n = S.n;
  constexpr int kS1Pass = 8;
  const double l_m = (double)S.L * 1e-6;
  const bool s1_ok = n <= 65536 && !(c->cfg.flags & LXY_FLAG_NO_STREAM1) &&
                     stream1_smem(P < kS1Pass ? (P <= 1 ? 1 : P <= 2 ? 2 : P <= 4 ? 4 : 8) : kS1Pass, n, 2) <= (size_t)c->max_smem_optin;
  const bool model_ok = s1_ok && S.wmode == W_UNIT && stream1p_applies(S.wmode, kS1Pass, n) &&
                        stream1p_smem(kS1Pass, n) <= (size_t)c->max_smem_optin;
  // ---- streaming regime: up to 8 candidates per pass (a lone Score() call is a batch of one) ----
  // Small populations - the reference's own GA runs 40 orderings (scaff.go:94) - take several passes
  // while the model says that wins: against 200 M records 16 orderings 2.27 -> 0.52 ms, 40 orderings
  // 2.15 -> 1.27 ms, 96 orderings 3.13 -> 3.03 ms; at 5 M records passes win up to 16 orderings.
  if (s1_ok && P <= kS1Pass) return scaff_enqueue_passes(c, perm_dev, P, out_dev, x100_dev, true, true);
  if (model_ok && P <= 128) {
    if (pass_cost_us(P, l_m) < group_cost_us(P, l_m)) return scaff_enqueue_passes(c, perm_dev, P, out_dev, x100_dev, true, true);
  }
  // ---- whole groups of 128 through the batch kernel; what is left over (P mod 128: a GA generation
  // makes ~P +- 30 children) goes through passes when those cost less than one more group, which would
  // score 128 - rem identity paddings at full price (cfg3, 1,028 children: 9 groups -> 8 groups + one
  // pass of 4) ----
  const int32_t rem = P % 128;
  if (model_ok && P > 128 && rem != 0 && pass_cost_us(rem, l_m) < kNextGroupUsPerM * l_m) {
    const int32_t full = P - rem;
    LXY_TRY(scaff_enqueue_batch(c, perm_dev, full, out_dev, x100_dev, true, false));
    return scaff_enqueue_passes(c, perm_dev + (size_t)full * n, rem, out_dev + full,
                                x100_dev ? x100_dev + full : nullptr, false, true);
  }
  return scaff_enqueue_batch(c, perm_dev, P, out_dev, x100_dev, true, true);
}

static int scaff_enqueue_batch(lxy_ctx* c, const int32_t* perm_dev, int32_t P, double* out_dev,
                               int64_t* x100_dev, bool first, bool last) {
  ScaffLinks& S = c->scaff;
  cudaStream_t st = c->stream;
  const int n = S.n;
  // table mode: 0 = 16-bit entries (pos + 21 | -pos) summed modulo 2^16 (n <= 65515, see eval_record),
  // 128 candidates in the 128 KB of tables; 1 = 32-bit entries, 64 candidates.
  static const int force_tm = getenv("LXY_TABLE_MODE") ? atoi(getenv("LXY_TABLE_MODE")) : -1;  // experiments
  const int tm = (n <= 65536 - (kBias + 1) && force_tm != 1) ? 0 : 1;
  const bool wide = (tm == 1);
  // candidates per lane: 4 (16-bit tables) / 2 (32-bit) fill 128 KB of tables; small populations use less
  int cpl = wide ? 2 : 4;
  while (cpl > 1 && P <= 32 * (cpl / 2)) cpl /= 2;
  const int cand = 32 * cpl;
  const int groups = (P + cand - 1) / cand;
  const int Ppad = groups * cand;
  const size_t es = wide ? 4 : 2;
  LXY_TRY(c->d_posA.reserve((size_t)Ppad * n * es));
  LXY_TRY(c->d_posB.reserve((size_t)Ppad * n * es));
  LXY_TRY(c->d_flag.reserve(64));
  LXY_TRY(c->d_corr.reserve(8 * (size_t)P));
  LXY_TRY(prefer_max_carveout(c));
  if (first) LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, st));
  int launches = 0;

  if (first) LXY_CUDA(record_timing_event(c, 0));
  // The head-of-ordering look-ups only need the population and the link store: they run on a side
  // stream next to the layout and streaming kernels and are joined before k_scaff_finish.
  {
    cudaStream_t s2 = c->stream2;
    LXY_CUDA(cudaEventRecord(c->ev_fork, st));
    LXY_CUDA(cudaStreamWaitEvent(s2, c->ev_fork, 0));
    const int hb = (P * 32 + 255) / 256;
    const uint16_t* rec = S.rec.as<uint16_t>();
    const int64_t* ts = S.tile_start.as<int64_t>();
    if (S.wmode == W_UNIT) k_scaff_head<W_UNIT><<<hb, 256, 0, s2>>>(perm_dev, P, n, S.nblk, rec, S.w.p, ts, c->d_corr.p);
    else if (S.wmode == W_INT32) k_scaff_head<W_INT32><<<hb, 256, 0, s2>>>(perm_dev, P, n, S.nblk, rec, S.w.p, ts, c->d_corr.p);
    else k_scaff_head<W_F64><<<hb, 256, 0, s2>>>(perm_dev, P, n, S.nblk, rec, S.w.p, ts, c->d_corr.p);
    LXY_CUDA(cudaGetLastError());
    LXY_CUDA(cudaEventRecord(c->ev_join, s2));
  }
  const int lg = grid_for((int64_t)Ppad * n, 256, c->sm_count);
  static const bool old_layout = getenv("LXY_OLD_LAYOUT") != nullptr;   // experiments
  if (tm == 0 && !old_layout && 2 * (size_t)n + 64 <= (size_t)c->max_smem_optin) {
    // two coalesced passes: inverse permutation per candidate in shared memory (+ validation), transpose
    LXY_TRY(c->d_start.reserve(2 * (size_t)P * n + 64));
    const size_t inv_smem = 2 * (size_t)n;
    if (inv_smem > 48 * 1024) {
      static size_t configured[64] = {};
      std::lock_guard<std::mutex> cfg_lock(config_mutex());
      if (configured[c->cfg.device & 63] < inv_smem) {
        LXY_CUDA(cudaFuncSetAttribute(k_scaff_inv16, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)inv_smem));
        configured[c->cfg.device & 63] = inv_smem;
      }
    }
    k_scaff_inv16<<<P, 512, inv_smem, st>>>(perm_dev, n, c->d_start.as<uint16_t>(),
                                           (c->cfg.flags & LXY_FLAG_NO_VALIDATE) ? 0 : 1, c->d_flag.as<int>());
    const dim3 tg((unsigned)((n + 63) / 64), (unsigned)groups);
    if (cand == 128) k_scaff_tables16<128><<<tg, 256, 0, st>>>(c->d_start.as<uint16_t>(), P, n, c->d_posA.as<uint16_t>(), c->d_posB.as<uint16_t>());
    else if (cand == 64) k_scaff_tables16<64><<<tg, 256, 0, st>>>(c->d_start.as<uint16_t>(), P, n, c->d_posA.as<uint16_t>(), c->d_posB.as<uint16_t>());
    else k_scaff_tables16<32><<<tg, 256, 0, st>>>(c->d_start.as<uint16_t>(), P, n, c->d_posA.as<uint16_t>(), c->d_posB.as<uint16_t>());
    launches += 2;
  } else {
    if (tm == 1) k_scaff_layout<1><<<lg, 256, 0, st>>>(perm_dev, P, Ppad, n, cand, c->d_posA.p, c->d_posB.p, c->d_flag.as<int>());
    else k_scaff_layout<0><<<lg, 256, 0, st>>>(perm_dev, P, Ppad, n, cand, c->d_posA.p, c->d_posB.p, c->d_flag.as<int>());
    ++launches;
    if (!(c->cfg.flags & LXY_FLAG_NO_VALIDATE)) {
      if (tm == 1) k_scaff_check<1><<<lg, 256, 0, st>>>(perm_dev, P, n, cand, c->d_posB.p, c->d_flag.as<int>());
      else k_scaff_check<0><<<lg, 256, 0, st>>>(perm_dev, P, n, cand, c->d_posB.p, c->d_flag.as<int>());
      ++launches;
    }
  }
  LXY_CUDA(cudaGetLastError());
  if (first) LXY_CUDA(record_timing_event(c, 1));

  // work units: groups x splits, about two waves of one CTA per SM
  const int64_t npk = (S.L + kPacket - 1) / kPacket;
  static const int waves = getenv("LXY_WAVES") ? atoi(getenv("LXY_WAVES")) : 2;   // experiments
  int splits = (waves * c->sm_count) / groups;  // two full waves of one CTA per SM, never a third
  if (splits < 1) splits = 1;
  const int64_t min_pk = 64;  // do not split below 512 records per unit
  if ((int64_t)splits * min_pk > npk) splits = (int)(npk / min_pk > 0 ? npk / min_pk : 1);
  const int units = groups * splits;
  const int grid = units < c->sm_count ? units : c->sm_count;
  LXY_TRY(c->d_partial.reserve(8 * (size_t)units * cand));
  ScoreParams prm;
  prm.rec = S.rec.as<uint16_t>();
  prm.w = S.w.p;
  prm.tile_start = S.tile_start.as<int64_t>();
  prm.tabA = c->d_posA.p;
  prm.tabB = c->d_posB.p;
  prm.partial = c->d_partial.p;
  prm.L = S.L;
  prm.n = n;
  prm.nblk = S.nblk;
  prm.groups = groups;
  prm.splits = splits;
  int rc;
  if (tm == 1) {
    rc = S.wmode == W_UNIT ? launch_score_cpl<1, W_UNIT>(c, cpl, prm, grid)
       : S.wmode == W_INT32 ? launch_score_cpl<1, W_INT32>(c, cpl, prm, grid)
                            : launch_score_cpl<1, W_F64>(c, cpl, prm, grid);
  } else {
    rc = S.wmode == W_UNIT ? launch_score_cpl<0, W_UNIT>(c, cpl, prm, grid)
       : S.wmode == W_INT32 ? launch_score_cpl<0, W_INT32>(c, cpl, prm, grid)
                            : launch_score_cpl<0, W_F64>(c, cpl, prm, grid);
  }
  LXY_TRY(rc);
  ++launches;
  if (last) LXY_CUDA(record_timing_event(c, 2));

  LXY_CUDA(cudaStreamWaitEvent(st, c->ev_join, 0));
  if (S.wmode == W_UNIT) k_scaff_finish<W_UNIT><<<(P + 7) / 8, 256, 0, st>>>(P, cand, splits, c->d_partial.p, c->d_corr.p, out_dev, x100_dev);
  else if (S.wmode == W_INT32) k_scaff_finish<W_INT32><<<(P + 7) / 8, 256, 0, st>>>(P, cand, splits, c->d_partial.p, c->d_corr.p, out_dev, x100_dev);
  else k_scaff_finish<W_F64><<<(P + 7) / 8, 256, 0, st>>>(P, cand, splits, c->d_partial.p, c->d_corr.p, out_dev, x100_dev);
  launches += 2;
  LXY_CUDA(cudaGetLastError());
  if (last) LXY_CUDA(record_timing_event(c, 3));

  c->timing.launches = launches;
  c->timing.grid = grid;
  c->timing.link_bytes = (int64_t)groups * S.L * (2 + (S.wmode == W_INT32 ? 4 : S.wmode == W_F64 ? 8 : 0));
  c->timing.table_bytes = 0;
  return LXY_OK;
}

int scaff_positions(lxy_ctx* c, const int32_t* perm_dev, int32_t P, int32_t n, int32_t* pos_dev) {
  LXY_TRY(c->d_flag.reserve(64));
  LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, c->stream));
  k_positions<<<grid_for((int64_t)P * n, 256, c->sm_count), 256, 0, c->stream>>>(perm_dev, P, n, pos_dev, c->d_flag.as<int>());
  LXY_CUDA(cudaGetLastError());
  return LXY_OK;
}

int scaff_link_distances(lxy_ctx* c, int32_t n, int64_t L, const uint32_t* a_dev,
                         const uint32_t* b_dev, const int32_t* perm_dev, int64_t* dist_dev) {
  LXY_TRY(c->d_start.reserve(sizeof(int32_t) * (size_t)n));
  LXY_TRY(scaff_positions(c, perm_dev, 1, n, c->d_start.as<int32_t>()));
  if (L > 0) {
    k_link_dist<<<grid_for(L, 256, c->sm_count), 256, 0, c->stream>>>(L, a_dev, b_dev, c->d_start.as<int32_t>(), n, dist_dev, c->d_flag.as<int>());
    LXY_CUDA(cudaGetLastError());
  }
  return LXY_OK;
}

}  // namespace lxy

#ifdef LXY_S1_TRACE
extern "C" int lxy_debug_s1_trace(unsigned long long* out, int n) {
  return (int)cudaMemcpyFromSymbol(out, lxy::g_s1_trace, sizeof(unsigned long long) * (size_t)n);
}
#endif
