// scaff_stream1.cuh — the streaming regime of the scaffolding scorer: 1..8 candidates per call.
// (included by scaff_kernels.cu inside its anonymous namespace)
//
// This is the reference's literal boundary: Genome.Score() hands ONE ordering to score()
// (/root/reference/src/lxy/scaff/genome.go:153-159 -> scaff.go:246-310).  With a single candidate
// there is no cross-candidate reuse of a record, so the pass is bound by HBM: every 2-byte record
// is read once and scored once.  The mapping is therefore turned around with respect to
// k_scaff_score: lane = record (8 consecutive records per lane per 128-bit shared-memory load; the
// records arrive through a per-warp ring of 2 KB bulk copies), and the whole placement table pos[] of each candidate (u16, n <= 65536) sits in
// shared memory for the lifetime of the CTA, built in the kernel from the ordering itself
// (pos[perm[i]] = i: the exclusive prefix scan of unit lengths is the ordering index).
//   per record:  d = |pos[a] - pos[b]|,  acc += 100*T[min(d, 21)]      (T of scaff.go:252-302)
// The `i-k > 0` quirk (scaff.go:256) is applied by k_scaff_head / k_scaff_finish as in the batch
// path (fitness = -(2S - H)/100), so both paths return identical integers.
#pragma once

#ifndef LXY_S1_THREADS
#define LXY_S1_THREADS 512
#endif
constexpr int kS1Threads = LXY_S1_THREADS;
constexpr int kS1MinBlocks = kS1Threads > 512 ? 1 : 2;
constexpr int kS1Warps = kS1Threads / 32;
// Ring geometry of the streaming kernels: every warp owns kS1Sub packets per lane (2 KB) of each chunk and
// keeps `stages` such slices in flight, each filled by ONE bulk copy (UBLKCP) counted on the warp's own
// mbarrier of that stage: warps stay independent of each other (a ring stage shared by the CTA couples
// them to the slowest, measured 100 us against 82), and the copy engine sees 2 KB requests (512-byte
// ones cost 158 us: its cost is per request).
constexpr int kS1Sub = 4;
constexpr int kS1WarpStage = kS1Sub * 512;             // bytes per warp and ring stage
// Unit weights: the two records of a 32-bit word share ONE decay look-up, lut2[v0 * 43 + v1] =
// 100*T[v0] + 100*T[v1] (<= 200: a byte).  Saves one of the two look-ups, one add and moves the index
// arithmetic to an IMAD (the FMA pipe idles in this kernel): 25 -> 21 shared-pipe wavefronts per packet.
constexpr int kS1LutSide = kMiss + 2;                  // 43 values of v
constexpr int kS1Lut2Len = kS1LutSide * kS1LutSide;    // 1849
constexpr int kS1Lut2Bytes = 1856;


// The 43-entry decay table (padded to 64 bytes) followed by the pair sums lut2, built at compile time and
// copied into shared memory as 120 16-byte pieces (one load per thread, requested with the other prologue
// loads): no constant-bank reads with divergent indices, no barrier between "table in shared memory" and
// "pair sums built from it".
struct S1Luts { uint8_t v[64 + kS1Lut2Bytes]; };
constexpr uint8_t s1_lut100_at(int v) {   // c_lut100[v], v = d + kBias + 1
  const int d = v - (kBias + 1), ad = d < 0 ? -d : d;
  return ad == 1 ? 100 : ad == 2 ? 50 : ad == 3 ? 33 : ad == 5 ? 20 : ad == 11 ? 10 : ad == 20 ? 5 : 0;
}
constexpr S1Luts s1_make_luts() {
  S1Luts t{};
  for (int i = 0; i < kS1LutSide; ++i) t.v[i] = s1_lut100_at(i);
  for (int i = 0; i < kS1Lut2Len; ++i) t.v[64 + i] = (uint8_t)(s1_lut100_at(i / kS1LutSide) + s1_lut100_at(i % kS1LutSide));
  return t;
}
__device__ const S1Luts g_s1_luts = s1_make_luts();
constexpr int kS1LutPieces = (64 + kS1Lut2Bytes) / 16;   // 120
static_assert((64 + kS1Lut2Bytes) % 16 == 0, "lut pieces");
constexpr int kS1Pre = 8192 / kS1Threads;                // genes per thread requested ahead of the record ring

// Ordered loads: volatile asm statements keep their order among themselves, so these are issued before
// the cp.async copies that follow them in the source (a plain __ldg may be scheduled below them).
__device__ __forceinline__ int32_t ldg_s32_ordered(const int32_t* p) {
  int32_t v;
  asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ int64_t ldg_s64_ordered(const int64_t* p) {
  int64_t v;
  asm volatile("ld.global.nc.s64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint4 ldg_v4_ordered(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

struct Stream1Params {
  const uint16_t* rec;
  const void* w;
  const int64_t* tile_start;
  const int32_t* perm;  // [P][n]
  void* partial;        // [grid][NC] int64 / double
  int* flag;
  int64_t L;
  int32_t n, nblk, P;
  int32_t validate;
  int32_t stages;       // ring stages per warp (k_scaff_stream1: 2 or 3)
};

// Table entries are tab[x] = 32767 - pos[x] as int16 (pos in [0, 65535]), so that
//   pos a - pos b = tab[b] - tab[a]   and   v = clamp(tab[b] + (21 - tab[a]), 0, 42)
// is one VIADDMNMX.RELU per record; v indexes the 43-entry decay table c_lut100 (v = d + 21; the two
// far entries 0 and 42 hold 0).  Every contig block of the table starts on a 512-byte boundary of
// the shared window, so a record's row address is one LOP3: (2 * local id) | block base.
__device__ __forceinline__ int32_t lds_s16(uint32_t addr) {
  int32_t v;
  asm("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(addr)
               : "memory");
  return r;
}

// The stream phase was bound by the ALU pipe (80 % busy: LOP3, SHF, VIADDMNMX, IADD3) while the FMA pipe idled
// at 14 %.  Two changes move work across: the pair-table index is one multiply-add in front of an
// address-only LDS (was IMAD + a three-input IADD3), and the row address of a word's second record is
// PRMT + IMAD (byte * entry size + base, the entry size read from the constant bank so that it stays a
// multiply) instead of SHF + LOP3: 77.8 -> 73.7 us for one ordering against 200 M records.  Right shifts as
// IMAD.HI by 2^(32 - s) were measured too (half-rate on that pipe): 79.9 us, not kept.
__constant__ uint32_t c_s1_mul[6] = {0u, 0u, 0u, 0u, 2u, 4u};   // [4], [5]: table entry sizes (k_scaff_stream1, k_scaff_stream1p)
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

template <int WM>
__device__ __forceinline__ void s1_acc(int32_t t, int32_t wi, double wd, int32_t& a32, int64_t& a64,
                                       double& ad) {
  if (WM == W_UNIT) a32 += t;
  else if (WM == W_INT32) a64 += (int64_t)t * (int64_t)wi;
  else ad += (double)t * wd;
}

// One record (general path: boundary packets, packets that span two a's).
template <int NC, int WM>
__device__ __forceinline__ void s1_eval(uint32_t baseA, uint32_t baseB, uint32_t cstride,
                                        const uint8_t* lut, uint32_t r, int32_t wi, double wd,
                                        int32_t (&a32)[NC], int64_t (&a64)[NC], double (&ad)[NC]) {
  const uint32_t oa = (r >> 7) & 0x1FEu, ob = (r << 1) & 0x1FEu;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int32_t ta = lds_s16((baseA + c * cstride) | oa), tb = lds_s16((baseB + c * cstride) | ob);
    const int32_t v = __viaddmin_s32_relu(tb, (kBias + 1) - ta, kMiss + 1);
    s1_acc<WM>(lut[v], wi, wd, a32[c], a64[c], ad[c]);
  }
}

// Eight records of one packet, all inside the segment and all with the same a (the common case:
// records are sorted by (a, b) inside a tile): tab[a] is fetched once.
template <int NC, int WM>
__device__ __forceinline__ void s1_packet_same_a(const uint4 v, uint32_t baseA, uint32_t baseB,
                                                 uint32_t cstride, const uint8_t* lut, const uint8_t* lut2,
                                                 const int32_t (&wi)[kPacket], const double (&wd)[kPacket],
                                                 int32_t (&a32)[NC], int64_t (&a64)[NC], double (&ad)[NC]) {
  const uint32_t words[4] = {v.x, v.y, v.z, v.w};
  const uint32_t oa = (v.x >> 7) & 0x1FEu;
  const uint32_t lut2_s = smem_u32(lut2);
  int32_t pterm[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) pterm[c] = (kBias + 1) - lds_s16((baseA + c * cstride) | oa);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t o0 = (words[q] << 1) & 0x1FEu;
    const uint32_t b1 = __byte_perm(words[q], 0u, 0x4442u);   // b of the second record, zero-extended (PRMT)
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const uint32_t bb = baseB + c * cstride;
      const int32_t v0 = __viaddmin_s32_relu(lds_s16(bb | o0), pterm[c], kMiss + 1);
      const int32_t v1 = __viaddmin_s32_relu(lds_s16(b1 * c_s1_mul[4] + bb), pterm[c], kMiss + 1);
      if (WM == W_UNIT) {
        a32[c] += (int32_t)lds_u8(lut2_s + ((uint32_t)v0 * kS1LutSide + (uint32_t)v1));
      } else {
        s1_acc<WM>(lut[v0], wi[2 * q], wd[2 * q], a32[c], a64[c], ad[c]);
        s1_acc<WM>(lut[v1], wi[2 * q + 1], wd[2 * q + 1], a32[c], a64[c], ad[c]);
      }
    }
  }
}

// Eight records inside the segment whose a's are two NEIGHBOURING ids (a packet that straddles the
// change from a to a + 1; with ~250 records per a, as at n = 20,000 / 200 M records, two thirds of
// the warps hold such a lane): both rows of table A are fetched, every record selects by comparing
// its a with the first one.  Sorted records: a_last - a_first <= 1 means no third value.
template <int NC, int WM>
__device__ __forceinline__ void s1_packet_two_a(const uint4 v, uint32_t baseA, uint32_t baseB,
                                                uint32_t cstride, const uint8_t* lut, const uint8_t* lut2,
                                                const int32_t (&wi)[kPacket], const double (&wd)[kPacket],
                                                int32_t (&a32)[NC], int64_t (&a64)[NC], double (&ad)[NC]) {
  const uint32_t words[4] = {v.x, v.y, v.z, v.w};
  const uint32_t oa0 = (v.x >> 7) & 0x1FEu, oa1 = (v.w >> 23) & 0x1FEu;
  const uint32_t rep = (v.x & 0xFF00u) * 0x00010001u;   // the first a at bits 8..15 and 24..31
  const uint32_t lut2_s = smem_u32(lut2);
  int32_t p0[NC], p1[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    p0[c] = (kBias + 1) - lds_s16((baseA + c * cstride) | oa0);
    p1[c] = (kBias + 1) - lds_s16((baseA + c * cstride) | oa1);
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t x = words[q] ^ rep;
    const bool f0 = (x & 0xFF00u) == 0u, f1 = (x & 0xFF000000u) == 0u;
    const uint32_t o0 = (words[q] << 1) & 0x1FEu;
    const uint32_t b1 = __byte_perm(words[q], 0u, 0x4442u);
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const uint32_t bb = baseB + c * cstride;
      const int32_t v0 = __viaddmin_s32_relu(lds_s16(bb | o0), f0 ? p0[c] : p1[c], kMiss + 1);
      const int32_t v1 = __viaddmin_s32_relu(lds_s16(b1 * c_s1_mul[4] + bb), f1 ? p0[c] : p1[c], kMiss + 1);
      if (WM == W_UNIT) {
        a32[c] += (int32_t)lds_u8(lut2_s + ((uint32_t)v0 * kS1LutSide + (uint32_t)v1));
      } else {
        s1_acc<WM>(lut[v0], wi[2 * q], wd[2 * q], a32[c], a64[c], ad[c]);
        s1_acc<WM>(lut[v1], wi[2 * q + 1], wd[2 * q + 1], a32[c], a64[c], ad[c]);
      }
    }
  }
}

__device__ __forceinline__ void s1_bulk_fill(uint32_t dst_smem, const void* src, uint32_t bar_smem) {
  asm volatile(
      "mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %3;\n\t"
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%1], [%2], %3, [%0];" ::"r"(bar_smem),
      "r"(dst_smem), "l"(src), "n"(kS1WarpStage)
      : "memory");
}
__device__ __forceinline__ void s1_bar_wait(uint32_t bar_smem, uint32_t parity) {
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

struct S1Tile {  // a tile cursor: tile t = (I, J) holds records [.., end)
  int64_t t, end;
  int32_t I, J;
  __device__ __forceinline__ void advance_to(int64_t r, const int64_t* __restrict__ tile_start, int32_t nblk) {
    while (r >= end) {  // skips empty tiles too
      ++t; ++J;
      if (J == nblk) { ++I; J = I; }
      end = tile_start[t + 1];
    }
  }
};

#ifdef LXY_S1_TRACE
__device__ unsigned long long g_s1_trace[1024 * 8 + 8 * 32];
__device__ __forceinline__ unsigned long long s1_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define S1_STAMP(k) do { if (threadIdx.x == 0 && blockIdx.x < 1024) g_s1_trace[blockIdx.x * 8 + (k)] = s1_now(); } while (0)
#else
#define S1_STAMP(k) do { } while (0)
#endif

template <int NC, int WM>
__global__ void __launch_bounds__(kS1Threads + (kS1MinBlocks == 2 && NC <= 2 ? 64 : 0), kS1MinBlocks) k_scaff_stream1(Stream1Params prm) {
  typedef typename Acc<WM>::type acc_t;
  extern __shared__ __align__(128) uint8_t smem_raw1[];
  // contig blocks of the table must start on 512-byte boundaries of the shared window
  uint8_t* smem = smem_raw1 + ((512u - (smem_u32(smem_raw1) & 511u)) & 511u);
  const int32_t n = prm.n;
  const int32_t nstride = (n + kBlk - 1) & ~(kBlk - 1);      // whole contig blocks: 512-byte aligned
  int16_t* sPos = (int16_t*)smem;                            // [NC][nstride], 32767 - pos
  uint8_t* s_lut = smem + (size_t)NC * nstride * 2;          // kLutLen bytes
  uint8_t* s_lut2 = s_lut + 64;                              // kS1Lut2Bytes: pair sums (unit weights)
  acc_t* s_red = (acc_t*)(s_lut2 + kS1Lut2Bytes);            // [kS1Warps][NC]
  int64_t* s_tile = (int64_t*)(s_red + kS1Warps * NC);
  const uint32_t sbase = smem_u32(sPos), cstride = (uint32_t)nstride * 2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  S1_STAMP(0);
#ifdef LXY_S1_TRACE
  if (threadIdx.x == 0 && blockIdx.x < 1024) { unsigned sm; asm("mov.u32 %0, %smid;" : "=r"(sm)); g_s1_trace[blockIdx.x * 8 + 4] = sm; }
#endif

  // ---- this CTA's share of the record stream: an even split of the 8-record packets, walked in
  // chunks of kS1Sub packets per thread (32 KB; a warp owns 1,024 consecutive records = 2 KB of each
  // chunk).  The warp's slices arrive in its own ring of `stages` 2 KB slots, each filled by one bulk
  // copy that lane 0 requests and whose completion the warp awaits on the slot's mbarrier: the L1 data
  // pipe, which bounds the stream phase (one wavefront per cycle; per warp-packet 4 for the packet's own
  // LDS.128, 1 + 8 table look-ups, 4 pair look-ups), no longer carries the 4 wavefronts of 32 lanes'
  // LDGSTS writes, and the three `@!PT LDS` issue slots ptxas pads every LDGSTS with are gone. ----
  const int64_t npk_all = (prm.L + kPacket - 1) / kPacket;
  const int64_t lo = (npk_all * blockIdx.x / gridDim.x) * kPacket;
  const int64_t hi = min((npk_all * (blockIdx.x + 1) / gridDim.x) * kPacket, prm.L);
  constexpr int32_t kSub = 32 * kPacket;                            // records per warp-packet
  constexpr int32_t kSlice = kSub * kS1Sub;                         // records per warp per chunk
  constexpr int32_t kChunk = kS1Warps * kSlice;                     // records per CTA per chunk
  const int32_t nchunks = (int32_t)((hi - lo + kChunk - 1) / kChunk);
  const int32_t stages = prm.stages;
  const uint32_t ring_w = smem_u32(s_tile + 2) + (uint32_t)warp * (uint32_t)(stages * kS1WarpStage);   // this warp's slots
  const uint32_t bar_w = smem_u32(s_tile + 2) + (uint32_t)(kS1Warps * stages * kS1WarpStage) + (uint32_t)warp * (uint32_t)(stages * 8);
  const uint16_t* rec_w = prm.rec + lo + (int64_t)warp * kSlice;    // the warp's slice of chunk 0
  const int32_t hi_rel = (int32_t)(hi - lo);                        // a share is far below 2^31 records

  // ---- prologue loads, requested BEFORE the ring is filled: behind 64 KB of record copies per CTA
  // (19 MB over the GPU) they waited 2 us in the memory system's queues (per-CTA time stamps: prologue
  // 3.2 us against 1 M records, 5.2 us against 200 M).  The first kS1Pre genes per thread of ordering 0,
  // the first two tile boundaries per thread, one 16-byte piece of the decay tables. ----
  const int64_t ntiles = (int64_t)prm.nblk * (prm.nblk + 1) / 2;
  int32_t pv[kS1Pre];
#pragma unroll
  for (int k = 0; k < kS1Pre; ++k) {
    const int i = tid + k * kS1Threads;
    pv[k] = i < n ? ldg_s32_ordered(prm.perm + i) : 0;
  }
  int64_t ts0 = 0, ts1 = 0;
  if (tid < ntiles) { ts0 = ldg_s64_ordered(prm.tile_start + tid); ts1 = ldg_s64_ordered(prm.tile_start + tid + 1); }
  uint4 lutv = make_uint4(0, 0, 0, 0);
  if (tid < kS1LutPieces) lutv = ldg_v4_ordered(g_s1_luts.v + tid * 16);
  if (lane == 0) {
    for (int q = 0; q < stages; ++q)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_w + q * 8) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int q = 0; q < stages; ++q)
      if (warp * kSlice + (int64_t)q * kChunk < hi_rel)
        s1_bulk_fill(ring_w + q * kS1WarpStage, rec_w + (int64_t)q * kChunk, bar_w + q * 8);
  }
  __syncwarp();
  // the tile holding record `lo` (the only t with tile_start[t] <= lo < tile_start[t+1])
  if (lo < hi) {
    if (tid < ntiles && ts0 <= lo && lo < ts1) { s_tile[0] = tid; s_tile[1] = ts1; }
    for (int64_t t = tid + kS1Threads; t < ntiles; t += kS1Threads)
      if (prm.tile_start[t] <= lo) {
        const int64_t e = prm.tile_start[t + 1];
        if (lo < e) { s_tile[0] = t; s_tile[1] = e; }
      }
  }
  if (tid < kS1LutPieces) reinterpret_cast<uint4*>(s_lut)[tid] = lutv;   // decay table | pair sums
  int bad = 0;
  // validation slice of this CTA: CTA k checks the contig IDS of slice k.  Every gene is in [0, n)
  // (checked during the build), so a row is a permutation <=> every id was written <=> for every id x
  // the position p found in the table is in range and perm[p] == x.  An id that occurs twice leaves
  // another one unwritten; the unwritten slot holds whatever the shared memory held, and whichever
  // position that decodes to, perm[p] != x because x is nowhere in the row.  The outcome therefore
  // does not depend on which of two racing stores to a duplicated id wins (a check by ordering
  // position, slice by slice, did: each CTA could see its own store win).
  const int32_t vi0 = (int32_t)((int64_t)n * blockIdx.x / gridDim.x), vi1 = (int32_t)((int64_t)n * (blockIdx.x + 1) / gridDim.x);
  constexpr int kVPre = NC <= 2 ? NC : 0;   // orderings whose first id per thread is compared at the end of the kernel
  int32_t vgot[kVPre > 0 ? kVPre : 1];
#pragma unroll
  for (int k = 0; k < kS1Pre; ++k) {   // ordering 0 (P >= 1), the genes requested above
    const int i = tid + k * kS1Threads;
    if (i < n) {
      if ((uint32_t)pv[k] >= (uint32_t)n) bad |= 1;
      else sPos[pv[k]] = (int16_t)(32767 - i);
    }
  }
#pragma unroll 1
  for (int c = 0; c < NC; ++c) {
    const int32_t* row = prm.perm + (size_t)c * n;
    for (int i = tid + (c == 0 ? kS1Pre * kS1Threads : 0); i < n; i += kS1Threads) {
      const int32_t cid = (c < prm.P) ? row[i] : i;
      if ((uint32_t)cid >= (uint32_t)n) bad |= 1;
      else sPos[c * nstride + cid] = (int16_t)(32767 - i);
    }
  }
  __syncthreads();
  if (prm.validate) {
    // the gather perm[p] of the first id per thread stays in flight while the records are scored
#pragma unroll
    for (int c = 0; c < kVPre; ++c) {
      vgot[c] = vi0 + tid;
      if (c < prm.P && vi0 + tid < vi1) {
        const uint32_t p = (uint32_t)(32767 - (int32_t)sPos[c * nstride + vi0 + tid]);
        vgot[c] = p < (uint32_t)n ? prm.perm[(size_t)c * n + p] : -1;
      }
    }
    for (int c = 0; c < prm.P && c < NC; ++c) {
      const int32_t* row = prm.perm + (size_t)c * n;
      for (int x = vi0 + tid + (c < kVPre ? kS1Threads : 0); x < vi1; x += kS1Threads) {
        const uint32_t p = (uint32_t)(32767 - (int32_t)sPos[c * nstride + x]);
        if (p >= (uint32_t)n || row[p] != x) bad |= 2;
      }
    }
  }
  if (bad) atomicOr(prm.flag, bad);
  S1_STAMP(1);

  int32_t a32[NC];
  int64_t a64[NC];
  double ad[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) { a32[c] = 0; a64[c] = 0; ad[c] = 0.0; }

  if (lo < hi) {
    // warp-uniform tile cursor: the tile of the first record of this warp's slice
    S1Tile cur;
    cur.t = s_tile[0];
    {
      int32_t a = 0, b = prm.nblk;
      while (b - a > 1) {
        const int32_t m = (a + b) >> 1;
        if (tile_row_start(m, prm.nblk) <= cur.t) a = m; else b = m;
      }
      cur.I = a;
    }
    cur.J = cur.I + (int32_t)(cur.t - tile_row_start(cur.I, prm.nblk));
    cur.end = s_tile[1];
    int32_t w_rel = warp * kSlice;                            // first record of the warp's slice (relative to lo)
    int32_t lim_rel = 0;                                      // end of (share ∩ current tile); 0 forces the set-up
    uint32_t baseA = sbase, baseB = sbase;
    int32_t st = 0;                                           // ring stage of chunk c, and its mbarrier phase
    uint32_t ph = 0;
    // Hand the stage back: lane 0 requests the warp's slice of chunk c + stages into it.
    auto refill = [&]() {
      if (lane == 0 && w_rel + stages * kChunk < hi_rel)
        s1_bulk_fill(ring_w + (uint32_t)st * kS1WarpStage, rec_w + (int64_t)w_rel - (int64_t)warp * kSlice + (int64_t)stages * kChunk,
                     bar_w + (uint32_t)st * 8);
    };
    // One warp-packet whose 256 records lie in the current tile: one a for the whole packet (sorted by
    // (a, b): first a == last a) is the common case, two neighbouring a's the next.
    auto score_packet = [&](const uint4 v, const int32_t (&wi)[kPacket], const double (&wd)[kPacket]) {
      if (__all_sync(0xffffffffu, ((v.x ^ (v.w >> 16)) & 0xFF00u) == 0u)) {
        s1_packet_same_a<NC, WM>(v, baseA, baseB, cstride, s_lut, s_lut2, wi, wd, a32, a64, ad);
      } else if (__all_sync(0xffffffffu, (v.w >> 24) - ((v.x >> 8) & 0xFFu) <= 1u)) {
        s1_packet_two_a<NC, WM>(v, baseA, baseB, cstride, s_lut, s_lut2, wi, wd, a32, a64, ad);
      } else {
        const uint32_t words[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < kPacket; ++k) {
          const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
          s1_eval<NC, WM>(baseA, baseB, cstride, s_lut, r, WM == W_INT32 ? wi[k] : 0,
                          WM == W_F64 ? wd[k] : 0.0, a32, a64, ad);
        }
      }
    };
    // Chunks are walked in RUNS: as long as the warp's slice lies inside (share ∩ current tile) nothing
    // but the ring and the packets themselves are touched.  A slice that crosses a tile boundary or the
    // end of the share takes the per-lane path, one chunk, and the next run starts behind it.
    int32_t c = 0;
#pragma unroll 1
    while (c < nchunks) {
      if (w_rel < hi_rel && w_rel + kSlice > lim_rel) {            // next tile (or the first one)
        cur.advance_to(lo + w_rel, prm.tile_start, prm.nblk);
        lim_rel = (int32_t)(min(hi, cur.end) - lo);
        baseA = sbase + (uint32_t)cur.I * (kBlk * 2);
        baseB = sbase + (uint32_t)cur.J * (kBlk * 2);
      }
      int32_t nrun = 0;                                            // warp-uniform
      if (w_rel + kSlice <= lim_rel) nrun = min((lim_rel - kSlice - w_rel) / kChunk + 1, nchunks - c);
      if (nrun == 0) {
        // The slice crosses a tile boundary or the end of the share.  Its packets are taken one by one: a
        // warp-packet that lies inside one tile and the share goes through the fast path with that tile's
        // rows, only the one that straddles takes a per-lane tile cursor.
        if (w_rel < hi_rel) {                                      // warp-uniform: a copy was requested
          s1_bar_wait(bar_w + (uint32_t)st * 8, ph);
          const uint32_t slot = ring_w + (uint32_t)st * kS1WarpStage + (uint32_t)lane * 16;
#pragma unroll 1
          for (int j = 0; j < kS1Sub; ++j) {
            const int32_t ws = w_rel + j * kSub;                   // first record of the warp-packet
            if (ws >= hi_rel) break;
            if (ws >= lim_rel) {                                   // the packet starts in a later tile
              cur.advance_to(lo + ws, prm.tile_start, prm.nblk);
              lim_rel = (int32_t)(min(hi, cur.end) - lo);
              baseA = sbase + (uint32_t)cur.I * (kBlk * 2);
              baseB = sbase + (uint32_t)cur.J * (kBlk * 2);
            }
            const int32_t rr = ws + lane * kPacket;
            const uint4 v = lds_v4(slot + (uint32_t)j * 512);
            int32_t wi[kPacket];
            double wd[kPacket];
            if (ws + kSub <= lim_rel) {
              if (WM != W_UNIT) load_weights<WM>(prm.w, lo + rr, wi, wd);
              score_packet(v, wi, wd);
            } else if (rr < hi_rel) {
              if (WM != W_UNIT) load_weights<WM>(prm.w, lo + rr, wi, wd);
              S1Tile mine = cur;
              const uint32_t words[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
              for (int k = 0; k < kPacket; ++k) {  // unrolled: words[] stays in registers
                if (rr + k < hi_rel) {
                  mine.advance_to(lo + rr + k, prm.tile_start, prm.nblk);
                  const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
                  s1_eval<NC, WM>(sbase + (uint32_t)mine.I * (kBlk * 2), sbase + (uint32_t)mine.J * (kBlk * 2),
                                  cstride, s_lut, r, WM == W_INT32 ? wi[k] : 0, WM == W_F64 ? wd[k] : 0.0,
                                  a32, a64, ad);
                }
              }
            }
          }
          fence_proxy_async();   // every lane's reads of the slot are complete before it is handed back
          __syncwarp();
          refill();
        }
        ++c;
        w_rel += kChunk;
        if (++st == stages) { st = 0; ph ^= 1u; }
        continue;
      }
      const int32_t c_end = c + nrun;
#pragma unroll 1
      for (; c < c_end; ++c) {
        s1_bar_wait(bar_w + (uint32_t)st * 8, ph);
        const uint32_t slot = ring_w + (uint32_t)st * kS1WarpStage + (uint32_t)lane * 16;
#pragma unroll
        for (int j = 0; j < kS1Sub; ++j) {
          const uint4 v = lds_v4(slot + (uint32_t)j * 512);
          // one a for the whole packet (sorted by (a, b): first a == last a)
          const bool one_a = __all_sync(0xffffffffu, ((v.x ^ (v.w >> 16)) & 0xFF00u) == 0u);
          // the last packet of the slot is in registers (the vote consumed it): the slot can be refilled
          if (j == kS1Sub - 1) refill();
          int32_t wi[kPacket];
          double wd[kPacket];
          if (WM != W_UNIT) load_weights<WM>(prm.w, lo + w_rel + j * kSub + lane * kPacket, wi, wd);
          if (one_a) {
            s1_packet_same_a<NC, WM>(v, baseA, baseB, cstride, s_lut, s_lut2, wi, wd, a32, a64, ad);
          } else if (__all_sync(0xffffffffu, (v.w >> 24) - ((v.x >> 8) & 0xFFu) <= 1u)) {
            s1_packet_two_a<NC, WM>(v, baseA, baseB, cstride, s_lut, s_lut2, wi, wd, a32, a64, ad);
          } else {
            const uint32_t words[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < kPacket; ++k) {
              const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
              s1_eval<NC, WM>(baseA, baseB, cstride, s_lut, r, WM == W_INT32 ? wi[k] : 0,
                              WM == W_F64 ? wd[k] : 0.0, a32, a64, ad);
            }
          }
        }
        w_rel += kChunk;
        if (++st == stages) { st = 0; ph ^= 1u; }
      }
    }
  }
  S1_STAMP(2);
#ifdef LXY_S1_TRACE
  if (lane == 0 && blockIdx.x < 8) g_s1_trace[1024 * 8 + blockIdx.x * 32 + warp] = s1_now();
#endif
  if (kVPre > 0 && prm.validate) {
    bool miss = false;
#pragma unroll
    for (int c = 0; c < kVPre; ++c) miss |= (vgot[c] != vi0 + tid);
    if (miss) atomicOr(prm.flag, 2);
  }

  // ---- warp shuffle tree, then a fixed-order sum over the warps ----
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    acc_t v;
    if (WM == W_F64) v = (acc_t)ad[c];
    else if (WM == W_UNIT) v = (acc_t)a32[c];
    else v = (acc_t)a64[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      if (WM == W_F64) v += (acc_t)__shfl_xor_sync(0xffffffffu, (double)v, o);
      else v += (acc_t)__shfl_xor_sync(0xffffffffu, (long long)v, o);
    }
    if (lane == 0) s_red[warp * NC + c] = v;
  }
  __syncthreads();
  if (tid < NC) {
    acc_t sum = 0;
    for (int wq = 0; wq < kS1Warps; ++wq) sum += s_red[wq * NC + tid];
    ((acc_t*)prm.partial)[(size_t)blockIdx.x * NC + tid] = sum;
  }
  S1_STAMP(3);
}

template <int NC, int WM>
int launch_stream1(lxy_ctx* c, const Stream1Params& prm, int grid, size_t smem) {
  static size_t configured[64] = {};
  std::lock_guard<std::mutex> cfg_lock(config_mutex());
  if (configured[c->cfg.device & 63] < smem) {
    LXY_CUDA(cudaFuncSetAttribute(k_scaff_stream1<NC, WM>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)smem));
    LXY_CUDA(cudaFuncSetAttribute(k_scaff_stream1<NC, WM>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
    configured[c->cfg.device & 63] = smem;
  }
  k_scaff_stream1<NC, WM><<<grid, kS1Threads, smem, c->stream>>>(prm);
  LXY_CUDA(cudaGetLastError());
  return LXY_OK;
}

template <int WM>
int launch_stream1_nc(lxy_ctx* c, int nc, const Stream1Params& prm, int grid, size_t smem) {
  switch (nc) {
    case 1: return launch_stream1<1, WM>(c, prm, grid, smem);
    case 2: return launch_stream1<2, WM>(c, prm, grid, smem);
    case 4: return launch_stream1<4, WM>(c, prm, grid, smem);
    default: return launch_stream1<8, WM>(c, prm, grid, smem);
  }
}

inline size_t stream1_smem(int nc, int n, int stages = 3) {
  return (size_t)nc * ((n + kBlk - 1) & ~(kBlk - 1)) * 2 + 64 + kS1Lut2Bytes + (size_t)kS1Warps * nc * 8 + 16 + 512 +
         (size_t)kS1Warps * stages * (kS1WarpStage + 8);  // (alignment slack) table | lut | warp sums | start tile | record ring | ring barriers
}
