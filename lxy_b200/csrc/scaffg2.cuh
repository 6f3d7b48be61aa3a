// scaffg2.cuh — the G-profile scorer for unit-weight records, 128 candidates per group.
// (included by scaffg_kernels.cu inside its anonymous namespace; same semantics as k_scaffg_score —
// the oracle's scaffg_* functions define them, no reference counterpart — and the same skeleton:
// tile walk, TMA-staged table of block J, 8-record packets, fixed-order reduction.)
//
// What changed, and why it is ~2x the evaluations per second of k_scaffg_score:
//  * 4 candidates per lane instead of 2: the per-record work that does not depend on the candidate
//    (packet loads, row addresses, the "first contig" test, loop control) is shared by twice as many
//    evaluations.  Two 128-candidate origin tables (2 x 128 KB) do not fit the SM, so only the table of
//    block J lives in shared memory; row a — constant over the run of records with the same first
//    contig, records being sorted by (a, b) inside a tile — is read from global memory (L2) into
//    registers, one packet ahead.
//  * records and offsets (10 bytes per record in three arrays) arrive through a per-warp ring of bulk
//    copies and are read from shared memory two records at a time: 64 registers, 32 warps per SM
//    (the first version kept two whole packets per warp in registers: 128 registers, 16 warps, 6 % slower).
//  * per evaluation: table entry e = -start (forward) or start + len - 1 (reversed), so that
//    coord = |x - e| is ONE VABSDIFF (the absolute value handles the orientation), d = |coord a - coord b|
//    a second one; d is clamped to 22 * unit - 1 (VIMNMX) and divided by one IMAD.HI + shift:
//    bucket = (d * m) >> (32 + sh) with m = ceil(2^(32 + sh) / unit), exact for every d <= 22 * unit - 1
//    because d * (m * unit - 2^(32 + sh)) < 2^(32 + sh) (g2_magic picks the smallest such sh).
//    k_scaffg_score spent IADD + IABS per coordinate and two IMAD.WIDE + adds on the 64-bit reciprocal.
//  * two candidates share one decay look-up (pair table lut2[q1][q0] = 100 T[q0] | 100 T[q1] << 16, one
//    LDS.32 and one IADD into a packed pair of 16-bit sums, flushed every 64 packets), as in k_scaff_score.
//  * the `i-k > 0` multiplicity: a byte per contig row says whether ANY candidate of the group has the
//    contig first; only then (warp-uniform, <= 128 of n rows) the 128 first-bits are consulted.
// Domain: unit weights, total length < 2^31 (signed coordinates), 2 <= unit < ~1.9e8; everything else
// stays on k_scaffg_score.
#pragma once

constexpr int kCand2 = 128;
constexpr int kRow2 = kCand2 * 4;          // bytes per contig row of the origin table
constexpr int kTab2 = kBlk * kRow2;        // 128 KB
constexpr int kFirstRow2 = 32;             // bytes per contig row: 4 words of first-bits, 1 any-first word, 3 pad
constexpr int kFirst2 = kBlk * kFirstRow2;
constexpr int kQ2 = 22;                    // bucket indices 0..21; 21 = far (T = 0)
constexpr int kLutG2Stride = 23;
constexpr int kLutG2Words = kQ2 * kLutG2Stride;
constexpr int kG2Flush = 64;               // packets between flushes of the packed sums: 64 * 8 * 100 < 2^16
constexpr int kG2Smem = kTab2 + kFirst2 + 4 * ((kLutG2Words + 3) & ~3) + 64;

struct G2Params {
  const uint16_t* rec;
  const uint32_t* xa;
  const uint32_t* xb;
  const int64_t* tile_start;
  const int32_t* origin;     // [groups][n][128]: -start | start + len - 1
  const uint32_t* firstw;    // [groups][n][8]: first-bits of the 128 candidates | any-first | pad
  int64_t* partial;          // [units][2][128]: S then H
  int64_t L;
  int32_t n, nblk;
  int32_t groups, splits;
  uint32_t K, m, sh;         // clamp, magic multiplier, post-shift
  uint32_t shmul;            // 2^(32 - sh) (sh > 0)
};

// Smallest sh with m = ceil(2^(32 + sh) / unit) < 2^32 and K * (m * unit - 2^(32 + sh)) < 2^(32 + sh).
inline bool g2_magic(uint32_t unit, uint32_t& K, uint32_t& m, uint32_t& sh) {
  if (unit < 2) return false;
  const uint64_t k64 = 22ull * unit - 1;
  K = k64 > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)k64;
  for (sh = 0; sh < 31; ++sh) {
    const unsigned __int128 two = (unsigned __int128)1 << (32 + sh);
    const unsigned __int128 mm = (two + unit - 1) / unit;
    if (mm >> 32) return false;
    const unsigned __int128 e = mm * unit - two;
    if ((unsigned __int128)K * e < two) { m = (uint32_t)mm; return true; }
  }
  return false;
}

// One warp per candidate: prefix scan of the permuted lengths, tables in the 128-candidate layout.
__global__ void k_scaffg2_layout(const int32_t* __restrict__ perm, const uint8_t* __restrict__ ori,
                                 const uint32_t* __restrict__ len, int32_t P, int32_t Ppad, int32_t n,
                                 int32_t* __restrict__ origin, uint32_t* __restrict__ firstw,
                                 int32_t* __restrict__ pos, int* __restrict__ flag) {
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= Ppad) return;
  const int g = p / kCand2, pc = p - g * kCand2;
  uint32_t carry = 0;
  for (int i0 = 0; i0 < n; i0 += 32) {
    const int i = i0 + lane;
    int32_t cid = -1;
    uint32_t l = 0;
    if (i < n) {
      cid = (p < P) ? perm[(size_t)p * n + i] : i;  // pad candidates: identity, forward
      if ((uint32_t)cid >= (uint32_t)n) { atomicOr(flag, 1); cid = -1; }
      else l = len[cid];
    }
    uint32_t s = l;   // inclusive warp scan of the lengths (Hillis-Steele over shuffle-up)
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += v;
    }
    const uint32_t start = carry + s - l;            // exclusive
    carry += __shfl_sync(0xffffffffu, s, 31);
    if (cid >= 0) {
      const bool rev = (p < P) && ori && ori[(size_t)p * n + i] != 0;
      origin[((size_t)g * n + cid) * kCand2 + pc] = rev ? (int32_t)(start + l - 1) : -(int32_t)start;
      if (i == 0) {
        atomicOr(firstw + ((size_t)g * n + cid) * 8 + (pc >> 5), 1u << (pc & 31));
        firstw[((size_t)g * n + cid) * 8 + 4] = 1u;
      }
      if (p < P) pos[(size_t)p * n + cid] = i;
    }
  }
}

__device__ __forceinline__ uint32_t lds_u32g(uint32_t addr) {
  uint32_t v;
  asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

struct G2RowA {
  int4 e;          // the four origins of this lane's candidates
  uint32_t fw;     // first-bits word holding this lane's candidates
  uint32_t any;    // any candidate of the group has the contig first (warp-uniform)
};

__device__ __forceinline__ G2RowA g2_load_row_a(const G2Params& prm, size_t row, int lane) {
  G2RowA R;
  R.e = __ldg(reinterpret_cast<const int4*>(prm.origin + row * kCand2) + lane);
  R.any = __ldg(prm.firstw + row * 8 + 4);
  R.fw = __ldg(prm.firstw + row * 8 + (lane >> 3));
  return R;
}

struct G2Acc {
  uint32_t p[2], hp[2];    // packed pairs of 16-bit sums (S and H)
  int32_t s32[4], h32[4];  // per unit; the host sizes the units so that these cannot overflow
  __device__ __forceinline__ void clear() {
    p[0] = p[1] = hp[0] = hp[1] = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) s32[j] = h32[j] = 0;
  }
  __device__ __forceinline__ void flush16() {
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      s32[2 * q] += (int32_t)(p[q] & 0xFFFFu); s32[2 * q + 1] += (int32_t)(p[q] >> 16);
      h32[2 * q] += (int32_t)(hp[q] & 0xFFFFu); h32[2 * q + 1] += (int32_t)(hp[q] >> 16);
      p[q] = hp[q] = 0;
    }
  }
};

// One record against the four candidates of this lane.  SH0: the magic needs no post-shift.
template <bool SH0>
__device__ __forceinline__ void g2_eval(const G2RowA& A, uint32_t b_loc, uint32_t xa, uint32_t xb,
                                        const uint8_t* sB, const uint8_t* sFirstB,
                                        uint32_t lut, int lane, const G2Params& prm, G2Acc& acc) {
  const int4 eb = *reinterpret_cast<const int4*>(sB + b_loc * kRow2 + lane * 16);
  const int ea[4] = {A.e.x, A.e.y, A.e.z, A.e.w}, ebv[4] = {eb.x, eb.y, eb.z, eb.w};
  uint32_t q[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t ca = __sad((int)xa, ea[j], 0u);    // |x - e|: start + x, or start + len - 1 - x
    const uint32_t cb = __sad((int)xb, ebv[j], 0u);
    const uint32_t d = __usad(ca, cb, 0u);
    // bucket, 21 when far; the post-shift is a second multiply-high by 2^(32 - sh): IMAD.HI runs on the
    // FMA pipe, which idles here, SHF on the ALU pipe, which is the busy one (VABSDIFF, VIMNMX)
    const uint32_t h = __umulhi(min(d, prm.K), prm.m);
    q[j] = SH0 ? h : __umulhi(h, prm.shmul);
  }
  const uint32_t e01 = lds_u32g(lut + q[0] * 4u + q[1] * (4u * kLutG2Stride));
  const uint32_t e23 = lds_u32g(lut + q[2] * 4u + q[3] * (4u * kLutG2Stride));
  acc.p[0] += e01;
  acc.p[1] += e23;
  if (A.any | *reinterpret_cast<const uint32_t*>(sFirstB + b_loc * kFirstRow2 + 16)) {
    // rare and warp-uniform: some candidate of the group has a or b first
    const uint32_t fw = A.fw | *reinterpret_cast<const uint32_t*>(sFirstB + b_loc * kFirstRow2 + (lane >> 3) * 4);
    const uint32_t bits = (fw >> (4 * (lane & 7))) & 15u;
    acc.hp[0] += e01 & (((bits & 1u) ? 0xFFFFu : 0u) | ((bits & 2u) ? 0xFFFF0000u : 0u));
    acc.hp[1] += e23 & (((bits & 4u) ? 0xFFFFu : 0u) | ((bits & 8u) ? 0xFFFF0000u : 0u));
  }
}

struct G2Packet {
  uint4 rec, xa0, xa1, xb0, xb1;
};
__device__ __forceinline__ G2Packet g2_load_packet(const G2Params& prm, int64_t base) {
  G2Packet p;
  p.rec = ldg_stream_v4(prm.rec + base);
  p.xa0 = ldg_stream_v4(prm.xa + base); p.xa1 = ldg_stream_v4(prm.xa + base + 4);
  p.xb0 = ldg_stream_v4(prm.xb + base); p.xb1 = ldg_stream_v4(prm.xb + base + 4);
  return p;
}

// Eight records with one a: row A comes in registers.
template <bool SH0>
__device__ __forceinline__ void g2_eval_packet(const G2Packet& pk, const G2RowA& A, const uint8_t* sB,
                                               const uint8_t* sFirstB, uint32_t lut,
                                               int lane, const G2Params& prm, G2Acc& acc) {
  const uint32_t words[4] = {pk.rec.x, pk.rec.y, pk.rec.z, pk.rec.w};
  const uint32_t xav[8] = {pk.xa0.x, pk.xa0.y, pk.xa0.z, pk.xa0.w, pk.xa1.x, pk.xa1.y, pk.xa1.z, pk.xa1.w};
  const uint32_t xbv[8] = {pk.xb0.x, pk.xb0.y, pk.xb0.z, pk.xb0.w, pk.xb1.x, pk.xb1.y, pk.xb1.z, pk.xb1.w};
#pragma unroll
  for (int k = 0; k < kPacket; ++k) {
    const uint32_t b_loc = (k & 1) ? ((words[k >> 1] >> 16) & 0xFFu) : (words[k >> 1] & 0xFFu);
    g2_eval<SH0>(A, b_loc, xav[k], xbv[k], sB, sFirstB, lut, lane, prm, acc);
  }
}

// Records k in [k0, k1) of a packet, row A fetched per record (boundary packets, packets that span
// several a's).
template <bool SH0>
__device__ __forceinline__ void g2_eval_records(const G2Packet& pk, int k0, int k1, size_t rowA0,
                                                const uint8_t* sB, const uint8_t* sFirstB,
                                                uint32_t lut, int lane,
                                                const G2Params& prm, G2Acc& acc) {
  const uint32_t words[4] = {pk.rec.x, pk.rec.y, pk.rec.z, pk.rec.w};
  const uint32_t xav[8] = {pk.xa0.x, pk.xa0.y, pk.xa0.z, pk.xa0.w, pk.xa1.x, pk.xa1.y, pk.xa1.z, pk.xa1.w};
  const uint32_t xbv[8] = {pk.xb0.x, pk.xb0.y, pk.xb0.z, pk.xb0.w, pk.xb1.x, pk.xb1.y, pk.xb1.z, pk.xb1.w};
  uint32_t a_prev = 0xFFFFFFFFu;
  G2RowA A;
  A.e = make_int4(0, 0, 0, 0); A.fw = 0; A.any = 0;
#pragma unroll
  for (int k = 0; k < kPacket; ++k) {
    if (k < k0 || k >= k1) continue;
    const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
    const uint32_t a_loc = r >> 8;
    if (a_loc != a_prev) { A = g2_load_row_a(prm, rowA0 + a_loc, lane); a_prev = a_loc; }   // warp-uniform
    g2_eval<SH0>(A, r & 0xFFu, xav[k], xbv[k], sB, sFirstB, lut, lane, prm, acc);
  }
}

// ---- k_scaffg2r_score: the scorer; records and offsets arrive through a per-warp ring of bulk copies ----
#ifndef LXY_G2R_THREADS
#define LXY_G2R_THREADS 1024
#endif
constexpr int kG2RThreads = LXY_G2R_THREADS;
constexpr int kG2RWarps = kG2RThreads / 32;
constexpr int kG2RingPackets = 16;                         // packets per ring slot
constexpr int kG2SlotBytes = kG2RingPackets * 80;          // 16 B of records + 2 x 32 B of offsets per packet
constexpr int kG2RSmem = kG2Smem + kG2RWarps * (2 * kG2SlotBytes + 16);
__device__ __forceinline__ uint32_t g2_lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint2 g2_lds_v2(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint4 g2_lds_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void g2_bar_wait(uint32_t bar_smem, uint32_t parity) {
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

template <bool SH0>
__global__ void __launch_bounds__(kG2RThreads, 1) k_scaffg2r_score(G2Params prm) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sB = smem;
  uint8_t* sFirstB = smem + kTab2;
  uint32_t* s_lut2 = (uint32_t*)(sFirstB + kFirst2);
  uint64_t* bar = (uint64_t*)(s_lut2 + ((kLutG2Words + 3) & ~3));
  const uint32_t lut = smem_u32(s_lut2);
  // per-warp record ring: two slots of kG2RingPackets packets (records | offsets a | offsets b) and their mbarriers
  const uint32_t ring_w = smem_u32(bar + 8) + (uint32_t)(threadIdx.x >> 5) * (2 * kG2SlotBytes);
  const uint32_t ring_bar = smem_u32(bar + 8) + (uint32_t)kG2RWarps * (2 * kG2SlotBytes) + (uint32_t)(threadIdx.x >> 5) * 16;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < kLutG2Words; i += kG2RThreads) {
    const int q1 = i / kLutG2Stride, q0 = i - q1 * kLutG2Stride;
    s_lut2[i] = (q0 < kQ2 ? (uint32_t)c_lutg[q0] : 0u) | ((uint32_t)c_lutg[q1] << 16);
  }
  if (tid == 0) mbar_init(bar, 1);
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ring_bar) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(ring_bar + 8) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  uint32_t ring_ph = 0;   // bit s: phase of slot s's mbarrier
  __syncthreads();
  uint32_t parity = 0;

  const int units = prm.groups * prm.splits;
  const int64_t npk_all = (prm.L + kPacket - 1) / kPacket;
  const int64_t ntiles = (int64_t)prm.nblk * (prm.nblk + 1) / 2;
  for (int u = blockIdx.x; u < units; u += gridDim.x) {
    const int g = u / prm.splits, s = u - g * prm.splits;
    const int64_t lo = (npk_all * s / prm.splits) * kPacket;
    const int64_t hi = min((npk_all * (s + 1) / prm.splits) * kPacket, prm.L);
    G2Acc acc;
    acc.clear();
    int32_t loadedJ = -1;
    if (lo < hi) {
      int64_t tl = 0, th = ntiles;
      while (th - tl > 1) {
        const int64_t mid = (tl + th) >> 1;
        if (prm.tile_start[mid] <= lo) tl = mid; else th = mid;
      }
      int64_t t = tl;
      int32_t I = 0;
      {
        int32_t a = 0, b = prm.nblk;
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
        if (tend <= cur) {
          ++t; ++J;
          if (J == prm.nblk) { ++I; J = I; }
          continue;
        }
        const int64_t seg_end = min(hi, tend);
        if (J != loadedJ) {   // stage block J of the origin table, its first-bits and any-first bytes
          __syncthreads();
          if (tid == 0) {
            const int rowsB = min(kBlk, prm.n - J * kBlk);
            const size_t rb = (size_t)g * prm.n + (size_t)J * kBlk;
            fence_proxy_async();
            mbar_expect_tx(bar, (uint32_t)rowsB * (kRow2 + kFirstRow2));
            tma_bulk_g2s(sB, prm.origin + rb * kCand2, (uint32_t)rowsB * kRow2, bar);
            tma_bulk_g2s(sFirstB, prm.firstw + rb * 8, (uint32_t)rowsB * kFirstRow2, bar);
          }
          loadedJ = J;
          mbar_wait(bar, parity);
          parity ^= 1;
        }
        const size_t rowA0 = (size_t)g * prm.n + (size_t)I * kBlk;

        const int64_t p_first = cur / kPacket, p_last = (seg_end - 1) / kPacket;
        const bool head_partial = (p_first * kPacket < cur);
        const bool tail_partial = ((p_last + 1) * kPacket > seg_end);
        if (head_partial && warp == 0) {   // boundary packets: masked, one warp each
          const int64_t base = p_first * kPacket;
          const G2Packet pk = g2_load_packet(prm, base);
          g2_eval_records<SH0>(pk, (int)(cur - base), (int)min((int64_t)kPacket, seg_end - base), rowA0, sB, sFirstB,
                          lut, lane, prm, acc);
          acc.flush16();
        }
        if (tail_partial && warp == 1 && !(head_partial && p_last == p_first)) {
          const int64_t base = p_last * kPacket;
          const G2Packet pk = g2_load_packet(prm, base);
          g2_eval_records<SH0>(pk, 0, (int)(seg_end - base), rowA0, sB, sFirstB, lut, lane, prm, acc);
          acc.flush16();
        }
        const int64_t f0 = p_first + (head_partial ? 1 : 0);
        const int64_t f1 = p_last - (tail_partial ? 1 : 0);  // inclusive
        const int64_t nfull = f1 - f0 + 1;
        if (nfull > 0) {
          // Every warp takes a contiguous run of the segment's packets, in slices of kG2RingPackets that arrive
          // in its own two-slot ring (three bulk copies per slice: records, offsets a, offsets b, counted on
          // the slot's mbarrier).  The packets are read from shared memory as they are scored, two records at a
          // time (one word of records, two words of each offset: 5 registers), so a thread needs ~60
          // registers instead of the 128 that two whole packets in flight took, and the SM holds 32 warps.
          const int64_t w_lo = nfull * warp / kG2RWarps, w_hi = nfull * (warp + 1) / kG2RWarps;
          const int64_t nsl = (w_hi - w_lo + kG2RingPackets - 1) / kG2RingPackets;
          auto fill = [&](int64_t sl) {   // lane 0
            const int64_t p0 = f0 + w_lo + sl * kG2RingPackets;               // first packet of the slice (absolute)
            const uint32_t np = (uint32_t)min((int64_t)kG2RingPackets, w_hi - w_lo - sl * kG2RingPackets);
            const uint32_t b = ring_bar + (uint32_t)(sl & 1) * 8, dst = ring_w + (uint32_t)(sl & 1) * kG2SlotBytes;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(np * 80u) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                         "l"(prm.rec + p0 * kPacket), "r"(np * 16u), "r"(b) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst + kG2RingPackets * 16),
                         "l"(prm.xa + p0 * kPacket), "r"(np * 32u), "r"(b) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst + kG2RingPackets * 48),
                         "l"(prm.xb + p0 * kPacket), "r"(np * 32u), "r"(b) : "memory");
          };
          if (lane == 0) {
            if (nsl > 0) fill(0);
            if (nsl > 1) fill(1);
          }
          int32_t since = 0;
#pragma unroll 1
          for (int64_t sl = 0; sl < nsl; ++sl) {
            const uint32_t sb = (uint32_t)(sl & 1);
            g2_bar_wait(ring_bar + sb * 8, (ring_ph >> sb) & 1u);
            ring_ph ^= 1u << sb;
            const int32_t cnt = (int32_t)min((int64_t)kG2RingPackets, w_hi - w_lo - sl * kG2RingPackets);
            const uint32_t s_rec = ring_w + sb * kG2SlotBytes, s_xa = s_rec + kG2RingPackets * 16, s_xb = s_rec + kG2RingPackets * 48;
            // row a of a packet is requested one packet ahead (its index is the packet's first word, already in
            // shared memory); the first packet of a slice waits for it
            G2RowA RA = g2_load_row_a(prm, rowA0 + ((g2_lds_u32(s_rec) >> 8) & 0xFFu), lane);
#pragma unroll 1
            for (int32_t k = 0; k < cnt; ++k) {
              const uint4 w = g2_lds_v4(s_rec + (uint32_t)k * 16);
              G2RowA RN = RA;
              if (k + 1 < cnt) RN = g2_load_row_a(prm, rowA0 + ((g2_lds_u32(s_rec + (uint32_t)(k + 1) * 16) >> 8) & 0xFFu), lane);
              const uint32_t words[4] = {w.x, w.y, w.z, w.w};
              if (((w.x >> 8) & 0xFFu) == (w.w >> 24)) {   // one a for the whole packet (warp-uniform)
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                  const uint2 xa2 = g2_lds_v2(s_xa + (uint32_t)k * 32 + (uint32_t)h * 8);
                  const uint2 xb2 = g2_lds_v2(s_xb + (uint32_t)k * 32 + (uint32_t)h * 8);
                  g2_eval<SH0>(RA, words[h] & 0xFFu, xa2.x, xb2.x, sB, sFirstB, lut, lane, prm, acc);
                  g2_eval<SH0>(RA, (words[h] >> 16) & 0xFFu, xa2.y, xb2.y, sB, sFirstB, lut, lane, prm, acc);
                }
              } else {                                     // several a's: row a per record (L2)
                uint32_t a_prev = 0xFFFFFFFFu;
                G2RowA A;
                A.e = make_int4(0, 0, 0, 0); A.fw = 0; A.any = 0;
#pragma unroll 1
                for (int r8 = 0; r8 < kPacket; ++r8) {
                  const uint32_t r = (r8 & 1) ? (words[r8 >> 1] >> 16) : (words[r8 >> 1] & 0xFFFFu);
                  const uint32_t a_loc = r >> 8;
                  if (a_loc != a_prev) { A = g2_load_row_a(prm, rowA0 + a_loc, lane); a_prev = a_loc; }
                  g2_eval<SH0>(A, r & 0xFFu, g2_lds_u32(s_xa + (uint32_t)k * 32 + (uint32_t)r8 * 4),
                               g2_lds_u32(s_xb + (uint32_t)k * 32 + (uint32_t)r8 * 4), sB, sFirstB, lut, lane, prm, acc);
                }
              }
              RA = RN;
              if (++since == kG2Flush) { acc.flush16(); since = 0; }
            }
            // the slot's packets have all been read (their values were consumed): hand it back
            __syncwarp();
            if (lane == 0 && sl + 2 < nsl) fill(sl + 2);
          }
          acc.flush16();
        }
        cur = seg_end;
      }
    }
    acc.flush16();
    __syncthreads();
    loadedJ = -1;                        // the table area is reused by the reduction
    int64_t* red = (int64_t*)smem;       // [kG2RWarps][2][128]
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      red[(warp * 2 + 0) * kCand2 + lane * 4 + j] = acc.s32[j];
      red[(warp * 2 + 1) * kCand2 + lane * 4 + j] = acc.h32[j];
    }
    __syncthreads();
    if (tid < 2 * kCand2) {
      int64_t sum = 0;
      for (int wq = 0; wq < kG2RWarps; ++wq) sum += red[wq * 2 * kCand2 + tid];
      prm.partial[(size_t)u * 2 * kCand2 + tid] = sum;
    }
  }
}

__global__ void k_scaffg2_finish(int32_t P, int32_t splits, const int64_t* __restrict__ partial,
                                 double* __restrict__ out, int64_t* __restrict__ x100) {
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= P) return;
  const int g = p / kCand2, pc = p - g * kCand2;
  int64_t S = 0, H = 0;
  for (int s = lane; s < splits; s += 32) {
    const int64_t* row = partial + ((size_t)g * splits + s) * 2 * kCand2;
    S += row[pc];
    H += row[kCand2 + pc];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    S += __shfl_xor_sync(0xffffffffu, (long long)S, o);
    H += __shfl_xor_sync(0xffffffffu, (long long)H, o);
  }
  if (lane != 0) return;
  const int64_t tot = S + S - H;
  out[p] = -((double)tot / 100.0);
  if (x100) x100[p] = tot;
}
