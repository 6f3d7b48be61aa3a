// scaffg_kernels.cu — scaffolding fitness, G-profile: the generalisation BASELINE.json's north_star
// describes (contig lengths, orientations, per-link in-contig offsets, genomic distance -> decay).
// It has no counterpart in the reference; the CPU checker under oracle/ (its scaffg_* functions) defines it, constrained
// to collapse to scaff.go:246-310 bit for bit when len == 1, x == 0, unit == 1.
//
//   k_scaffg_layout  one warp per candidate: warp prefix scan (shuffle-up) of the permuted contig
//                    lengths -> placement offsets; writes the origin table
//                    origin[g][contig][cand] = start (forward) | start + len - 1 (reversed) and the
//                    per-contig bit rows {reversed?, first in the ordering?} of the 64 candidates.
//   k_scaffg_score   same streaming structure as k_scaff_score (tiles of 256 x 256 contigs, tables
//                    staged by TMA bulk copies, 2-byte records + their two in-contig offsets
//                    streamed with 128-bit loads, lane = candidate):
//                    coord = origin -/+ x, d = |coord a - coord b|, bucket = d / unit by a 64-bit
//                    multiply-high (exact for 32-bit d), decay table, integer / fp64 sums.  The
//                    `i-k > 0` multiplicity (pairs touching the first contig count once) is kept as
//                    a second running sum that is only touched when a record's row has a "first"
//                    bit set for some candidate of the group (a warp-uniform test).
//   k_scaffg_finish  fixed-order reduction, fitness = -(2 S - H) / 100.
#include "common.cuh"
#include "device_utils.cuh"

namespace lxy {
namespace {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kPacket = 8;
constexpr int kCandG = 64;                 // candidates per group: lane l owns l and l + 32
constexpr int kRowG = kCandG * 4;          // bytes per contig row of the origin table
constexpr int kTabG = kBlk * kRowG;        // 64 KB
constexpr int kBitsG = kBlk * 16;          // uint4 {rev[0..31], rev[32..63], first[0..31], first[32..63]}
constexpr int kLutG = 24;

__constant__ uint8_t c_lutg[kLutG] = {0, 100, 50, 33, 0, 20, 0, 0, 0, 0, 0, 10,
                                      0, 0,   0,  0,  0, 0,  0, 0, 5, 0, 0, 0};

// ---- layout ---------------------------------------------------------------------------------
__global__ void k_scaffg_layout(const int32_t* __restrict__ perm, const uint8_t* __restrict__ ori,
                                const uint32_t* __restrict__ len, int32_t P, int32_t Ppad, int32_t n,
                                int sgn, uint32_t* __restrict__ origin, uint32_t* __restrict__ bits,
                                int32_t* __restrict__ pos, int* __restrict__ flag) {
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= Ppad) return;
  const int g = p / kCandG, pc = p - g * kCandG;
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
    // inclusive warp scan of the lengths (Hillis-Steele over shuffle-up)
    uint32_t s = l;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += v;
    }
    const uint32_t start = carry + s - l;            // exclusive
    carry += __shfl_sync(0xffffffffu, s, 31);
    if (cid >= 0) {
      const bool rev = (p < P) && ori && ori[(size_t)p * n + i] != 0;
      // sgn: orientation in the sign (coord = |v + x|); else unsigned origin + orientation bit
      const uint32_t e = start + l - 1;
      origin[((size_t)g * n + cid) * kCandG + pc] = rev ? (sgn ? (uint32_t)(-(int32_t)e) : e) : start;
      uint32_t* row = bits + ((size_t)g * n + cid) * 4;
      if (rev) atomicOr(row + (pc >> 5), 1u << (pc & 31));
      if (i == 0) atomicOr(row + 2 + (pc >> 5), 1u << (pc & 31));
      if (p < P) pos[(size_t)p * n + cid] = i;
    }
  }
}

__global__ void k_scaffg_check(const int32_t* __restrict__ perm, int32_t P, int32_t n,
                               const int32_t* __restrict__ pos, int* __restrict__ flag) {
  int64_t total = (int64_t)P * n;
  int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; idx < total; idx += stride) {
    const int32_t p = (int32_t)(idx / n), i = (int32_t)(idx - (int64_t)p * n);
    const int32_t cid = perm[idx];
    if ((uint32_t)cid >= (uint32_t)n) continue;
    if (pos[(size_t)p * n + cid] != i) atomicOr(flag, 2);
  }
}

// 64-bit layout for the API's integer check (lxy_scaffg_layout): one warp per candidate.
__global__ void k_scaffg_layout64(const int32_t* __restrict__ perm, const uint32_t* __restrict__ len,
                                  int32_t P, int32_t n, uint64_t* __restrict__ start_by_pos,
                                  int32_t* __restrict__ pos_by_contig, int* __restrict__ flag) {
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= P) return;
  uint64_t carry = 0;
  for (int i0 = 0; i0 < n; i0 += 32) {
    const int i = i0 + lane;
    int32_t cid = -1;
    uint64_t l = 0;
    if (i < n) {
      cid = perm[(size_t)p * n + i];
      if ((uint32_t)cid >= (uint32_t)n) { atomicOr(flag, 1); cid = -1; }
      else l = len[cid];
    }
    uint64_t s = l;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t v = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += v;
    }
    if (i < n) start_by_pos[(size_t)p * n + i] = carry + s - l;
    if (cid >= 0) pos_by_contig[(size_t)p * n + cid] = i;
    carry += __shfl_sync(0xffffffffu, s, 31);
  }
}

__global__ void k_scaffg_link_dist(int64_t L, const uint32_t* __restrict__ a, const uint32_t* __restrict__ xa,
                                   const uint32_t* __restrict__ b, const uint32_t* __restrict__ xb,
                                   const uint32_t* __restrict__ len, const uint64_t* __restrict__ start,
                                   const int32_t* __restrict__ pos, const uint8_t* __restrict__ ori,
                                   int32_t n, uint64_t* __restrict__ dist, int* __restrict__ flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < L; i += stride) {
    const uint32_t x = a[i], y = b[i];
    if (x >= (uint32_t)n || y >= (uint32_t)n) { atomicOr(flag, 1); continue; }
    const int32_t pa = pos[x], pb = pos[y];
    const bool ra = ori && ori[pa], rb = ori && ori[pb];
    const uint64_t ca = start[pa] + (ra ? (uint64_t)len[x] - 1 - xa[i] : (uint64_t)xa[i]);
    const uint64_t cb = start[pb] + (rb ? (uint64_t)len[y] - 1 - xb[i] : (uint64_t)xb[i]);
    dist[i] = ca > cb ? ca - cb : cb - ca;
  }
}

// ---- score ----------------------------------------------------------------------------------
struct GParams {
  const uint16_t* rec;
  const uint32_t* xa;
  const uint32_t* xb;
  const void* w;
  const int64_t* tile_start;
  const uint32_t* origin;     // [groups][n][64]
  const uint32_t* bits;       // [groups][n][4]
  void* partial;              // [units][2][64]: S then H
  int64_t L;
  int32_t n, nblk;
  int32_t groups, splits;
  uint32_t unit;
  uint32_t m_lo, m_hi;        // floor(2^64 / unit) + 1 (general units)
  uint32_t shift;             // log2(unit) when unit is a power of two
};

template <int WM>
struct GLane {
  int32_t s32[2], h32[2];
  int64_t s64[2], h64[2];
  double sd[2], hd[2];
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int j = 0; j < 2; ++j) { s32[j] = h32[j] = 0; s64[j] = h64[j] = 0; sd[j] = hd[j] = 0.0; }
  }
  __device__ __forceinline__ void flush() {
    if (WM == W_UNIT) {
#pragma unroll
      for (int j = 0; j < 2; ++j) { s64[j] += s32[j]; h64[j] += h32[j]; s32[j] = h32[j] = 0; }
    }
  }
};

// Row A of a record (or of a whole packet when all its records share one a): the two origins of
// this lane's candidates and the bit words of the row.
struct GRowA {
  uint32_t o[2], rev[2], first[2];
};

template <bool SGN>
__device__ __forceinline__ void g_load_row(const uint8_t* tab, const uint8_t* bits, uint32_t loc, int lane,
                                           GRowA& R) {
  const uint8_t* row = tab + loc * kRowG + lane * 4;
  R.o[0] = *(const uint32_t*)row;
  R.o[1] = *(const uint32_t*)(row + 128);
  if (SGN) {  // orientation lives in the sign of the origin: only the "first" words are needed
    const uint2 f = *(const uint2*)(bits + loc * 16 + 8);   // warp-uniform address: broadcast
    R.first[0] = f.x; R.first[1] = f.y; R.rev[0] = R.rev[1] = 0;
  } else {
    const uint4 b = *(const uint4*)(bits + loc * 16);
    R.rev[0] = b.x; R.rev[1] = b.y; R.first[0] = b.z; R.first[1] = b.w;
  }
}

// One record against the two candidates of this lane.
//   SGN  (total length < 2^31): the table holds v = start (forward) or -(start + len - 1) (reversed),
//        so coord = |v + x| and d = |coord a - coord b|: four IADD/IABS, no orientation bits.
//   !SGN (total length < 2^32): unsigned origins, orientation from the bit rows.
//   POW2: unit is a power of two -> bucket = d >> shift; otherwise bucket = (d * M) >> 64 with
//        M = floor(2^64 / unit) + 1 (two IMAD.HI.U32), exact for every 32-bit d.
template <int WM, bool SGN, bool POW2>
__device__ __forceinline__ void g_eval(const GRowA& A, const GRowA& B, uint32_t xa, uint32_t xb, int32_t wi,
                                       double wd, const uint8_t* lut, int lane, const GParams& prm,
                                       GLane<WM>& L) {
  uint32_t t[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    uint32_t d;
    if (SGN) {
      const int ca = abs((int)A.o[j] + (int)xa), cb = abs((int)B.o[j] + (int)xb);
      d = (uint32_t)abs(ca - cb);
    } else {
      const uint32_t ca = ((A.rev[j] >> lane) & 1u) ? A.o[j] - xa : A.o[j] + xa;
      const uint32_t cb = ((B.rev[j] >> lane) & 1u) ? B.o[j] - xb : B.o[j] + xb;
      d = ca > cb ? ca - cb : cb - ca;
    }
    uint32_t q;
    if (POW2) {
      q = d >> prm.shift;
    } else {
      const uint64_t x = (uint64_t)d * prm.m_lo;
      const uint64_t y = (uint64_t)d * prm.m_hi + (x >> 32);
      q = (uint32_t)(y >> 32);
    }
    t[j] = lut[min(q, 21u)];
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    if (WM == W_UNIT) L.s32[j] += (int32_t)t[j];
    else if (WM == W_INT32) L.s64[j] += (int64_t)(int32_t)t[j] * (int64_t)wi;
    else L.sd[j] += (double)t[j] * wd;
  }
  const uint32_t f0 = A.first[0] | B.first[0], f1 = A.first[1] | B.first[1];
  if ((f0 | f1) != 0) {   // rare and warp-uniform: some candidate of the group has this contig first
    const uint32_t f[2] = {f0, f1};
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      if ((f[j] >> lane) & 1u) {
        if (WM == W_UNIT) L.h32[j] += (int32_t)t[j];
        else if (WM == W_INT32) L.h64[j] += (int64_t)(int32_t)t[j] * (int64_t)wi;
        else L.hd[j] += (double)t[j] * wd;
      }
    }
  }
}

struct GPacket {
  uint4 rec, xa0, xa1, xb0, xb1;
};
__device__ __forceinline__ GPacket g_load_packet(const GParams& prm, int64_t base) {
  GPacket p;
  p.rec = ldg_stream_v4(prm.rec + base);
  p.xa0 = ldg_stream_v4(prm.xa + base); p.xa1 = ldg_stream_v4(prm.xa + base + 4);
  p.xb0 = ldg_stream_v4(prm.xb + base); p.xb1 = ldg_stream_v4(prm.xb + base + 4);
  return p;
}

// Records k in [k0, k1) of one packet; SAME_A: one a for the whole packet (sorted records).
template <int WM, bool SGN, bool POW2, bool SAME_A>
__device__ __forceinline__ void g_eval_packet(const GPacket& pk, const int32_t* wi, const double* wd, int k0,
                                              int k1, const uint8_t* sA, const uint8_t* sB,
                                              const uint8_t* sBitsA, const uint8_t* sBitsB, const uint8_t* lut,
                                              int lane, const GParams& prm, GLane<WM>& L) {
  const uint32_t words[4] = {pk.rec.x, pk.rec.y, pk.rec.z, pk.rec.w};
  const uint32_t xav[8] = {pk.xa0.x, pk.xa0.y, pk.xa0.z, pk.xa0.w, pk.xa1.x, pk.xa1.y, pk.xa1.z, pk.xa1.w};
  const uint32_t xbv[8] = {pk.xb0.x, pk.xb0.y, pk.xb0.z, pk.xb0.w, pk.xb1.x, pk.xb1.y, pk.xb1.z, pk.xb1.w};
  GRowA A, B;
  if (SAME_A) g_load_row<SGN>(sA, sBitsA, (words[0] >> 8) & 0xFFu, lane, A);
#pragma unroll
  for (int k = 0; k < kPacket; ++k) {
    if (k < k0 || k >= k1) continue;
    const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
    if (!SAME_A) g_load_row<SGN>(sA, sBitsA, r >> 8, lane, A);
    g_load_row<SGN>(sB, sBitsB, r & 0xFFu, lane, B);
    g_eval<WM, SGN, POW2>(A, B, xav[k], xbv[k], WM == W_INT32 ? wi[k] : 0, WM == W_F64 ? wd[k] : 0.0, lut,
                          lane, prm, L);
  }
}

template <int WM>
__device__ __forceinline__ void g_load_weights(const void* w, int64_t base, int32_t (&wi)[kPacket],
                                               double (&wd)[kPacket]) {
  if (WM == W_INT32) {
    const uint4* wp = reinterpret_cast<const uint4*>((const int32_t*)w + base);
    const uint4 x = __ldg(wp), y = __ldg(wp + 1);
    wi[0] = x.x; wi[1] = x.y; wi[2] = x.z; wi[3] = x.w; wi[4] = y.x; wi[5] = y.y; wi[6] = y.z; wi[7] = y.w;
  } else if (WM == W_F64) {
    const double2* wp = reinterpret_cast<const double2*>((const double*)w + base);
#pragma unroll
    for (int q = 0; q < 4; ++q) { const double2 d2 = __ldg(wp + q); wd[2 * q] = d2.x; wd[2 * q + 1] = d2.y; }
  }
}

template <int WM, bool SGN, bool POW2>
__global__ void __launch_bounds__(kThreads, 1) k_scaffg_score(GParams prm) {
  typedef typename Acc<WM>::type acc_t;
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + kTabG;
  uint8_t* sBitsA = smem + 2 * kTabG;
  uint8_t* sBitsB = sBitsA + kBitsG;
  uint8_t* s_lut = sBitsB + kBitsG;
  uint64_t* bar = (uint64_t*)(s_lut + 32);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < kLutG) s_lut[tid] = c_lutg[tid];
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t parity = 0;

  const int units = prm.groups * prm.splits;
  const int64_t npk_all = (prm.L + kPacket - 1) / kPacket;
  const int64_t ntiles = (int64_t)prm.nblk * (prm.nblk + 1) / 2;
  for (int u = blockIdx.x; u < units; u += gridDim.x) {
    const int g = u / prm.splits, s = u - g * prm.splits;
    const int64_t lo = (npk_all * s / prm.splits) * kPacket;
    const int64_t hi = min((npk_all * (s + 1) / prm.splits) * kPacket, prm.L);
    GLane<WM> acc;
    acc.clear();
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
        __syncthreads();
        if (tid == 0) {
          const int rowsA = min(kBlk, prm.n - I * kBlk), rowsB = min(kBlk, prm.n - J * kBlk);
          const size_t ra = (size_t)g * prm.n + (size_t)I * kBlk, rb = (size_t)g * prm.n + (size_t)J * kBlk;
          fence_proxy_async();
          mbar_expect_tx(bar, (uint32_t)(rowsA + rowsB) * (kRowG + 16));
          tma_bulk_g2s(sA, prm.origin + ra * kCandG, (uint32_t)rowsA * kRowG, bar);
          tma_bulk_g2s(sB, prm.origin + rb * kCandG, (uint32_t)rowsB * kRowG, bar);
          tma_bulk_g2s(sBitsA, prm.bits + ra * 4, (uint32_t)rowsA * 16, bar);
          tma_bulk_g2s(sBitsB, prm.bits + rb * 4, (uint32_t)rowsB * 16, bar);
        }
        mbar_wait(bar, parity);
        parity ^= 1;

        const int64_t p_first = cur / kPacket, p_last = (seg_end - 1) / kPacket;
        const bool head_partial = (p_first * kPacket < cur);
        const bool tail_partial = ((p_last + 1) * kPacket > seg_end);
        int32_t wi[kPacket];
        double wd[kPacket];
        if (head_partial && warp == 0) {   // boundary packets: masked, one warp each
          const int64_t base = p_first * kPacket;
          const GPacket pk = g_load_packet(prm, base);
          g_load_weights<WM>(prm.w, base, wi, wd);
          const int k0 = (int)(cur - base), k1 = (int)min((int64_t)kPacket, seg_end - base);
#pragma unroll 1
          for (int k = k0; k < k1; ++k)
            g_eval_packet<WM, SGN, POW2, false>(pk, wi, wd, k, k + 1, sA, sB, sBitsA, sBitsB, s_lut, lane, prm, acc);
        }
        if (tail_partial && warp == 1 && !(head_partial && p_last == p_first)) {
          const int64_t base = p_last * kPacket;
          const GPacket pk = g_load_packet(prm, base);
          g_load_weights<WM>(prm.w, base, wi, wd);
          const int k1 = (int)(seg_end - base);
#pragma unroll 1
          for (int k = 0; k < k1; ++k)
            g_eval_packet<WM, SGN, POW2, false>(pk, wi, wd, k, k + 1, sA, sB, sBitsA, sBitsB, s_lut, lane, prm, acc);
        }
        const int64_t f0 = p_first + (head_partial ? 1 : 0);
        const int64_t f1 = p_last - (tail_partial ? 1 : 0);  // inclusive
        const int64_t nfull = f1 - f0 + 1;
        if (nfull > 0) {
          for (int64_t c0 = 0; c0 < nfull; c0 += (int64_t)kWarps << 16) {
            const int32_t cnt = (int32_t)min((int64_t)kWarps << 16, nfull - c0);
            const int64_t pbase = (f0 + c0) * kPacket;
            int32_t it = warp;
            GPacket nx;
            if (it < cnt) nx = g_load_packet(prm, pbase + (int64_t)it * kPacket);
#pragma unroll 1
            for (; it < cnt; it += kWarps) {
              const GPacket pk = nx;
              if (it + kWarps < cnt) nx = g_load_packet(prm, pbase + (int64_t)(it + kWarps) * kPacket);
              g_load_weights<WM>(prm.w, pbase + (int64_t)it * kPacket, wi, wd);
              if (((pk.rec.x >> 8) & 0xFFu) == (pk.rec.w >> 24))
                g_eval_packet<WM, SGN, POW2, true>(pk, wi, wd, 0, kPacket, sA, sB, sBitsA, sBitsB, s_lut, lane, prm, acc);
              else
                g_eval_packet<WM, SGN, POW2, false>(pk, wi, wd, 0, kPacket, sA, sB, sBitsA, sBitsB, s_lut, lane, prm, acc);
            }
            acc.flush();
          }
        }
        cur = seg_end;
      }
    }
    acc.flush();
    __syncthreads();
    acc_t* red = (acc_t*)smem;  // [kWarps][2][64]
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      acc_t sv, hv;
      if (WM == W_F64) { sv = (acc_t)acc.sd[j]; hv = (acc_t)acc.hd[j]; }
      else { sv = (acc_t)acc.s64[j]; hv = (acc_t)acc.h64[j]; }
      red[(warp * 2 + 0) * kCandG + lane + 32 * j] = sv;
      red[(warp * 2 + 1) * kCandG + lane + 32 * j] = hv;
    }
    __syncthreads();
    if (tid < 2 * kCandG) {
      acc_t sum = 0;
      for (int wq = 0; wq < kWarps; ++wq) sum += red[wq * 2 * kCandG + tid];
      ((acc_t*)prm.partial)[(size_t)u * 2 * kCandG + tid] = sum;
    }
  }
}

template <int WM>
__global__ void k_scaffg_finish(int32_t P, int32_t splits, const void* __restrict__ partial,
                                double* __restrict__ out, int64_t* __restrict__ x100) {
  typedef typename Acc<WM>::type acc_t;
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= P) return;
  const int g = p / kCandG, pc = p - g * kCandG;
  acc_t S = 0, H = 0;
  for (int s = lane; s < splits; s += 32) {
    const acc_t* row = (const acc_t*)partial + ((size_t)g * splits + s) * 2 * kCandG;
    S += row[pc];
    H += row[kCandG + pc];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    if (WM == W_F64) {
      S += (acc_t)__shfl_xor_sync(0xffffffffu, (double)S, o);
      H += (acc_t)__shfl_xor_sync(0xffffffffu, (double)H, o);
    } else {
      S += (acc_t)__shfl_xor_sync(0xffffffffu, (long long)S, o);
      H += (acc_t)__shfl_xor_sync(0xffffffffu, (long long)H, o);
    }
  }
  if (lane != 0) return;
  const acc_t tot = S + S - H;
  if (WM == W_F64) {
    out[p] = -((double)tot / 100.0);
    if (x100) x100[p] = INT64_MIN;
  } else {
    out[p] = -((double)(int64_t)tot / 100.0);
    if (x100) x100[p] = (int64_t)tot;
  }
}

template <int WM, bool SGN, bool POW2>
int launch_g3(lxy_ctx* c, const GParams& prm, int grid) {
  constexpr int smem = 2 * kTabG + 2 * kBitsG + 32 + 16;
  static bool configured[64] = {};
  std::lock_guard<std::mutex> cfg_lock(config_mutex());
  if (!configured[c->cfg.device & 63]) {
    LXY_CUDA(cudaFuncSetAttribute(k_scaffg_score<WM, SGN, POW2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured[c->cfg.device & 63] = true;
  }
  k_scaffg_score<WM, SGN, POW2><<<grid, kThreads, smem, c->stream>>>(prm);
  LXY_CUDA(cudaGetLastError());
  return LXY_OK;
}
#include "scaffg2.cuh"

template <int WM>
int launch_g(lxy_ctx* c, const GParams& prm, int grid, bool sgn, bool pow2) {
  if (sgn) return pow2 ? launch_g3<WM, true, true>(c, prm, grid) : launch_g3<WM, true, false>(c, prm, grid);
  return pow2 ? launch_g3<WM, false, true>(c, prm, grid) : launch_g3<WM, false, false>(c, prm, grid);
}

}  // namespace

// Unit-weight records, signed 32-bit coordinates, a unit the one-multiply division covers: scaffg2.cuh
static int scaffg2_score(lxy_ctx* c, const int32_t* perm_dev, const uint8_t* ori_dev, int32_t P, double* out_dev,
                         int64_t* x100_dev, uint32_t K, uint32_t m, uint32_t sh) {
  ScaffLinks& S = c->scaff;
  cudaStream_t st = c->stream;
  const int n = S.n;
  const int groups = (P + kCand2 - 1) / kCand2;
  const int Ppad = groups * kCand2;
  LXY_TRY(c->d_posA.reserve(4 * (size_t)Ppad * n));           // origin table
  LXY_TRY(c->d_posB.reserve(32 * (size_t)groups * n + 32));   // first-bit rows
  LXY_TRY(c->d_start.reserve(4 * (size_t)P * n));             // pos (validation)
  LXY_TRY(c->d_flag.reserve(64));
  LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, st));
  LXY_CUDA(cudaEventRecord(c->ev[0], st));
  LXY_CUDA(cudaMemsetAsync(c->d_posB.p, 0, 32 * (size_t)groups * n, st));
  int launches = 0;
  k_scaffg2_layout<<<(Ppad * 32 + 255) / 256, 256, 0, st>>>(perm_dev, ori_dev, S.len.as<uint32_t>(), P, Ppad, n,
                                                           c->d_posA.as<int32_t>(), c->d_posB.as<uint32_t>(),
                                                           c->d_start.as<int32_t>(), c->d_flag.as<int>());
  ++launches;
  if (!(c->cfg.flags & LXY_FLAG_NO_VALIDATE)) {
    k_scaffg_check<<<grid_for((int64_t)P * n, 256, c->sm_count), 256, 0, st>>>(perm_dev, P, n, c->d_start.as<int32_t>(), c->d_flag.as<int>());
    ++launches;
  }
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(c->ev[1], st));
  const int64_t npk = (S.L + kPacket - 1) / kPacket;
  int splits = (2 * c->sm_count) / groups;  // two full waves of one CTA per SM, never a third
  if (splits < 1) splits = 1;
  if ((int64_t)splits * 64 > npk) splits = (int)(npk / 64 > 0 ? npk / 64 : 1);
  // a warp keeps int32 sums per unit: at most 2^31 / 100 records per warp and unit (kept far below)
  while (S.L / splits > ((int64_t)1 << 28)) splits *= 2;
  const int units = groups * splits;
  const int grid = units < c->sm_count ? units : c->sm_count;
  LXY_TRY(c->d_partial.reserve(8 * (size_t)units * 2 * kCand2));
  G2Params prm;
  prm.rec = S.rec.as<uint16_t>();
  prm.xa = S.xa.as<uint32_t>();
  prm.xb = S.xb.as<uint32_t>();
  prm.tile_start = S.tile_start.as<int64_t>();
  prm.origin = c->d_posA.as<int32_t>();
  prm.firstw = c->d_posB.as<uint32_t>();
  prm.partial = c->d_partial.as<int64_t>();
  prm.L = S.L;
  prm.n = n;
  prm.nblk = S.nblk;
  prm.groups = groups;
  prm.splits = splits;
  prm.K = K; prm.m = m; prm.sh = sh;
  if (S.L > 0) {
    static bool configured[64] = {};
    {
      std::lock_guard<std::mutex> cfg_lock(config_mutex());
      if (!configured[c->cfg.device & 63]) {
        LXY_CUDA(cudaFuncSetAttribute(k_scaffg2r_score<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kG2RSmem));
        LXY_CUDA(cudaFuncSetAttribute(k_scaffg2r_score<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kG2RSmem));
        configured[c->cfg.device & 63] = true;
      }
    }
    prm.shmul = sh ? (1u << (32 - sh)) : 0u;
    if (sh == 0) k_scaffg2r_score<true><<<grid, kG2RThreads, kG2RSmem, st>>>(prm);
    else k_scaffg2r_score<false><<<grid, kG2RThreads, kG2RSmem, st>>>(prm);
    LXY_CUDA(cudaGetLastError());
    ++launches;
  } else {
    LXY_CUDA(cudaMemsetAsync(c->d_partial.p, 0, 8 * (size_t)units * 2 * kCand2, st));
  }
  LXY_CUDA(cudaEventRecord(c->ev[2], st));
  k_scaffg2_finish<<<(P + 7) / 8, 256, 0, st>>>(P, splits, c->d_partial.as<int64_t>(), out_dev, x100_dev);
  ++launches;
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(c->ev[3], st));
  c->timing.launches = launches;
  c->timing.grid = grid;
  c->timing.link_bytes = (int64_t)groups * S.L * 10;
  c->timing.table_bytes = 0;
  return LXY_OK;
}

int scaffg_score(lxy_ctx* c, const int32_t* perm_dev, const uint8_t* ori_dev, int32_t P,
                 double* out_dev, int64_t* x100_dev) {
  ScaffLinks& S = c->scaff;
  if (!S.ready || !S.gprofile) { set_error("scaffg score: G-profile links not set"); return LXY_ERR_STATE; }
  if (P <= 0) return LXY_OK;
  {
    uint32_t K, m, sh;
    static const bool g2_off = getenv("LXY_G2_OFF") != nullptr;   // experiments: the 64-candidate kernel everywhere
    if (!g2_off && S.wmode == W_UNIT && S.total_len < 0x7FFFFFF0ull && g2_magic(S.unit, K, m, sh))
      return scaffg2_score(c, perm_dev, ori_dev, P, out_dev, x100_dev, K, m, sh);
  }
  cudaStream_t st = c->stream;
  const int n = S.n;
  const int groups = (P + kCandG - 1) / kCandG;
  const int Ppad = groups * kCandG;
  LXY_TRY(c->d_posA.reserve(4 * (size_t)Ppad * n));           // origin table
  LXY_TRY(c->d_posB.reserve(16 * (size_t)groups * n));        // bit rows
  LXY_TRY(c->d_start.reserve(4 * (size_t)P * n));             // pos (validation)
  LXY_TRY(c->d_flag.reserve(64));
  LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, st));
  LXY_CUDA(cudaEventRecord(c->ev[0], st));
  LXY_CUDA(cudaMemsetAsync(c->d_posB.p, 0, 16 * (size_t)groups * n, st));
  int launches = 0;
  const bool sgn = S.total_len < 0x7FFFFFF0ull;               // coordinates fit a signed 32-bit word
  const bool pow2 = (S.unit & (S.unit - 1)) == 0;
  k_scaffg_layout<<<(Ppad * 32 + 255) / 256, 256, 0, st>>>(perm_dev, ori_dev, S.len.as<uint32_t>(), P, Ppad, n, sgn ? 1 : 0,
                                                          c->d_posA.as<uint32_t>(), c->d_posB.as<uint32_t>(),
                                                          c->d_start.as<int32_t>(), c->d_flag.as<int>());
  ++launches;
  if (!(c->cfg.flags & LXY_FLAG_NO_VALIDATE)) {
    k_scaffg_check<<<grid_for((int64_t)P * n, 256, c->sm_count), 256, 0, st>>>(perm_dev, P, n, c->d_start.as<int32_t>(), c->d_flag.as<int>());
    ++launches;
  }
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(c->ev[1], st));

  const int64_t npk = (S.L + kPacket - 1) / kPacket;
  int splits = (2 * c->sm_count) / groups;  // two full waves of one CTA per SM, never a third
  if (splits < 1) splits = 1;
  if ((int64_t)splits * 64 > npk) splits = (int)(npk / 64 > 0 ? npk / 64 : 1);
  const int units = groups * splits;
  const int grid = units < c->sm_count ? units : c->sm_count;
  LXY_TRY(c->d_partial.reserve(8 * (size_t)units * 2 * kCandG));
  GParams prm;
  prm.rec = S.rec.as<uint16_t>();
  prm.xa = S.xa.as<uint32_t>();
  prm.xb = S.xb.as<uint32_t>();
  prm.w = S.w.p;
  prm.tile_start = S.tile_start.as<int64_t>();
  prm.origin = c->d_posA.as<uint32_t>();
  prm.bits = c->d_posB.as<uint32_t>();
  prm.partial = c->d_partial.p;
  prm.L = S.L;
  prm.n = n;
  prm.nblk = S.nblk;
  prm.groups = groups;
  prm.splits = splits;
  prm.unit = S.unit;
  uint64_t M = 0;
  if (S.unit > 1) M = (~(uint64_t)0) / S.unit + 1;  // floor((2^64 - 1) / u) + 1 == floor(2^64 / u) + 1 unless u | 2^64
  if (S.unit > 1 && (S.unit & (S.unit - 1)) == 0) M = ((uint64_t)1 << 63) / (S.unit >> 1);  // power of two: exact 2^64 / u
  prm.m_lo = (uint32_t)M;
  prm.m_hi = (uint32_t)(M >> 32);
  prm.shift = 0;
  while (pow2 && (1u << prm.shift) < S.unit) ++prm.shift;
  if (S.L > 0) {
    int rc = S.wmode == W_UNIT ? launch_g<W_UNIT>(c, prm, grid, sgn, pow2)
           : S.wmode == W_INT32 ? launch_g<W_INT32>(c, prm, grid, sgn, pow2)
                                : launch_g<W_F64>(c, prm, grid, sgn, pow2);
    LXY_TRY(rc);
    ++launches;
  } else {
    LXY_CUDA(cudaMemsetAsync(c->d_partial.p, 0, 8 * (size_t)units * 2 * kCandG, st));
  }
  LXY_CUDA(cudaEventRecord(c->ev[2], st));
  if (S.wmode == W_UNIT) k_scaffg_finish<W_UNIT><<<(P + 7) / 8, 256, 0, st>>>(P, splits, c->d_partial.p, out_dev, x100_dev);
  else if (S.wmode == W_INT32) k_scaffg_finish<W_INT32><<<(P + 7) / 8, 256, 0, st>>>(P, splits, c->d_partial.p, out_dev, x100_dev);
  else k_scaffg_finish<W_F64><<<(P + 7) / 8, 256, 0, st>>>(P, splits, c->d_partial.p, out_dev, x100_dev);
  ++launches;
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(c->ev[3], st));
  c->timing.launches = launches;
  c->timing.grid = grid;
  c->timing.link_bytes = (int64_t)groups * S.L * (10 + (S.wmode == W_INT32 ? 4 : S.wmode == W_F64 ? 8 : 0));
  c->timing.table_bytes = 0;
  return LXY_OK;
}

int scaffg_layout64(lxy_ctx* c, const int32_t* perm_dev, int32_t P, uint64_t* start_dev, int32_t* pos_dev) {
  ScaffLinks& S = c->scaff;
  if (!S.ready || !S.gprofile) { set_error("scaffg layout: G-profile links not set"); return LXY_ERR_STATE; }
  LXY_TRY(c->d_flag.reserve(64));
  LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, c->stream));
  k_scaffg_layout64<<<(P * 32 + 255) / 256, 256, 0, c->stream>>>(perm_dev, S.len.as<uint32_t>(), P, S.n, start_dev, pos_dev, c->d_flag.as<int>());
  LXY_CUDA(cudaGetLastError());
  return LXY_OK;
}

int scaffg_link_distances(lxy_ctx* c, int64_t L, const uint32_t* a, const uint32_t* xa, const uint32_t* b,
                          const uint32_t* xb, const uint64_t* start_dev, const int32_t* pos_dev,
                          const uint8_t* ori_dev, uint64_t* dist_dev) {
  ScaffLinks& S = c->scaff;
  if (L > 0) {
    k_scaffg_link_dist<<<grid_for(L, 256, c->sm_count), 256, 0, c->stream>>>(L, a, xa, b, xb, S.len.as<uint32_t>(), start_dev, pos_dev, ori_dev, S.n, dist_dev, c->d_flag.as<int>());
    LXY_CUDA(cudaGetLastError());
  }
  return LXY_OK;
}

}  // namespace lxy
