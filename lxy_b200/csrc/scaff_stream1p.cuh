// scaff_stream1p.cuh — the streaming-regime kernel for 2..8 orderings per pass with unit weights:
// the placement tables of TWO orderings are interleaved as one 32-bit word per contig.
// (included by scaff_kernels.cu inside its anonymous namespace, after scaff_stream1.cuh)
//
// Same boundary, mapping (lane = record, per-warp ring of bulk copies, runs of chunks) and integers as
// k_scaff_stream1 (scaff.go:246-310 through DESIGN.md §2); what changes is the cost per evaluation.
//   table word of pair j at contig x:  (-pos_{2j}[x] mod 2^16) | (-pos_{2j+1}[x] mod 2^16) << 16
//   per packet:   A_j = (21 | 21 << 16) - word_j[a]      (per half-word: pos a + 21)
//   per record:   m = VIADDMNMX.U16x2(A_j, word_j[b], 43|43<<16)   -> v = min(pos a - pos b + 21 mod 2^16, 43)
//                 e = lutc[m.lo + 45 * m.hi]  (one DP2A for the address: 100*T[v_lo] | 100*T[v_hi] << 16)
//                 p_j += e                    (two 16-bit sums in one add, flushed every 64 packets)
// i.e. one table look-up, one min-add, one look-up of the decay pair and one add serve TWO orderings
// (k_scaff_stream1: one of each per ordering) - the batch kernel's per-lane arithmetic (eval_record)
// with the roles of lanes and records exchanged.  Exact while n <= 65,515 (a difference cannot alias
// into [0, 42] modulo 2^16); larger n and weighted records stay on k_scaff_stream1.
#pragma once

constexpr uint32_t kS1pCap = (uint32_t)(kMiss + 2) | ((uint32_t)(kMiss + 2) << 16);   // 43 | 43 << 16
constexpr uint32_t kS1pBias = (uint32_t)(kBias + 1) | ((uint32_t)(kBias + 1) << 16);  // 21 | 21 << 16
constexpr uint32_t kS1pDp = 4u | ((4u * kLut2Stride) << 8);                             // byte strides of lutc
constexpr int kS1pFlush = 64;  // packets between flushes of the packed sums: 64 * 8 * 100 < 2^16

__device__ __forceinline__ uint32_t s1p_pair(uint32_t A, uint32_t Bw, uint32_t lutc) {
  const uint32_t m = __viaddmin_u16x2(A, Bw, kS1pCap);
  return lds_u32(__dp2a_lo(m, kS1pDp, lutc));
}

// Eight records, one a (see s1_packet_same_a).
template <int NP>
__device__ __forceinline__ void s1p_packet_same_a(const uint4 v, uint32_t baseA, uint32_t baseB, uint32_t pstride,
                                                  uint32_t lutc, uint32_t (&p32)[NP]) {
  const uint32_t words[4] = {v.x, v.y, v.z, v.w};
  const uint32_t oa = (v.x >> 6) & 0x3FCu;
  uint32_t A[NP];
#pragma unroll
  for (int j = 0; j < NP; ++j) A[j] = __vsub2(kS1pBias, lds_u32((baseA + j * pstride) | oa));
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t o0 = (words[q] << 2) & 0x3FCu;
    const uint32_t b1 = __byte_perm(words[q], 0u, 0x4442u);   // PRMT + IMAD instead of SHF + LOP3 (see scaff_stream1.cuh)
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const uint32_t bb = baseB + j * pstride;
      p32[j] += s1p_pair(A[j], lds_u32(bb | o0), lutc);
      p32[j] += s1p_pair(A[j], lds_u32(b1 * c_s1_mul[5] + bb), lutc);
    }
  }
}

// Eight records whose a's are two neighbouring ids (see s1_packet_two_a).
template <int NP>
__device__ __forceinline__ void s1p_packet_two_a(const uint4 v, uint32_t baseA, uint32_t baseB, uint32_t pstride,
                                                 uint32_t lutc, uint32_t (&p32)[NP]) {
  const uint32_t words[4] = {v.x, v.y, v.z, v.w};
  const uint32_t oa0 = (v.x >> 6) & 0x3FCu, oa1 = (v.w >> 22) & 0x3FCu;
  const uint32_t rep = (v.x & 0xFF00u) * 0x00010001u;   // the first a at bits 8..15 and 24..31
  uint32_t A0[NP], A1[NP];
#pragma unroll
  for (int j = 0; j < NP; ++j) {
    A0[j] = __vsub2(kS1pBias, lds_u32((baseA + j * pstride) | oa0));
    A1[j] = __vsub2(kS1pBias, lds_u32((baseA + j * pstride) | oa1));
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t x = words[q] ^ rep;
    const bool f0 = (x & 0xFF00u) == 0u, f1 = (x & 0xFF000000u) == 0u;
    const uint32_t o0 = (words[q] << 2) & 0x3FCu;
    const uint32_t b1 = __byte_perm(words[q], 0u, 0x4442u);
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const uint32_t bb = baseB + j * pstride;
      p32[j] += s1p_pair(f0 ? A0[j] : A1[j], lds_u32(bb | o0), lutc);
      p32[j] += s1p_pair(f1 ? A0[j] : A1[j], lds_u32(b1 * c_s1_mul[5] + bb), lutc);
    }
  }
}

// One record (boundary packets, packets with three or more a's): unpacked sums.
template <int NP>
__device__ __forceinline__ void s1p_eval(uint32_t baseA, uint32_t baseB, uint32_t pstride, const uint8_t* lut,
                                         uint32_t r, int32_t (&a32)[2 * NP]) {
  const uint32_t oa = (r >> 6) & 0x3FCu, ob = (r << 2) & 0x3FCu;
#pragma unroll
  for (int j = 0; j < NP; ++j) {
    const uint32_t A = __vsub2(kS1pBias, lds_u32((baseA + j * pstride) | oa));
    const uint32_t m = __viaddmin_u16x2(A, lds_u32((baseB + j * pstride) | ob), kS1pCap);
    a32[2 * j] += lut[m & 0xFFFFu];
    a32[2 * j + 1] += lut[m >> 16];
  }
}

template <int NP>
__device__ __forceinline__ void s1p_flush(uint32_t (&p32)[NP], int32_t (&a32)[2 * NP]) {
#pragma unroll
  for (int j = 0; j < NP; ++j) {
    a32[2 * j] += (int32_t)(p32[j] & 0xFFFFu);
    a32[2 * j + 1] += (int32_t)(p32[j] >> 16);
    p32[j] = 0;
  }
}

// The decay table (64 bytes) and the pair table lutc[v1][v0] = 100 T[v0] | 100 T[v1] << 16, built at
// compile time and copied into shared memory as 16-byte pieces requested with the other prologue loads.
struct S1pLuts { uint8_t lut[64]; uint32_t lutc[kLut2Words]; };
constexpr S1pLuts s1p_make_luts() {
  S1pLuts t{};
  for (int i = 0; i < kLutLen; ++i) t.lut[i] = s1_lut100_at(i);
  for (int i = 0; i < kLut2Words; ++i) {
    const int v0 = i % kLut2Stride, v1 = i / kLut2Stride;
    t.lutc[i] = (v0 < kLutLen ? (uint32_t)s1_lut100_at(v0) : 0u) | ((uint32_t)s1_lut100_at(v1) << 16);
  }
  return t;
}
__device__ const S1pLuts g_s1p_luts = s1p_make_luts();
constexpr int kS1PreP = 5120 / kS1Threads;                    // genes per thread and ordering requested ahead of the ring
constexpr int kS1pLutPieces = (int)(sizeof(S1pLuts) / 16);   // 544
static_assert(sizeof(S1pLuts) % 16 == 0 && kS1pLutPieces <= 2 * kS1Threads, "lut pieces");

__device__ __forceinline__ uint32_t s1_atom_inc(uint32_t addr_smem) {
  uint32_t old;
  asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(addr_smem) : "memory");
  return old;
}

template <int NC>
__global__ void __launch_bounds__(kS1Threads + (kS1MinBlocks == 2 && NC <= 2 ? 64 : 0), kS1MinBlocks) k_scaff_stream1p(Stream1Params prm) {
  typedef typename Acc<W_UNIT>::type acc_t;
  constexpr int NP = NC / 2;
  extern __shared__ __align__(128) uint8_t smem_raw1p[];
  // contig blocks of the table (256 words) must start on 1024-byte boundaries of the shared window
  uint8_t* smem = smem_raw1p + ((1024u - (smem_u32(smem_raw1p) & 1023u)) & 1023u);
  const int32_t n = prm.n;
  const int32_t nstride = (n + kBlk - 1) & ~(kBlk - 1);
  uint16_t* sNeg = (uint16_t*)smem;                          // [NP][nstride][2]: -pos mod 2^16 of orderings 2j, 2j+1
  uint8_t* s_lut = smem + (size_t)NP * nstride * 4;          // kLutLen bytes
  uint32_t* s_lutc = (uint32_t*)(s_lut + 64);                // [kLutLen][kLut2Stride]: decay values of a pair
  acc_t* s_red = (acc_t*)(s_lutc + kLut2Words);              // [kS1Warps][NC]
  int64_t* s_tile = (int64_t*)(s_red + kS1Warps * NC);
  const uint32_t sbase = smem_u32(sNeg), pstride = (uint32_t)nstride * 4, lutc = smem_u32(s_lutc);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  S1_STAMP(0);
#ifdef LXY_S1_TRACE
  if (threadIdx.x == 0 && blockIdx.x < 1024) { unsigned sm; asm("mov.u32 %0, %smid;" : "=r"(sm)); g_s1_trace[blockIdx.x * 8 + 4] = sm; }
#endif

  // share of the record stream, per-warp ring of 2 KB bulk copies: as in k_scaff_stream1
  const int64_t npk_all = (prm.L + kPacket - 1) / kPacket;
  const int64_t lo = (npk_all * blockIdx.x / gridDim.x) * kPacket;
  const int64_t hi = min((npk_all * (blockIdx.x + 1) / gridDim.x) * kPacket, prm.L);
  constexpr int32_t kSub = 32 * kPacket;
  constexpr int32_t kSlice = kSub * kS1Sub;
  constexpr int32_t kChunk = kS1Warps * kSlice;
  const int32_t nchunks = (int32_t)((hi - lo + kChunk - 1) / kChunk);
  const int32_t stages = prm.stages;
  const uint32_t ring_w = smem_u32(s_tile + 2) + (uint32_t)warp * (uint32_t)(stages * kS1WarpStage);
  const uint32_t bar_w = smem_u32(s_tile + 2) + (uint32_t)(kS1Warps * stages * kS1WarpStage) + (uint32_t)warp * (uint32_t)(stages * 8);
  const uint16_t* rec_w = prm.rec + lo + (int64_t)warp * kSlice;
  const uint32_t s_next = smem_u32(s_tile + 2) + (uint32_t)(kS1Warps * stages * (kS1WarpStage + 8));   // the CTA's slice counter
  const int32_t hi_rel = (int32_t)(hi - lo);
  const int64_t ntiles = (int64_t)prm.nblk * (prm.nblk + 1) / 2;
  auto fill_ring = [&]() {
    if (lane == 0) {
      for (int q = 0; q < stages; ++q)
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_w + q * 8) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      for (int q = 0; q < stages; ++q)
        if (warp * kSlice + (int64_t)q * kChunk < hi_rel)
          s1_bulk_fill(ring_w + q * kS1WarpStage, rec_w + (int64_t)q * kChunk, bar_w + q * 8);
      if (warp == 0) asm volatile("st.shared.u32 [%0], %1;" ::"r"(s_next), "r"(kS1Warps * stages) : "memory");
    }
    __syncwarp();
  };

  // ---- prologue loads ahead of the record copies (see k_scaff_stream1): with two orderings the first
  // kS1Pre genes per thread of both; with more the tables are built before the ring is filled ----
  constexpr int kPreC = NC <= 2 ? NC : 0;
  int32_t pv[kPreC > 0 ? kPreC : 1][kS1PreP];
#pragma unroll
  for (int c = 0; c < kPreC; ++c)
#pragma unroll
    for (int k = 0; k < kS1PreP; ++k) {
      const int i = tid + k * kS1Threads;
      pv[c][k] = (i < n && c < prm.P) ? ldg_s32_ordered(prm.perm + (size_t)c * n + i) : i;
    }
  int64_t ts0 = 0, ts1 = 0;
  if (tid < ntiles) { ts0 = ldg_s64_ordered(prm.tile_start + tid); ts1 = ldg_s64_ordered(prm.tile_start + tid + 1); }
  uint4 lutv0 = ldg_v4_ordered(reinterpret_cast<const uint8_t*>(&g_s1p_luts) + tid * 16), lutv1 = make_uint4(0, 0, 0, 0);
  if (tid + kS1Threads < kS1pLutPieces) lutv1 = ldg_v4_ordered(reinterpret_cast<const uint8_t*>(&g_s1p_luts) + (tid + kS1Threads) * 16);
  if (kPreC > 0) fill_ring();
  S1_STAMP(5);
  if (lo < hi) {
    if (tid < ntiles && ts0 <= lo && lo < ts1) { s_tile[0] = tid; s_tile[1] = ts1; }
    for (int64_t t = tid + kS1Threads; t < ntiles; t += kS1Threads)
      if (prm.tile_start[t] <= lo) {
        const int64_t e = prm.tile_start[t + 1];
        if (lo < e) { s_tile[0] = t; s_tile[1] = e; }
      }
  }
  reinterpret_cast<uint4*>(s_lut)[tid] = lutv0;                                   // decay table | pair table
  if (tid + kS1Threads < kS1pLutPieces) reinterpret_cast<uint4*>(s_lut)[tid + kS1Threads] = lutv1;
  S1_STAMP(6);
  // ---- placement tables: half-word (c & 1) of word [c >> 1][perm_c[i]] = -i (pad orderings: identity) ----
  int bad = 0;
#pragma unroll
  for (int c = 0; c < kPreC; ++c)
#pragma unroll
    for (int k = 0; k < kS1PreP; ++k) {
      const int i = tid + k * kS1Threads;
      if (i < n) {
        if ((uint32_t)pv[c][k] >= (uint32_t)n) bad |= 1;
        else sNeg[((size_t)(c >> 1) * nstride + pv[c][k]) * 2 + (c & 1)] = (uint16_t)(0 - i);
      }
    }
  if (kPreC > 0) {
#pragma unroll 1
    for (int c = 0; c < NC; ++c) {   // what the requests above did not cover (n > 5,120)
      const int32_t* row = prm.perm + (size_t)c * n;
      for (int i = tid + kS1PreP * kS1Threads; i < n; i += kS1Threads) {
        const int32_t cid = (c < prm.P) ? row[i] : i;
        if ((uint32_t)cid >= (uint32_t)n) bad |= 1;
        else sNeg[((size_t)(c >> 1) * nstride + cid) * 2 + (c & 1)] = (uint16_t)(0 - i);
      }
    }
  } else {
    // all orderings as ONE sequence of (ordering, gene) items, 16 loads in flight per thread: ordering by
    // ordering the loop waited one memory round trip per ordering and four genes (12 us for four orderings)
    constexpr int kB = 16;
    const int K = (n + kS1Threads - 1) / kS1Threads, total = NC * K;
    int c_ld = 0, k_ld = 0, c_st = 0, k_st = 0;
#pragma unroll 1
    for (int q0 = 0; q0 < total; q0 += kB) {
      int32_t buf[kB];
#pragma unroll
      for (int j = 0; j < kB; ++j) {
        const int i = tid + k_ld * kS1Threads;
        buf[j] = (q0 + j < total && i < n && c_ld < prm.P) ? ldg_s32_ordered(prm.perm + (size_t)c_ld * n + i) : i;
        if (++k_ld == K) { k_ld = 0; ++c_ld; }
      }
#pragma unroll
      for (int j = 0; j < kB; ++j) {
        const int i = tid + k_st * kS1Threads;
        if (q0 + j < total && i < n) {
          if ((uint32_t)buf[j] >= (uint32_t)n) bad |= 1;
          else sNeg[((size_t)(c_st >> 1) * nstride + buf[j]) * 2 + (c_st & 1)] = (uint16_t)(0 - i);
        }
        if (++k_st == K) { k_st = 0; ++c_st; }
      }
    }
    fill_ring();
  }
  S1_STAMP(7);
  __syncthreads();
  // CTA k checks the contig IDS of slice k: the position p decoded from the table entry of id x must be in
  // range with perm[p] == x.  Independent of which store to a duplicated id won (see k_scaff_stream1): a
  // duplicate leaves another id unwritten, and an unwritten slot cannot point at its own id.  The gathers
  // perm[p] of the first id per thread are requested for all orderings at once; with two orderings they
  // stay in flight while the records are scored.
  const int32_t i0 = (int32_t)((int64_t)n * blockIdx.x / gridDim.x), i1 = (int32_t)((int64_t)n * (blockIdx.x + 1) / gridDim.x);
  int32_t vgot[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) vgot[c] = i0 + tid;
  if (prm.validate) {
    if (i0 + tid < i1) {
#pragma unroll
      for (int c = 0; c < NC; ++c)
        if (c < prm.P) {
          // the whole word through an opaque 32-bit load, decoded in 32 bits
          const uint32_t word = lds_u32(sbase + (uint32_t)(c >> 1) * pstride + (uint32_t)(i0 + tid) * 4u);
          const uint32_t p = (0u - ((c & 1) ? (word >> 16) : (word & 0xFFFFu))) & 0xFFFFu;
          vgot[c] = p < (uint32_t)n ? prm.perm[(size_t)c * n + p] : -1;
        }
    }
    for (int c = 0; c < prm.P && c < NC; ++c) {
      const int32_t* row = prm.perm + (size_t)c * n;
      for (int x = i0 + tid + kS1Threads; x < i1; x += kS1Threads) {
        const uint32_t word = lds_u32(sbase + (uint32_t)(c >> 1) * pstride + (uint32_t)x * 4u);
        const uint32_t p = (0u - ((c & 1) ? (word >> 16) : (word & 0xFFFFu))) & 0xFFFFu;
        if (p >= (uint32_t)n || row[p] != x) bad |= 2;
      }
    }
    if (NC > 2) {
#pragma unroll
      for (int c = 0; c < NC; ++c) if (vgot[c] != i0 + tid) bad |= 2;
    }
  }
  if (bad) atomicOr(prm.flag, bad);
  S1_STAMP(1);

  int32_t a32[NC];
  uint32_t p32[NP];
#pragma unroll
  for (int c = 0; c < NC; ++c) a32[c] = 0;
#pragma unroll
  for (int j = 0; j < NP; ++j) p32[j] = 0;

  if (lo < hi) {
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
    int32_t w_rel = warp * kSlice;
    int32_t lim_rel = 0;
    uint32_t baseA = sbase, baseB = sbase;
    int32_t st = 0;
    uint32_t ph = 0;
    int32_t since = 0;  // packets added to the packed sums since the last flush
    auto refill = [&]() {
      if (lane == 0 && w_rel + stages * kChunk < hi_rel)
        s1_bulk_fill(ring_w + (uint32_t)st * kS1WarpStage, rec_w + (int64_t)w_rel - (int64_t)warp * kSlice + (int64_t)stages * kChunk,
                     bar_w + (uint32_t)st * 8);
    };
    auto score_packet = [&](const uint4 v) {
      if (__all_sync(0xffffffffu, ((v.x ^ (v.w >> 16)) & 0xFF00u) == 0u)) {
        s1p_packet_same_a<NP>(v, baseA, baseB, pstride, lutc, p32);
      } else if (__all_sync(0xffffffffu, (v.w >> 24) - ((v.x >> 8) & 0xFFu) <= 1u)) {
        s1p_packet_two_a<NP>(v, baseA, baseB, pstride, lutc, p32);
      } else {
        const uint32_t words[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < kPacket; ++k) {
          const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
          s1p_eval<NP>(baseA, baseB, pstride, s_lut, r, a32);
        }
      }
      if (++since == kS1pFlush) {  // warp-uniform
        s1p_flush<NP>(p32, a32);
        since = 0;
      }
    };
    // Eight orderings (one CTA per SM, 16 warps): slices of the share are taken from the CTA's counter, so
    // that all warps finish within one slice of each other (statically assigned the last warp finished up
    // to 14 us after the first: 280 -> 258 us per pass).  With two CTAs per SM the run loop below is 4 us
    // per pass faster than this per-slice loop and the warps' pace evens out by itself.
    constexpr bool kDyn = NC >= 8;
    if (kDyn) {
      // slices of the share taken from the CTA's counter, as in k_scaff_stream1
      const int32_t nslices = (hi_rel + kSlice - 1) / kSlice;
      int32_t q0 = warp, q1 = warp + kS1Warps, q2 = warp + 2 * kS1Warps, q3 = warp + 3 * kS1Warps;
      int32_t newest = warp + (stages - 1) * kS1Warps;
      constexpr bool kDyn = NC >= 8;
      uint32_t claimed = 0;
      auto claim = [&]() {   // asked at the start of a slice, read when the slot is handed back
        if (kDyn && lane == 0) claimed = s1_atom_inc(s_next);
      };
      auto hand_back = [&]() {   // whole warp, converged
        const int32_t nxt = !kDyn ? newest + kS1Warps : (int32_t)__shfl_sync(0xffffffffu, claimed, 0);
        if (lane == 0 && nxt < nslices)
          s1_bulk_fill(ring_w + (uint32_t)st * kS1WarpStage, prm.rec + lo + (int64_t)nxt * kSlice, bar_w + (uint32_t)st * 8);
        newest = nxt;
      };
  #pragma unroll 1
      while (q0 < nslices) {
        claim();
        const int32_t w_rel = q0 * kSlice;
        if (w_rel >= lim_rel) {   // the slice starts in a later tile (or is the first)
          cur.advance_to(lo + w_rel, prm.tile_start, prm.nblk);
          lim_rel = (int32_t)(min(hi, cur.end) - lo);
          baseA = sbase + (uint32_t)cur.I * (kBlk * 4);
          baseB = sbase + (uint32_t)cur.J * (kBlk * 4);
        }
        s1_bar_wait(bar_w + (uint32_t)st * 8, ph);
        const uint32_t slot = ring_w + (uint32_t)st * kS1WarpStage + (uint32_t)lane * 16;
        if (w_rel + kSlice <= lim_rel) {
  #pragma unroll
          for (int j = 0; j < kS1Sub; ++j) {
            const uint4 v = lds_v4(slot + (uint32_t)j * 512);
            const bool one_a = __all_sync(0xffffffffu, ((v.x ^ (v.w >> 16)) & 0xFF00u) == 0u);
            if (j == kS1Sub - 1) hand_back();   // the slot's last packet is in registers (the vote consumed it)
            if (one_a) {
              s1p_packet_same_a<NP>(v, baseA, baseB, pstride, lutc, p32);
            } else if (__all_sync(0xffffffffu, (v.w >> 24) - ((v.x >> 8) & 0xFFu) <= 1u)) {
              s1p_packet_two_a<NP>(v, baseA, baseB, pstride, lutc, p32);
            } else {
              const uint32_t words[4] = {v.x, v.y, v.z, v.w};
  #pragma unroll
              for (int k = 0; k < kPacket; ++k) {
                const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
                s1p_eval<NP>(baseA, baseB, pstride, s_lut, r, a32);
              }
            }
          }
          since += kS1Sub;
          if (since >= kS1pFlush) {  // warp-uniform
            s1p_flush<NP>(p32, a32);
            since = 0;
          }
        } else {  // the slice crosses a tile boundary or the end of the share: packet by packet
  #pragma unroll 1
          for (int j = 0; j < kS1Sub; ++j) {
            const int32_t ws = w_rel + j * kSub;
            if (ws >= hi_rel) break;
            if (ws >= lim_rel) {
              cur.advance_to(lo + ws, prm.tile_start, prm.nblk);
              lim_rel = (int32_t)(min(hi, cur.end) - lo);
              baseA = sbase + (uint32_t)cur.I * (kBlk * 4);
              baseB = sbase + (uint32_t)cur.J * (kBlk * 4);
            }
            const int32_t rr = ws + lane * kPacket;
            const uint4 v = lds_v4(slot + (uint32_t)j * 512);
            if (ws + kSub <= lim_rel) {
              score_packet(v);
            } else if (rr < hi_rel) {
              S1Tile mine = cur;
              const uint32_t words[4] = {v.x, v.y, v.z, v.w};
  #pragma unroll
              for (int k = 0; k < kPacket; ++k) {
                if (rr + k < hi_rel) {
                  mine.advance_to(lo + rr + k, prm.tile_start, prm.nblk);
                  const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
                  s1p_eval<NP>(sbase + (uint32_t)mine.I * (kBlk * 4), sbase + (uint32_t)mine.J * (kBlk * 4), pstride, s_lut,
                               r, a32);
                }
              }
            }
          }
          fence_proxy_async();
          __syncwarp();
          hand_back();
        }
        q0 = q1;
        q1 = stages > 2 ? q2 : newest;
        q2 = stages > 3 ? q3 : newest;
        q3 = newest;
        if (++st == stages) { st = 0; ph ^= 1u; }
      }

    } else {
      int32_t c = 0;
  #pragma unroll 1
      while (c < nchunks) {
        if (w_rel < hi_rel && w_rel + kSlice > lim_rel) {   // next tile (or the first one)
          cur.advance_to(lo + w_rel, prm.tile_start, prm.nblk);
          lim_rel = (int32_t)(min(hi, cur.end) - lo);
          baseA = sbase + (uint32_t)cur.I * (kBlk * 4);
          baseB = sbase + (uint32_t)cur.J * (kBlk * 4);
        }
        int32_t nrun = 0;
        if (w_rel + kSlice <= lim_rel) nrun = min((lim_rel - kSlice - w_rel) / kChunk + 1, nchunks - c);
        if (nrun == 0) {  // the slice crosses a tile boundary or the end of the share: packet by packet
          if (w_rel < hi_rel) {
            s1_bar_wait(bar_w + (uint32_t)st * 8, ph);
            const uint32_t slot = ring_w + (uint32_t)st * kS1WarpStage + (uint32_t)lane * 16;
  #pragma unroll 1
            for (int j = 0; j < kS1Sub; ++j) {
              const int32_t ws = w_rel + j * kSub;
              if (ws >= hi_rel) break;
              if (ws >= lim_rel) {
                cur.advance_to(lo + ws, prm.tile_start, prm.nblk);
                lim_rel = (int32_t)(min(hi, cur.end) - lo);
                baseA = sbase + (uint32_t)cur.I * (kBlk * 4);
                baseB = sbase + (uint32_t)cur.J * (kBlk * 4);
              }
              const int32_t rr = ws + lane * kPacket;
              const uint4 v = lds_v4(slot + (uint32_t)j * 512);
              if (ws + kSub <= lim_rel) {
                score_packet(v);
              } else if (rr < hi_rel) {
                S1Tile mine = cur;
                const uint32_t words[4] = {v.x, v.y, v.z, v.w};
  #pragma unroll
                for (int k = 0; k < kPacket; ++k) {
                  if (rr + k < hi_rel) {
                    mine.advance_to(lo + rr + k, prm.tile_start, prm.nblk);
                    const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
                    s1p_eval<NP>(sbase + (uint32_t)mine.I * (kBlk * 4), sbase + (uint32_t)mine.J * (kBlk * 4), pstride, s_lut,
                                 r, a32);
                  }
                }
              }
            }
            fence_proxy_async();
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
            const bool one_a = __all_sync(0xffffffffu, ((v.x ^ (v.w >> 16)) & 0xFF00u) == 0u);
            if (j == kS1Sub - 1) refill();   // the slot's last packet is in registers (the vote consumed it)
            if (one_a) {
              s1p_packet_same_a<NP>(v, baseA, baseB, pstride, lutc, p32);
            } else if (__all_sync(0xffffffffu, (v.w >> 24) - ((v.x >> 8) & 0xFFu) <= 1u)) {
              s1p_packet_two_a<NP>(v, baseA, baseB, pstride, lutc, p32);
            } else {
              const uint32_t words[4] = {v.x, v.y, v.z, v.w};
  #pragma unroll
              for (int k = 0; k < kPacket; ++k) {
                const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
                s1p_eval<NP>(baseA, baseB, pstride, s_lut, r, a32);
              }
            }
          }
          since += kS1Sub;
          if (since >= kS1pFlush) {  // warp-uniform; kS1pFlush is a multiple of kS1Sub away from overflow
            s1p_flush<NP>(p32, a32);
            since = 0;
          }
          w_rel += kChunk;
          if (++st == stages) { st = 0; ph ^= 1u; }
        }
      }

    }
  }
  S1_STAMP(2);
  s1p_flush<NP>(p32, a32);
  if (NC <= 2 && prm.validate) {
    bool miss = false;
#pragma unroll
    for (int c = 0; c < NC; ++c) miss |= (vgot[c] != i0 + tid);
    if (miss) atomicOr(prm.flag, 2);
  }

  // ---- warp shuffle tree, then a fixed-order sum over the warps ----
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    acc_t v = (acc_t)a32[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += (acc_t)__shfl_xor_sync(0xffffffffu, (long long)v, o);
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

inline size_t stream1p_smem(int nc, int n, int stages = 2) {
  return (size_t)nc * ((n + kBlk - 1) & ~(kBlk - 1)) * 2 + 64 + (size_t)kLut2Words * 4 + (size_t)kS1Warps * nc * 8 + 16 +
         1024 + (size_t)kS1Warps * stages * (kS1WarpStage + 8) + 16;  // (alignment slack) table | lut | pair lut | warp sums | start tile | ring | ring barriers | slice counter
}

template <int NC>
int launch_stream1p(lxy_ctx* c, const Stream1Params& prm, int grid, size_t smem) {
  static size_t configured[64] = {};
  std::lock_guard<std::mutex> cfg_lock(config_mutex());
  if (configured[c->cfg.device & 63] < smem) {
    LXY_CUDA(cudaFuncSetAttribute(k_scaff_stream1p<NC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    LXY_CUDA(cudaFuncSetAttribute(k_scaff_stream1p<NC>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                  cudaSharedmemCarveoutMaxShared));
    configured[c->cfg.device & 63] = smem;
  }
  k_scaff_stream1p<NC><<<grid, kS1Threads, smem, c->stream>>>(prm);
  LXY_CUDA(cudaGetLastError());
  return LXY_OK;
}

inline int launch_stream1p_nc(lxy_ctx* c, int nc, const Stream1Params& prm, int grid, size_t smem) {
  switch (nc) {
    case 2: return launch_stream1p<2>(c, prm, grid, smem);
    case 4: return launch_stream1p<4>(c, prm, grid, smem);
    default: return launch_stream1p<8>(c, prm, grid, smem);
  }
}

// The packed kernel's domain: unit weights, 2..8 orderings, differences that cannot alias modulo 2^16
// (n <= 65,515; the tables of two orderings fit the shared memory up to n ~ 38,000).  The permutation
// check reads whole table words through lds_u32: written as a comparison of the uint16_t element with
// (uint16_t)(0 - i), nvcc 12.9 produced code that reported a mismatch for exactly i = 32,768 (the value
// whose negation modulo 2^16 is itself), found with a device printf on the box.
inline bool stream1p_applies(int wmode, int nc, int n) {
  static const int max_n = getenv("LXY_S1P_MAX_N") ? atoi(getenv("LXY_S1P_MAX_N")) : 65536 - (kBias + 1);  // experiments
  return wmode == W_UNIT && nc >= 2 && n <= max_n;
}
