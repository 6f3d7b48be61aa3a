// scaff_stream1p.cuh — the streaming-regime kernel for 2..8 orderings per pass with unit weights:
// the placement tables of TWO orderings are interleaved as one 32-bit word per contig.
// (included by scaff_kernels.cu inside its anonymous namespace, after scaff_stream1.cuh)
//
// Same boundary, mapping (lane = record, cp.async ring, runs of chunks) and integers as
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
    const uint32_t o0 = (words[q] << 2) & 0x3FCu, o1 = (words[q] >> 14) & 0x3FCu;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const uint32_t bb = baseB + j * pstride;
      p32[j] += s1p_pair(A[j], lds_u32(bb | o0), lutc);
      p32[j] += s1p_pair(A[j], lds_u32(bb | o1), lutc);
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
    const uint32_t o0 = (words[q] << 2) & 0x3FCu, o1 = (words[q] >> 14) & 0x3FCu;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
      const uint32_t bb = baseB + j * pstride;
      p32[j] += s1p_pair(f0 ? A0[j] : A1[j], lds_u32(bb | o0), lutc);
      p32[j] += s1p_pair(f1 ? A0[j] : A1[j], lds_u32(bb | o1), lutc);
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

template <int NC>
__global__ void __launch_bounds__(kS1Threads, kS1MinBlocks) k_scaff_stream1p(Stream1Params prm) {
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

  // share of the record stream and the first ring fill: as in k_scaff_stream1
  const int64_t npk_all = (prm.L + kPacket - 1) / kPacket;
  const int64_t lo = (npk_all * blockIdx.x / gridDim.x) * kPacket;
  const int64_t hi = min((npk_all * (blockIdx.x + 1) / gridDim.x) * kPacket, prm.L);
  constexpr int64_t kSlice = 32 * kPacket;
  constexpr int64_t kChunk = (int64_t)kS1Threads * kPacket;
  const int32_t nchunks = (int32_t)((hi - lo + kChunk - 1) / kChunk);
  const uint32_t my_slot = smem_u32(s_tile + 2) + (uint32_t)tid * 16;
  const int64_t r0 = lo + (int64_t)tid * kPacket;
#pragma unroll
  for (int q = 0; q < kS1Depth; ++q) {
    if (r0 + q * kChunk < hi) cp_async16(my_slot + q * kS1Stage, prm.rec + r0 + q * kChunk);
    cp_async_commit();
  }
  if (lo < hi) {
    const int64_t ntiles = (int64_t)prm.nblk * (prm.nblk + 1) / 2;
    for (int64_t t = tid; t < ntiles; t += kS1Threads)
      if (prm.tile_start[t] <= lo) {
        const int64_t e = prm.tile_start[t + 1];
        if (lo < e) { s_tile[0] = t; s_tile[1] = e; }
      }
  }
  // ---- placement tables: half-word (c & 1) of word [c >> 1][perm_c[i]] = -i (pad orderings: identity) ----
  if (tid < kLutLen) s_lut[tid] = c_lut100[tid];
  int bad = 0;
#pragma unroll 1
  for (int c = 0; c < NC; ++c) {
    const int32_t* row = prm.perm + (size_t)c * n;
    for (int i = tid; i < n; i += kS1Threads) {
      const int32_t cid = (c < prm.P) ? row[i] : i;
      if ((uint32_t)cid >= (uint32_t)n) bad |= 1;
      else sNeg[((size_t)(c >> 1) * nstride + cid) * 2 + (c & 1)] = (uint16_t)(0 - i);
    }
  }
  __syncthreads();
  for (int i = tid; i < kLut2Words; i += kS1Threads) {  // from the shared copy of the decay table
    const int v0 = i % kLut2Stride, v1 = i / kLut2Stride;
    s_lutc[i] = (v0 < kLutLen ? (uint32_t)s_lut[v0] : 0u) | ((uint32_t)s_lut[v1] << 16);
  }
  if (prm.validate) {
    // CTA k checks the contig IDS of slice k: the position p decoded from the table entry of id x must be
    // in range with perm[p] == x.  Independent of which store to a duplicated id won (see k_scaff_stream1):
    // a duplicate leaves another id unwritten, and an unwritten slot cannot point at its own id.
    const int32_t i0 = (int32_t)((int64_t)n * blockIdx.x / gridDim.x), i1 = (int32_t)((int64_t)n * (blockIdx.x + 1) / gridDim.x);
    for (int c = 0; c < prm.P && c < NC; ++c) {
      const int32_t* row = prm.perm + (size_t)c * n;
      for (int x = i0 + tid; x < i1; x += kS1Threads) {
        // the whole word through an opaque 32-bit load, decoded in 32 bits
        const uint32_t word = lds_u32(sbase + (uint32_t)(c >> 1) * pstride + (uint32_t)x * 4u);
        const uint32_t p = (0u - ((c & 1) ? (word >> 16) : (word & 0xFFFFu))) & 0xFFFFu;
        if (p >= (uint32_t)n || row[p] != x) bad |= 2;
      }
    }
  }
  if (bad) atomicOr(prm.flag, bad);
  S1_STAMP(1);
  __syncthreads();

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
    const int32_t hi_rel = (int32_t)(hi - lo);
    int32_t w_rel = warp * (int32_t)kSlice;
    int32_t r_rel = tid * kPacket;
    int32_t lim_rel = 0;
    uint32_t baseA = sbase, baseB = sbase;
    const uint16_t* pre = prm.rec + lo + r_rel + (int64_t)kS1Depth * kChunk;
    int32_t c = 0, since = 0;  // since: packets added to the packed sums since the last flush
#pragma unroll 1
    while (c < nchunks) {
      if (w_rel < hi_rel && w_rel + (int32_t)kSlice > lim_rel) {   // next tile (or the first one)
        cur.advance_to(lo + w_rel, prm.tile_start, prm.nblk);
        lim_rel = (int32_t)(min(hi, cur.end) - lo);
        baseA = sbase + (uint32_t)cur.I * (kBlk * 4);
        baseB = sbase + (uint32_t)cur.J * (kBlk * 4);
      }
      int32_t nrun = 0;
      if (w_rel + (int32_t)kSlice <= lim_rel) nrun = min((lim_rel - (int32_t)kSlice - w_rel) / (int32_t)kChunk + 1, nchunks - c);
      if (nrun == 0) {  // the slice crosses a tile boundary or the end of the share: per-lane tile cursor
        cp_async_wait<kS1Depth - 1>();
        const uint32_t slot = my_slot + (uint32_t)(c & (kS1Depth - 1)) * kS1Stage;
        const uint4 v = lds_v4(slot);
        cp_async16_zfill(slot, pre, r_rel + kS1Depth * (int32_t)kChunk < hi_rel);
        cp_async_commit();
        if (r_rel < hi_rel) {
          S1Tile mine = cur;
          const uint32_t words[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int k = 0; k < kPacket; ++k) {
            if (r_rel + k < hi_rel) {
              mine.advance_to(lo + r_rel + k, prm.tile_start, prm.nblk);
              const uint32_t r = (k & 1) ? (words[k >> 1] >> 16) : (words[k >> 1] & 0xFFFFu);
              s1p_eval<NP>(sbase + (uint32_t)mine.I * (kBlk * 4), sbase + (uint32_t)mine.J * (kBlk * 4), pstride, s_lut,
                           r, a32);
            }
          }
        }
        ++c;
        w_rel += (int32_t)kChunk;
        r_rel += (int32_t)kChunk;
        pre += kChunk;
        continue;
      }
      const int32_t c_end = c + nrun;
      const int32_t r_pf_end = hi_rel - kS1Depth * (int32_t)kChunk;
#pragma unroll 1
      for (; c < c_end; ++c) {
        cp_async_wait<kS1Depth - 1>();
        const uint32_t slot = my_slot + (uint32_t)(c & (kS1Depth - 1)) * kS1Stage;
        const uint4 v = lds_v4(slot);
        const bool one_a = __all_sync(0xffffffffu, ((v.x ^ (v.w >> 16)) & 0xFF00u) == 0u);
        cp_async16_zfill(slot, pre, r_rel < r_pf_end);
        cp_async_commit();
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
        if (++since == kS1pFlush) {  // warp-uniform
          s1p_flush<NP>(p32, a32);
          since = 0;
        }
        r_rel += (int32_t)kChunk;
        pre += kChunk;
      }
      w_rel += nrun * (int32_t)kChunk;
    }
  }
  cp_async_wait<0>();
  S1_STAMP(2);
  s1p_flush<NP>(p32, a32);

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

inline size_t stream1p_smem(int nc, int n) {
  return (size_t)nc * ((n + kBlk - 1) & ~(kBlk - 1)) * 2 + 64 + (size_t)kLut2Words * 4 + (size_t)kS1Warps * nc * 8 + 16 +
         1024 + (size_t)kS1Depth * kS1Stage;  // (alignment slack) table | lut | pair lut | warp sums | start tile | ring
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
