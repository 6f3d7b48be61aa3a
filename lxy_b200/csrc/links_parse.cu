// links_parse.cu — LoadLinks on the device: a `.links` text file becomes the link table (names -> ids
// by first appearance, one record per pair with the value of the pair's LAST row) without the host
// ever tokenising the text.  Same results as the host loader lxy_links_load (links_io.cu), which
// follows /root/reference/src/sequtil/links.go:427-491:
//   bufio.Scanner lines (split at '\n', one trailing '\r' dropped)            :441-446
//   lines whose first byte is '#' are skipped; an empty line panics (line[0]) :450   -> LXY_ERR_ARG
//   strings.Split(line, " "): every single space separates, empty fields kept :454   (arr[2] missing
//                                                                              panics -> LXY_ERR_ARG)
//   ids by first appearance of a name, dense from 0                            :458-470, :51-62
//   strconv.ParseFloat with the error ignored (malformed -> 0)                 :473
//   Links.Set: (a,b) == (b,a), the last row of a pair wins                     :477, :113-145
//   error when nothing was loaded                                              :485-487
//
// Device pipeline (all byte / integer work, HBM-bound; the text is read three times, 16 bytes per
// thread, coalesced; every table is a flat array in HBM):
//   1. k_newline_count / k_scan_counts / k_newline_pos  — positions of all '\n' in file order
//   2. k_parse_lines    — one thread per line: three field boundaries, a seeded 64-bit hash of each
//                         name, the value as an exactly rounded double when the token is a plain
//                         decimal with <= 19 digits, mantissa < 2^53 and |exponent| <= 22 (one IEEE
//                         multiply or divide of two exact doubles: Clinger's fast path — what `%f`
//                         output always is).  Any other token (inf, nan, hex, underscores, 20+ digits,
//                         garbage) is put on a list and resolved by the host loader's own
//                         parse_float_or_zero, so both loaders agree on every value by construction.
//   3. k_names_insert   — open-addressing table keyed by the name hash; atomicMin keeps the first
//                         token index of every name.  k_names_verify compares every token with the
//                         first token of its slot byte for byte: a 64-bit hash collision is DETECTED
//                         (and the file re-hashed with another seed), never silently merged.
//   4. slots sorted by first token index (cub radix sort over the distinct names only) -> ids.
//   5. k_pairs_insert   — table keyed by the exact (lo, hi) id pair; atomicMin / atomicMax keep the
//                         first and last data line of every pair.  Sorted by first line -> records in
//                         the host loader's order, each with the value of its last line.
// The record arrays stay on the device for lxy_ctx_load_links (they feed the link store directly).
//
// ScaffoldLinksFromSam (/root/reference/src/sequtil/alignment.go:624-676) runs through the same
// pipeline with another front end (mode kSam): k_sam_fields splits every line at tabs (read id,
// contig name), the data lines are compacted (cub select), k_sam_starts marks the lines whose read
// id differs from the data line before (exact byte compare), and k_sam_emit turns every group of
// exactly two hits that is FOLLOWED by another group into one record (contig of hit 1, contig of
// hit 2) — the reference adds a group only when the next read id arrives and never flushes the last
// one.  Names then get their ids by first appearance over the records, and a pair's value is the
// number of its records (Links.Add(.., 1) per read pair, alignment.go:656-660).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"
#include "links_table.h"

using namespace lxy;

namespace {

constexpr int kThreads = 256;
constexpr int kIters = 8;
constexpr int64_t kChunk = (int64_t)kThreads * 16 * kIters;  // text bytes per CTA in passes 1a / 1b
constexpr uint32_t kInvalid = 0xFFFFFFFFu;
constexpr uint64_t kFnvPrime = 0x100000001b3ull;

#define P_CUDA(expr)                                                                                \
  do {                                                                                              \
    cudaError_t e_ = (expr);                                                                        \
    if (e_ != cudaSuccess) {                                                                        \
      set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__);        \
      return LXY_ERR_CUDA;                                                                          \
    }                                                                                               \
  } while (0)

// bit k set when byte k of the 16 loaded bytes is '\n'
__device__ __forceinline__ uint32_t newline_mask(uint4 v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  uint32_t m = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t x = (__vcmpeq4(w[q], 0x0a0a0a0au) >> 7) & 0x01010101u;  // one bit per equal byte
    m |= (((x * 0x00204081u) >> 21) & 0xFu) << (4 * q);                   // bits 0,8,16,24 -> 0..3
  }
  return m;
}

__device__ __forceinline__ uint32_t load_mask(const uint8_t* text, int64_t T, int64_t off) {
  if (off >= T) return 0;
  uint32_t m = newline_mask(__ldg(reinterpret_cast<const uint4*>(text + off)));
  if (off + 16 > T) m &= (1u << (int)(T - off)) - 1u;
  return m;
}

// exclusive scan of one int per thread over the CTA (kThreads = 8 warps); *total = CTA sum
__device__ __forceinline__ int block_excl_scan(int c, int* total, int* wsum) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = c;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int up = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += up;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  int base = 0, all = 0;
#pragma unroll
  for (int k = 0; k < kThreads / 32; ++k) {
    const int s = wsum[k];
    if (k < warp) base += s;
    all += s;
  }
  __syncthreads();
  *total = all;
  return base + incl - c;
}

__global__ void __launch_bounds__(kThreads) k_newline_count(const uint8_t* __restrict__ text, int64_t T,
                                                            uint32_t* __restrict__ counts) {
  __shared__ int wsum[kThreads / 32];
  const int64_t base = (int64_t)blockIdx.x * kChunk;
  int c = 0;
#pragma unroll
  for (int it = 0; it < kIters; ++it)
    c += __popc(load_mask(text, T, base + ((int64_t)it * kThreads + threadIdx.x) * 16));
  int total;
  block_excl_scan(c, &total, wsum);
  if (threadIdx.x == 0) counts[blockIdx.x] = (uint32_t)total;
}

// one CTA: exclusive prefix sums of the per-chunk counts (64-bit), grand total to *total
__global__ void __launch_bounds__(1024) k_scan_counts(const uint32_t* __restrict__ counts, int64_t n,
                                                      int64_t* __restrict__ base, int64_t* __restrict__ total) {
  __shared__ long long wsum[32];
  __shared__ long long carry_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int64_t i0 = 0; i0 < n; i0 += 1024) {
    const int64_t i = i0 + threadIdx.x;
    const long long v = i < n ? (long long)counts[i] : 0;
    long long incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const long long up = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += up;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    long long wb = 0, all = 0;
    for (int k = 0; k < 32; ++k) {
      const long long s = wsum[k];
      if (k < warp) wb += s;
      all += s;
    }
    const long long carry = carry_s;
    if (i < n) base[i] = carry + wb + incl - v;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + all;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry_s;
}

__global__ void __launch_bounds__(kThreads) k_newline_pos(const uint8_t* __restrict__ text, int64_t T,
                                                          const int64_t* __restrict__ chunk_base,
                                                          int64_t* __restrict__ nl_pos) {
  __shared__ int wsum[kThreads / 32];
  const int64_t base = (int64_t)blockIdx.x * kChunk;
  int64_t out = chunk_base[blockIdx.x];
  for (int it = 0; it < kIters; ++it) {
    const int64_t off = base + ((int64_t)it * kThreads + threadIdx.x) * 16;
    uint32_t m = load_mask(text, T, off);
    int total;
    int64_t o = out + block_excl_scan(__popc(m), &total, wsum);
    while (m) {
      const int k = __ffs(m) - 1;
      m &= m - 1;
      nl_pos[o++] = off + k;
    }
    out += total;
  }
}

__device__ __forceinline__ uint64_t fmix64(uint64_t k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return k;
}

__constant__ double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                  1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

// A plain decimal `[+-]?(d+[.d*]|.d+)([eE][+-]?d+)?` whose value one IEEE operation rounds exactly.
// Returns false for everything else (the host resolves those tokens).
__device__ bool parse_decimal_exact(const uint8_t* __restrict__ p, int64_t s, int64_t e, double* out) {
  if (s >= e) return false;
  int64_t i = s;
  bool neg = false;
  uint8_t c = p[i];
  if (c == '+' || c == '-') { neg = c == '-'; ++i; }
  uint64_t mant = 0;
  int nd = 0, exp10 = 0;
  bool sawdigits = false, sawdot = false;
  for (; i < e; ++i) {
    c = p[i];
    if (c == '.') {
      if (sawdot) return false;
      sawdot = true;
      continue;
    }
    if (c < '0' || c > '9') break;
    sawdigits = true;
    if (c == '0' && mant == 0) {  // leading zeros carry no digit
      if (sawdot) --exp10;
      continue;
    }
    if (nd >= 19) return false;
    mant = mant * 10 + (uint64_t)(c - '0');
    ++nd;
    if (sawdot) --exp10;
  }
  if (!sawdigits) return false;
  if (i < e) {
    if ((p[i] | 0x20) != 'e') return false;
    ++i;
    if (i >= e) return false;
    bool eneg = false;
    c = p[i];
    if (c == '+' || c == '-') { eneg = c == '-'; ++i; }
    if (i >= e) return false;
    int ex = 0;
    for (; i < e; ++i) {
      c = p[i];
      if (c < '0' || c > '9') return false;
      if (ex < 100000) ex = ex * 10 + (c - '0');
    }
    exp10 += eneg ? -ex : ex;
  }
  if (mant == 0) { *out = neg ? -0.0 : 0.0; return true; }
  if (mant >> 53) return false;
  if (exp10 < -22 || exp10 > 22) return false;
  double v = (double)(long long)mant;
  v = exp10 < 0 ? v / kPow10[-exp10] : v * kPow10[exp10];
  *out = neg ? -v : v;
  return true;
}

struct ParseArgs {
  const uint8_t* text;
  int64_t T;
  const int64_t* nl_pos;
  int64_t NL, lines;
  uint64_t seed, hmask;
  uint64_t* h;        // [2 * lines]
  uint32_t* tl;       // [2 * lines] token lengths; kInvalid on lines without a record
  double* val;        // [lines]
  unsigned long long* err;  // min over (line << 2 | kind)
  uint32_t* slow_cnt;
  int64_t* slow_off;
  uint32_t* slow_len;
  uint32_t* slow_line;
  unsigned long long* ndata;  // data lines
};

__global__ void __launch_bounds__(kThreads) k_parse_lines(const ParseArgs A) {
  const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  if (i >= A.lines) return;
  const uint8_t* __restrict__ p = A.text;
  int64_t s = i ? A.nl_pos[i - 1] + 1 : 0;
  int64_t e = i < A.NL ? A.nl_pos[i] : A.T;
  A.tl[2 * i] = kInvalid;
  A.tl[2 * i + 1] = kInvalid;
  if (e > s && p[e - 1] == '\r') --e;
  if (e == s) { atomicMin(A.err, ((unsigned long long)i << 2) | 1ull); return; }
  if (e - s >= 0x7fffffffll) { atomicMin(A.err, ((unsigned long long)i << 2) | 3ull); return; }
  if (p[s] == '#') return;
  const uint64_t h_init = 0xcbf29ce484222325ull ^ A.seed;
  int64_t j = s;
  uint64_t h0 = h_init;
  for (; j < e; ++j) {
    const uint8_t c = p[j];
    if (c == ' ') break;
    h0 = (h0 ^ c) * kFnvPrime;
  }
  if (j >= e) { atomicMin(A.err, ((unsigned long long)i << 2) | 2ull); return; }
  const uint32_t len0 = (uint32_t)(j - s);
  const int64_t f1 = ++j;
  uint64_t h1 = h_init;
  for (; j < e; ++j) {
    const uint8_t c = p[j];
    if (c == ' ') break;
    h1 = (h1 ^ c) * kFnvPrime;
  }
  if (j >= e) { atomicMin(A.err, ((unsigned long long)i << 2) | 2ull); return; }
  const uint32_t len1 = (uint32_t)(j - f1);
  const int64_t f2 = ++j;
  for (; j < e; ++j)
    if (p[j] == ' ') break;
  double v = 0.0;
  if (j > f2 && !parse_decimal_exact(p, f2, j, &v)) {  // the empty token is a syntax error: 0
    const uint32_t k = atomicAdd(A.slow_cnt, 1u);
    A.slow_off[k] = f2;
    A.slow_len[k] = (uint32_t)(j - f2);
    A.slow_line[k] = (uint32_t)i;
  }
  h0 = fmix64(h0 ^ len0) & A.hmask;
  h1 = fmix64(h1 ^ len1) & A.hmask;
  A.h[2 * i] = h0 ? h0 : 1;
  A.h[2 * i + 1] = h1 ? h1 : 1;
  A.tl[2 * i] = len0;
  A.tl[2 * i + 1] = len1;
  A.val[i] = v;
  const unsigned peers = __activemask();  // one counter update per warp, not per line
  if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(A.ndata, (unsigned long long)__popc(peers));
}

// ---- SAM front end (ScaffoldLinksFromSam, alignment.go:624-676) -----------------------------------
struct SamArgs {
  const uint8_t* text;
  int64_t T;
  const int64_t* nl_pos;
  int64_t NL, lines;
  uint64_t seed, hmask;
  uint32_t* tl;        // [2 * lines] reset to kInvalid here; k_sam_emit fills the record lines
  uint8_t* is_data;    // [lines] 1 on lines that do not start with '@'
  uint32_t* qlen;      // [lines] length of arr[0] (the read id starts the line)
  int64_t* r_off;      // [lines] arr[2]
  uint32_t* r_len;
  uint64_t* r_h;
  unsigned long long* err;  // min over (line << 2 | kind)
};

// one thread per line: strings.Split(line, "\t") as far as arr[0] and arr[2] (alignment.go:650)
__global__ void __launch_bounds__(kThreads) k_sam_fields(const SamArgs A) {
  const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  if (i >= A.lines) return;
  const uint8_t* __restrict__ p = A.text;
  int64_t s = i ? A.nl_pos[i - 1] + 1 : 0;
  int64_t e = i < A.NL ? A.nl_pos[i] : A.T;
  A.tl[2 * i] = kInvalid;
  A.tl[2 * i + 1] = kInvalid;
  A.is_data[i] = 0;
  if (e > s && p[e - 1] == '\r') --e;
  if (e == s) { atomicMin(A.err, ((unsigned long long)i << 2) | 1ull); return; }   // line[0] panics (:648)
  if (e - s >= 0x7fffffffll) { atomicMin(A.err, ((unsigned long long)i << 2) | 3ull); return; }
  if (p[s] == '@') return;
  int64_t j = s;
  for (; j < e; ++j)
    if (p[j] == '\t') break;
  A.qlen[i] = (uint32_t)(j - s);
  if (j < e) {
    for (++j; j < e; ++j)
      if (p[j] == '\t') break;
  }
  if (j >= e) { atomicMin(A.err, ((unsigned long long)i << 2) | 2ull); return; }   // arr[2] panics (:668)
  const int64_t f2 = ++j;
  uint64_t h = 0xcbf29ce484222325ull ^ A.seed;
  for (; j < e; ++j) {
    const uint8_t c = p[j];
    if (c == '\t') break;
    h = (h ^ c) * kFnvPrime;
  }
  const uint32_t len = (uint32_t)(j - f2);
  h = fmix64(h ^ len) & A.hmask;
  A.r_off[i] = f2;
  A.r_len[i] = len;
  A.r_h[i] = h ? h : 1;
  A.is_data[i] = 1;  // only complete lines reach the later stages (a bad line ends the load anyway)
}

// start[d] = the d-th data line opens a group: its read id differs from the data line before
// (alignment.go:652; currentID always equals the previous data line's arr[0]).  start[0] = 1: the
// first group begins there whether or not its id equals the initial "".
__global__ void __launch_bounds__(kThreads) k_sam_starts(const uint8_t* __restrict__ text, const int64_t* __restrict__ nl_pos,
                                                         const uint32_t* __restrict__ data_idx,
                                                         const uint32_t* __restrict__ ndata_lines,
                                                         const uint32_t* __restrict__ qlen, uint8_t* __restrict__ start) {
  const uint32_t d = blockIdx.x * kThreads + threadIdx.x;
  const uint32_t D = *ndata_lines;
  if (d >= D) return;
  uint8_t st = 1;
  if (d > 0) {
    const uint32_t l1 = data_idx[d], l0 = data_idx[d - 1];
    const uint32_t n = qlen[l1];
    if (qlen[l0] == n) {
      const uint8_t* x = text + (l1 ? nl_pos[l1 - 1] + 1 : 0);
      const uint8_t* y = text + (l0 ? nl_pos[l0 - 1] + 1 : 0);
      uint32_t k = 0;
      while (k < n && x[k] == y[k]) ++k;
      st = k < n;
    }
  }
  start[d] = st;
}

// a group of exactly two hits that is followed by another group -> one record on its first line:
// tokens (arr[2] of hit 1, arr[2] of hit 2), in the order ID() sees them (alignment.go:654-658)
__global__ void __launch_bounds__(kThreads) k_sam_emit(const uint32_t* __restrict__ data_idx,
                                                       const uint32_t* __restrict__ ndata_lines,
                                                       const uint8_t* __restrict__ start, const int64_t* __restrict__ r_off,
                                                       const uint32_t* __restrict__ r_len, const uint64_t* __restrict__ r_h,
                                                       uint64_t* __restrict__ h, uint32_t* __restrict__ tl,
                                                       int64_t* __restrict__ tok_off, double* __restrict__ val,
                                                       unsigned long long* nrec) {
  const uint32_t d = blockIdx.x * kThreads + threadIdx.x;
  const uint32_t D = *ndata_lines;
  if (d >= D || (uint64_t)d + 2 >= D) return;
  if (!(start[d] && !start[d + 1] && start[d + 2])) return;
  const uint32_t l0 = data_idx[d], l1 = data_idx[d + 1];
  h[2 * (size_t)l0] = r_h[l0];
  h[2 * (size_t)l0 + 1] = r_h[l1];
  tl[2 * (size_t)l0] = r_len[l0];
  tl[2 * (size_t)l0 + 1] = r_len[l1];
  tok_off[2 * (size_t)l0] = r_off[l0];
  tok_off[2 * (size_t)l0 + 1] = r_off[l1];
  val[l0] = 1.0;
  atomicAdd(nrec, 1ull);
}

__global__ void k_scatter_vals(const uint32_t* __restrict__ line, const double* __restrict__ v, uint32_t n,
                               double* __restrict__ val) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) val[line[k]] = v[k];
}

// tok_off (SAM mode) holds the offsets explicitly; in a .links file they follow from the line start
__device__ __forceinline__ int64_t token_offset(const int64_t* __restrict__ nl_pos, const uint32_t* __restrict__ tl,
                                                const int64_t* __restrict__ tok_off, uint32_t t) {
  if (tok_off) return tok_off[t];
  const uint32_t line = t >> 1;
  const int64_t s = line ? nl_pos[line - 1] + 1 : 0;
  return (t & 1) ? s + tl[t - 1] + 1 : s;
}

// name table: keys[slot] = hash (0 = empty), first[slot] = smallest token index with that hash
__global__ void __launch_bounds__(kThreads) k_names_insert(const uint64_t* __restrict__ h, const uint32_t* __restrict__ tl,
                                                           uint32_t ntok, unsigned long long* keys, uint32_t* first,
                                                           uint32_t mask, uint32_t* __restrict__ tok_slot, uint32_t* used,
                                                           uint32_t limit, int* overflow) {
  const uint32_t t = blockIdx.x * kThreads + threadIdx.x;
  if (t >= ntok || tl[t] == kInvalid) return;
  const unsigned long long key = h[t];
  uint32_t s = (uint32_t)key & mask;
  for (uint32_t probe = 0; probe <= mask; ++probe) {
    unsigned long long k = *(volatile unsigned long long*)(keys + s);
    if (k == 0) {
      k = atomicCAS(keys + s, 0ull, key);
      if (k == 0) {
        if (atomicAdd(used, 1u) >= limit) *overflow = 1;
        k = key;
      }
    }
    if (k == key) {
      if (*(volatile uint32_t*)(first + s) > t) atomicMin(first + s, t);
      tok_slot[t] = s;
      return;
    }
    s = (s + 1) & mask;
  }
  *overflow = 1;
}

__global__ void __launch_bounds__(kThreads) k_names_verify(const uint8_t* __restrict__ text, const int64_t* __restrict__ nl_pos,
                                                           const uint32_t* __restrict__ tl,
                                                           const int64_t* __restrict__ tok_off, uint32_t ntok,
                                                           const uint32_t* __restrict__ first,
                                                           const uint32_t* __restrict__ tok_slot, const int* overflow,
                                                           int* collision) {
  const uint32_t t = blockIdx.x * kThreads + threadIdx.x;
  if (t >= ntok || tl[t] == kInvalid || *overflow) return;  // after an overflow tok_slot is incomplete
  const uint32_t ft = first[tok_slot[t]];
  if (ft == t) return;
  const uint32_t len = tl[t];
  if (tl[ft] != len) { *collision = 1; return; }
  const uint8_t* x = text + token_offset(nl_pos, tl, tok_off, t);
  const uint8_t* y = text + token_offset(nl_pos, tl, tok_off, ft);
  for (uint32_t k = 0; k < len; ++k)
    if (x[k] != y[k]) { *collision = 1; return; }
}

// occupied slots of a table -> (first, slot) list
__global__ void __launch_bounds__(kThreads) k_collect_slots(const unsigned long long* __restrict__ keys,
                                                            const uint32_t* __restrict__ first, uint32_t nslots,
                                                            uint32_t* __restrict__ list_first, uint32_t* __restrict__ list_slot,
                                                            uint32_t* cnt) {
  const uint32_t s = blockIdx.x * kThreads + threadIdx.x;
  if (s >= nslots || keys[s] == 0) return;
  const uint32_t k = atomicAdd(cnt, 1u);
  list_first[k] = first[s];
  list_slot[k] = s;
}

__global__ void __launch_bounds__(kThreads) k_assign_ids(const uint32_t* __restrict__ first_sorted,
                                                         const uint32_t* __restrict__ slot_sorted, uint32_t n,
                                                         const int64_t* __restrict__ nl_pos, const uint32_t* __restrict__ tl,
                                                         const int64_t* __restrict__ tok_off,
                                                         uint32_t* __restrict__ slot_id, int64_t* __restrict__ name_off,
                                                         uint32_t* __restrict__ name_len) {
  const uint32_t i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= n) return;
  const uint32_t t = first_sorted[i];
  slot_id[slot_sorted[i]] = i;
  name_off[i] = token_offset(nl_pos, tl, tok_off, t);
  name_len[i] = tl[t];
}

// pair table: keys[slot] = (lo << 32 | hi) + 1, first / last data line of the pair
__global__ void __launch_bounds__(kThreads) k_pairs_insert(const uint32_t* __restrict__ tl, const uint32_t* __restrict__ tok_slot,
                                                           const uint32_t* __restrict__ slot_id, uint32_t lines,
                                                           unsigned long long* keys, uint32_t* first, uint32_t* last,
                                                           uint32_t* count, uint32_t mask, int* overflow) {
  const uint32_t l = blockIdx.x * kThreads + threadIdx.x;
  if (l >= lines || tl[2 * l] == kInvalid) return;
  const uint32_t ia = slot_id[tok_slot[2 * l]], ib = slot_id[tok_slot[2 * l + 1]];
  const uint32_t lo = min(ia, ib), hi = max(ia, ib);
  const unsigned long long key = (((unsigned long long)lo << 32) | hi) + 1ull;
  uint32_t s = (uint32_t)fmix64(key) & mask;
  for (uint32_t probe = 0; probe <= mask; ++probe) {
    unsigned long long k = *(volatile unsigned long long*)(keys + s);
    if (k == 0) {
      k = atomicCAS(keys + s, 0ull, key);
      if (k == 0) k = key;
    }
    if (k == key) {
      if (*(volatile uint32_t*)(first + s) > l) atomicMin(first + s, l);
      if (*(volatile uint32_t*)(last + s) < l) atomicMax(last + s, l);
      if (count) atomicAdd(count + s, 1u);  // Links.Add(.., 1) per record (SAM mode)
      return;
    }
    s = (s + 1) & mask;
  }
  *overflow = 1;
}

__global__ void __launch_bounds__(kThreads) k_pairs_emit(const uint32_t* __restrict__ slot_sorted, uint32_t m,
                                                         const unsigned long long* __restrict__ keys,
                                                         const uint32_t* __restrict__ last, const double* __restrict__ val,
                                                         const uint32_t* __restrict__ count, uint32_t* __restrict__ a, uint32_t* __restrict__ b, double* __restrict__ w) {
  const uint32_t i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= m) return;
  const uint32_t s = slot_sorted[i];
  const unsigned long long key = keys[s] - 1ull;
  a[i] = (uint32_t)(key >> 32);
  b[i] = (uint32_t)key;
  w[i] = count ? (double)count[s] : val[last[s]];
}

// Device memory of one load: one cudaMalloc per stage instead of one per array (26 cudaMalloc /
// cudaFree pairs cost several times what all the kernels take); everything is freed on every exit path.
struct Arena {
  std::vector<void*> slabs;
  char* cur = nullptr;
  size_t left = 0;
  int open(size_t bytes) {
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      set_error("links_load: cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
      return LXY_ERR_NOMEM;
    }
    slabs.push_back(p);
    cur = (char*)p;
    left = bytes;
    return LXY_OK;
  }
  template <class T> T* take(size_t count) {
    const size_t b = (count * sizeof(T) + 255) & ~(size_t)255;
    if (b > left) return nullptr;  // sizing bug; callers check through ok()
    T* r = reinterpret_cast<T*>(cur);
    cur += b;
    left -= b;
    return r;
  }
  ~Arena() {
    for (void* p : slabs) cudaFree(p);
  }
};
inline size_t r256(size_t b) { return (b + 255) & ~(size_t)255; }

struct Handles {
  cudaStream_t st = nullptr;
  cudaEvent_t ev[3] = {};
  ~Handles() {
    for (auto& e : ev)
      if (e) cudaEventDestroy(e);
    if (st) cudaStreamDestroy(st);
  }
};

struct Parsed {
  int32_t n = 0;
  int64_t m = 0;
  void* slab = nullptr;  // device records a | b | w, order of first appearance
  uint32_t *a = nullptr, *b = nullptr;
  double* w = nullptr;
  std::vector<std::string> names;
  ~Parsed() { if (slab) cudaFree(slab); }
};

// The file as one read-only byte range: mapped (the page cache is the buffer; nothing is copied on the
// host before the DMA staging) or, for things that cannot be mapped, read into memory.
struct FileText {
  const char* p = nullptr;
  size_t n = 0;
  bool mapped = false;
  int fd = -1;  // kept open for a mapped file: the staging threads pread() from it
  std::vector<char> buf;
  ~FileText() {
    if (mapped) munmap(const_cast<char*>(p), n);
    if (fd >= 0) ::close(fd);
  }
};

int read_file(const char* path, FileText& ft) {
  const int fd = ::open(path, O_RDONLY);
  if (fd < 0) { set_error("links_load: cannot open %s: %s", path, std::strerror(errno)); return LXY_ERR_ARG; }
  struct stat sb;
  if (fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size > 0) {
    void* m = mmap(nullptr, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);
    if (m != MAP_FAILED) {
      ft.p = (const char*)m;
      ft.n = (size_t)sb.st_size;
      ft.mapped = true;
      ft.fd = fd;
      return LXY_OK;
    }
  }
  char chunk[1 << 16];
  for (;;) {
    const ssize_t r = ::read(fd, chunk, sizeof(chunk));
    if (r < 0 && errno == EINTR) continue;
    if (r < 0) { set_error("links_load: cannot read %s: %s", path, std::strerror(errno)); ::close(fd); return LXY_ERR_ARG; }
    if (r == 0) break;
    ft.buf.insert(ft.buf.end(), chunk, chunk + r);
  }
  ::close(fd);
  ft.p = ft.buf.data();
  ft.n = ft.buf.size();
  return LXY_OK;
}

// Host -> device copy of a large text through a page-locked ring.  A pageable cudaMemcpy of the mapped
// file runs at ~11 GB/s (the driver's own staging, one thread); here four threads pread() 4 MB pieces
// straight from the page cache into their two ring slots each and queue the DMA of a piece as soon as
// it is staged, so the copy engine overlaps the other threads' reads.  The ring (32 MB) is allocated
// once per process on the first large load and kept (loading one file per chromosome re-uses it);
// smaller texts take the plain copy.  Returns LXY_OK with *done = false when the ring is unavailable.
constexpr size_t kRingSlot = (size_t)4 << 20;
constexpr int kRingThreads = 4, kRingSlotsPer = 2;
constexpr size_t kRingMinBytes = (size_t)48 << 20;

int h2d_through_ring(uint8_t* d_dst, const FileText& ft, int device, bool* done) {
  static std::mutex mu;
  static char* ring = nullptr;
  static bool ring_failed = false;
  *done = false;
  size_t min_bytes = kRingMinBytes;
  if (const char* e = std::getenv("LXY_PARSE_RING_MIN_BYTES")) min_bytes = (size_t)std::strtoull(e, nullptr, 10);  // tests
  if (ft.n < min_bytes || std::getenv("LXY_PARSE_NO_RING")) return LXY_OK;
  std::lock_guard<std::mutex> lk(mu);  // one load at a time owns the ring
  if (!ring && !ring_failed) {
    if (cudaHostAlloc((void**)&ring, kRingSlot * kRingThreads * kRingSlotsPer, cudaHostAllocDefault) != cudaSuccess) {
      (void)cudaGetLastError();
      ring = nullptr;
      ring_failed = true;
    }
  }
  if (!ring) return LXY_OK;
  const size_t nchunks = (ft.n + kRingSlot - 1) / kRingSlot;
  std::atomic<int> err(0);
  auto work = [&](int t) {
    if (cudaSetDevice(device) != cudaSuccess) { err = 1; return; }
    cudaStream_t s = nullptr;
    cudaEvent_t ev[kRingSlotsPer] = {};
    if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) { err = 1; return; }
    for (auto& e : ev)
      if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) err = 1;
    int use = 0;
    for (size_t i = (size_t)t; i < nchunks && !err; i += kRingThreads, ++use) {
      const int k = use % kRingSlotsPer;
      char* slot = ring + ((size_t)t * kRingSlotsPer + k) * kRingSlot;
      if (use >= kRingSlotsPer && cudaEventSynchronize(ev[k]) != cudaSuccess) { err = 1; break; }
      const size_t off = i * kRingSlot, len = std::min(kRingSlot, ft.n - off);
      if (ft.fd >= 0) {
        size_t got = 0;
        while (got < len) {
          const ssize_t r = ::pread(ft.fd, slot + got, len - got, (off_t)(off + got));
          if (r < 0 && errno == EINTR) continue;
          if (r <= 0) { err = 2; break; }
          got += (size_t)r;
        }
        if (err) break;
      } else {
        std::memcpy(slot, ft.p + off, len);
      }
      if (cudaMemcpyAsync(d_dst + off, slot, len, cudaMemcpyHostToDevice, s) != cudaSuccess ||
          cudaEventRecord(ev[k], s) != cudaSuccess) { err = 1; break; }
    }
    if (cudaStreamSynchronize(s) != cudaSuccess) err = 1;
    for (auto& e : ev)
      if (e) cudaEventDestroy(e);
    cudaStreamDestroy(s);
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < kRingThreads; ++t) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
  if (err == 2) { set_error("links_load: the file changed while it was read"); return LXY_ERR_ARG; }
  if (err) { (void)cudaGetLastError(); set_error("links_load: staged host -> device copy failed"); return LXY_ERR_CUDA; }
  *done = true;
  return LXY_OK;
}

uint32_t pow2_at_least(uint64_t x) {
  uint64_t c = 1024;
  while (c < x && c < (1ull << 31)) c <<= 1;
  return (uint32_t)c;
}

// (first, slot) pairs -> sorted by first
struct SortBufs {
  uint32_t *lf = nullptr, *ls = nullptr, *lf2 = nullptr, *ls2 = nullptr;
  void* tmp = nullptr;
  size_t tmp_bytes = 0;
};
size_t sort_tmp_bytes(uint32_t n) {
  size_t tmp = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, (int64_t)n, 0, 32, (cudaStream_t) nullptr);
  return tmp + 256;
}
size_t sort_bytes(uint32_t n) { return 4 * r256((size_t)n * 4 + 4) + r256(sort_tmp_bytes(n)); }
bool sort_take(Arena& A, uint32_t n, SortBufs& B) {
  B.lf = A.take<uint32_t>((size_t)n + 1);
  B.ls = A.take<uint32_t>((size_t)n + 1);
  B.lf2 = A.take<uint32_t>((size_t)n + 1);
  B.ls2 = A.take<uint32_t>((size_t)n + 1);
  B.tmp_bytes = sort_tmp_bytes(n);
  B.tmp = A.take<char>(B.tmp_bytes);
  return B.lf && B.ls && B.lf2 && B.ls2 && B.tmp;
}
int sort_by_first(const SortBufs& B, uint32_t n, cudaStream_t st) {
  size_t tmp = B.tmp_bytes;
  P_CUDA(cub::DeviceRadixSort::SortPairs(B.tmp, tmp, (const uint32_t*)B.lf, B.lf2, (const uint32_t*)B.ls, B.ls2,
                                         (int64_t)n, 0, 32, st));
  return LXY_OK;
}

double ms_since(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

// flags block layout (one small device allocation, zeroed per attempt)
struct Flags {
  unsigned long long err;     // min (line << 2 | kind); ~0 = none
  unsigned long long ndata;
  int64_t nl_total;
  uint32_t slow_cnt, used, cnt;
  int overflow, collision, pad;
};

struct NameTable {  // arrays whose size follows the table capacity
  unsigned long long* keys = nullptr;
  uint32_t *first = nullptr, *slot_id = nullptr, *name_len = nullptr;
  int64_t* name_off = nullptr;
  SortBufs sort;
};
size_t name_table_bytes(uint32_t cap) {
  return r256((size_t)cap * 8) + 2 * r256((size_t)cap * 4) + r256((size_t)cap / 2 * 4) + r256((size_t)cap / 2 * 8) +
         sort_bytes(cap / 2);
}
bool name_table_take(Arena& A, uint32_t cap, NameTable& N) {
  N.keys = A.take<unsigned long long>(cap);
  N.first = A.take<uint32_t>(cap);
  N.slot_id = A.take<uint32_t>(cap);
  N.name_len = A.take<uint32_t>(cap / 2);
  N.name_off = A.take<int64_t>(cap / 2);
  return N.keys && N.first && N.slot_id && N.name_len && N.name_off && sort_take(A, cap / 2, N.sort);
}

#define P_SIZED(ok) do { if (!(ok)) { set_error("links_load: arena sizing (internal)"); return LXY_ERR_STATE; } } while (0)

enum Mode { kLinks = 0, kSam = 1 };  // LoadLinks (links.go:427-491) / ScaffoldLinksFromSam (alignment.go:624-676)

size_t select_tmp_bytes(int64_t lines) {
  size_t tmp = 0;
  cub::DeviceSelect::Flagged(nullptr, tmp, thrust::counting_iterator<uint32_t>(0), (const uint8_t*)nullptr,
                             (uint32_t*)nullptr, (uint32_t*)nullptr, lines, (cudaStream_t) nullptr);
  return tmp + 256;
}

int parse_links_on_device(const char* path, int device, Mode mode, Parsed& out, lxy_load_stats* stats) {
  const char* who = mode == kSam ? "links_from_sam" : "links_load";
  const auto t_start = std::chrono::steady_clock::now();
  lxy_load_stats stt{};
  FileText ft;
  LXY_TRY(read_file(path, ft));
  const char* text = ft.p;
  stt.ms_read = (float)ms_since(t_start);
  const int64_t T = (int64_t)ft.n;
  stt.bytes = T;
  if (T == 0) {
    if (mode == kSam) {  // no read pairs: the reference writes the empty table
      stt.ms_total = (float)ms_since(t_start);
      if (stats) *stats = stt;
      return LXY_OK;
    }
    set_error("links_load: Loaded empty links object...");
    return LXY_ERR_STATE;
  }

  P_CUDA(cudaSetDevice(device));
  Handles H;
  Arena A;
  P_CUDA(cudaStreamCreateWithFlags(&H.st, cudaStreamNonBlocking));
  for (auto& e : H.ev) P_CUDA(cudaEventCreate(&e));
  cudaStream_t st = H.st;
  const bool trace = std::getenv("LXY_PARSE_TRACE") != nullptr;  // phase times on stderr (adds a sync per phase)
  auto mark = [&](const char* what) {
    if (!trace) return;
    const double t_host = ms_since(t_start);
    cudaStreamSynchronize(st);
    std::fprintf(stderr, "[%s] host %8.3f ms, stream idle %8.3f ms  %s\n", who, t_host, ms_since(t_start), what);
  };

  // stage A: the text and the per-chunk newline counts
  const int64_t nchunks = (T + kChunk - 1) / kChunk;
  LXY_TRY(A.open(r256((size_t)T + 48) + r256((size_t)nchunks * 4) + r256((size_t)nchunks * 8) + r256(sizeof(Flags))));
  uint8_t* d_text_w = A.take<uint8_t>((size_t)T + 48);
  uint32_t* d_counts = A.take<uint32_t>((size_t)nchunks);
  int64_t* d_chunk_base = A.take<int64_t>((size_t)nchunks);
  Flags* dF = A.take<Flags>(1);
  P_SIZED(d_text_w && d_counts && d_chunk_base && dF);
  const uint8_t* d_text = d_text_w;
  P_CUDA(cudaEventRecord(H.ev[0], st));
  P_CUDA(cudaMemsetAsync(d_text_w + T, 0, 48, st));
  {
    bool staged = false;
    LXY_TRY(h2d_through_ring(d_text_w, ft, device, &staged));
    if (!staged) P_CUDA(cudaMemcpyAsync(d_text_w, text, (size_t)T, cudaMemcpyHostToDevice, st));
  }
  P_CUDA(cudaEventRecord(H.ev[1], st));

  // 1. newline positions
  P_CUDA(cudaMemsetAsync(dF, 0, sizeof(Flags), st));
  k_newline_count<<<(unsigned)nchunks, kThreads, 0, st>>>(d_text, T, d_counts);
  k_scan_counts<<<1, 1024, 0, st>>>(d_counts, nchunks, d_chunk_base, &dF->nl_total);
  int64_t NL = 0;
  P_CUDA(cudaMemcpyAsync(&NL, &dF->nl_total, 8, cudaMemcpyDeviceToHost, st));
  P_CUDA(cudaStreamSynchronize(st));
  mark("text resident, newlines counted");
  const int64_t lines = NL + (text[(size_t)T - 1] != '\n' ? 1 : 0);
  if (lines >= (1ll << 31)) { set_error("%s: %lld lines exceed the device loader's 2^31", who, (long long)lines); return LXY_ERR_RANGE; }
  stt.lines = lines;

  // stage B: everything sized by the line count, and the first name table
  const uint32_t ntok = (uint32_t)(2 * lines);
  uint32_t cap = pow2_at_least(1u << 16);
  const size_t sel_bytes = mode == kSam ? select_tmp_bytes(lines) : 0;
  const size_t sam_bytes = mode == kSam ? 2 * r256((size_t)lines + 2) + 3 * r256((size_t)lines * 4) +
                                              2 * r256((size_t)lines * 8) + r256((size_t)ntok * 8) + r256(sel_bytes)
                                        : 0;
  mark("before stage B allocation");
  LXY_TRY(A.open(r256((size_t)(NL + 1) * 8) + r256((size_t)ntok * 8) + 2 * r256((size_t)ntok * 4) +
                 3 * r256((size_t)lines * 8) + 2 * r256((size_t)lines * 4) + sam_bytes + name_table_bytes(cap)));
  int64_t* d_nl = A.take<int64_t>((size_t)NL + 1);
  uint64_t* d_h = A.take<uint64_t>(ntok);
  uint32_t* d_tl = A.take<uint32_t>(ntok);
  uint32_t* d_tok_slot = A.take<uint32_t>(ntok);
  double* d_val = A.take<double>((size_t)lines);
  int64_t* d_slow_off = A.take<int64_t>((size_t)lines);
  double* d_slow_val = A.take<double>((size_t)lines);
  uint32_t* d_slow_len = A.take<uint32_t>((size_t)lines);
  uint32_t* d_slow_line = A.take<uint32_t>((size_t)lines);
  uint8_t *d_is_data = nullptr, *d_start = nullptr;
  uint32_t *d_data_idx = nullptr, *d_qlen = nullptr, *d_rlen = nullptr;
  int64_t *d_roff = nullptr, *d_tok_off = nullptr;
  uint64_t* d_rh = nullptr;
  void* d_sel_tmp = nullptr;
  if (mode == kSam) {
    d_is_data = A.take<uint8_t>((size_t)lines + 2);
    d_start = A.take<uint8_t>((size_t)lines + 2);
    d_data_idx = A.take<uint32_t>((size_t)lines);
    d_qlen = A.take<uint32_t>((size_t)lines);
    d_rlen = A.take<uint32_t>((size_t)lines);
    d_roff = A.take<int64_t>((size_t)lines);
    d_rh = A.take<uint64_t>((size_t)lines);
    d_tok_off = A.take<int64_t>(ntok);
    d_sel_tmp = A.take<char>(sel_bytes);
    P_SIZED(d_is_data && d_start && d_data_idx && d_qlen && d_rlen && d_roff && d_rh && d_tok_off && d_sel_tmp);
  }
  NameTable NT;
  P_SIZED(d_nl && d_h && d_tl && d_tok_slot && d_val && d_slow_off && d_slow_val && d_slow_len && d_slow_line &&
          name_table_take(A, cap, NT));
  mark("stage B allocated");
  k_newline_pos<<<(unsigned)nchunks, kThreads, 0, st>>>(d_text, T, d_chunk_base, d_nl);
  P_CUDA(cudaGetLastError());
  mark("newline positions");

  // 2..3: parse + name table; a detected hash collision re-runs both with another seed
  const unsigned line_grid = (unsigned)((lines + kThreads - 1) / kThreads);
  const unsigned tok_grid = (unsigned)(((int64_t)ntok + kThreads - 1) / kThreads);
  uint64_t hmask0 = ~0ull;  // test hook: a narrow first hash forces the collision path
  if (const char* hb = std::getenv("LXY_PARSE_HASH_BITS")) {
    const int bits = std::atoi(hb);
    if (bits > 0 && bits < 64) hmask0 = (1ull << bits) - 1;
  }
  uint32_t N = 0;
  Flags hF{};
  for (int attempt = 0;; ++attempt) {
    if (attempt == 4) { set_error("%s: name hash collisions persisted over 4 seeds", who); return LXY_ERR_STATE; }
    stt.attempts = attempt + 1;
    P_CUDA(cudaMemsetAsync(dF, 0, sizeof(Flags), st));
    P_CUDA(cudaMemsetAsync(&dF->err, 0xFF, 8, st));
    const uint64_t seed = 0x9e3779b97f4a7c15ull * (uint64_t)attempt, hmask = attempt == 0 ? hmask0 : ~0ull;
    if (mode == kLinks) {
      ParseArgs PA{d_text, T, d_nl, NL, lines, seed, hmask, d_h, d_tl, d_val, &dF->err, &dF->slow_cnt, d_slow_off,
                   d_slow_len, d_slow_line, &dF->ndata};
      k_parse_lines<<<line_grid, kThreads, 0, st>>>(PA);
    } else {
      SamArgs SA{d_text, T, d_nl, NL, lines, seed, hmask, d_tl, d_is_data, d_qlen, d_roff, d_rlen, d_rh, &dF->err};
      k_sam_fields<<<line_grid, kThreads, 0, st>>>(SA);
      size_t tmp = sel_bytes;  // data lines in file order; their count stays on the device (dF->cnt)
      P_CUDA(cub::DeviceSelect::Flagged(d_sel_tmp, tmp, thrust::counting_iterator<uint32_t>(0), (const uint8_t*)d_is_data,
                                        d_data_idx, &dF->cnt, lines, st));
      k_sam_starts<<<line_grid, kThreads, 0, st>>>(d_text, d_nl, d_data_idx, &dF->cnt, d_qlen, d_start);
      k_sam_emit<<<line_grid, kThreads, 0, st>>>(d_data_idx, &dF->cnt, d_start, d_roff, d_rlen, d_rh, d_h, d_tl, d_tok_off,
                                                 d_val, &dF->ndata);
    }
    mark("lines parsed");
    // the name table is filled on the assumption that the file is well formed (checked right after)
    bool first_pass = true;
    for (;;) {
      P_CUDA(cudaMemsetAsync(NT.keys, 0, (size_t)cap * 8, st));
      P_CUDA(cudaMemsetAsync(NT.first, 0xFF, (size_t)cap * 4, st));
      P_CUDA(cudaMemsetAsync(&dF->used, 0, 4, st));
      P_CUDA(cudaMemsetAsync(&dF->overflow, 0, 8, st));  // overflow + collision
      k_names_insert<<<tok_grid, kThreads, 0, st>>>(d_h, d_tl, ntok, NT.keys, NT.first, cap - 1, d_tok_slot, &dF->used,
                                                    cap / 2, &dF->overflow);
      k_names_verify<<<tok_grid, kThreads, 0, st>>>(d_text, d_nl, d_tl, d_tok_off, ntok, NT.first, d_tok_slot,
                                                    &dF->overflow, &dF->collision);
      P_CUDA(cudaMemcpyAsync(&hF, dF, sizeof(Flags), cudaMemcpyDeviceToHost, st));
      P_CUDA(cudaStreamSynchronize(st));
      if (first_pass) {
        first_pass = false;
        if (hF.err != ~0ull) {
          const unsigned long long ln = (hF.err >> 2) + 1;
          const int kind = (int)(hF.err & 3);
          if (kind == 1) set_error("%s: %s:%llu: empty line (the reference panics here)", who, path, ln);
          else if (kind == 2) set_error("%s: %s:%llu: fewer than three fields (the reference panics here)", who, path, ln);
          else set_error("%s: %s:%llu: line longer than 2 GiB", who, path, ln);
          return LXY_ERR_ARG;
        }
        stt.records = (int64_t)hF.ndata;
        if (hF.ndata == 0) {
          if (mode == kSam) {  // no group of two was followed by another read id: the empty table
            cudaEventElapsedTime(&stt.ms_h2d, H.ev[0], H.ev[1]);
            stt.ms_total = (float)ms_since(t_start);
            if (stats) *stats = stt;
            return LXY_OK;
          }
          set_error("links_load: Loaded empty links object...");
          return LXY_ERR_STATE;
        }
        stt.slow_values = (int32_t)hF.slow_cnt;
      }
      if (!hF.overflow) break;
      if (cap >= (1u << 31)) { set_error("%s: more than 2^30 distinct names", who); return LXY_ERR_RANGE; }
      cap = cap >= (1u << 28) ? cap * 2 : cap * 8;  // grown until the names fill at most half of it
      LXY_TRY(A.open(name_table_bytes(cap)));
      P_SIZED(name_table_take(A, cap, NT));
    }
    if (hF.slow_cnt) {  // tokens outside the exact fast path: the host loader's own function decides
      const uint32_t k = hF.slow_cnt;
      std::vector<int64_t> off(k);
      std::vector<uint32_t> len(k);
      std::vector<double> v(k);
      P_CUDA(cudaMemcpyAsync(off.data(), d_slow_off, (size_t)k * 8, cudaMemcpyDeviceToHost, st));
      P_CUDA(cudaMemcpyAsync(len.data(), d_slow_len, (size_t)k * 4, cudaMemcpyDeviceToHost, st));
      P_CUDA(cudaStreamSynchronize(st));
      for (uint32_t q = 0; q < k; ++q) v[q] = parse_float_or_zero(text + off[q], len[q]);
      P_CUDA(cudaMemcpyAsync(d_slow_val, v.data(), (size_t)k * 8, cudaMemcpyHostToDevice, st));
      k_scatter_vals<<<(k + 255) / 256, 256, 0, st>>>(d_slow_line, d_slow_val, k, d_val);
      P_CUDA(cudaStreamSynchronize(st));  // v must outlive the copy
    }
    if (!hF.collision) { N = hF.used; break; }
  }

  mark("name table");
  // 4. ids = rank of a name's first token
  P_CUDA(cudaMemsetAsync(&dF->cnt, 0, 4, st));
  k_collect_slots<<<(cap + kThreads - 1) / kThreads, kThreads, 0, st>>>(NT.keys, NT.first, cap, NT.sort.lf, NT.sort.ls, &dF->cnt);
  LXY_TRY(sort_by_first(NT.sort, N, st));
  k_assign_ids<<<(N + kThreads - 1) / kThreads, kThreads, 0, st>>>(NT.sort.lf2, NT.sort.ls2, N, d_nl, d_tl, d_tok_off,
                                                                  NT.slot_id, NT.name_off, NT.name_len);
  std::vector<int64_t> noff(N);
  std::vector<uint32_t> nlen(N);
  P_CUDA(cudaMemcpyAsync(noff.data(), NT.name_off, (size_t)N * 8, cudaMemcpyDeviceToHost, st));
  P_CUDA(cudaMemcpyAsync(nlen.data(), NT.name_len, (size_t)N * 4, cudaMemcpyDeviceToHost, st));

  mark("ids assigned");
  // stage C / 5. pairs: first / last data line of every (lo, hi)
  const uint64_t max_pairs = std::min<uint64_t>((uint64_t)stt.records, (uint64_t)N * ((uint64_t)N + 1) / 2);
  const uint32_t pcap = pow2_at_least(2 * max_pairs);
  LXY_TRY(A.open(r256((size_t)pcap * 8) + 3 * r256((size_t)pcap * 4) + sort_bytes((uint32_t)max_pairs)));
  unsigned long long* d_pkeys = A.take<unsigned long long>(pcap);
  uint32_t* d_pfirst = A.take<uint32_t>(pcap);
  uint32_t* d_plast = A.take<uint32_t>(pcap);
  uint32_t* d_pcount = mode == kSam ? A.take<uint32_t>(pcap) : nullptr;  // records per pair = Add(.., 1) each
  SortBufs PS;
  P_SIZED(d_pkeys && d_pfirst && d_plast && (mode != kSam || d_pcount) && sort_take(A, (uint32_t)max_pairs, PS));
  if (d_pcount) P_CUDA(cudaMemsetAsync(d_pcount, 0, (size_t)pcap * 4, st));
  mark("ids assigned, stage C allocated");
  P_CUDA(cudaMemsetAsync(d_pkeys, 0, (size_t)pcap * 8, st));
  P_CUDA(cudaMemsetAsync(d_pfirst, 0xFF, (size_t)pcap * 4, st));
  P_CUDA(cudaMemsetAsync(d_plast, 0, (size_t)pcap * 4, st));
  P_CUDA(cudaMemsetAsync(&dF->cnt, 0, 4, st));
  P_CUDA(cudaMemsetAsync(&dF->overflow, 0, 4, st));
  k_pairs_insert<<<line_grid, kThreads, 0, st>>>(d_tl, d_tok_slot, NT.slot_id, (uint32_t)lines, d_pkeys, d_pfirst, d_plast,
                                                 d_pcount, pcap - 1, &dF->overflow);
  k_collect_slots<<<(pcap + kThreads - 1) / kThreads, kThreads, 0, st>>>(d_pkeys, d_pfirst, pcap, PS.lf, PS.ls, &dF->cnt);
  P_CUDA(cudaMemcpyAsync(&hF, dF, sizeof(Flags), cudaMemcpyDeviceToHost, st));
  P_CUDA(cudaStreamSynchronize(st));
  if (hF.overflow) { set_error("%s: pair table overflow (internal)", who); return LXY_ERR_STATE; }
  mark("pair table");
  const uint32_t M = hF.cnt;
  LXY_TRY(sort_by_first(PS, M, st));
  {
    const size_t ab = r256((size_t)M * 4 + 4);
    cudaError_t e = cudaMalloc(&out.slab, 2 * ab + (size_t)M * 8 + 8);
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      out.slab = nullptr;
      set_error("%s: cudaMalloc of the records failed: %s", who, cudaGetErrorString(e));
      return LXY_ERR_NOMEM;
    }
    out.a = reinterpret_cast<uint32_t*>(out.slab);
    out.b = reinterpret_cast<uint32_t*>((char*)out.slab + ab);
    out.w = reinterpret_cast<double*>((char*)out.slab + 2 * ab);
  }
  k_pairs_emit<<<(M + kThreads - 1) / kThreads, kThreads, 0, st>>>(PS.ls2, M, d_pkeys, d_plast, d_val, d_pcount, out.a, out.b,
                                                                  out.w);
  P_CUDA(cudaEventRecord(H.ev[2], st));
  P_CUDA(cudaStreamSynchronize(st));
  P_CUDA(cudaGetLastError());

  mark("records emitted");
  out.n = (int32_t)N;
  out.m = (int64_t)M;
  out.names.resize(N);
  for (uint32_t i = 0; i < N; ++i) out.names[i].assign(text + noff[i], nlen[i]);
  stt.names = (int32_t)N;
  stt.pairs = (int64_t)M;
  cudaEventElapsedTime(&stt.ms_h2d, H.ev[0], H.ev[1]);
  cudaEventElapsedTime(&stt.ms_device, H.ev[1], H.ev[2]);
  stt.ms_total = (float)ms_since(t_start);
  if (stats) *stats = stt;
  return LXY_OK;
}

// names (and, when wanted, the records) into a host table object
int to_host_table(const Parsed& P, bool with_records, lxy_links** out) {
  lxy_links* l = new (std::nothrow) lxy_links();
  if (!l) { set_error("out of memory"); return LXY_ERR_NOMEM; }
  l->name_of = P.names;
  l->id_of.reserve(P.names.size());
  for (size_t i = 0; i < P.names.size(); ++i) l->id_of.emplace(P.names[i], (int32_t)i);
  if (with_records && P.m) {
    l->a.resize((size_t)P.m);
    l->b.resize((size_t)P.m);
    l->w.resize((size_t)P.m);
    cudaError_t e = cudaMemcpy(l->a.data(), P.a, (size_t)P.m * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(l->b.data(), P.b, (size_t)P.m * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(l->w.data(), P.w, (size_t)P.m * 8, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
      set_error("links_load: device -> host copy failed: %s", cudaGetErrorString(e));
      delete l;
      return LXY_ERR_CUDA;
    }
    l->slot_ready = false;  // built by the first Set / Add / Get
  }
  *out = l;
  return LXY_OK;
}

}  // namespace

extern "C" {

int lxy_links_load_gpu(const char* path, int32_t device, lxy_links** out, lxy_load_stats* stats) {
  if (!path || !out) { set_error("links_load_gpu: null argument"); return LXY_ERR_ARG; }
  *out = nullptr;
  Parsed P;
  LXY_TRY(parse_links_on_device(path, device, kLinks, P, stats));
  return to_host_table(P, true, out);
}

// ScaffoldLinksFromSam's table (alignment.go:624-672), built on the device
int lxy_links_from_sam_gpu(const char* path, int32_t device, lxy_links** out, lxy_load_stats* stats) {
  if (!path || !out) { set_error("links_from_sam_gpu: null argument"); return LXY_ERR_ARG; }
  *out = nullptr;
  Parsed P;
  LXY_TRY(parse_links_on_device(path, device, kSam, P, stats));
  return to_host_table(P, true, out);
}

// ScaffoldLinksFromSam(samPath, outPath) (alignment.go:624-676): the table, then links.Write
int lxy_scaffold_links_from_sam(const char* sam_path, const char* out_path, int32_t device, lxy_load_stats* stats) {
  if (!sam_path || !out_path) { set_error("scaffold_links_from_sam: null argument"); return LXY_ERR_ARG; }
  lxy_links* l = nullptr;
  LXY_TRY(lxy_links_from_sam_gpu(sam_path, device, &l, stats));
  const int rc = lxy_links_write(l, out_path);
  lxy_links_free(l);
  return rc;
}

// SAM -> link store without the .links file in between: the records never leave the device
int lxy_ctx_load_sam(lxy_ctx* c, const char* path, lxy_links** out, lxy_load_stats* stats) {
  if (!c || !path) { set_error("ctx_load_sam: null argument"); return LXY_ERR_ARG; }
  if (out) *out = nullptr;
  Parsed P;
  LXY_TRY(parse_links_on_device(path, c->cfg.device, kSam, P, stats));
  if (P.n == 0) { set_error("ctx_load_sam: %s holds no read pair to add (empty link table)", path); return LXY_ERR_STATE; }
  lxy_links* l = nullptr;
  if (out) LXY_TRY(to_host_table(P, true, &l));
  const int rc = lxy_ctx_set_links_dev(c, P.n, P.m, P.a, P.b, P.w);
  if (rc != LXY_OK) delete l;
  else if (out) *out = l;
  return rc;
}

int lxy_ctx_load_links(lxy_ctx* c, const char* path, int32_t kind, lxy_links** out, lxy_load_stats* stats) {
  if (!c || !path) { set_error("ctx_load_links: null argument"); return LXY_ERR_ARG; }
  if (kind != 0 && kind != 1) { set_error("ctx_load_links: kind must be 0 (scaffolding) or 1 (phasing)"); return LXY_ERR_ARG; }
  if (out) *out = nullptr;
  Parsed P;
  LXY_TRY(parse_links_on_device(path, c->cfg.device, kLinks, P, stats));
  lxy_links* l = nullptr;
  if (out || kind == 1) LXY_TRY(to_host_table(P, true, &l));
  int rc;
  if (kind == 0) {  // the device records feed the link store as they are
    rc = lxy_ctx_set_links_dev(c, P.n, P.m, P.a, P.b, P.w);
  } else {          // the phasing store is built from host arrays (deduplicated pairs: small)
    rc = lxy_ctx_set_phase_links(c, P.n, P.m, l->a.data(), l->b.data(), l->w.data());
  }
  if (rc != LXY_OK || !out) delete l;
  else *out = l;
  return rc;
}

}  // extern "C"
