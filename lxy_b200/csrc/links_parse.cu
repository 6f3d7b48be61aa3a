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
#include <cub/device/device_radix_sort.cuh>

#include <cerrno>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
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
  atomicAdd(A.ndata, 1ull);
}

__global__ void k_scatter_vals(const uint32_t* __restrict__ line, const double* __restrict__ v, uint32_t n,
                               double* __restrict__ val) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) val[line[k]] = v[k];
}

__device__ __forceinline__ int64_t token_offset(const int64_t* __restrict__ nl_pos, const uint32_t* __restrict__ tl,
                                                uint32_t t) {
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
                                                           const uint32_t* __restrict__ tl, uint32_t ntok,
                                                           const uint32_t* __restrict__ first,
                                                           const uint32_t* __restrict__ tok_slot, int* collision) {
  const uint32_t t = blockIdx.x * kThreads + threadIdx.x;
  if (t >= ntok || tl[t] == kInvalid) return;
  const uint32_t ft = first[tok_slot[t]];
  if (ft == t) return;
  const uint32_t len = tl[t];
  if (tl[ft] != len) { *collision = 1; return; }
  const uint8_t* x = text + token_offset(nl_pos, tl, t);
  const uint8_t* y = text + token_offset(nl_pos, tl, ft);
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
                                                         uint32_t* __restrict__ slot_id, int64_t* __restrict__ name_off,
                                                         uint32_t* __restrict__ name_len) {
  const uint32_t i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= n) return;
  const uint32_t t = first_sorted[i];
  slot_id[slot_sorted[i]] = i;
  name_off[i] = token_offset(nl_pos, tl, t);
  name_len[i] = tl[t];
}

// pair table: keys[slot] = (lo << 32 | hi) + 1, first / last data line of the pair
__global__ void __launch_bounds__(kThreads) k_pairs_insert(const uint32_t* __restrict__ tl, const uint32_t* __restrict__ tok_slot,
                                                           const uint32_t* __restrict__ slot_id, uint32_t lines,
                                                           unsigned long long* keys, uint32_t* first, uint32_t* last,
                                                           uint32_t mask, int* overflow) {
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
      return;
    }
    s = (s + 1) & mask;
  }
  *overflow = 1;
}

__global__ void __launch_bounds__(kThreads) k_pairs_emit(const uint32_t* __restrict__ slot_sorted, uint32_t m,
                                                         const unsigned long long* __restrict__ keys,
                                                         const uint32_t* __restrict__ last, const double* __restrict__ val,
                                                         uint32_t* __restrict__ a, uint32_t* __restrict__ b, double* __restrict__ w) {
  const uint32_t i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= m) return;
  const uint32_t s = slot_sorted[i];
  const unsigned long long key = keys[s] - 1ull;
  a[i] = (uint32_t)(key >> 32);
  b[i] = (uint32_t)key;
  w[i] = val[last[s]];
}

struct Scratch {  // every device buffer of one load, freed on every exit path
  DevBuf text, counts, chunk_base, nl_pos, h, tl, val, tok_slot, flags, slow_off, slow_len, slow_line, slow_val,
      nkeys, nfirst, lf, ls, lf2, ls2, slot_id, name_off, name_len, pkeys, pfirst, plast, sort_tmp;
  cudaStream_t st = nullptr;
  cudaEvent_t ev[3] = {};
  ~Scratch() {
    DevBuf* all[] = {&text, &counts, &chunk_base, &nl_pos, &h, &tl, &val, &tok_slot, &flags, &slow_off, &slow_len,
                     &slow_line, &slow_val, &nkeys, &nfirst, &lf, &ls, &lf2, &ls2, &slot_id, &name_off, &name_len,
                     &pkeys, &pfirst, &plast, &sort_tmp};
    for (DevBuf* b : all) b->release();
    for (auto& e : ev)
      if (e) cudaEventDestroy(e);
    if (st) cudaStreamDestroy(st);
  }
};

struct Parsed {
  int32_t n = 0;
  int64_t m = 0;
  DevBuf a, b, w;  // device records, order of first appearance (released by the caller)
  std::vector<std::string> names;
  ~Parsed() { a.release(); b.release(); w.release(); }
};

uint32_t pow2_at_least(uint64_t x) {
  uint64_t c = 1024;
  while (c < x && c < (1ull << 31)) c <<= 1;
  return (uint32_t)c;
}

// (first, slot) pairs sorted by first; results in lf2 / ls2
int sort_by_first(Scratch& S, uint32_t n) {
  LXY_TRY(S.lf2.reserve((size_t)n * 4 + 4));
  LXY_TRY(S.ls2.reserve((size_t)n * 4 + 4));
  size_t tmp = 0;
  P_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp, S.lf.as<uint32_t>(), S.lf2.as<uint32_t>(), S.ls.as<uint32_t>(),
                                         S.ls2.as<uint32_t>(), (int64_t)n, 0, 32, S.st));
  LXY_TRY(S.sort_tmp.reserve(tmp + 16));
  P_CUDA(cub::DeviceRadixSort::SortPairs(S.sort_tmp.p, tmp, S.lf.as<uint32_t>(), S.lf2.as<uint32_t>(), S.ls.as<uint32_t>(),
                                         S.ls2.as<uint32_t>(), (int64_t)n, 0, 32, S.st));
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

int parse_links_on_device(const char* path, int device, Parsed& out, lxy_load_stats* stats) {
  const auto t_start = std::chrono::steady_clock::now();
  lxy_load_stats stt{};
  FILE* f = std::fopen(path, "rb");
  if (!f) { set_error("links_load: cannot open %s: %s", path, std::strerror(errno)); return LXY_ERR_ARG; }
  std::vector<char> text;
  {
    std::fseek(f, 0, SEEK_END);
    const long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (sz > 0) text.resize((size_t)sz);
    size_t got = 0;
    while (got < text.size()) {
      const size_t r = std::fread(text.data() + got, 1, text.size() - got, f);
      if (r == 0) break;
      got += r;
    }
    text.resize(got);
    char buf[1 << 12];  // a file that grew, or a stream without a size
    size_t r;
    while ((r = std::fread(buf, 1, sizeof(buf), f)) > 0) text.insert(text.end(), buf, buf + r);
  }
  std::fclose(f);
  stt.ms_read = (float)ms_since(t_start);
  const int64_t T = (int64_t)text.size();
  stt.bytes = T;
  if (T == 0) { set_error("links_load: Loaded empty links object..."); return LXY_ERR_STATE; }

  P_CUDA(cudaSetDevice(device));
  Scratch S;
  P_CUDA(cudaStreamCreateWithFlags(&S.st, cudaStreamNonBlocking));
  for (auto& e : S.ev) P_CUDA(cudaEventCreate(&e));
  LXY_TRY(S.text.reserve((size_t)T + 48));
  LXY_TRY(S.flags.reserve(sizeof(Flags)));
  Flags* dF = S.flags.as<Flags>();
  P_CUDA(cudaEventRecord(S.ev[0], S.st));
  P_CUDA(cudaMemsetAsync(S.text.as<uint8_t>() + T, 0, 48, S.st));
  P_CUDA(cudaMemcpyAsync(S.text.p, text.data(), (size_t)T, cudaMemcpyHostToDevice, S.st));
  P_CUDA(cudaEventRecord(S.ev[1], S.st));
  const uint8_t* d_text = S.text.as<uint8_t>();

  // 1. newline positions
  const int64_t nchunks = (T + kChunk - 1) / kChunk;
  LXY_TRY(S.counts.reserve((size_t)nchunks * 4));
  LXY_TRY(S.chunk_base.reserve((size_t)nchunks * 8));
  P_CUDA(cudaMemsetAsync(dF, 0, sizeof(Flags), S.st));
  k_newline_count<<<(unsigned)nchunks, kThreads, 0, S.st>>>(d_text, T, S.counts.as<uint32_t>());
  k_scan_counts<<<1, 1024, 0, S.st>>>(S.counts.as<uint32_t>(), nchunks, S.chunk_base.as<int64_t>(), &dF->nl_total);
  int64_t NL = 0;
  P_CUDA(cudaMemcpyAsync(&NL, &dF->nl_total, 8, cudaMemcpyDeviceToHost, S.st));
  P_CUDA(cudaStreamSynchronize(S.st));
  const int64_t lines = NL + (text[(size_t)T - 1] != '\n' ? 1 : 0);
  if (lines >= (1ll << 31)) { set_error("links_load: %lld lines exceed the device loader's 2^31", (long long)lines); return LXY_ERR_RANGE; }
  stt.lines = lines;
  LXY_TRY(S.nl_pos.reserve((size_t)(NL + 1) * 8));
  k_newline_pos<<<(unsigned)nchunks, kThreads, 0, S.st>>>(d_text, T, S.chunk_base.as<int64_t>(), S.nl_pos.as<int64_t>());
  P_CUDA(cudaGetLastError());

  // 2..3: parse + name table; a detected hash collision re-runs both with another seed
  const uint32_t ntok = (uint32_t)(2 * lines);
  LXY_TRY(S.h.reserve((size_t)ntok * 8));
  LXY_TRY(S.tl.reserve((size_t)ntok * 4));
  LXY_TRY(S.val.reserve((size_t)lines * 8));
  LXY_TRY(S.tok_slot.reserve((size_t)ntok * 4));
  LXY_TRY(S.slow_off.reserve((size_t)lines * 8));
  LXY_TRY(S.slow_len.reserve((size_t)lines * 4));
  LXY_TRY(S.slow_line.reserve((size_t)lines * 4));
  const unsigned line_grid = (unsigned)((lines + kThreads - 1) / kThreads);
  const unsigned tok_grid = (unsigned)(((int64_t)ntok + kThreads - 1) / kThreads);
  uint64_t hmask0 = ~0ull;  // test hook: a narrow first hash forces the collision path
  if (const char* hb = std::getenv("LXY_PARSE_HASH_BITS")) {
    const int bits = std::atoi(hb);
    if (bits > 0 && bits < 64) hmask0 = (1ull << bits) - 1;
  }
  uint32_t cap = pow2_at_least(1u << 16);
  uint32_t N = 0;
  Flags hF{};
  for (int attempt = 0;; ++attempt) {
    if (attempt == 4) { set_error("links_load: name hash collisions persisted over 4 seeds"); return LXY_ERR_STATE; }
    stt.attempts = attempt + 1;
    P_CUDA(cudaMemsetAsync(dF, 0, sizeof(Flags), S.st));
    P_CUDA(cudaMemsetAsync(&dF->err, 0xFF, 8, S.st));
    ParseArgs A{d_text, T, S.nl_pos.as<int64_t>(), (int64_t)NL, lines, 0x9e3779b97f4a7c15ull * (uint64_t)attempt,
                attempt == 0 ? hmask0 : ~0ull, S.h.as<uint64_t>(), S.tl.as<uint32_t>(), S.val.as<double>(),
                &dF->err, &dF->slow_cnt, S.slow_off.as<int64_t>(), S.slow_len.as<uint32_t>(),
                S.slow_line.as<uint32_t>(), &dF->ndata};
    k_parse_lines<<<line_grid, kThreads, 0, S.st>>>(A);
    P_CUDA(cudaMemcpyAsync(&hF, dF, sizeof(Flags), cudaMemcpyDeviceToHost, S.st));
    P_CUDA(cudaStreamSynchronize(S.st));
    if (hF.err != ~0ull) {
      const unsigned long long ln = (hF.err >> 2) + 1;
      const int kind = (int)(hF.err & 3);
      if (kind == 1) set_error("links_load: %s:%llu: empty line (the reference panics here)", path, ln);
      else if (kind == 2) set_error("links_load: %s:%llu: fewer than three fields (the reference panics here)", path, ln);
      else set_error("links_load: %s:%llu: line longer than 2 GiB", path, ln);
      return LXY_ERR_ARG;
    }
    stt.records = (int64_t)hF.ndata;
    if (hF.ndata == 0) { set_error("links_load: Loaded empty links object..."); return LXY_ERR_STATE; }
    stt.slow_values = (int32_t)hF.slow_cnt;
    if (hF.slow_cnt) {  // tokens outside the exact fast path: the host loader's own function decides
      const uint32_t k = hF.slow_cnt;
      std::vector<int64_t> off(k);
      std::vector<uint32_t> len(k);
      std::vector<double> v(k);
      P_CUDA(cudaMemcpyAsync(off.data(), S.slow_off.p, (size_t)k * 8, cudaMemcpyDeviceToHost, S.st));
      P_CUDA(cudaMemcpyAsync(len.data(), S.slow_len.p, (size_t)k * 4, cudaMemcpyDeviceToHost, S.st));
      P_CUDA(cudaStreamSynchronize(S.st));
      for (uint32_t q = 0; q < k; ++q) v[q] = parse_float_or_zero(text.data() + off[q], len[q]);
      LXY_TRY(S.slow_val.reserve((size_t)k * 8));
      P_CUDA(cudaMemcpyAsync(S.slow_val.p, v.data(), (size_t)k * 8, cudaMemcpyHostToDevice, S.st));
      k_scatter_vals<<<(k + 255) / 256, 256, 0, S.st>>>(S.slow_line.as<uint32_t>(), S.slow_val.as<double>(), k,
                                                         S.val.as<double>());
      P_CUDA(cudaStreamSynchronize(S.st));  // v must outlive the copy
    }
    // name table, grown until the distinct names fill at most half of it
    bool collided = false;
    for (;;) {
      LXY_TRY(S.nkeys.reserve((size_t)cap * 8));
      LXY_TRY(S.nfirst.reserve((size_t)cap * 4));
      P_CUDA(cudaMemsetAsync(S.nkeys.p, 0, (size_t)cap * 8, S.st));
      P_CUDA(cudaMemsetAsync(S.nfirst.p, 0xFF, (size_t)cap * 4, S.st));
      P_CUDA(cudaMemsetAsync(&dF->used, 0, 4, S.st));
      P_CUDA(cudaMemsetAsync(&dF->overflow, 0, 8, S.st));  // overflow + collision
      k_names_insert<<<tok_grid, kThreads, 0, S.st>>>(S.h.as<uint64_t>(), S.tl.as<uint32_t>(), ntok,
                                                      S.nkeys.as<unsigned long long>(), S.nfirst.as<uint32_t>(), cap - 1,
                                                      S.tok_slot.as<uint32_t>(), &dF->used, cap / 2, &dF->overflow);
      P_CUDA(cudaMemcpyAsync(&hF, dF, sizeof(Flags), cudaMemcpyDeviceToHost, S.st));
      P_CUDA(cudaStreamSynchronize(S.st));
      if (!hF.overflow) break;
      if (cap >= (1u << 31)) { set_error("links_load: more than 2^30 distinct names"); return LXY_ERR_RANGE; }
      cap = cap >= (1u << 28) ? cap * 2 : cap * 8;
    }
    k_names_verify<<<tok_grid, kThreads, 0, S.st>>>(d_text, S.nl_pos.as<int64_t>(), S.tl.as<uint32_t>(), ntok,
                                                    S.nfirst.as<uint32_t>(), S.tok_slot.as<uint32_t>(), &dF->collision);
    P_CUDA(cudaMemcpyAsync(&hF, dF, sizeof(Flags), cudaMemcpyDeviceToHost, S.st));
    P_CUDA(cudaStreamSynchronize(S.st));
    collided = hF.collision != 0;
    if (!collided) { N = hF.used; break; }
  }

  // 4. ids = rank of a name's first token
  LXY_TRY(S.lf.reserve((size_t)N * 4 + 4));
  LXY_TRY(S.ls.reserve((size_t)N * 4 + 4));
  LXY_TRY(S.slot_id.reserve((size_t)cap * 4));
  LXY_TRY(S.name_off.reserve((size_t)N * 8));
  LXY_TRY(S.name_len.reserve((size_t)N * 4));
  P_CUDA(cudaMemsetAsync(&dF->cnt, 0, 4, S.st));
  k_collect_slots<<<(cap + kThreads - 1) / kThreads, kThreads, 0, S.st>>>(S.nkeys.as<unsigned long long>(),
                                                                         S.nfirst.as<uint32_t>(), cap, S.lf.as<uint32_t>(),
                                                                         S.ls.as<uint32_t>(), &dF->cnt);
  LXY_TRY(sort_by_first(S, N));
  k_assign_ids<<<(N + kThreads - 1) / kThreads, kThreads, 0, S.st>>>(S.lf2.as<uint32_t>(), S.ls2.as<uint32_t>(), N,
                                                                    S.nl_pos.as<int64_t>(), S.tl.as<uint32_t>(),
                                                                    S.slot_id.as<uint32_t>(), S.name_off.as<int64_t>(),
                                                                    S.name_len.as<uint32_t>());
  std::vector<int64_t> noff(N);
  std::vector<uint32_t> nlen(N);
  P_CUDA(cudaMemcpyAsync(noff.data(), S.name_off.p, (size_t)N * 8, cudaMemcpyDeviceToHost, S.st));
  P_CUDA(cudaMemcpyAsync(nlen.data(), S.name_len.p, (size_t)N * 4, cudaMemcpyDeviceToHost, S.st));

  // 5. pairs: first / last data line of every (lo, hi)
  const uint64_t max_pairs = std::min<uint64_t>((uint64_t)stt.records, (uint64_t)N * ((uint64_t)N + 1) / 2);
  const uint32_t pcap = pow2_at_least(2 * max_pairs);
  LXY_TRY(S.pkeys.reserve((size_t)pcap * 8));
  LXY_TRY(S.pfirst.reserve((size_t)pcap * 4));
  LXY_TRY(S.plast.reserve((size_t)pcap * 4));
  P_CUDA(cudaMemsetAsync(S.pkeys.p, 0, (size_t)pcap * 8, S.st));
  P_CUDA(cudaMemsetAsync(S.pfirst.p, 0xFF, (size_t)pcap * 4, S.st));
  P_CUDA(cudaMemsetAsync(S.plast.p, 0, (size_t)pcap * 4, S.st));
  P_CUDA(cudaMemsetAsync(&dF->cnt, 0, 4, S.st));
  P_CUDA(cudaMemsetAsync(&dF->overflow, 0, 4, S.st));
  k_pairs_insert<<<line_grid, kThreads, 0, S.st>>>(S.tl.as<uint32_t>(), S.tok_slot.as<uint32_t>(), S.slot_id.as<uint32_t>(),
                                                   (uint32_t)lines, S.pkeys.as<unsigned long long>(),
                                                   S.pfirst.as<uint32_t>(), S.plast.as<uint32_t>(), pcap - 1, &dF->overflow);
  LXY_TRY(S.lf.reserve((size_t)max_pairs * 4 + 4));
  LXY_TRY(S.ls.reserve((size_t)max_pairs * 4 + 4));
  k_collect_slots<<<(pcap + kThreads - 1) / kThreads, kThreads, 0, S.st>>>(S.pkeys.as<unsigned long long>(),
                                                                          S.pfirst.as<uint32_t>(), pcap, S.lf.as<uint32_t>(),
                                                                          S.ls.as<uint32_t>(), &dF->cnt);
  P_CUDA(cudaMemcpyAsync(&hF, dF, sizeof(Flags), cudaMemcpyDeviceToHost, S.st));
  P_CUDA(cudaStreamSynchronize(S.st));
  if (hF.overflow) { set_error("links_load: pair table overflow (internal)"); return LXY_ERR_STATE; }
  const uint32_t M = hF.cnt;
  LXY_TRY(sort_by_first(S, M));
  LXY_TRY(out.a.reserve((size_t)M * 4 + 4));
  LXY_TRY(out.b.reserve((size_t)M * 4 + 4));
  LXY_TRY(out.w.reserve((size_t)M * 8 + 8));
  k_pairs_emit<<<(M + kThreads - 1) / kThreads, kThreads, 0, S.st>>>(S.ls2.as<uint32_t>(), M, S.pkeys.as<unsigned long long>(),
                                                                    S.plast.as<uint32_t>(), S.val.as<double>(),
                                                                    out.a.as<uint32_t>(), out.b.as<uint32_t>(), out.w.as<double>());
  P_CUDA(cudaEventRecord(S.ev[2], S.st));
  P_CUDA(cudaStreamSynchronize(S.st));
  P_CUDA(cudaGetLastError());

  out.n = (int32_t)N;
  out.m = (int64_t)M;
  out.names.resize(N);
  for (uint32_t i = 0; i < N; ++i) out.names[i].assign(text.data() + noff[i], nlen[i]);
  stt.names = (int32_t)N;
  stt.pairs = (int64_t)M;
  cudaEventElapsedTime(&stt.ms_h2d, S.ev[0], S.ev[1]);
  cudaEventElapsedTime(&stt.ms_device, S.ev[1], S.ev[2]);
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
    cudaError_t e = cudaMemcpy(l->a.data(), P.a.p, (size_t)P.m * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(l->b.data(), P.b.p, (size_t)P.m * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(l->w.data(), P.w.p, (size_t)P.m * 8, cudaMemcpyDeviceToHost);
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
  LXY_TRY(parse_links_on_device(path, device, P, stats));
  return to_host_table(P, true, out);
}

int lxy_ctx_load_links(lxy_ctx* c, const char* path, int32_t kind, lxy_links** out, lxy_load_stats* stats) {
  if (!c || !path) { set_error("ctx_load_links: null argument"); return LXY_ERR_ARG; }
  if (kind != 0 && kind != 1) { set_error("ctx_load_links: kind must be 0 (scaffolding) or 1 (phasing)"); return LXY_ERR_ARG; }
  if (out) *out = nullptr;
  Parsed P;
  LXY_TRY(parse_links_on_device(path, c->cfg.device, P, stats));
  lxy_links* l = nullptr;
  if (out || kind == 1) LXY_TRY(to_host_table(P, true, &l));
  int rc;
  if (kind == 0) {  // the device records feed the link store as they are
    rc = lxy_ctx_set_links_dev(c, P.n, P.m, P.a.as<uint32_t>(), P.b.as<uint32_t>(), P.w.as<double>());
  } else {          // the phasing store is built from host arrays (deduplicated pairs: small)
    rc = lxy_ctx_set_phase_links(c, P.n, P.m, l->a.data(), l->b.data(), l->w.data());
  }
  if (rc != LXY_OK || !out) delete l;
  else *out = l;
  return rc;
}

}  // extern "C"
