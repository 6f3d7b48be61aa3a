// phase_kernels.cu — phasing fitness on sm_100a (/root/reference/src/lxy/phase/phase.go:171-207).
//
// Link-centric form: the gene index IS the link id (phase.go:185), so for a record
// (i < j, w) both the step test (j - i in {1..11,15,20,25,30}, phase.go:177) and the multiplicity
// m = 1 + [i > 0] (the `i-v > 0` test, phase.go:192) are candidate-independent.  They are applied
// once when the links are set: records off the step list are dropped and w' = w * m is stored.
//   fitness(p) = - sum w' * (g[i] == g[j] ? +1 : -1) = -(sumW - 2 * D(p)),  D(p) = sum w' * (g[i]^g[j])
// Everything is integer; int64 accumulation is exact.
//
//   k_phase_bits    genes[P][n] bytes -> bit-sliced words bits[g][i] (bit c = candidate 32g + c)
//   k_phase_score   one warp per (block of up to four 32-candidate groups, link chunk); lane = candidate
//                   of each group; records and gene words are warp-uniform loads (L1-resident window)
//   k_phase_finish  out = -(sumW - 2 D)
#include <algorithm>
#include <cmath>
#include <vector>

#include "common.cuh"

namespace lxy {
namespace {

constexpr int kChunk = 2048;  // records per warp work item

// bits[q][i][G]: word (q, i, s) holds the gene bits of group g = q * G + s (bit c = candidate 32g + c),
// so the G words a warp needs for variant i are one aligned G * 4-byte load.  Groups past the
// population read as zero.
__global__ void k_phase_bits(const uint8_t* __restrict__ genes, int32_t P, int32_t n, int32_t groups,
                             int32_t G, uint32_t* __restrict__ bits) {
  // one warp per (group, 32 consecutive variants); lane = variant
  const int64_t warp_id = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int64_t vchunks = (n + 31) / 32;
  if (warp_id >= vchunks * groups) return;
  const int g = (int)(warp_id / vchunks);
  const int32_t i = (int32_t)(warp_id - (int64_t)g * vchunks) * 32 + lane;
  uint32_t word = 0;
  if (i < n) {
#pragma unroll 8
    for (int cnd = 0; cnd < 32; ++cnd) {
      const int p = g * 32 + cnd;
      if (p < P && genes[(size_t)p * n + i] != 0) word |= (1u << cnd);
    }
    bits[((size_t)(g / G) * n + i) * G + (g % G)] = word;
  }
}

// Same, four variants per lane (n % 4 == 0 and a 4-byte aligned population): a warp reads 128
// consecutive gene bytes of each of its 32 candidates with one coalesced 32-bit load per lane, 32
// independent loads in flight per lane.
__global__ void k_phase_bits4(const uint8_t* __restrict__ genes, int32_t P, int32_t n, int32_t groups,
                              int32_t G, uint32_t* __restrict__ bits) {
  const int64_t warp_id = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int64_t vchunks = (n + 127) / 128;
  if (warp_id >= vchunks * groups) return;
  const int g = (int)(warp_id / vchunks);
  const int32_t i0 = (int32_t)(warp_id - (int64_t)g * vchunks) * 128 + lane * 4;
  if (i0 >= n) return;
  uint32_t v[32];
#pragma unroll
  for (int cnd = 0; cnd < 32; ++cnd) {
    const int p = g * 32 + cnd;
    v[cnd] = p < P ? __ldg(reinterpret_cast<const uint32_t*>(genes + (size_t)p * n + i0)) : 0u;
  }
  uint32_t word[4] = {0, 0, 0, 0};
#pragma unroll
  for (int cnd = 0; cnd < 32; ++cnd) {
#pragma unroll
    for (int t = 0; t < 4; ++t) word[t] |= ((v[cnd] >> (8 * t)) & 0xFFu) ? (1u << cnd) : 0u;
  }
  uint32_t* out = bits + ((size_t)(g / G) * n + i0) * G + (g % G);
#pragma unroll
  for (int t = 0; t < 4; ++t) out[(size_t)t * G] = word[t];
}


// One butterfly stage of the 32 x 32 bit transpose (LSB first): rows k and k + J exchange the bits whose
// index has bit J set / clear.
template <int J>
__device__ __forceinline__ void transpose_stage(uint32_t (&a)[32], uint32_t m) {
#pragma unroll
  for (int h = 0; h < 16; ++h) {
    const int k = ((h & ~(J - 1)) << 1) | (h & (J - 1));   // the 16 rows with bit J clear
    const uint32_t t = ((a[k] >> J) ^ a[k + J]) & m;
    a[k + J] ^= t;
    a[k] ^= t << J;
  }
}

// Row-packed populations (lxy_phase_score_batch_bits): words[P][W], W = ceil(n / 32), bit (i & 31) of
// word [p][i >> 5] = gene i of candidate p.  One CTA per (block q of G groups, span of 32 words = 1,024
// variants); warp s takes group q * G + s: lane L reads word w0 + L of each of the group's 32 rows
// (128 contiguous bytes per row), transposes its 32 x 32 bit block in registers (five butterfly
// stages), and the CTA writes the [variant][G] words of the span as one contiguous run.
template <int G>
__global__ void __launch_bounds__(32 * G) k_phase_bits_packed(const uint32_t* __restrict__ words, int32_t P,
                                                              int32_t n, int32_t W, int32_t groups,
                                                              uint32_t* __restrict__ bits) {
  constexpr int kRow = 32 * G + 1;              // words of one lane's 32 variants, padded against bank conflicts
  __shared__ uint32_t tile[32 * kRow];
  const int s = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = blockIdx.y, g = q * G + s;
  const int w0 = blockIdx.x * 32;
  uint32_t a[32];
#pragma unroll
  for (int r = 0; r < 32; ++r) {
    const int p = g * 32 + r;
    a[r] = (g < groups && p < P && w0 + lane < W) ? __ldg(words + (size_t)p * W + w0 + lane) : 0u;
  }
  // a[t] bit r  <-  a[r] bit t
  transpose_stage<16>(a, 0x0000FFFFu);
  transpose_stage<8>(a, 0x00FF00FFu);
  transpose_stage<4>(a, 0x0F0F0F0Fu);
  transpose_stage<2>(a, 0x33333333u);
  transpose_stage<1>(a, 0x55555555u);
#pragma unroll
  for (int t = 0; t < 32; ++t) tile[lane * kRow + t * G + s] = a[t];
  __syncthreads();
  const int64_t i0 = (int64_t)w0 * 32;
  uint32_t* out = bits + ((size_t)q * n + (size_t)i0) * G;
  for (int k = threadIdx.x; k < 1024 * G; k += 32 * G) {
    const int vloc = k / G, ss = k - vloc * G;
    if (i0 + vloc < n) out[k] = tile[(vloc >> 5) * kRow + (vloc & 31) * G + ss];
  }
}

template <int G> struct GeneWords;
template <> struct GeneWords<1> { typedef uint32_t type; };
template <> struct GeneWords<2> { typedef uint2 type; };
template <> struct GeneWords<4> { typedef uint4 type; };
__device__ __forceinline__ uint32_t gw_get(uint32_t v, int) { return v; }
__device__ __forceinline__ uint32_t gw_get(const uint2& v, int s) { return s ? v.y : v.x; }
__device__ __forceinline__ uint32_t gw_get(const uint4& v, int s) { return s == 0 ? v.x : s == 1 ? v.y : s == 2 ? v.z : v.w; }

// One warp per (block of G 32-candidate groups, chunk of kChunk records); lane = candidate c of each of
// its G groups.  Every load is warp-uniform, and what bounds the kernel is the L1 data pipe that
// returns those loads to the 32 lanes (128 B per cycle per SM: a uniform 16-byte load costs 4
// wavefronts): the record (i, j, w) is 12 bytes per lane whatever the number of candidates, so each
// warp scores G = 4 groups per record (one 16-byte load for the gene words of j; the word of i is kept
// across the run of records with the same i).  Per group the test is one LOP3 ((bits i ^ bits j) &
// lane bit) + SEL + half an IADD3; addresses are IMAD.WIDE on the FMA pipe.
// NARROW: |w| * kChunk < 2^31, so a chunk's sum fits 32 bits.
template <int G, bool NARROW>
__global__ void __launch_bounds__(256) k_phase_score(const uint2* __restrict__ rec,
                                                     const int32_t* __restrict__ w, int64_t L,
                                                     const uint32_t* __restrict__ bits, int32_t n,
                                                     int32_t qgroups, int64_t chunks, int32_t chunk,
                                                     unsigned long long* __restrict__ D) {
  typedef typename GeneWords<G>::type words_t;
  const int lane = threadIdx.x & 31;
  const uint32_t lanebit = 1u << lane;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  int64_t unit = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  for (; unit < chunks * qgroups; unit += warps) {
    const int q = (int)(unit / chunks);
    const int64_t ch = unit - (int64_t)q * chunks;
    const int64_t lo = ch * chunk, hi = min(L, lo + chunk);
    const words_t* gb = reinterpret_cast<const words_t*>(bits) + (size_t)q * n;
    asm volatile("" : "+l"(gb));   // keep the row base one 64-bit value: gb + i is then ONE IMAD.WIDE.U32
    long long acc[G];
    int32_t acc32[G];
#pragma unroll
    for (int s = 0; s < G; ++s) { acc[s] = 0; acc32[s] = 0; }
    uint32_t prev_i = 0xFFFFFFFFu;   // records are sorted by i: its gene words are loaded once per run
    words_t word_i = words_t();
#pragma unroll 2
    for (int64_t base = lo; base < hi; base += 4) {
      // 4 records per iteration (the store is padded to a multiple of 4 with i = j = 0, w = 0)
      const uint4 r0 = __ldg(reinterpret_cast<const uint4*>(rec + base));
      const uint4 r1 = __ldg(reinterpret_cast<const uint4*>(rec + base + 2));
      const int4 ww = __ldg(reinterpret_cast<const int4*>(w + base));
      const uint32_t ii[4] = {r0.x, r0.z, r1.x, r1.z}, jj[4] = {r0.y, r0.w, r1.y, r1.w};
      const int32_t wv[4] = {ww.x, ww.y, ww.z, ww.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (ii[k] != prev_i) { word_i = __ldg(gb + ii[k]); prev_i = ii[k]; }   // warp-uniform
        const words_t word_j = __ldg(gb + jj[k]);
#pragma unroll
        for (int s = 0; s < G; ++s) {
          const uint32_t x = (gw_get(word_i, s) ^ gw_get(word_j, s)) & lanebit;
          if (NARROW) acc32[s] += x ? wv[k] : 0;
          else acc[s] += x ? (long long)wv[k] : 0ll;
        }
      }
    }
#pragma unroll
    for (int s = 0; s < G; ++s) {
      if (NARROW) acc[s] = acc32[s];
      const int p = (q * G + s) * 32 + lane;
      if (acc[s] != 0) atomicAdd(D + p, (unsigned long long)acc[s]);
    }
  }
}

__global__ void k_phase_finish(int32_t P, long long sum_w, const long long* __restrict__ D,
                               int64_t* __restrict__ out_i64, double* __restrict__ out_f64) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const long long total = sum_w - 2 * D[p];
  if (out_i64) out_i64[p] = (int64_t)(-total);
  if (out_f64) out_f64[p] = (double)(-total);
}

inline bool phase_step(int64_t d) {  // phase.go:177
  return (d >= 1 && d <= 11) || d == 15 || d == 20 || d == 25 || d == 30;
}

}  // namespace

int build_phase_links(lxy_ctx* c, int32_t n, int64_t L, const uint32_t* a, const uint32_t* b,
                      const double* w) {
  PhaseLinks& S = c->phase;
  S.ready = false;
  if (n <= 0 || L < 0 || (L > 0 && (!a || !b || !w))) { set_error("set_phase_links: bad argument"); return LXY_ERR_ARG; }
  if (n >= (1 << 27)) { set_error("set_phase_links: n_ids %d exceeds 2^27", n); return LXY_ERR_RANGE; }
  // Candidate-independent work happens once, here, on the host (one pass + a counting sort).
  std::vector<int64_t> count((size_t)n + 1, 0);
  int64_t kept = 0;
  for (int64_t l = 0; l < L; ++l) {
    if (a[l] >= (uint32_t)n || b[l] >= (uint32_t)n) { set_error("set_phase_links: a link id is outside [0, %d)", n); return LXY_ERR_ARG; }
    const double v = w[l];
    if (!(std::fabs(v) < 1073741824.0) || v != std::nearbyint(v)) {
      set_error("set_phase_links: weight %g of record %lld is not an integer below 2^30", v, (long long)l);
      return LXY_ERR_RANGE;
    }
    const int64_t i = std::min(a[l], b[l]), j = std::max(a[l], b[l]);
    if (!phase_step(j - i)) continue;
    ++count[(size_t)i + 1];
    ++kept;
  }
  for (int32_t i = 0; i < n; ++i) count[(size_t)i + 1] += count[(size_t)i];
  const int64_t padded = (kept + 3) / 4 * 4;
  std::vector<uint32_t> rec(2 * ((size_t)padded + 4), 0);   // (i, j) per record
  std::vector<int32_t> ww((size_t)padded + 4, 0);
  int64_t sum_w = 0, max_abs = 0;
  for (int64_t l = 0; l < L; ++l) {
    const int64_t i = std::min(a[l], b[l]), j = std::max(a[l], b[l]);
    if (!phase_step(j - i)) continue;
    const int64_t at = count[(size_t)i]++;
    const int64_t wm = (int64_t)w[l] * (i > 0 ? 2 : 1);  // 1 + [min(i,j) > 0], phase.go:192
    rec[2 * (size_t)at] = (uint32_t)i;
    rec[2 * (size_t)at + 1] = (uint32_t)j;
    ww[(size_t)at] = (int32_t)wm;
    sum_w += wm;
    max_abs = std::max<int64_t>(max_abs, wm < 0 ? -wm : wm);
  }
  S.n = n;
  S.L = padded;
  S.L_in = L;
  S.sum_w = sum_w;
  S.narrow = max_abs * kChunk < ((int64_t)1 << 31);
  LXY_TRY(S.rec.reserve(8 * (size_t)(padded + 4 + 128)));   // + the prefetch distance of k_phase_score
  LXY_TRY(S.w.reserve(4 * (size_t)(padded + 4 + 128)));
  LXY_CUDA(cudaMemcpyAsync(S.rec.p, rec.data(), 8 * (size_t)(padded + 4), cudaMemcpyHostToDevice, c->stream));
  LXY_CUDA(cudaMemcpyAsync(S.w.p, ww.data(), 4 * (size_t)(padded + 4), cudaMemcpyHostToDevice, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  S.ready = true;
  return LXY_OK;
}

int phase_score(lxy_ctx* c, const uint8_t* genes_dev, int32_t P, int64_t* out_i64_dev,
                double* out_f64_dev) {
  return phase_score_any(c, genes_dev, nullptr, P, out_i64_dev, out_f64_dev);
}

// genes_dev: one byte per gene, [P][n]; or packed_dev: row-packed bits, [P][ceil(n / 32)] words.
int phase_score_any(lxy_ctx* c, const uint8_t* genes_dev, const uint32_t* packed_dev, int32_t P,
                    int64_t* out_i64_dev, double* out_f64_dev) {
  PhaseLinks& S = c->phase;
  if (!S.ready) { set_error("phase score: links not set"); return LXY_ERR_STATE; }
  cudaStream_t st = c->stream;
  const int groups = (P + 31) / 32;
  const int G = groups >= 4 ? 4 : groups >= 2 ? 2 : 1;   // groups scored per warp (k_phase_score)
  const int qgroups = (groups + G - 1) / G;
  const int n = S.n;
  const size_t bits_bytes = 4 * (size_t)qgroups * G * n;
  LXY_TRY(c->d_gbits.reserve(bits_bytes));
  LXY_TRY(c->d_corr.reserve(8 * (size_t)qgroups * G * 32));
  LXY_TRY(c->d_flag.reserve(64));
  LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, st));
  LXY_CUDA(cudaEventRecord(c->ev[0], st));
  LXY_CUDA(cudaMemsetAsync(c->d_corr.p, 0, 8 * (size_t)qgroups * G * 32, st));
  if (groups % G) LXY_CUDA(cudaMemsetAsync(c->d_gbits.p, 0, bits_bytes, st));   // words of the absent groups
  if (packed_dev) {
    const int W = (n + 31) / 32;
    const dim3 pg((unsigned)((W + 31) / 32), (unsigned)qgroups);
    if (G == 4) k_phase_bits_packed<4><<<pg, 128, 0, st>>>(packed_dev, P, n, W, groups, c->d_gbits.as<uint32_t>());
    else if (G == 2) k_phase_bits_packed<2><<<pg, 64, 0, st>>>(packed_dev, P, n, W, groups, c->d_gbits.as<uint32_t>());
    else k_phase_bits_packed<1><<<pg, 32, 0, st>>>(packed_dev, P, n, W, groups, c->d_gbits.as<uint32_t>());
  } else if ((n & 3) == 0 && ((uintptr_t)genes_dev & 3) == 0) {
    const int64_t bwarps = (int64_t)((n + 127) / 128) * groups;
    k_phase_bits4<<<(unsigned)((bwarps * 32 + 255) / 256), 256, 0, st>>>(genes_dev, P, n, groups, G, c->d_gbits.as<uint32_t>());
  } else {
    const int64_t bwarps = (int64_t)((n + 31) / 32) * groups;
    k_phase_bits<<<(unsigned)((bwarps * 32 + 255) / 256), 256, 0, st>>>(genes_dev, P, n, groups, G, c->d_gbits.as<uint32_t>());
  }
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(c->ev[1], st));
  int launches = 1, grid = 0;
  if (S.L > 0) {
    // records per warp work item: kChunk, less when that would leave the SMs short of warps
    int64_t chunk = S.L * qgroups / ((int64_t)c->sm_count * 64 * 2);
    chunk = std::min<int64_t>(kChunk, std::max<int64_t>(128, chunk)) / 4 * 4;
    const int64_t chunks = (S.L + chunk - 1) / chunk;
    const int64_t units = chunks * qgroups;
    int64_t blocks = (units + 7) / 8;
    const int64_t cap = (int64_t)c->sm_count * 8;
    if (blocks > cap) blocks = cap;
    grid = (int)blocks;
    const uint2* rec = S.rec.as<uint2>();
    const int32_t* w = S.w.as<int32_t>();
    const uint32_t* bits = c->d_gbits.as<uint32_t>();
    unsigned long long* D = c->d_corr.as<unsigned long long>();
#define LXY_PHASE_LAUNCH(GG)                                                                              \
    do {                                                                                                   \
      if (S.narrow) k_phase_score<GG, true><<<grid, 256, 0, st>>>(rec, w, S.L, bits, n, qgroups, chunks, (int32_t)chunk, D); \
      else k_phase_score<GG, false><<<grid, 256, 0, st>>>(rec, w, S.L, bits, n, qgroups, chunks, (int32_t)chunk, D);        \
    } while (0)
    if (G == 4) LXY_PHASE_LAUNCH(4);
    else if (G == 2) LXY_PHASE_LAUNCH(2);
    else LXY_PHASE_LAUNCH(1);
#undef LXY_PHASE_LAUNCH
    LXY_CUDA(cudaGetLastError());
    ++launches;
  }
  LXY_CUDA(cudaEventRecord(c->ev[2], st));
  k_phase_finish<<<(P + 127) / 128, 128, 0, st>>>(P, (long long)S.sum_w, c->d_corr.as<long long>(), out_i64_dev, out_f64_dev);
  LXY_CUDA(cudaGetLastError());
  ++launches;
  LXY_CUDA(cudaEventRecord(c->ev[3], st));
  c->timing.launches = launches;
  c->timing.grid = grid;
  c->timing.link_bytes = (int64_t)qgroups * S.L * 12;
  c->timing.table_bytes = 0;
  return LXY_OK;
}

}  // namespace lxy

// ---- allocation-free variant of the packed phasing scorer -------------------------------------------
// lxy_phase_score_batch_bits_noalloc_dev: the launches of phase_score_any's packed branch with the two
// scratch buffers (the per-candidate accumulator D and the bit-sliced gene words) carved out of ONE
// file-scope device array instead of the context's growable buffers.  No reference counterpart; added
// for callers that may not allocate (graph capture, serving loops).  Every GPU that loads the library
// holds kPhaseStaticBytes until the process exits.  All calls on one device share the array: it keeps
// the previous call's contents (D is cleared on the stream at the start of every call; the gene words
// are rewritten in full), and two calls that overlap on the GPU (two contexts = two streams) race.
constexpr size_t kPhaseStaticBytes = (size_t)8 << 20;
static __device__ __align__(256) unsigned char g_phase_static[kPhaseStaticBytes];

namespace lxy {
namespace {
constexpr int kStaticMaxDev = 64;
std::mutex g_static_mu;
unsigned char* g_static_addr[kStaticMaxDev];   // device address of g_phase_static, per device ordinal

// In host code `g_phase_static` is only a handle: the device address comes from cudaGetSymbolAddress,
// asked once per device with that device current (the caller has selected it).
int phase_static_scratch(int device, unsigned char** out) {
  if (device < 0 || device >= kStaticMaxDev) { set_error("phase score (noalloc): device %d beyond %d", device, kStaticMaxDev); return LXY_ERR_RANGE; }
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_addr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_phase_static);
    if (e != cudaSuccess || !p) {
      set_error("phase score (noalloc): cudaGetSymbolAddress failed on device %d: %s", device, cudaGetErrorString(e));
      (void)cudaGetLastError();
      return LXY_ERR_CUDA;
    }
    g_static_addr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_static_addr[device];
  return LXY_OK;
}
}  // namespace
}  // namespace lxy

extern "C" int lxy_phase_score_batch_bits_noalloc_dev(lxy_ctx* c, const uint32_t* words_dev, int32_t P,
                                                      int64_t* out_i64_dev, double* out_f64_dev) {
  using namespace lxy;
  if (!c) { set_error("null context"); return LXY_ERR_ARG; }
  std::lock_guard<std::mutex> lk(c->mu);
  if (cudaSetDevice(c->cfg.device) != cudaSuccess) { set_error("cudaSetDevice(%d) failed", c->cfg.device); return LXY_ERR_CUDA; }
  if (P < 0 || (P > 0 && (!words_dev || (!out_i64_dev && !out_f64_dev)))) { set_error("phase_score_batch_bits_noalloc_dev: bad argument"); return LXY_ERR_ARG; }
  PhaseLinks& S = c->phase;
  if (!S.ready) { set_error("phase_score_batch_bits_noalloc_dev: links not set"); return LXY_ERR_STATE; }
  if (P == 0) return LXY_OK;
  const int groups = (P + 31) / 32;
  const int G = groups >= 4 ? 4 : groups >= 2 ? 2 : 1;
  const int qgroups = (groups + G - 1) / G;
  const int n = S.n;
  const size_t d_bytes = (8 * (size_t)qgroups * G * 32 + 255) & ~(size_t)255;
  const size_t bits_bytes = 4 * (size_t)qgroups * G * n;
  if (d_bytes + bits_bytes > kPhaseStaticBytes) {
    set_error("phase_score_batch_bits_noalloc_dev: %d candidates x %d variants need %zu bytes of scratch, the static "
              "array holds %zu (n_ids * P / 8 + 8 * P, P rounded up to %d); use lxy_phase_score_batch_bits_dev",
              P, n, d_bytes + bits_bytes, kPhaseStaticBytes, 32 * G);
    return LXY_ERR_RANGE;
  }
  unsigned char* scratch = nullptr;
  LXY_TRY(phase_static_scratch(c->cfg.device, &scratch));
  unsigned long long* D = reinterpret_cast<unsigned long long*>(scratch);
  uint32_t* bits = reinterpret_cast<uint32_t*>(scratch + d_bytes);
  cudaStream_t st = c->stream;
  LXY_CUDA(cudaEventRecord(c->ev[0], st));
  LXY_CUDA(cudaMemsetAsync(D, 0, d_bytes, st));   // the accumulator must start at zero: the last call's sums are still here
  const int W = (n + 31) / 32;
  const dim3 pg((unsigned)((W + 31) / 32), (unsigned)qgroups);   // writes every word of bits[qgroups][n][G]
  if (G == 4) k_phase_bits_packed<4><<<pg, 128, 0, st>>>(words_dev, P, n, W, groups, bits);
  else if (G == 2) k_phase_bits_packed<2><<<pg, 64, 0, st>>>(words_dev, P, n, W, groups, bits);
  else k_phase_bits_packed<1><<<pg, 32, 0, st>>>(words_dev, P, n, W, groups, bits);
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(c->ev[1], st));
  int launches = 1, grid = 0;
  if (S.L > 0) {
    int64_t chunk = S.L * qgroups / ((int64_t)c->sm_count * 64 * 2);   // as phase_score_any
    chunk = std::min<int64_t>(kChunk, std::max<int64_t>(128, chunk)) / 4 * 4;
    const int64_t chunks = (S.L + chunk - 1) / chunk;
    const int64_t blocks = std::min<int64_t>((chunks * qgroups + 7) / 8, (int64_t)c->sm_count * 8);
    grid = (int)blocks;
    const uint2* rec = S.rec.as<uint2>();
    const int32_t* w = S.w.as<int32_t>();
#define LXY_PHASE_LAUNCH(GG)                                                                              \
    do {                                                                                                   \
      if (S.narrow) k_phase_score<GG, true><<<grid, 256, 0, st>>>(rec, w, S.L, bits, n, qgroups, chunks, (int32_t)chunk, D); \
      else k_phase_score<GG, false><<<grid, 256, 0, st>>>(rec, w, S.L, bits, n, qgroups, chunks, (int32_t)chunk, D);        \
    } while (0)
    if (G == 4) LXY_PHASE_LAUNCH(4);
    else if (G == 2) LXY_PHASE_LAUNCH(2);
    else LXY_PHASE_LAUNCH(1);
#undef LXY_PHASE_LAUNCH
    LXY_CUDA(cudaGetLastError());
    ++launches;
  }
  LXY_CUDA(cudaEventRecord(c->ev[2], st));
  k_phase_finish<<<(P + 127) / 128, 128, 0, st>>>(P, (long long)S.sum_w, reinterpret_cast<const long long*>(D), out_i64_dev, out_f64_dev);
  LXY_CUDA(cudaGetLastError());
  ++launches;
  LXY_CUDA(cudaEventRecord(c->ev[3], st));
  c->timing.launches = launches;
  c->timing.grid = grid;
  c->timing.link_bytes = (int64_t)qgroups * S.L * 12;
  c->timing.table_bytes = 0;
  c->timing_pending = 1;
  return LXY_OK;   // asynchronous on the context's stream, like lxy_phase_score_batch_bits_dev
}
