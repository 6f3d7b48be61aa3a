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
//   k_phase_score   one warp per (32-candidate group, link chunk); lane = candidate; the two gene
//                   words of a record are warp-uniform loads from a window that stays in L1
//   k_phase_finish  out = -(sumW - 2 D)
#include <algorithm>
#include <cmath>
#include <vector>

#include "common.cuh"

namespace lxy {
namespace {

constexpr int kChunk = 2048;  // records per warp work item

__global__ void k_phase_bits(const uint8_t* __restrict__ genes, int32_t P, int32_t n, int32_t groups,
                             uint32_t* __restrict__ bits) {
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
    bits[(size_t)g * n + i] = word;
  }
}

__global__ void __launch_bounds__(256) k_phase_score(const uint32_t* __restrict__ rec,
                                                     const int32_t* __restrict__ w, int64_t L,
                                                     const uint32_t* __restrict__ bits, int32_t n,
                                                     int32_t groups, int64_t chunks,
                                                     unsigned long long* __restrict__ D) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  int64_t unit = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  for (; unit < chunks * groups; unit += warps) {
    const int g = (int)(unit / chunks);
    const int64_t ch = unit - (int64_t)g * chunks;
    const int64_t lo = ch * kChunk, hi = min(L, lo + kChunk);
    const uint32_t* gb = bits + (size_t)g * n;
    long long acc = 0;
    for (int64_t base = lo; base < hi; base += 4) {
      // 4 records per 128-bit load (the store is padded to a multiple of 4)
      const uint4 r = *reinterpret_cast<const uint4*>(rec + base);
      const int4 ww = *reinterpret_cast<const int4*>(w + base);
      const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
      const int32_t wv[4] = {ww.x, ww.y, ww.z, ww.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t i = rr[k] >> 5, j = i + (rr[k] & 31u);
        const uint32_t x = (__ldg(gb + i) ^ __ldg(gb + j)) >> lane;
        acc += (x & 1u) ? (long long)wv[k] : 0ll;   // pad records carry w = 0
      }
    }
    const int p = g * 32 + lane;
    if (acc != 0) atomicAdd(D + p, (unsigned long long)acc);
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
  std::vector<uint32_t> rec((size_t)padded + 4, 0);
  std::vector<int32_t> ww((size_t)padded + 4, 0);
  int64_t sum_w = 0;
  for (int64_t l = 0; l < L; ++l) {
    const int64_t i = std::min(a[l], b[l]), j = std::max(a[l], b[l]);
    if (!phase_step(j - i)) continue;
    const int64_t at = count[(size_t)i]++;
    const int64_t wm = (int64_t)w[l] * (i > 0 ? 2 : 1);  // 1 + [min(i,j) > 0], phase.go:192
    rec[(size_t)at] = ((uint32_t)i << 5) | (uint32_t)(j - i);
    ww[(size_t)at] = (int32_t)wm;
    sum_w += wm;
  }
  S.n = n;
  S.L = padded;
  S.L_in = L;
  S.sum_w = sum_w;
  LXY_TRY(S.rec.reserve(4 * (size_t)(padded + 4)));
  LXY_TRY(S.w.reserve(4 * (size_t)(padded + 4)));
  LXY_CUDA(cudaMemcpyAsync(S.rec.p, rec.data(), 4 * (size_t)(padded + 4), cudaMemcpyHostToDevice, c->stream));
  LXY_CUDA(cudaMemcpyAsync(S.w.p, ww.data(), 4 * (size_t)(padded + 4), cudaMemcpyHostToDevice, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  S.ready = true;
  return LXY_OK;
}

int phase_score(lxy_ctx* c, const uint8_t* genes_dev, int32_t P, int64_t* out_i64_dev,
                double* out_f64_dev) {
  PhaseLinks& S = c->phase;
  if (!S.ready) { set_error("phase score: links not set"); return LXY_ERR_STATE; }
  cudaStream_t st = c->stream;
  const int groups = (P + 31) / 32;
  const int n = S.n;
  LXY_TRY(c->d_gbits.reserve(4 * (size_t)groups * n));
  LXY_TRY(c->d_corr.reserve(8 * (size_t)groups * 32));
  LXY_TRY(c->d_flag.reserve(64));
  LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, st));
  LXY_CUDA(cudaEventRecord(c->ev[0], st));
  LXY_CUDA(cudaMemsetAsync(c->d_corr.p, 0, 8 * (size_t)groups * 32, st));
  const int64_t bwarps = (int64_t)((n + 31) / 32) * groups;
  k_phase_bits<<<(unsigned)((bwarps * 32 + 255) / 256), 256, 0, st>>>(genes_dev, P, n, groups, c->d_gbits.as<uint32_t>());
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(c->ev[1], st));
  int launches = 1, grid = 0;
  if (S.L > 0) {
    const int64_t chunks = (S.L + kChunk - 1) / kChunk;
    const int64_t units = chunks * groups;
    int64_t blocks = (units + 7) / 8;
    const int64_t cap = (int64_t)c->sm_count * 8;
    if (blocks > cap) blocks = cap;
    grid = (int)blocks;
    k_phase_score<<<grid, 256, 0, st>>>(S.rec.as<uint32_t>(), S.w.as<int32_t>(), S.L, c->d_gbits.as<uint32_t>(), n, groups, chunks, c->d_corr.as<unsigned long long>());
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
  c->timing.link_bytes = (int64_t)groups * S.L * 8;
  c->timing.table_bytes = 0;
  return LXY_OK;
}

}  // namespace lxy
