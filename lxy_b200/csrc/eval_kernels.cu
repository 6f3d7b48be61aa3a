// eval_kernels.cu — the reference's solution-quality metrics for whole populations on the device.
//
//   EvalScaffolding  /root/reference/src/lxy/scaff/scaff.go:329-384   (called once per reported
//                    intermediate, scaff/cmd.go:165,195-198).  The reference walks all n^3 index
//                    triples; for i < j < k it counts the triples whose key ranks ascend (and, with the
//                    reversed key, descend), plus the same over adjacent triples, and returns the
//                    better direction's pair (all-triple fraction, adjacent-triple fraction).
//                    With r = key rank along the inferred order (a permutation) and
//                    L_j = #{i < j : r_i < r_j}:
//                       ascending  triples through j = L_j * ((n-1-r_j) - (j-L_j))
//                       descending triples through j = (j-L_j) * (r_j - L_j)
//                    so one O(n^2) count of L replaces the O(n^3) walk, in exact integers.
//   EvalPhasing      /root/reference/src/lxy/phase/phase.go:211-263.  Over all ORDERED pairs (i, j),
//                    i == j included, a pair matches when phasing and key agree on "same / different".
//                    With x_i = phasing_i xor key_i: matches = c0^2 + c1^2; the matches at distance d
//                    are 2 * #{i : x_i == x_(i+d)}.  The reference adds the comparisons at distance
//                    1, 2 AND 3 to one counter and leaves the other two at zero (phase.go:225-235);
//                    that is reproduced, since these are the numbers its callers log.
//   EvalPhasingDev   /root/reference/src/lxy/phase/phase.go:266-325, the metric `lxy phase infer`
//                    logs (phase/cmd.go:152).  As above over all ordered pairs, plus the pairs whose
//                    name indices (the integer after '_' in a block name, :279-280) differ by at most
//                    2 and at most 9.  With the blocks sorted by index, the partners of block s are
//                    one contiguous range [lo, hi) per window (two binary searches, done once per
//                    call), and with S = prefix count of x along that order
//                       matches_w = sum_s (x_s ? S[hi]-S[lo] : (hi-lo) - (S[hi]-S[lo]))
//                    so a candidate costs O(n), the whole call O(n log n + P n), in exact integers
//                    (duplicate or unparsable indices - Atoi errors read as 0 - included).
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>

#include "common.cuh"
#include "device_utils.cuh"

namespace lxy {
namespace {

__device__ __forceinline__ unsigned long long block_sum(unsigned long long v, unsigned long long* s_red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  unsigned long long t = 0;
  for (int w = 0; w < nw; ++w) t += s_red[w];
  return t;
}

// One CTA per ordering.  rank[] (key rank of the contig at each ordering index) is staged in shared
// memory as u16 (n <= 65536) and every thread counts L_j for its j's.
__global__ void __launch_bounds__(512) k_eval_scaff(const int32_t* __restrict__ perm, const int32_t* __restrict__ key_rank,
                                                    int32_t n, double* __restrict__ out, int* __restrict__ flag) {
  extern __shared__ uint16_t s_r[];
  __shared__ unsigned long long s_red[16];
  const int32_t* g = perm + (size_t)blockIdx.x * n;
  for (int32_t i = threadIdx.x; i < n; i += blockDim.x) {
    const int32_t c = g[i];
    if ((uint32_t)c >= (uint32_t)n) { atomicOr(flag, 1); s_r[i] = 0; }
    else s_r[i] = (uint16_t)key_rank[c];
  }
  __syncthreads();
  unsigned long long asc = 0, desc = 0, nasc = 0, ndesc = 0;
  for (int32_t j = threadIdx.x; j < n; j += blockDim.x) {
    const uint32_t rj = s_r[j];
    uint32_t L = 0;
    for (int32_t i = 0; i < j; ++i) L += (s_r[i] < rj) ? 1u : 0u;
    const long long before_gt = (long long)j - L;            // i < j with r_i > r_j
    const long long after_gt = (long long)(n - 1 - rj) - before_gt;
    const long long after_lt = (long long)rj - L;
    asc += (unsigned long long)((long long)L * after_gt);
    desc += (unsigned long long)(before_gt * after_lt);
    if (j >= 1 && j + 1 < n) {
      const uint32_t a = s_r[j - 1], b = s_r[j + 1];
      nasc += (a < rj && rj < b) ? 1 : 0;
      ndesc += (a > rj && rj > b) ? 1 : 0;
    }
  }
  asc = block_sum(asc, s_red);
  desc = block_sum(desc, s_red);
  nasc = block_sum(nasc, s_red);
  ndesc = block_sum(ndesc, s_red);
  if (threadIdx.x == 0) {
    // the reference accumulates float64 counters by +1: exact below 2^53, as are these conversions
    const double comparisons = (double)((unsigned long long)n * (unsigned long long)(n - 1) / 2ull *
                                        (unsigned long long)(n - 2) / 3ull);
    const double neighbors = n >= 3 ? (double)(n - 2) : 0.0;
    const double qf = (double)asc / comparisons, qr = (double)desc / comparisons;
    double* o = out + 2 * (size_t)blockIdx.x;
    if (qf > qr) { o[0] = qf; o[1] = (double)nasc / neighbors; }       // scaff.go:378-382
    else { o[0] = qr; o[1] = (double)ndesc / neighbors; }
  }
}

// One CTA per phasing.
__global__ void __launch_bounds__(256) k_eval_phase(const uint8_t* __restrict__ genes, const uint8_t* __restrict__ key,
                                                    int32_t n, double* __restrict__ out) {
  __shared__ unsigned long long s_red[8];
  const uint8_t* g = genes + (size_t)blockIdx.x * n;
  unsigned long long c1 = 0, m1 = 0, m2 = 0, m3 = 0;
  for (int32_t i = threadIdx.x; i < n; i += blockDim.x) {
    const int x = (g[i] != 0) != (key[i] != 0);
    c1 += x;
    if (i + 1 < n) m1 += (((g[i + 1] != 0) != (key[i + 1] != 0)) == x);
    if (i + 2 < n) m2 += (((g[i + 2] != 0) != (key[i + 2] != 0)) == x);
    if (i + 3 < n) m3 += (((g[i + 3] != 0) != (key[i + 3] != 0)) == x);
  }
  c1 = block_sum(c1, s_red);
  m1 = block_sum(m1, s_red);
  m2 = block_sum(m2, s_red);
  m3 = block_sum(m3, s_red);
  if (threadIdx.x == 0) {
    const double N = (double)n, c0 = N - (double)c1;
    const double matches = c0 * c0 + (double)c1 * (double)c1, comparisons = N * N;
    const double nm = 2.0 * (double)m1, snm = 2.0 * (double)m2, tnm = 2.0 * (double)m3;
    // phase.go:225-235: distances 1, 2 and 3 all count into neighbor_comparisons
    const double nc = 2.0 * ((n > 1 ? n - 1 : 0) + (n > 2 ? n - 2 : 0) + (n > 3 ? n - 3 : 0));
    double* o = out + 4 * (size_t)blockIdx.x;
    o[0] = matches / comparisons;
    o[1] = nm / nc;
    o[2] = (nm + snm) / (nc + 0.0);
    o[3] = (nm + snm + tnm) / (nc + 0.0 + 0.0);
  }
}

__global__ void k_evdev_iota(uint32_t* __restrict__ out, int32_t n) {
  const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (uint32_t)i;
}

// EvalPhasingDev, step 1: partner ranges of every sorted position for the two windows; the comparison
// counts (candidate independent) are summed into cmp[0] (<= 2) and cmp[1] (<= 9).
__global__ void __launch_bounds__(256) k_evdev_ranges(const int32_t* __restrict__ si, int32_t n, uint32_t* __restrict__ lo2,
                                                      uint32_t* __restrict__ hi2, uint32_t* __restrict__ lo9,
                                                      uint32_t* __restrict__ hi9, unsigned long long* __restrict__ cmp) {
  __shared__ unsigned long long s_red[8];
  const int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long c2 = 0, c9 = 0;
  if (s < n) {
    const long long v = si[s];
    auto lower = [&](long long x) {  // first position with si >= x
      int32_t a = 0, b = n;
      while (a < b) { const int32_t m = (a + b) >> 1; if ((long long)si[m] < x) a = m + 1; else b = m; }
      return (uint32_t)a;
    };
    const uint32_t l2 = lower(v - 2), h2 = lower(v + 3), l9 = lower(v - 9), h9 = lower(v + 10);
    lo2[s] = l2; hi2[s] = h2; lo9[s] = l9; hi9[s] = h9;
    c2 = h2 - l2;
    c9 = h9 - l9;
  }
  c2 = block_sum(c2, s_red);
  c9 = block_sum(c9, s_red);
  if (threadIdx.x == 0) { atomicAdd(cmp, c2); atomicAdd(cmp + 1, c9); }
}

// step 2: one CTA per phasing; S (n + 1 prefix counts, this CTA's slice of scratch) is written and then
// read back by the same CTA.
__global__ void __launch_bounds__(512) k_evdev_score(const uint8_t* __restrict__ genes, const uint8_t* __restrict__ key,
                                                     const uint32_t* __restrict__ order, int32_t n,
                                                     const uint32_t* __restrict__ lo2, const uint32_t* __restrict__ hi2,
                                                     const uint32_t* __restrict__ lo9, const uint32_t* __restrict__ hi9,
                                                     const unsigned long long* __restrict__ cmp, uint32_t* __restrict__ scratch,
                                                     double* __restrict__ out) {
  __shared__ unsigned long long s_red[16];
  __shared__ uint32_t s_w[16];
  __shared__ uint32_t s_carry;
  const uint8_t* g = genes + (size_t)blockIdx.x * n;
  uint32_t* S = scratch + (size_t)blockIdx.x * ((size_t)n + 1);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int32_t base = 0; base < n; base += 512) {
    const int32_t s = base + threadIdx.x;
    uint32_t x = 0;
    if (s < n) { const uint32_t i = order[s]; x = (g[i] != 0) != (key[i] != 0); }
    uint32_t incl = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += up;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    uint32_t wb = 0, all = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) { const uint32_t t = s_w[k]; if (k < warp) wb += t; all += t; }
    const uint32_t carry = s_carry;
    if (s < n) S[s] = carry + wb + incl - x;
    __syncthreads();
    if (threadIdx.x == 0) s_carry = carry + all;
    __syncthreads();
  }
  if (threadIdx.x == 0) S[n] = s_carry;
  __syncthreads();
  unsigned long long m2 = 0, m9 = 0;
  for (int32_t s = threadIdx.x; s < n; s += 512) {
    const uint32_t x = S[s + 1] - S[s];
    const uint32_t l2 = lo2[s], h2 = hi2[s], l9 = lo9[s], h9 = hi9[s];
    const uint32_t o2 = S[h2] - S[l2], o9 = S[h9] - S[l9];
    m2 += x ? o2 : (h2 - l2) - o2;
    m9 += x ? o9 : (h9 - l9) - o9;
  }
  m2 = block_sum(m2, s_red);
  m9 = block_sum(m9, s_red);
  if (threadIdx.x == 0) {
    const double N = (double)n, c1 = (double)S[n], c0 = N - c1;
    double* o = out + 3 * (size_t)blockIdx.x;
    o[0] = (c0 * c0 + c1 * c1) / (N * N);
    o[1] = (double)m2 / (double)cmp[0];
    o[2] = (double)m9 / (double)cmp[1];
  }
}

__global__ void k_check_perm(const int32_t* __restrict__ key_rank, int32_t n, int32_t* __restrict__ seen,
                             int* __restrict__ flag) {
  const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t r = key_rank[i];
  if ((uint32_t)r >= (uint32_t)n) { atomicOr(flag, 1); return; }
  if (atomicAdd(&seen[r], 1) != 0) atomicOr(flag, 2);
}

}  // namespace
}  // namespace lxy

using namespace lxy;

extern "C" {

int lxy_eval_scaffolding(lxy_ctx* c, const int32_t* perm, const int32_t* key_rank, int32_t n, int32_t P,
                         double* out) {
  if (!c) { set_error("null context"); return LXY_ERR_ARG; }
  std::lock_guard<std::mutex> lk(c->mu);
  LXY_CUDA(cudaSetDevice(c->cfg.device));
  if (n <= 0 || P < 0 || (P > 0 && (!perm || !key_rank || !out))) { set_error("eval_scaffolding: bad argument"); return LXY_ERR_ARG; }
  if (n > 65536) { set_error("eval_scaffolding: n > 65536 is not supported (u16 rank table)"); return LXY_ERR_RANGE; }
  if (P == 0) return LXY_OK;
  cudaStream_t st = c->stream;
  const size_t pbytes = 4 * (size_t)n * (size_t)P;
  LXY_TRY(c->d_perm.reserve(pbytes));
  LXY_TRY(c->d_tmp0.reserve(4 * (size_t)n));
  LXY_TRY(c->d_tmp1.reserve(4 * (size_t)n));
  LXY_TRY(c->d_out.reserve(16 * (size_t)P));
  LXY_TRY(c->d_flag.reserve(64));
  LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, st));
  LXY_CUDA(cudaMemsetAsync(c->d_tmp1.p, 0, 4 * (size_t)n, st));
  LXY_CUDA(cudaMemcpyAsync(c->d_perm.p, perm, pbytes, cudaMemcpyHostToDevice, st));
  LXY_CUDA(cudaMemcpyAsync(c->d_tmp0.p, key_rank, 4 * (size_t)n, cudaMemcpyHostToDevice, st));
  k_check_perm<<<(n + 255) / 256, 256, 0, st>>>(c->d_tmp0.as<int32_t>(), n, c->d_tmp1.as<int32_t>(), c->d_flag.as<int>());
  const int smem = 2 * n;
  static bool configured[64] = {};
  std::lock_guard<std::mutex> cfg_lock(config_mutex());
  if (!configured[c->cfg.device & 63]) {
    LXY_CUDA(cudaFuncSetAttribute(k_eval_scaff, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 65536));
    configured[c->cfg.device & 63] = true;
  }
  k_eval_scaff<<<P, 512, smem, st>>>(c->d_perm.as<int32_t>(), c->d_tmp0.as<int32_t>(), n, c->d_out.as<double>(), c->d_flag.as<int>());
  LXY_CUDA(cudaGetLastError());
  int flag = 0;
  LXY_CUDA(cudaMemcpyAsync(&flag, c->d_flag.p, sizeof(int), cudaMemcpyDeviceToHost, st));
  LXY_CUDA(cudaMemcpyAsync(out, c->d_out.p, 16 * (size_t)P, cudaMemcpyDeviceToHost, st));
  LXY_CUDA(cudaStreamSynchronize(st));
  if (flag) { set_error("eval_scaffolding: the key ranks or an ordering are not permutations of 0..n-1"); return LXY_ERR_BAD_GENE; }
  return LXY_OK;
}

int lxy_eval_phasing(lxy_ctx* c, const uint8_t* genes, const uint8_t* key, int32_t n, int32_t P, double* out) {
  if (!c) { set_error("null context"); return LXY_ERR_ARG; }
  std::lock_guard<std::mutex> lk(c->mu);
  LXY_CUDA(cudaSetDevice(c->cfg.device));
  if (n <= 0 || P < 0 || (P > 0 && (!genes || !key || !out))) { set_error("eval_phasing: bad argument"); return LXY_ERR_ARG; }
  if (P == 0) return LXY_OK;
  cudaStream_t st = c->stream;
  LXY_TRY(c->d_genes.reserve((size_t)n * (size_t)P));
  LXY_TRY(c->d_tmp0.reserve((size_t)n));
  LXY_TRY(c->d_out.reserve(32 * (size_t)P));
  LXY_CUDA(cudaMemcpyAsync(c->d_genes.p, genes, (size_t)n * (size_t)P, cudaMemcpyHostToDevice, st));
  LXY_CUDA(cudaMemcpyAsync(c->d_tmp0.p, key, (size_t)n, cudaMemcpyHostToDevice, st));
  k_eval_phase<<<P, 256, 0, st>>>(c->d_genes.as<uint8_t>(), c->d_tmp0.as<uint8_t>(), n, c->d_out.as<double>());
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaMemcpyAsync(out, c->d_out.p, 32 * (size_t)P, cudaMemcpyDeviceToHost, st));
  LXY_CUDA(cudaStreamSynchronize(st));
  return LXY_OK;
}

int lxy_eval_phasing_dev(lxy_ctx* c, const uint8_t* genes, const uint8_t* key, const int32_t* index, int32_t n,
                         int32_t P, double* out) {
  if (!c) { set_error("null context"); return LXY_ERR_ARG; }
  std::lock_guard<std::mutex> lk(c->mu);
  LXY_CUDA(cudaSetDevice(c->cfg.device));
  if (n <= 0 || P < 0 || (P > 0 && (!genes || !key || !index || !out))) { set_error("eval_phasing_dev: bad argument"); return LXY_ERR_ARG; }
  if (n > (1 << 26)) { set_error("eval_phasing_dev: n > 2^26 (the pair counts would leave the exact range of a double)"); return LXY_ERR_RANGE; }
  if (P == 0) return LXY_OK;
  cudaStream_t st = c->stream;
  const size_t N = (size_t)n;
  // index -> sorted index + original position
  LXY_TRY(c->d_tmp0.reserve(4 * N * 4));      // index in, index sorted, iota in, order out
  LXY_TRY(c->d_tmp1.reserve(4 * N * 4));      // lo2, hi2, lo9, hi9
  LXY_TRY(c->d_tmp2.reserve(N));              // key
  LXY_TRY(c->d_flag.reserve(64));
  int32_t* d_idx = c->d_tmp0.as<int32_t>();
  int32_t* d_si = d_idx + N;
  uint32_t* d_iota = reinterpret_cast<uint32_t*>(d_idx + 2 * N);
  uint32_t* d_order = d_iota + N;
  uint32_t* d_lo2 = c->d_tmp1.as<uint32_t>();
  unsigned long long* d_cmp = c->d_flag.as<unsigned long long>();
  LXY_CUDA(cudaMemsetAsync(d_cmp, 0, 16, st));
  LXY_CUDA(cudaMemcpyAsync(d_idx, index, 4 * N, cudaMemcpyHostToDevice, st));
  LXY_CUDA(cudaMemcpyAsync(c->d_tmp2.p, key, N, cudaMemcpyHostToDevice, st));
  k_evdev_iota<<<(n + 255) / 256, 256, 0, st>>>(d_iota, n);
  size_t tmp = 0;
  LXY_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const int32_t*)d_idx, d_si, (const uint32_t*)d_iota, d_order, n, 0, 32, st));
  LXY_TRY(c->d_tmp3.reserve(tmp + 16));
  LXY_CUDA(cub::DeviceRadixSort::SortPairs(c->d_tmp3.p, tmp, (const int32_t*)d_idx, d_si, (const uint32_t*)d_iota, d_order, n, 0, 32, st));
  k_evdev_ranges<<<(n + 255) / 256, 256, 0, st>>>(d_si, n, d_lo2, d_lo2 + N, d_lo2 + 2 * N, d_lo2 + 3 * N, d_cmp);
  LXY_CUDA(cudaGetLastError());
  // candidates in batches bounded by the prefix-count scratch (256 MB)
  const int32_t batch = (int32_t)std::max<size_t>(1, std::min<size_t>((size_t)P, ((size_t)64 << 20) / (N + 1)));
  LXY_TRY(c->d_genes.reserve(N * (size_t)batch));
  LXY_TRY(c->d_partial.reserve(4 * (N + 1) * (size_t)batch));
  LXY_TRY(c->d_out.reserve(24 * (size_t)batch));
  for (int32_t p0 = 0; p0 < P; p0 += batch) {
    const int32_t pb = std::min(batch, P - p0);
    LXY_CUDA(cudaMemcpyAsync(c->d_genes.p, genes + (size_t)p0 * N, N * (size_t)pb, cudaMemcpyHostToDevice, st));
    k_evdev_score<<<pb, 512, 0, st>>>(c->d_genes.as<uint8_t>(), c->d_tmp2.as<uint8_t>(), d_order, n, d_lo2, d_lo2 + N,
                                      d_lo2 + 2 * N, d_lo2 + 3 * N, d_cmp, c->d_partial.as<uint32_t>(), c->d_out.as<double>());
    LXY_CUDA(cudaGetLastError());
    LXY_CUDA(cudaMemcpyAsync(out + 3 * (size_t)p0, c->d_out.p, 24 * (size_t)pb, cudaMemcpyDeviceToHost, st));
    LXY_CUDA(cudaStreamSynchronize(st));
  }
  return LXY_OK;
}

}  // extern "C"
