// varblock_links.cu — TabulateVariantBlockLinks (/root/reference/src/sequtil/links.go:372-420), the
// accumulation step of BlockLinksFromSam (src/sequtil/alignment.go:461-555, the producer of every
// phasing `.links` file — the step directly in front of phase.score), for a batch of read pairs.
//
// Input per read pair: the two maps GetVariantBlocks built for its mates (alignment.go:265-323), as CSR
// rows {block number, #ref, #alt, #other}.  For every (k1 of mate 1, k2 of mate 2) with k1 > k2 and
// neither entry ambiguous (#other == 0) the pair ("chrom_k1", "chrom_k2") receives +1 when both blocks
// were read ref-major or both alt-major (strictly: a tie is out of phase, links.go:394-395), else -1;
// ct counts the emissions, balance sums them (the function's two return values).
//
// Device pipeline (one launch each): k_vb_count (emissions per read pair) -> exclusive scan ->
// k_vb_emit (order-preserving 64-bit key (k1, k2), value +-1) -> cub radix sort by key -> reduce by key.
// The unique (k1, k2, sum) triples come back to the host, which names them and adds them to the table.
// The host loop (device < 0) builds the same triples with a std::map.  Both then apply them in
// ascending (k1, k2) order; a name that is new to the table receives its id at that moment, id1
// before id2 (links.go:402 / :409 call ID(id1), ID(id2) in that order).  The reference ranges over Go
// maps, so its id order is random from run to run: tables compare by name.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_reduce.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include <algorithm>

#include <map>
#include <string>
#include <vector>

#include "common.cuh"
#include "links_table.h"

using namespace lxy;

namespace {

struct VbEntry { int32_t block, ref, alt, other; };

__host__ __device__ inline int vb_sign(const VbEntry& x, const VbEntry& y) {
  const bool one = x.ref > x.alt && y.ref > y.alt;   // inPhaseCaseOne  links.go:394
  const bool two = x.ref < x.alt && y.ref < y.alt;   // inPhaseCaseTwo  links.go:395
  return (one || two) ? 1 : -1;
}
__host__ __device__ inline bool vb_emits(const VbEntry& x, const VbEntry& y) {
  return x.block > y.block && x.other == 0 && y.other == 0;   // links.go:385, :392
}
// (k1, k2) as one unsigned key that sorts like the signed pair
__host__ __device__ inline uint64_t vb_key(int32_t k1, int32_t k2) {
  return ((uint64_t)((uint32_t)k1 ^ 0x80000000u) << 32) | (uint64_t)((uint32_t)k2 ^ 0x80000000u);
}

__global__ void k_vb_count(int64_t n_pairs, const int64_t* __restrict__ off1, const VbEntry* __restrict__ f1,
                           const int64_t* __restrict__ off2, const VbEntry* __restrict__ f2,
                           int64_t* __restrict__ count) {
  int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; p < n_pairs; p += stride) {
    int64_t c = 0;
    for (int64_t i = off1[p]; i < off1[p + 1]; ++i) {
      const VbEntry x = f1[i];
      for (int64_t j = off2[p]; j < off2[p + 1]; ++j) c += vb_emits(x, f2[j]) ? 1 : 0;
    }
    count[p] = c;
  }
}

__global__ void k_vb_emit(int64_t n_pairs, const int64_t* __restrict__ off1, const VbEntry* __restrict__ f1,
                          const int64_t* __restrict__ off2, const VbEntry* __restrict__ f2,
                          const int64_t* __restrict__ at, uint64_t* __restrict__ keys, int32_t* __restrict__ vals) {
  int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; p < n_pairs; p += stride) {
    int64_t o = at[p];
    for (int64_t i = off1[p]; i < off1[p + 1]; ++i) {
      const VbEntry x = f1[i];
      for (int64_t j = off2[p]; j < off2[p + 1]; ++j) {
        const VbEntry y = f2[j];
        if (!vb_emits(x, y)) continue;
        keys[o] = vb_key(x.block, y.block);
        vals[o] = vb_sign(x, y);
        ++o;
      }
    }
  }
}

struct Widen { __host__ __device__ int64_t operator()(int32_t v) const { return (int64_t)v; } };
struct Triple { int32_t k1, k2; int64_t sum; };

int apply_triples(lxy_links* l, const std::string& chrom, const std::vector<Triple>& ts) {
  for (const Triple& t : ts) {
    const std::string n1 = chrom + "_" + std::to_string(t.k1), n2 = chrom + "_" + std::to_string(t.k2);
    const int32_t id1 = lxy_links_id(l, n1.c_str(), 1);
    const int32_t id2 = lxy_links_id(l, n2.c_str(), 1);
    LXY_TRY(lxy_links_add(l, id1, id2, (double)t.sum));   // the +-1 of every read pair, summed (Links.Add)
  }
  return LXY_OK;
}

int tabulate_host(int64_t n_pairs, const int64_t* off1, const VbEntry* f1, const int64_t* off2,
                  const VbEntry* f2, std::vector<Triple>& out, int64_t& ct, int64_t& balance) {
  std::map<uint64_t, Triple> acc;
  for (int64_t p = 0; p < n_pairs; ++p)
    for (int64_t i = off1[p]; i < off1[p + 1]; ++i)
      for (int64_t j = off2[p]; j < off2[p + 1]; ++j) {
        if (!vb_emits(f1[i], f2[j])) continue;
        const int s = vb_sign(f1[i], f2[j]);
        Triple& t = acc[vb_key(f1[i].block, f2[j].block)];
        t.k1 = f1[i].block; t.k2 = f2[j].block; t.sum += s;
        ct += 1;
        balance += s;
      }
  out.reserve(acc.size());
  for (auto& kv : acc) out.push_back(kv.second);
  return LXY_OK;
}

int tabulate_device(int32_t device, int64_t n_pairs, const int64_t* off1, const VbEntry* f1,
                    const int64_t* off2, const VbEntry* f2, std::vector<Triple>& out, int64_t& ct,
                    int64_t& balance) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device >= ndev) {
    (void)cudaGetLastError();
    set_error("tabulate_varblocks: no CUDA device %d (liblxyb200 has no CPU fallback for the device path)", device);
    return LXY_ERR_CUDA;
  }
  LXY_CUDA(cudaSetDevice(device));
  const int64_t e1 = off1[n_pairs], e2 = off2[n_pairs];
  DevBuf d_off1, d_off2, d_f1, d_f2, d_cnt, d_at, d_keys, d_keys2, d_vals, d_vals2, d_tmp, d_ukeys, d_sums, d_runs;
  DevBuf* all[] = {&d_off1, &d_off2, &d_f1, &d_f2, &d_cnt, &d_at, &d_keys, &d_keys2, &d_vals, &d_vals2, &d_tmp, &d_ukeys, &d_sums, &d_runs};
  auto cleanup = [&]() { for (DevBuf* b : all) b->release(); };
#define V_TRY(expr) do { int rc_ = (expr); if (rc_ != LXY_OK) { cleanup(); return rc_; } } while (0)
#define V_CUDA(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { set_error("%s failed: %s", #expr, cudaGetErrorString(e_)); cleanup(); return LXY_ERR_CUDA; } } while (0)
  cudaStream_t st = nullptr;   // the legacy default stream: a one-off, host-synchronous step
  V_TRY(d_off1.reserve(8 * (size_t)(n_pairs + 1)));
  V_TRY(d_off2.reserve(8 * (size_t)(n_pairs + 1)));
  V_TRY(d_f1.reserve(16 * (size_t)e1 + 16));
  V_TRY(d_f2.reserve(16 * (size_t)e2 + 16));
  V_TRY(d_cnt.reserve(8 * (size_t)(n_pairs + 1)));
  V_TRY(d_at.reserve(8 * (size_t)(n_pairs + 1)));
  V_CUDA(cudaMemcpyAsync(d_off1.p, off1, 8 * (size_t)(n_pairs + 1), cudaMemcpyHostToDevice, st));
  V_CUDA(cudaMemcpyAsync(d_off2.p, off2, 8 * (size_t)(n_pairs + 1), cudaMemcpyHostToDevice, st));
  if (e1) V_CUDA(cudaMemcpyAsync(d_f1.p, f1, 16 * (size_t)e1, cudaMemcpyHostToDevice, st));
  if (e2) V_CUDA(cudaMemcpyAsync(d_f2.p, f2, 16 * (size_t)e2, cudaMemcpyHostToDevice, st));
  const int grid = (int)std::min<int64_t>((n_pairs + 255) / 256, 148 * 16);
  V_CUDA(cudaMemsetAsync(d_cnt.p, 0, 8 * (size_t)(n_pairs + 1), st));
  k_vb_count<<<grid, 256, 0, st>>>(n_pairs, d_off1.as<int64_t>(), d_f1.as<VbEntry>(), d_off2.as<int64_t>(),
                                  d_f2.as<VbEntry>(), d_cnt.as<int64_t>());
  V_CUDA(cudaGetLastError());
  size_t tb = 0;
  V_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, d_cnt.as<int64_t>(), d_at.as<int64_t>(), n_pairs + 1, st));
  V_TRY(d_tmp.reserve(tb));
  V_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tb, d_cnt.as<int64_t>(), d_at.as<int64_t>(), n_pairs + 1, st));
  int64_t E = 0;
  V_CUDA(cudaMemcpyAsync(&E, d_at.as<int64_t>() + n_pairs, 8, cudaMemcpyDeviceToHost, st));
  V_CUDA(cudaStreamSynchronize(st));
  ct = E;
  balance = 0;
  if (E == 0) { cleanup(); return LXY_OK; }
  V_TRY(d_keys.reserve(8 * (size_t)E));
  V_TRY(d_keys2.reserve(8 * (size_t)E));
  V_TRY(d_vals.reserve(4 * (size_t)E));
  V_TRY(d_vals2.reserve(4 * (size_t)E));
  k_vb_emit<<<grid, 256, 0, st>>>(n_pairs, d_off1.as<int64_t>(), d_f1.as<VbEntry>(), d_off2.as<int64_t>(),
                                 d_f2.as<VbEntry>(), d_at.as<int64_t>(), d_keys.as<uint64_t>(), d_vals.as<int32_t>());
  V_CUDA(cudaGetLastError());
  cub::DoubleBuffer<uint64_t> kb(d_keys.as<uint64_t>(), d_keys2.as<uint64_t>());
  cub::DoubleBuffer<int32_t> vb(d_vals.as<int32_t>(), d_vals2.as<int32_t>());
  V_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, kb, vb, E, 0, 64, st));
  V_TRY(d_tmp.reserve(tb));
  V_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp.p, tb, kb, vb, E, 0, 64, st));
  V_TRY(d_ukeys.reserve(8 * (size_t)E));
  V_TRY(d_sums.reserve(8 * (size_t)E));
  V_TRY(d_runs.reserve(64));
  // sums in int64: a pair can collect more than 2^31 read pairs in principle
  cub::TransformInputIterator<int64_t, Widen, const int32_t*> wide(vb.Current(), Widen());
  V_CUDA(cub::DeviceReduce::ReduceByKey(nullptr, tb, kb.Current(), d_ukeys.as<uint64_t>(), wide, d_sums.as<int64_t>(),
                                        d_runs.as<int64_t>(), cub::Sum(), E, st));
  V_TRY(d_tmp.reserve(tb));
  V_CUDA(cub::DeviceReduce::ReduceByKey(d_tmp.p, tb, kb.Current(), d_ukeys.as<uint64_t>(), wide, d_sums.as<int64_t>(),
                                        d_runs.as<int64_t>(), cub::Sum(), E, st));
  int64_t runs = 0;
  V_CUDA(cudaMemcpyAsync(&runs, d_runs.p, 8, cudaMemcpyDeviceToHost, st));
  V_CUDA(cudaStreamSynchronize(st));
  std::vector<uint64_t> hk((size_t)runs);
  std::vector<int64_t> hs((size_t)runs);
  V_CUDA(cudaMemcpy(hk.data(), d_ukeys.p, 8 * (size_t)runs, cudaMemcpyDeviceToHost));
  V_CUDA(cudaMemcpy(hs.data(), d_sums.p, 8 * (size_t)runs, cudaMemcpyDeviceToHost));
  cleanup();
#undef V_TRY
#undef V_CUDA
  out.resize((size_t)runs);
  for (int64_t r = 0; r < runs; ++r) {
    out[(size_t)r].k1 = (int32_t)((uint32_t)(hk[(size_t)r] >> 32) ^ 0x80000000u);
    out[(size_t)r].k2 = (int32_t)((uint32_t)hk[(size_t)r] ^ 0x80000000u);
    out[(size_t)r].sum = hs[(size_t)r];
    balance += hs[(size_t)r];
  }
  return LXY_OK;
}

}  // namespace

extern "C" int lxy_links_tabulate_varblocks(lxy_links* l, const char* chrom, int64_t n_pairs,
                                            const int64_t* off1, const int32_t* blk1, const int64_t* off2,
                                            const int32_t* blk2, int32_t device, int64_t* ct, int64_t* balance) {
  if (!l || !chrom || n_pairs < 0 || !off1 || !off2) { set_error("tabulate_varblocks: null argument"); return LXY_ERR_ARG; }
  for (int64_t p = 0; p < n_pairs; ++p)
    if (off1[p + 1] < off1[p] || off2[p + 1] < off2[p] || off1[0] != 0 || off2[0] != 0) {
      set_error("tabulate_varblocks: offsets must start at 0 and not decrease (read pair %lld)", (long long)p);
      return LXY_ERR_ARG;
    }
  if ((off1[n_pairs] > 0 && !blk1) || (off2[n_pairs] > 0 && !blk2)) { set_error("tabulate_varblocks: null entries"); return LXY_ERR_ARG; }
  std::vector<Triple> ts;
  int64_t c = 0, b = 0;
  if (n_pairs == 0) {
    if (ct) *ct = 0;
    if (balance) *balance = 0;
    return LXY_OK;
  }
  if (device < 0) LXY_TRY(tabulate_host(n_pairs, off1, (const VbEntry*)blk1, off2, (const VbEntry*)blk2, ts, c, b));
  else LXY_TRY(tabulate_device(device, n_pairs, off1, (const VbEntry*)blk1, off2, (const VbEntry*)blk2, ts, c, b));
  LXY_TRY(apply_triples(l, chrom, ts));
  if (ct) *ct = c;
  if (balance) *balance = b;
  return LXY_OK;
}
