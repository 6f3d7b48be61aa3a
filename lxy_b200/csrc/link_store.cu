// link_store.cu — builds the device link stores (one-off, per lxy_ctx_set_*links call).
//
// Scaffolding: the reference keeps sequtil.Links as map[int]map[int]float64 keyed by the smaller
// id (/root/reference/src/sequtil/links.go:14-28,130-132) and the scorer probes it per ordering
// step (scaff.go:252-302).  Here the raw records are kept and streamed; to make every streamed
// byte reusable across many candidates they are bucketed into tiles of (kBlk x kBlk) contig
// blocks and, inside a tile, sorted by (a_local, b_local) — a blocked CSR.  A record is then two
// bytes.  The sort uses cub::DeviceRadixSort (CUDA toolkit library; build time, not the hot path).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_reduce.cuh>
#include <cub/iterator/constant_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include <algorithm>
#include <cmath>
#include <vector>

#include "common.cuh"

namespace lxy {

int DevBuf::reserve(size_t bytes) {
  if (bytes <= cap) return LXY_OK;
  if (p) cudaFree(p);
  p = nullptr;
  cap = 0;
  size_t want = bytes + (bytes >> 3) + 256;
  cudaError_t e = cudaMalloc(&p, want);
  if (e != cudaSuccess) {
    e = cudaMalloc(&p, bytes);
    want = bytes;
  }
  if (e != cudaSuccess) {
    p = nullptr;
    set_error("cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    return LXY_ERR_NOMEM;
  }
  cap = want;
  return LXY_OK;
}
void DevBuf::release() {
  if (p) cudaFree(p);
  p = nullptr;
  cap = 0;
}

namespace {

__host__ __device__ inline int64_t tile_row_start(int64_t I, int64_t nblk) {
  return I * nblk - I * (I - 1) / 2;  // tiles (I, J>=I) are stored row-major
}

// key = tile << 16 | a_loc << 8 | b_loc with a < b by id (links.go:130-132 orders the same way).
__global__ void k_make_keys(int64_t L, int32_t n, int32_t nblk, const uint32_t* __restrict__ a,
                            const uint32_t* __restrict__ b, uint64_t* __restrict__ keys,
                            int* __restrict__ flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < L; i += stride) {
    uint32_t x = a[i], y = b[i];
    if (x >= (uint32_t)n || y >= (uint32_t)n) {
      atomicOr(flag, 1);
      x = 0;
      y = 0;
    }
    uint32_t lo = min(x, y), hi = max(x, y);
    int64_t I = lo >> kBlkShift, J = hi >> kBlkShift;
    uint64_t tile = (uint64_t)(tile_row_start(I, nblk) + (J - I));
    keys[i] = (tile << 16) | ((uint64_t)(lo & (kBlk - 1)) << 8) | (uint64_t)(hi & (kBlk - 1));
  }
}

__global__ void k_keys_to_rec(int64_t L, const uint64_t* __restrict__ keys,
                              uint16_t* __restrict__ rec) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < L; i += stride) rec[i] = (uint16_t)(keys[i] & 0xFFFFu);
}

// tile_start[t] = first sorted record whose tile >= t  (t in [0, ntiles]).
__global__ void k_tile_starts(int64_t L, int64_t ntiles, const uint64_t* __restrict__ keys,
                              int64_t* __restrict__ tile_start) {
  int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t > ntiles) return;
  int64_t lo = 0, hi = L;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if ((int64_t)(keys[mid] >> 16) < t) lo = mid + 1; else hi = mid;
  }
  tile_start[t] = lo;
}

inline int grid_for(int64_t n, int threads, int sm) {
  int64_t g = (n + threads - 1) / threads;
  int64_t cap = (int64_t)sm * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

int bits_for(uint64_t v) {
  int b = 0;
  while (v) { ++b; v >>= 1; }
  return b < 1 ? 1 : b;
}

// bit0: some weight != 1; bit1: some weight is not an integer below 2^31
__global__ void k_weight_class(int64_t L, const double* __restrict__ w, int* __restrict__ flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int f = 0;
  for (; i < L; i += stride) {
    const double v = w[i];
    if (v != 1.0) f |= 1;
    if (!(fabs(v) < 2147483648.0) || v != rint(v)) f |= 2;
  }
  if (f) atomicOr(flag, f);
}

template <class T>
__global__ void k_gather_w(int64_t L, const uint32_t* __restrict__ idx, const double* __restrict__ w,
                           T* __restrict__ out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < L; i += stride) out[i] = w ? (T)w[idx[i]] : (T)1;
}

__global__ void k_gather_x(int64_t L, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ a,
                           const uint32_t* __restrict__ b, const uint32_t* __restrict__ xa,
                           const uint32_t* __restrict__ xb, uint32_t* __restrict__ oxa,
                           uint32_t* __restrict__ oxb) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < L; i += stride) {
    const uint32_t k = idx[i];
    const bool swap = a[k] > b[k];  // records are stored with the smaller id first
    oxa[i] = swap ? xb[k] : xa[k];
    oxb[i] = swap ? xa[k] : xb[k];
  }
}

// Duplicate summing (LXY_PATH_AUTO): weight of sorted record i as int64
struct WeightAt {
  const double* w;
  __host__ __device__ int64_t operator()(uint32_t k) const { return (int64_t)w[k]; }
};
// aggregated weights back to int32; bit 8 of *flag: a sum does not fit
__global__ void k_narrow_w(int64_t L, const int64_t* __restrict__ agg, int32_t* __restrict__ out, int* __restrict__ flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < L; i += stride) {
    const int64_t v = agg[i];
    if (v > 2147483647ll || v < -2147483648ll) atomicOr(flag, 8);
    out[i] = (int32_t)v;
  }
}

__global__ void k_iota(int64_t L, uint32_t* __restrict__ out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < L; i += stride) out[i] = (uint32_t)i;
}

}  // namespace

// in.on_device: a, b, w, xa, xb are device pointers (len is always a host pointer).
int build_scaff_links(lxy_ctx* c, const LinkInput& in) {
  ScaffLinks& S = c->scaff;
  S.ready = false;
  S.dense_ready = false;
  const int32_t n = in.n;
  const int64_t L = in.L;
  const bool G = (in.len != nullptr);
  if (n > (1 << 24)) { set_error("set_links: n_ids %d exceeds 2^24", n); return LXY_ERR_RANGE; }
  if (L >= (int64_t)1 << 32) { set_error("set_links: more than 2^32-1 records per context"); return LXY_ERR_RANGE; }
  if (n <= 0 || L < 0 || (L > 0 && (!in.a || !in.b))) { set_error("set_links: bad argument"); return LXY_ERR_ARG; }
  if (G && (L > 0 && (!in.xa || !in.xb))) { set_error("set_links_g: null offsets"); return LXY_ERR_ARG; }
  if (G && in.unit == 0) { set_error("set_links_g: unit must be > 0"); return LXY_ERR_ARG; }
  cudaStream_t st = c->stream;
  const int sm = c->sm_count;

  S.gprofile = G;
  S.n = n;
  S.L = L;
  S.L_in = L;
  S.path = LXY_PATH_STREAM;
  S.nblk = (n + kBlk - 1) / kBlk;
  S.ntiles = (int64_t)S.nblk * (S.nblk + 1) / 2;
  S.unit = in.unit;
  S.total_len = 0;
  LXY_TRY(S.tile_start.reserve(sizeof(int64_t) * (size_t)(S.ntiles + 1)));
  LXY_TRY(S.rec.reserve(sizeof(uint16_t) * (size_t)(L + 64 + 4096)));   // + the 8 KB bulk copy of a CTA's last chunk may run past L (scaff_stream1.cuh)
  LXY_TRY(c->d_flag.reserve(64));
  LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, st));
  if (G) {
    uint64_t tot = 0;
    for (int32_t i = 0; i < n; ++i) {
      if (in.len[i] == 0) { set_error("set_links_g: contig %d has length 0", i); return LXY_ERR_ARG; }
      tot += in.len[i];
    }
    S.total_len = tot;
    if (tot > 0xFFFFFFFFull) {
      set_error("set_links_g: total scaffold length %llu exceeds the 32-bit coordinate format",
                (unsigned long long)tot);
      return LXY_ERR_RANGE;
    }
    LXY_TRY(S.len.reserve(sizeof(uint32_t) * (size_t)n));
    LXY_CUDA(cudaMemcpyAsync(S.len.p, in.len, sizeof(uint32_t) * (size_t)n, cudaMemcpyHostToDevice, st));
  }

  DevBuf& da = c->d_tmp0; DevBuf& db = c->d_tmp1; DevBuf& dk = c->d_tmp2; DevBuf& dv = c->d_tmp3;
  DevBuf dk2, dv2, dscratch, dw, dxa, dxb;
  auto cleanup = [&]() { dk2.release(); dv2.release(); dscratch.release(); dw.release(); dxa.release(); dxb.release(); };
  int rc = LXY_OK;
#define B_TRY(expr) do { rc = (expr); if (rc != LXY_OK) { cleanup(); return rc; } } while (0)
#define B_CUDA(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { set_error("%s failed: %s", #expr, cudaGetErrorString(e_)); cleanup(); return LXY_ERR_CUDA; } } while (0)

  // ---- inputs on the device ----
  const uint32_t *a = in.a, *b = in.b, *xa = in.xa, *xb = in.xb;
  const double* w = in.w;
  if (!in.on_device && L > 0) {
    B_TRY(da.reserve(4 * (size_t)L));
    B_TRY(db.reserve(4 * (size_t)L));
    B_CUDA(cudaMemcpyAsync(da.p, in.a, 4 * (size_t)L, cudaMemcpyHostToDevice, st));
    B_CUDA(cudaMemcpyAsync(db.p, in.b, 4 * (size_t)L, cudaMemcpyHostToDevice, st));
    a = da.as<uint32_t>();
    b = db.as<uint32_t>();
    if (in.w) {
      B_TRY(dw.reserve(8 * (size_t)L));
      B_CUDA(cudaMemcpyAsync(dw.p, in.w, 8 * (size_t)L, cudaMemcpyHostToDevice, st));
      w = dw.as<double>();
    }
    if (G) {
      B_TRY(dxa.reserve(4 * (size_t)L));
      B_TRY(dxb.reserve(4 * (size_t)L));
      B_CUDA(cudaMemcpyAsync(dxa.p, in.xa, 4 * (size_t)L, cudaMemcpyHostToDevice, st));
      B_CUDA(cudaMemcpyAsync(dxb.p, in.xb, 4 * (size_t)L, cudaMemcpyHostToDevice, st));
      xa = dxa.as<uint32_t>();
      xb = dxb.as<uint32_t>();
    }
  }

  // ---- weight mode (links.go stores float64; every producer in the reference adds counts) ----
  int wmode = W_UNIT;
  if (w && L > 0) {
    int* wf = c->d_flag.as<int>() + 1;
    k_weight_class<<<grid_for(L, 256, sm), 256, 0, st>>>(L, w, wf);
    int f = 0;
    B_CUDA(cudaMemcpyAsync(&f, wf, sizeof(int), cudaMemcpyDeviceToHost, st));
    B_CUDA(cudaStreamSynchronize(st));
    wmode = (f & 2) ? W_F64 : ((f & 1) ? W_INT32 : W_UNIT);
  }
  if (c->cfg.acc_mode == LXY_ACC_INT && wmode == W_F64) {
    set_error("set_links: LXY_ACC_INT requires integral weights");
    cleanup();
    return LXY_ERR_RANGE;
  }
  if (c->cfg.acc_mode == LXY_ACC_F64) wmode = W_F64;
  S.wmode = wmode;

  if (L == 0) {
    B_CUDA(cudaMemsetAsync(S.tile_start.p, 0, sizeof(int64_t) * (size_t)(S.ntiles + 1), st));
    B_CUDA(cudaMemsetAsync(S.rec.p, 0, 64 * sizeof(uint16_t), st));
    B_CUDA(cudaStreamSynchronize(st));
    cleanup();
    S.ready = true;
    ++S.version;
    drop_score_graphs(c);
    return LXY_OK;
  }

  // ---- keys, sort ----
  B_TRY(dk.reserve(8 * (size_t)L));
  B_TRY(dk2.reserve(8 * (size_t)L));
  k_make_keys<<<grid_for(L, 256, sm), 256, 0, st>>>(L, n, S.nblk, a, b, dk.as<uint64_t>(), c->d_flag.as<int>());
  B_CUDA(cudaGetLastError());
  const int end_bit = 16 + bits_for((uint64_t)S.ntiles);
  cub::DoubleBuffer<uint64_t> kb(dk.as<uint64_t>(), dk2.as<uint64_t>());
  size_t tmp_bytes = 0;
  const bool need_idx = G || wmode != W_UNIT;
  const uint32_t* sorted_idx = nullptr;
  if (!need_idx) {
    B_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, kb, (int64_t)L, 0, end_bit, st));
    B_TRY(dscratch.reserve(tmp_bytes));
    B_CUDA(cub::DeviceRadixSort::SortKeys(dscratch.p, tmp_bytes, kb, (int64_t)L, 0, end_bit, st));
  } else {
    B_TRY(dv.reserve(4 * (size_t)L));
    B_TRY(dv2.reserve(4 * (size_t)L));
    k_iota<<<grid_for(L, 256, sm), 256, 0, st>>>(L, dv.as<uint32_t>());
    cub::DoubleBuffer<uint32_t> vb(dv.as<uint32_t>(), dv2.as<uint32_t>());
    B_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, kb, vb, (int64_t)L, 0, end_bit, st));
    B_TRY(dscratch.reserve(tmp_bytes));
    B_CUDA(cub::DeviceRadixSort::SortPairs(dscratch.p, tmp_bytes, kb, vb, (int64_t)L, 0, end_bit, st));
    sorted_idx = vb.Current();
  }
  const uint64_t* sorted = kb.Current();

  // ---- LXY_PATH_AUTO: the records of a pair summed into one (Links.Add, links.go:149-177) ----
  // R-profile records carry nothing beyond their pair, so a run of equal sorted keys is one record with
  // the run's summed weight; exact in integers (skipped for real weights, whose summation order the
  // raw stream defines).  Adopted when it at least halves the stream: the weighted kernel variants pay
  // for the int32 multiply, the unit-weight ones do not.
  bool aggregated = false;
  if (c->cfg.path == LXY_PATH_AUTO && !G && wmode != W_F64 && L >= 2) {
    uint64_t* ukeys = kb.Alternate();
    DevBuf dagg, druns;
    auto drop = [&]() { dagg.release(); druns.release(); };
    rc = dagg.reserve(8 * (size_t)L);
    if (rc == LXY_OK) rc = druns.reserve(64);
    if (rc != LXY_OK) { drop(); cleanup(); return rc; }
    size_t rb = 0;
    cudaError_t e;
    if (wmode == W_UNIT) {
      cub::ConstantInputIterator<int64_t> ones(1);
      e = cub::DeviceReduce::ReduceByKey(nullptr, rb, sorted, ukeys, ones, dagg.as<int64_t>(), druns.as<int64_t>(), cub::Sum(), (int64_t)L, st);
      if (e == cudaSuccess && dscratch.reserve(rb) != LXY_OK) e = cudaErrorMemoryAllocation;
      if (e == cudaSuccess)
        e = cub::DeviceReduce::ReduceByKey(dscratch.p, rb, sorted, ukeys, ones, dagg.as<int64_t>(), druns.as<int64_t>(), cub::Sum(), (int64_t)L, st);
    } else {
      cub::TransformInputIterator<int64_t, WeightAt, const uint32_t*> wit(sorted_idx, WeightAt{w});
      e = cub::DeviceReduce::ReduceByKey(nullptr, rb, sorted, ukeys, wit, dagg.as<int64_t>(), druns.as<int64_t>(), cub::Sum(), (int64_t)L, st);
      if (e == cudaSuccess && dscratch.reserve(rb) != LXY_OK) e = cudaErrorMemoryAllocation;
      if (e == cudaSuccess)
        e = cub::DeviceReduce::ReduceByKey(dscratch.p, rb, sorted, ukeys, wit, dagg.as<int64_t>(), druns.as<int64_t>(), cub::Sum(), (int64_t)L, st);
    }
    int64_t runs = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&runs, druns.p, 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { set_error("set_links: duplicate summing failed: %s", cudaGetErrorString(e)); drop(); cleanup(); return LXY_ERR_CUDA; }
    if (runs > 0 && runs * 2 <= L) {
      rc = S.w.reserve(4 * (size_t)runs + 256);
      if (rc != LXY_OK) { drop(); cleanup(); return rc; }
      int* nf = c->d_flag.as<int>() + 2;
      k_narrow_w<<<grid_for(runs, 256, sm), 256, 0, st>>>(runs, dagg.as<int64_t>(), S.w.as<int32_t>(), nf);
      int f = 0;
      e = cudaMemcpyAsync(&f, nf, sizeof(int), cudaMemcpyDeviceToHost, st);
      if (e == cudaSuccess) e = cudaStreamSynchronize(st);
      if (e != cudaSuccess) { set_error("set_links: duplicate summing failed: %s", cudaGetErrorString(e)); drop(); cleanup(); return LXY_ERR_CUDA; }
      if (!(f & 8)) {   // every sum fits int32: adopt
        aggregated = true;
        B_CUDA(cudaMemsetAsync(S.w.as<int32_t>() + runs, 0, 256, st));
        S.L = runs;
        S.wmode = wmode = W_INT32;
        k_keys_to_rec<<<grid_for(runs, 256, sm), 256, 0, st>>>(runs, ukeys, S.rec.as<uint16_t>());
        B_CUDA(cudaMemsetAsync(S.rec.as<uint16_t>() + runs, 0, 64 * sizeof(uint16_t), st));
        k_tile_starts<<<(int)((S.ntiles + 1 + 255) / 256), 256, 0, st>>>(runs, S.ntiles, ukeys, S.tile_start.as<int64_t>());
        B_CUDA(cudaGetLastError());
        B_CUDA(cudaStreamSynchronize(st));
      }
    }
    drop();
  }

  // ---- the resident store ----
  if (!aggregated) {
  k_keys_to_rec<<<grid_for(L, 256, sm), 256, 0, st>>>(L, sorted, S.rec.as<uint16_t>());
  B_CUDA(cudaMemsetAsync(S.rec.as<uint16_t>() + L, 0, 64 * sizeof(uint16_t), st));  // vector-load pad
  k_tile_starts<<<(int)((S.ntiles + 1 + 255) / 256), 256, 0, st>>>(L, S.ntiles, sorted, S.tile_start.as<int64_t>());
  }
  if (aggregated) {
    // weights already in place
  } else if (wmode == W_INT32) {
    B_TRY(S.w.reserve(4 * (size_t)L + 256));
    k_gather_w<int32_t><<<grid_for(L, 256, sm), 256, 0, st>>>(L, sorted_idx, w, S.w.as<int32_t>());
    B_CUDA(cudaMemsetAsync(S.w.as<int32_t>() + L, 0, 256, st));
  } else if (wmode == W_F64) {
    B_TRY(S.w.reserve(8 * (size_t)L + 512));
    k_gather_w<double><<<grid_for(L, 256, sm), 256, 0, st>>>(L, sorted_idx, w, S.w.as<double>());
    B_CUDA(cudaMemsetAsync(S.w.as<double>() + L, 0, 512, st));
  }
  if (G) {
    B_TRY(S.xa.reserve(4 * (size_t)L + 256));
    B_TRY(S.xb.reserve(4 * (size_t)L + 256));
    k_gather_x<<<grid_for(L, 256, sm), 256, 0, st>>>(L, sorted_idx, a, b, xa, xb, S.xa.as<uint32_t>(), S.xb.as<uint32_t>());
    B_CUDA(cudaMemsetAsync(S.xa.as<uint32_t>() + L, 0, 256, st));
    B_CUDA(cudaMemsetAsync(S.xb.as<uint32_t>() + L, 0, 256, st));
  }
  B_CUDA(cudaGetLastError());
  int flag = 0;
  B_CUDA(cudaMemcpyAsync(&flag, c->d_flag.p, sizeof(int), cudaMemcpyDeviceToHost, st));
  B_CUDA(cudaStreamSynchronize(st));
  cleanup();
#undef B_TRY
#undef B_CUDA
  if (flag & 1) {
    set_error("set_links: a link id is outside [0, %d)", n);
    return LXY_ERR_ARG;
  }
  S.ready = true;
  ++S.version;
  drop_score_graphs(c);
  if (G) return LXY_OK;
  // ---- the path of this store ----
  bool pairs = (c->cfg.path == LXY_PATH_PAIRS);
  if (c->cfg.path == LXY_PATH_AUTO && S.L_in >= (int64_t)256 * n) {
    // the dense table costs 6n gathers per candidate whatever the number of records (0.25 ms for 1,024
    // candidates at n = 5,000); the stream ~1.35e-7 ms per record and candidate: the table wins from a
    // few hundred records per contig on, when it fits next to everything else
    const size_t bytes = (size_t)n * (size_t)n * (S.wmode == W_F64 ? 8 : 4);
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { (void)cudaGetLastError(); free_b = 0; }
    pairs = bytes <= ((size_t)16 << 30) && bytes <= free_b / 4;
  }
  if (pairs) {
    const int prc = build_pair_table(c);
    if (prc != LXY_OK) {
      if (c->cfg.path == LXY_PATH_PAIRS) { S.ready = false; return prc; }
      S.dense.release();   // AUTO: the stream is always available
      S.dense_ready = false;
    } else {
      S.path = LXY_PATH_PAIRS;
    }
  }
  return LXY_OK;
}

// Links.Get (links.go:190-221): summed weight of the records of the pair, 0 when absent.
__global__ void k_links_get(uint32_t lo, uint32_t hi, int32_t nblk, int wmode,
                            const uint16_t* __restrict__ rec, const void* __restrict__ w,
                            const int64_t* __restrict__ tile_start, double* __restrict__ out) {
  const int64_t I = lo >> kBlkShift, J = hi >> kBlkShift;
  const int64_t t = tile_row_start(I, nblk) + (J - I);
  const uint32_t key = ((lo & (kBlk - 1)) << 8) | (hi & (kBlk - 1));
  int64_t a = tile_start[t], b = tile_start[t + 1];
  const int64_t end = b;
  while (a < b) {
    int64_t m = (a + b) >> 1;
    if (rec[m] < key) a = m + 1; else b = m;
  }
  double sum = 0.0;
  for (int64_t i = a; i < end && rec[i] == key; ++i)
    sum += wmode == W_UNIT ? 1.0 : wmode == W_INT32 ? (double)((const int32_t*)w)[i] : ((const double*)w)[i];
  *out = sum;
}

int links_get(lxy_ctx* c, int32_t id1, int32_t id2, double* out) {
  ScaffLinks& S = c->scaff;
  if (!S.ready) { set_error("links_get: links not set"); return LXY_ERR_STATE; }
  if (id1 < 0 || id2 < 0 || id1 >= S.n || id2 >= S.n) {  // links.go:193-200 returns -1 + error
    *out = -1.0;
    set_error("links_get: id outside [0, %d)", S.n);
    return LXY_ERR_ARG;
  }
  LXY_TRY(c->d_out.reserve(8));
  const uint32_t lo = (uint32_t)std::min(id1, id2), hi = (uint32_t)std::max(id1, id2);
  k_links_get<<<1, 1, 0, c->stream>>>(lo, hi, S.nblk, S.wmode, S.rec.as<uint16_t>(), S.w.p,
                                      S.tile_start.as<int64_t>(), c->d_out.as<double>());
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaMemcpyAsync(out, c->d_out.p, 8, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  return LXY_OK;
}

}  // namespace lxy
