// pair_kernels.cu — the pair-table path of the R-profile scorer (cfg.path = LXY_PATH_PAIRS).
//
// Records without in-contig offsets aggregate losslessly: the reference itself only ever sees the
// per-pair sums (Links.Add, /root/reference/src/sequtil/links.go:149-177) and scores a candidate
// with 12N-90 table look-ups (scaff.go:252-302), whatever the number of read pairs.  This path is
// that algorithm on the GPU: the sorted record store is reduced to a dense symmetric n x n table
// W (int32 counts, or fp64), and a candidate costs 6 gathers per ordering index,
//     x100(g) = sum_i (1 + [i > 0]) * sum_k [i + k < n] * T100[k] * W[g[i]][g[i + k]],
// each unordered pair taken once from its earlier member (the `i-k > 0` rule of scaff.go:256 is the
// factor 1 + [i > 0]).  Work is independent of the number of records; the bound is random 32-byte
// sector reads from L2 / HBM.
#include "common.cuh"
#include "device_utils.cuh"

namespace lxy {
namespace {

constexpr int kChunkI = 256;  // ordering indices per CTA
__constant__ int c_pstep[6] = {1, 2, 3, 5, 11, 20};
__constant__ int c_pw100[6] = {100, 50, 33, 20, 10, 5};

// One thread per record; the first record of every run of equal (tile, a, b) sums the run (fixed
// order => deterministic also in fp64) and writes both W[a][b] and W[b][a].
template <int WM>
__global__ void k_pairs_build(int64_t L, int32_t n, int32_t nblk, const uint16_t* __restrict__ rec,
                              const void* __restrict__ w, const int64_t* __restrict__ tile_start,
                              void* __restrict__ W, int* __restrict__ flag) {
  const int64_t ntiles = (int64_t)nblk * (nblk + 1) / 2;
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < L; i += stride) {
    int64_t tl = 0, th = ntiles;  // tile_start[tl] <= i < tile_start[th]
    while (th - tl > 1) {
      const int64_t mid = (tl + th) >> 1;
      if (tile_start[mid] <= i) tl = mid; else th = mid;
    }
    const uint32_t key = rec[i];
    if (i > tile_start[tl] && rec[i - 1] == key) continue;  // not the head of its run
    const int64_t tend = tile_start[tl + 1];
    int32_t I = 0;
    {
      int32_t a = 0, b = nblk;
      while (b - a > 1) {
        const int32_t m = (a + b) >> 1;
        if (tile_row_start(m, nblk) <= tl) a = m; else b = m;
      }
      I = a;
    }
    const int32_t J = I + (int32_t)(tl - tile_row_start(I, nblk));
    const size_t ga = (size_t)I * kBlk + (key >> 8), gb = (size_t)J * kBlk + (key & 0xFFu);
    if (WM == W_F64) {
      double sum = 0.0;
      for (int64_t j = i; j < tend && rec[j] == key; ++j) sum += ((const double*)w)[j];
      ((double*)W)[ga * n + gb] = sum;
      ((double*)W)[gb * n + ga] = sum;
    } else {
      long long sum = 0;
      for (int64_t j = i; j < tend && rec[j] == key; ++j) sum += (WM == W_UNIT) ? 1 : ((const int32_t*)w)[j];
      if (sum > 2147483647ll || sum < -2147483648ll) atomicOr(flag, 4);  // needs the fp64 table
      ((int32_t*)W)[ga * n + gb] = (int32_t)sum;
      ((int32_t*)W)[gb * n + ga] = (int32_t)sum;
    }
  }
}

// grid chunks * P: CTA scores ordering indices [c*256, c*256+256) of candidate p.
template <bool F64>
__global__ void __launch_bounds__(kChunkI) k_pairs_score(const int32_t* __restrict__ perm, int32_t n,
                                                        const void* __restrict__ W, int32_t chunks,
                                                        void* __restrict__ partial, int* __restrict__ flag) {
  __shared__ int32_t g[kChunkI + 20];
  __shared__ double redd[kChunkI / 32];
  __shared__ long long redi[kChunkI / 32];
  const int p = blockIdx.x / chunks, c = blockIdx.x - p * chunks, tid = threadIdx.x;
  const int i0 = c * kChunkI;
  const int32_t* row = perm + (size_t)p * n;
  for (int t = tid; t < kChunkI + 20; t += kChunkI) {
    int32_t v = -1;
    if (i0 + t < n) {
      v = row[i0 + t];
      if ((uint32_t)v >= (uint32_t)n) { atomicOr(flag, 1); v = -1; }
    }
    g[t] = v;
  }
  __syncthreads();
  const int i = i0 + tid;
  long long acci = 0;
  double accd = 0.0;
  const int32_t a = g[tid];
  if (i < n && a >= 0) {
    const size_t base = (size_t)a * n;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const int st = c_pstep[k];
      const int32_t b = g[tid + st];          // -1 beyond the end of the ordering (i + k >= n)
      if (b >= 0) {
        if (F64) accd += (double)c_pw100[k] * __ldg((const double*)W + base + b);
        else acci += (long long)c_pw100[k] * (long long)__ldg((const int32_t*)W + base + b);
      }
    }
    if (i > 0) { acci += acci; accd += accd; }   // visited from both ends unless the earlier is first
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    if (F64) accd += __shfl_xor_sync(0xffffffffu, accd, o);
    else acci += __shfl_xor_sync(0xffffffffu, acci, o);
  }
  if ((tid & 31) == 0) { redd[tid >> 5] = accd; redi[tid >> 5] = acci; }
  __syncthreads();
  if (tid == 0) {
    if (F64) {
      double s = 0.0;
      for (int q = 0; q < kChunkI / 32; ++q) s += redd[q];
      ((double*)partial)[(size_t)p * chunks + c] = s;
    } else {
      long long s = 0;
      for (int q = 0; q < kChunkI / 32; ++q) s += redi[q];
      ((long long*)partial)[(size_t)p * chunks + c] = s;
    }
  }
}

template <bool F64>
__global__ void k_pairs_finish(int32_t P, int32_t chunks, const void* __restrict__ partial,
                               double* __restrict__ out, int64_t* __restrict__ x100) {
  const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (p >= P) return;
  if (F64) {
    double s = 0.0;
    for (int c = lane; c < chunks; c += 32) s += ((const double*)partial)[(size_t)p * chunks + c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) { out[p] = -(s / 100.0); if (x100) x100[p] = INT64_MIN; }
  } else {
    long long s = 0;
    for (int c = lane; c < chunks; c += 32) s += ((const long long*)partial)[(size_t)p * chunks + c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) { out[p] = -((double)s / 100.0); if (x100) x100[p] = (int64_t)s; }
  }
}

// pos[p][g[i]] = i, then pos[p][g[i]] == i for all i  <=>  row p is a permutation
__global__ void k_pairs_pos(const int32_t* __restrict__ perm, int32_t P, int32_t n, int32_t* __restrict__ pos) {
  int64_t total = (int64_t)P * n;
  int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; idx < total; idx += stride) {
    const int32_t p = (int32_t)(idx / n), i = (int32_t)(idx - (int64_t)p * n);
    const int32_t cid = perm[idx];
    if ((uint32_t)cid < (uint32_t)n) pos[(size_t)p * n + cid] = i;
  }
}
__global__ void k_pairs_check(const int32_t* __restrict__ perm, int32_t P, int32_t n,
                              const int32_t* __restrict__ pos, int* __restrict__ flag) {
  int64_t total = (int64_t)P * n;
  int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; idx < total; idx += stride) {
    const int32_t p = (int32_t)(idx / n), i = (int32_t)(idx - (int64_t)p * n);
    const int32_t cid = perm[idx];
    if ((uint32_t)cid >= (uint32_t)n) { atomicOr(flag, 1); continue; }
    if (pos[(size_t)p * n + cid] != i) atomicOr(flag, 2);
  }
}

}  // namespace

int build_pair_table(lxy_ctx* c) {
  ScaffLinks& S = c->scaff;
  const size_t es = (S.wmode == W_F64) ? 8 : 4;
  const size_t bytes = (size_t)S.n * (size_t)S.n * es;
  if (bytes > ((size_t)48 << 30)) {
    set_error("pair table for %d ids needs %zu bytes; use the streaming path", S.n, bytes);
    return LXY_ERR_RANGE;
  }
  LXY_TRY(S.dense.reserve(bytes + 256));
  LXY_TRY(c->d_flag.reserve(64));
  LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, c->stream));
  LXY_CUDA(cudaMemsetAsync(S.dense.p, 0, bytes, c->stream));
  if (S.L > 0) {
    const int grid = grid_for(S.L, 256, c->sm_count);
    const uint16_t* rec = S.rec.as<uint16_t>();
    const int64_t* ts = S.tile_start.as<int64_t>();
    if (S.wmode == W_UNIT) k_pairs_build<W_UNIT><<<grid, 256, 0, c->stream>>>(S.L, S.n, S.nblk, rec, S.w.p, ts, S.dense.p, c->d_flag.as<int>());
    else if (S.wmode == W_INT32) k_pairs_build<W_INT32><<<grid, 256, 0, c->stream>>>(S.L, S.n, S.nblk, rec, S.w.p, ts, S.dense.p, c->d_flag.as<int>());
    else k_pairs_build<W_F64><<<grid, 256, 0, c->stream>>>(S.L, S.n, S.nblk, rec, S.w.p, ts, S.dense.p, c->d_flag.as<int>());
    LXY_CUDA(cudaGetLastError());
  }
  int flag = 0;
  LXY_CUDA(cudaMemcpyAsync(&flag, c->d_flag.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  if (flag & 4) {
    set_error("pair table: a summed pair weight exceeds int32; create the context with LXY_ACC_F64");
    return LXY_ERR_RANGE;
  }
  S.dense_ready = true;
  return LXY_OK;
}

int pairs_score(lxy_ctx* c, const int32_t* perm_dev, int32_t P, double* out_dev, int64_t* x100_dev) {
  ScaffLinks& S = c->scaff;
  if (!S.ready || !S.dense_ready) { set_error("pair-table score: table not built"); return LXY_ERR_STATE; }
  if (P <= 0) return LXY_OK;
  cudaStream_t st = c->stream;
  const int n = S.n;
  const int chunks = (n + kChunkI - 1) / kChunkI;
  LXY_TRY(c->d_flag.reserve(64));
  LXY_TRY(c->d_partial.reserve(8 * (size_t)P * chunks));
  LXY_CUDA(cudaMemsetAsync(c->d_flag.p, 0, 64, st));
  int launches = 0;
  LXY_CUDA(cudaEventRecord(c->ev[0], st));
  if (!(c->cfg.flags & LXY_FLAG_NO_VALIDATE)) {
    LXY_TRY(c->d_start.reserve(4 * (size_t)P * n));
    LXY_CUDA(cudaMemsetAsync(c->d_start.p, 0xFF, 4 * (size_t)P * n, st));
    const int g = grid_for((int64_t)P * n, 256, c->sm_count);
    k_pairs_pos<<<g, 256, 0, st>>>(perm_dev, P, n, c->d_start.as<int32_t>());
    k_pairs_check<<<g, 256, 0, st>>>(perm_dev, P, n, c->d_start.as<int32_t>(), c->d_flag.as<int>());
    launches += 2;
  }
  LXY_CUDA(cudaEventRecord(c->ev[1], st));
  const unsigned grid = (unsigned)chunks * (unsigned)P;
  if (S.wmode == W_F64) k_pairs_score<true><<<grid, kChunkI, 0, st>>>(perm_dev, n, S.dense.p, chunks, c->d_partial.p, c->d_flag.as<int>());
  else k_pairs_score<false><<<grid, kChunkI, 0, st>>>(perm_dev, n, S.dense.p, chunks, c->d_partial.p, c->d_flag.as<int>());
  ++launches;
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(c->ev[2], st));
  if (S.wmode == W_F64) k_pairs_finish<true><<<(P + 7) / 8, 256, 0, st>>>(P, chunks, c->d_partial.p, out_dev, x100_dev);
  else k_pairs_finish<false><<<(P + 7) / 8, 256, 0, st>>>(P, chunks, c->d_partial.p, out_dev, x100_dev);
  ++launches;
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(c->ev[3], st));
  c->timing.launches = launches;
  c->timing.grid = chunks * P;
  c->timing.link_bytes = 0;
  c->timing.table_bytes = (int64_t)P * n * 6 * 32;  // one 32-byte sector per gather
  return LXY_OK;
}

}  // namespace lxy
