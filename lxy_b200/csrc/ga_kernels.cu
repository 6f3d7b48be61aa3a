// ga_kernels.cu — the GA generation loop on the device (sm_100a): selection plan, crossover and
// mutation kernels, scoring of the children through the streaming scorers, survivor selection.
//
// Replaces, for one generation (all file:line relative to /root/reference/):
//   gao.Optimize(1) / gao.Best()         src/lxy/scaff/scaff.go:105-126, src/lxy/phase/phase.go:74-95
//   GAOrderedIntGenome.Crossover (PMX)   src/lxy/scaff/genome.go:47-91 (+ pmxmap :37-44, O(n) per
//                                        window element in the reference; O(1) here via the inverse
//                                        permutations of the population)
//   Splice / Switch / Invert / mutators  scaff/genome.go:93-97, 125-128, 161-177, 184-209
//   GAFixedBitstringGenome operators     src/lxy/phase/genome.go:37-46, 56-59, 95-109, 117-131
// The engine's random decisions are defined in include/lxy_ga_plan.h (shared with the oracle).
#include <vector>

#include "../../include/lxy_ga_plan.h"
#include "common.cuh"
#include "device_utils.cuh"

struct lxy_ga {
  lxy_ctx* ctx = nullptr;
  lxy_ga_params prm{};
  int32_t n = 0;            // gene length
  size_t esz = 4;           // bytes per gene element (int32 permutation entry / uint8 bit)
  uint64_t gen = 0;         // generations completed
  int cur = 0;              // which of the two population buffers is live
  bool ready = false;       // population scored and sorted
  bool in_step = false;     // between step_begin and step_end
  int32_t n_children = 0;
  int32_t shard_mode = -1;  // LXY_SHARD_* over the context's communicator, -1: score locally
  lxy::DevBuf pop[2], score[2], x100[2];
  lxy::DevBuf child, child_score, child_x100, plan, count, order, inv, scratch, breed;
  cudaEvent_t ev[4]{};
  lxy_ga_stats stats{};
};

namespace lxy {
namespace {

constexpr int kPlanThreads = 1024;

// ---- k_ga_plan ------------------------------------------------------------------------------------
// One CTA.  Slot p draws its (up to three) children with lxy_ga_plan_slot; a block-wide exclusive
// scan packs them in go-galib's append order (slot by slot: the two crossover children, then the
// mutation child).
__global__ void __launch_bounds__(kPlanThreads) k_ga_plan(lxy_ga_params prm, uint64_t gen, int32_t len,
                                                         lxy_ga_child* __restrict__ plan,
                                                         int32_t* __restrict__ count,
                                                         int32_t* __restrict__ breed) {
  __shared__ int32_t s_warp[kPlanThreads / 32];
  __shared__ int32_t s_base, s_nbreed;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) { s_base = 0; s_nbreed = 0; }
  __syncthreads();
  for (int32_t p0 = 0; p0 < prm.popsize; p0 += kPlanThreads) {
    const int32_t p = p0 + tid;
    lxy_ga_child k[3];
    k[0].kind = k[1].kind = k[2].kind = LXY_CHILD_NONE;
    int32_t cnt = 0;
    if (p < prm.popsize) {
      lxy_ga_plan_slot(prm.seed, gen, p, prm.popsize, len, prm.p_breed, prm.p_mutate, prm.tour_p,
                       prm.tour_k, prm.mutators, k);
      cnt = (k[0].kind != LXY_CHILD_NONE ? 2 : 0) + (k[2].kind != LXY_CHILD_NONE ? 1 : 0);
    }
    int32_t incl = cnt;  // warp inclusive scan
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int32_t w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    int32_t at = s_base + (warp ? s_warp[warp - 1] : 0) + incl - cnt;
    if (k[0].kind != LXY_CHILD_NONE) {
      breed[atomicAdd(&s_nbreed, 1)] = at;   // the crossover children, any order: k_ga_children launches over them
      plan[at++] = k[0]; plan[at++] = k[1];
    }
    if (k[2].kind != LXY_CHILD_NONE) plan[at++] = k[2];
    __syncthreads();
    if (tid == 0) s_base += s_warp[kPlanThreads / 32 - 1];
    __syncthreads();
  }
  if (tid == 0) { count[0] = s_base; count[1] = s_nbreed; }
}

// ---- k_ga_children ----------------------------------------------------------------------------------
// One CTA per child.  Mutations are an index map applied while copying the parent:
//   Switch(p1, p2)        out[p1] = in[p2], out[p2] = in[p1]                 (genome.go:125-128)
//   Shift                 out[p1] = in[p2], out[i] = in[i-1] for p1 < i <= p2  (two Splice calls, :93-97)
//   Invert, permutation   out[i] = in[p1 + p2 - i] for p1 <= i <= p2          (:161-177)
//   Invert, bits          out[i] = !in[i] for p1 <= i < p2                   (phase/genome.go:95-109)
// Crossover, bits: the windows [p1, p2] are exchanged (phase/genome.go:37-46).
// Crossover, permutations (PMX, scaff/genome.go:47-91): the windows are exchanged, then the
// reference walks i = p1..p2 and, per i, first asks whether ca[i] (= b[i]) also occurs outside the
// window of ca, then the same for cb[i] (= a[i]); hits are pushed on two stacks and, whenever both
// stacks are non-empty, the tops are popped and ca[i1] <-> cb[i2] swapped.  A swap never creates or
// removes a later hit (the values moved in come from the other parent's window), so hit k of ca is
// "inv_a[b[i]] outside the window" and its position is inv_a[b[i]]: both are known up front from the
// inverse permutations.  The CTA computes the two hit lists in parallel and one thread replays the
// stack discipline, which is what fixes the pairing.
// Largest index i in [lo, hi] with sgn * v[i] >= key, or lo - 1; eight independent loads per step (the
// one-at-a-time loop paid a shared-memory round trip per element).
__device__ __forceinline__ int32_t pmx_scan_back(const int16_t* v, int32_t hi, int32_t lo, int32_t sgn, int32_t key) {
  int32_t i = hi;
  while (i - 7 >= lo) {
    int32_t x[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) x[q] = sgn * (int32_t)v[i - q];
#pragma unroll
    for (int q = 0; q < 8; ++q)
      if (x[q] >= key) return i - q;
    i -= 8;
  }
  for (; i >= lo; --i)
    if (sgn * (int32_t)v[i] >= key) return i;
  return lo - 1;
}

template <class T, bool BITS>
__global__ void __launch_bounds__(1024) k_ga_children(const lxy_ga_child* __restrict__ plan,
                                                     const int32_t* __restrict__ count, int32_t n,
                                                     const T* __restrict__ pop, const int32_t* __restrict__ inv,
                                                     T* __restrict__ child, int32_t* __restrict__ scratch,
                                                     int use_smem, int which, const int32_t* __restrict__ breed) {
  int32_t c = blockIdx.x;
  if (which == 2) {   // launched over the crossover list (count[1] entries)
    if (c >= count[1]) return;
    c = breed[c];
  }
  if (c >= *count) return;
  const lxy_ga_child k = plan[c];
  const int tid = threadIdx.x;
  const int32_t p1 = k.p1, p2 = k.p2;
  if (k.kind == LXY_CHILD_BREED_B) return;  // written by the CTA of its sibling
  // which: 0 = every child; 1 = mutation children only; 2 = crossover children only.  The permutation
  // genome launches the two kinds separately: the ~P/5 crossover CTAs need 12 n bytes of shared memory
  // (3 CTAs per SM at n = 5,000) and tens of microseconds each, the ~3P/5 mutation CTAs are plain
  // copies - in one launch the copies queued behind the crossovers, wave after wave.
  if ((which == 1 && k.kind == LXY_CHILD_BREED_A) || (which == 2 && k.kind != LXY_CHILD_BREED_A)) return;
  const T* a = pop + (size_t)k.pa * n;
  T* out = child + (size_t)c * n;
  if (k.kind != LXY_CHILD_BREED_A) {
    for (int32_t i = tid; i < n; i += blockDim.x) {
      int32_t src = i;
      bool flip = false;
      if (k.kind == LXY_CHILD_SWITCH) src = (i == p1) ? p2 : (i == p2 ? p1 : i);
      else if (k.kind == LXY_CHILD_SHIFT) src = (i == p1) ? p2 : ((i > p1 && i <= p2) ? i - 1 : i);
      else if (BITS) flip = (i >= p1 && i < p2);
      else src = (i >= p1 && i <= p2) ? p1 + p2 - i : i;
      T v = a[src];
      if (BITS && flip) v = v ? 0 : 1;
      out[i] = v;
    }
    return;
  }
  const T* b = pop + (size_t)k.pb * n;
  T* outb = out + n;
  for (int32_t i = tid; i < n; i += blockDim.x) {
    const bool in = (i >= p1 && i <= p2);
    out[i] = in ? b[i] : a[i];
    outb[i] = in ? a[i] : b[i];
  }
  if (BITS) return;
  // ---- PMX repair ----
  const int32_t* inva = inv + (size_t)k.pa * n;
  const int32_t* invb = inv + (size_t)k.pb * n;
  const int32_t w = p2 - p1 + 1;
  // hit lists + stack: in shared memory when 3 n ints fit (the replay below is one thread walking
  // them: out of global memory every step paid a DRAM / L2 round trip), else in the global scratch
  extern __shared__ int32_t s_pmx[];
  int2* hit = reinterpret_cast<int2*>(use_smem ? s_pmx : scratch + (size_t)k.slot * 3 * n);  // (hit of ca, hit of cb) per window index: position outside the window, or -1
  int32_t* stack = reinterpret_cast<int32_t*>(hit + n);
  __shared__ int32_t s_matches;
  for (int32_t j = tid; j < w; j += blockDim.x) {
    const int32_t qa = inva[(int32_t)b[p1 + j]], qb = invb[(int32_t)a[p1 + j]];
    hit[j] = make_int2((qa < p1 || qa > p2) ? qa : -1, (qb < p1 || qb > p2) ? qb : -1);
  }
  __syncthreads();  // hit lists and both proto-children are complete
  if (use_smem) {
    // ---- the two-stack replay as a parallel bracket matching ----
    // Events in the reference's order: slot 2j = "ca[i] is duplicated outside the window" (+1), slot
    // 2j + 1 = the same for cb (-1), 0 where there is no hit; S = their prefix sum = the signed stack depth
    // (> 0: amap holds S entries, < 0: bmap holds -S).  An event that moves S towards 0 is a pop, and what
    // it pops is the most recent push that arrived at the level it leaves: the first event after the
    // LAST EARLIER SLOT u WITH S[u] == S[t] (S moves by at most 1 per slot, so everything between u and
    // t lies strictly on the far side of S[t]; u = -1, S = 0, when there is none).  Checked against the
    // stack replay on random event strings and, through the children, against the reference's O(n^2)
    // code as the CPU checker restates it (tests/).
    // Each thread owns a contiguous chunk of slots: chunk sums -> block scan -> S, chunk minima / maxima;
    // a pop searches backwards in its own chunk, then over the chunk extrema, then inside the chunk found.
    int16_t* S = reinterpret_cast<int16_t*>(hit + n);   // [2 n]; |S| <= w <= n < 2^15 on this path
    __shared__ int32_t s_wsum[32];
    __shared__ int16_t s_cmin[1024], s_cmax[1024], s_wmin[32], s_wmax[32];
    const int32_t NT = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarps = NT >> 5;
    const int32_t m2 = 2 * w, C = (m2 + NT - 1) / NT;
    const int32_t t0 = min(tid * C, m2), t1 = min(t0 + C, m2);
    auto delta = [&](int32_t t) -> int32_t {
      const int2 h = hit[t >> 1];
      return (t & 1) ? (h.y >= 0 ? -1 : 0) : (h.x >= 0 ? 1 : 0);
    };
    int32_t loc = 0;
    for (int32_t t = t0; t < t1; ++t) loc += delta(t);
    int32_t incl = loc;   // block exclusive scan of the chunk sums: shuffle scan per warp, then over the warps
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    int32_t run = incl - loc;
    for (int32_t q = 0; q < warp; ++q) run += s_wsum[q];
    int32_t mn = 32767, mx = -32768;
    for (int32_t t = t0; t < t1; ++t) {
      run += delta(t);
      S[t] = (int16_t)run;
      mn = min(mn, run);
      mx = max(mx, run);
    }
    s_cmin[tid] = (int16_t)mn;
    s_cmax[tid] = (int16_t)mx;
    const int32_t wmn = __reduce_min_sync(0xffffffffu, mn), wmx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0) { s_wmin[warp] = (int16_t)wmn; s_wmax[warp] = (int16_t)wmx; }
    __syncthreads();
    (void)nwarps;
    for (int32_t t = t0; t < t1; ++t) {
      const int32_t d = delta(t);
      if (d == 0) continue;
      const int32_t cur = S[t], prev = cur - d;
      if (!((d < 0 && prev > 0) || (d > 0 && prev < 0))) continue;   // a push
      // d < 0: leaving a positive level, everything since u is > cur (look for S[u] <= cur);
      // d > 0: leaving a negative level, everything since u is < cur (look for S[u] >= cur).
      // With sgn = d both read "sgn * S[u] >= sgn * cur".  Search order: own chunk, the earlier chunks of
      // the own warp (chunk extrema), the earlier warps (warp extrema), then down again.
      const int32_t sgn = d, key = sgn * cur;
      int32_t u = pmx_scan_back(S, t - 1, t0, sgn, key);
      if (u < t0) {
        const int16_t* ext = d < 0 ? s_cmin : s_cmax;
        int32_t c = pmx_scan_back(ext, tid - 1, warp * 32, sgn, key);
        if (c < warp * 32) {
          const int32_t wq = pmx_scan_back(d < 0 ? s_wmin : s_wmax, warp - 1, 0, sgn, key);
          c = wq < 0 ? -1 : pmx_scan_back(ext, wq * 32 + 31, wq * 32, sgn, key);
        }
        u = c < 0 ? -1 : pmx_scan_back(S, min((c + 1) * C, m2) - 1, c * C, sgn, key);
      }
      int32_t pslot = u + 1;
      while (delta(pslot) == 0) ++pslot;                  // the push: first event after u
      // ca[i1], cb[i2] = cb[i2], ca[i1]  with i1 the ca-side hit, i2 the cb-side hit (genome.go:73-85)
      const int32_t i1 = d > 0 ? hit[t >> 1].x : hit[pslot >> 1].x;
      const int32_t i2 = d > 0 ? hit[pslot >> 1].y : hit[t >> 1].y;
      out[i1] = b[i2];
      outb[i2] = a[i1];
    }
    return;
  }
  if (tid == 0) {
    // The replay only PAIRS the hits (which a-side hit meets which b-side hit is what the reference's
    // two stacks decide); the swaps themselves are independent of one another and are applied by the
    // whole CTA afterwards.  Match m is written over slot m of the hit list: after step j at most
    // j + 1 matches exist (a match consumes two of the <= 2 (j + 1) events seen), so the slot has been
    // read.  Eight steps' hits are fetched before any of them is replayed: the loads overlap, and the
    // chain that remains is the depth counter (the popped values only feed the match list).
    int32_t depth = 0;   // > 0: `depth` entries of amap on the stack; < 0: -depth entries of bmap
    int32_t m = 0;
    for (int32_t j0 = 0; j0 < w; j0 += 8) {
      int2 h[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) h[u] = (j0 + u < w) ? hit[j0 + u] : make_int2(-1, -1);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int32_t ha = h[u].x, hb = h[u].y;
        if (ha >= 0) {
          if (depth < 0) {                 // bmap non-empty: pop it, pair with this hit
            const int32_t i2 = stack[-depth - 1];
            ++depth;
            hit[m++] = make_int2(ha, i2);  // ca[i1], cb[i2] = cb[i2], ca[i1]
          } else {
            stack[depth++] = ha;
          }
        }
        if (hb >= 0) {
          if (depth > 0) {                 // amap non-empty
            const int32_t i1 = stack[depth - 1];
            --depth;
            hit[m++] = make_int2(i1, hb);
          } else {
            stack[-depth] = hb;
            --depth;
          }
        }
      }
    }
    s_matches = m;
  }
  __syncthreads();
  for (int32_t q = tid; q < s_matches; q += blockDim.x) {
    const int32_t i1 = hit[q].x, i2 = hit[q].y;
    out[i1] = b[i2];
    outb[i2] = a[i1];
  }
}

// GARandomInitializer: copy of `first` + Randomize() (n random transpositions, scaff/genome.go:130-138;
// n random bits, phase/genome.go:61-72).  Draws of individual p: generation ~0, slot p.
template <class T, bool BITS>
__global__ void k_ga_randomize(lxy_ga_params prm, int32_t n, const T* __restrict__ first, int randomize,
                               T* __restrict__ pop) {
  const int32_t p = blockIdx.x;
  T* g = pop + (size_t)p * n;
  for (int32_t i = threadIdx.x; i < n; i += blockDim.x) g[i] = first ? first[i] : (BITS ? (T)0 : (T)i);
  __syncthreads();
  if (!randomize) return;
  if (BITS) {
    for (int32_t i = threadIdx.x; i < n; i += blockDim.x)
      g[i] = (T)(lxy_ga_intn(lxy_ga_rand(prm.seed, ~0ull, (uint64_t)p, (uint64_t)i), 2) == 1);
  } else if (threadIdx.x == 0) {
    for (int32_t i = 0; i < n; ++i) {  // sequential by definition: each swap sees the previous ones
      const int32_t x = lxy_ga_intn(lxy_ga_rand(prm.seed, ~0ull, (uint64_t)p, 2ull * i), n);
      const int32_t y = lxy_ga_intn(lxy_ga_rand(prm.seed, ~0ull, (uint64_t)p, 2ull * i + 1), n);
      const T t = g[x]; g[x] = g[y]; g[y] = t;
    }
  }
}

// ---- survivor selection ---------------------------------------------------------------------------
// go-galib sorts the grown population by Score() and keeps the first popsize (lower is better,
// scaff.go:308).  rank(i) = #{j : s_j < s_i or (s_j == s_i and j < i)} over population ++ children
// (stable: ties keep append order); element i survives iff rank(i) < popsize.  M <= 4 * popsize,
// so the O(M^2) count is a few million comparisons out of shared memory.
__global__ void __launch_bounds__(256) k_ga_rank(int32_t P, const int32_t* __restrict__ count,
                                                 const double* __restrict__ s_pop,
                                                 const double* __restrict__ s_child,
                                                 int32_t* __restrict__ order) {
  __shared__ double tile[1024];
  const int32_t M = P + *count;
  const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const double si = i < P ? s_pop[i < P ? i : 0] : (i < M ? s_child[i - P] : 0.0);
  int32_t rank = 0;
  for (int32_t j0 = 0; j0 < M; j0 += 1024) {
    for (int32_t j = threadIdx.x; j < 1024; j += blockDim.x) {
      const int32_t jj = j0 + j;
      tile[j] = jj < P ? s_pop[jj] : (jj < M ? s_child[jj - P] : 0.0);
    }
    __syncthreads();
    const int32_t lim = min(1024, M - j0);
    for (int32_t j = 0; j < lim; ++j) {
      const double sj = tile[j];
      rank += (sj < si || (sj == si && j0 + j < i)) ? 1 : 0;
    }
    __syncthreads();
  }
  if (i < M && rank < P) order[rank] = i;
}

template <class T>
__global__ void __launch_bounds__(256) k_ga_gather(int32_t P, int32_t n, const int32_t* __restrict__ order,
                                                   const T* __restrict__ pop, const T* __restrict__ child,
                                                   const double* __restrict__ s_pop, const double* __restrict__ s_child,
                                                   const int64_t* __restrict__ x_pop, const int64_t* __restrict__ x_child,
                                                   T* __restrict__ npop, double* __restrict__ ns, int64_t* __restrict__ nx) {
  const int32_t r = blockIdx.x, src = order[r];
  const T* g = src < P ? pop + (size_t)src * n : child + (size_t)(src - P) * n;
  T* o = npop + (size_t)r * n;
  for (int32_t i = threadIdx.x; i < n; i += blockDim.x) o[i] = g[i];
  if (threadIdx.x == 0) {
    ns[r] = src < P ? s_pop[src] : s_child[src - P];
    nx[r] = src < P ? x_pop[src] : x_child[src - P];
  }
}

}  // namespace
}  // namespace lxy

using namespace lxy;

namespace {

struct GaGuard {  // the engine shares its context's mutex and device
  std::lock_guard<std::mutex> lk;
  bool ok;
  explicit GaGuard(lxy_ga* g) : lk(g->ctx->mu), ok(cudaSetDevice(g->ctx->cfg.device) == cudaSuccess) {
    if (!ok) set_error("cudaSetDevice(%d) failed", g->ctx->cfg.device);
  }
};
#define CHECK_GA(g)                                               \
  if (!(g)) { set_error("null GA engine"); return LXY_ERR_ARG; }  \
  GaGuard guard_(g);                                              \
  if (!guard_.ok) return LXY_ERR_CUDA;

// the local scorer on rows [first, first + count) of `genes`
int ga_score(lxy_ga* g, const void* genes, int32_t first, int32_t count, double* out, int64_t* x) {
  if (count <= 0) return LXY_OK;
  lxy_ctx* c = g->ctx;
  const char* rows = (const char*)genes + (size_t)first * g->n * g->esz;
  g->stats.scored += count;
  if (g->prm.kind == LXY_GA_PHASE) return phase_score(c, (const uint8_t*)rows, count, x + first, out + first);
  if (g->shard_mode >= 0)   // every rank holds the same rows; only the scores cross NVLink (comm.cu)
    return sharded_score_dev(c, g->shard_mode, (const int32_t*)rows, count, out + first, x + first, false);
  return scaff_dispatch(c, (const int32_t*)rows, count, out + first, x + first);
}

int ga_check_flags(lxy_ga* g, const char* what) {
  int flag = 0;
  lxy_ctx* c = g->ctx;
  LXY_CUDA(cudaMemcpyAsync(&flag, c->d_flag.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  if (flag & 3) { set_error("%s: a population row is not a valid genome", what); return LXY_ERR_BAD_GENE; }
  return LXY_OK;
}

// score the live population, sort it best-first (the plan reads ranks as indices)
int ga_settle(lxy_ga* g) {
  lxy_ctx* c = g->ctx;
  const int32_t P = g->prm.popsize;
  const int cur = g->cur, nxt = cur ^ 1;
  LXY_TRY(ga_score(g, g->pop[cur].p, 0, P, g->score[cur].as<double>(), g->x100[cur].as<int64_t>()));
  LXY_TRY(ga_check_flags(g, "GA population"));
  LXY_CUDA(cudaMemsetAsync(g->count.p, 0, sizeof(int32_t), c->stream));
  k_ga_rank<<<(P + 255) / 256, 256, 0, c->stream>>>(P, g->count.as<int32_t>(), g->score[cur].as<double>(),
                                                   g->child_score.as<double>(), g->order.as<int32_t>());
  if (g->prm.kind == LXY_GA_PHASE)
    k_ga_gather<uint8_t><<<P, 256, 0, c->stream>>>(P, g->n, g->order.as<int32_t>(), g->pop[cur].as<uint8_t>(), g->child.as<uint8_t>(),
        g->score[cur].as<double>(), g->child_score.as<double>(), g->x100[cur].as<int64_t>(), g->child_x100.as<int64_t>(),
        g->pop[nxt].as<uint8_t>(), g->score[nxt].as<double>(), g->x100[nxt].as<int64_t>());
  else
    k_ga_gather<int32_t><<<P, 256, 0, c->stream>>>(P, g->n, g->order.as<int32_t>(), g->pop[cur].as<int32_t>(), g->child.as<int32_t>(),
        g->score[cur].as<double>(), g->child_score.as<double>(), g->x100[cur].as<int64_t>(), g->child_x100.as<int64_t>(),
        g->pop[nxt].as<int32_t>(), g->score[nxt].as<double>(), g->x100[nxt].as<int64_t>());
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  g->cur = nxt;
  g->ready = true;
  return LXY_OK;
}

int ga_begin(lxy_ga* g) {
  lxy_ctx* c = g->ctx;
  if (!g->ready) { set_error("GA: population not initialised"); return LXY_ERR_STATE; }
  if (g->in_step) { set_error("GA: step_begin called twice"); return LXY_ERR_STATE; }
  const int32_t P = g->prm.popsize;
  cudaStream_t st = c->stream;
  LXY_CUDA(cudaEventRecord(g->ev[0], st));
  k_ga_plan<<<1, kPlanThreads, 0, st>>>(g->prm, g->gen, g->n, g->plan.as<lxy_ga_child>(), g->count.as<int32_t>(), g->breed.as<int32_t>());
  if (g->prm.kind == LXY_GA_PHASE) {
    k_ga_children<uint8_t, true><<<3 * P, 256, 0, st>>>(g->plan.as<lxy_ga_child>(), g->count.as<int32_t>(), g->n,
        g->pop[g->cur].as<uint8_t>(), nullptr, g->child.as<uint8_t>(), nullptr, 0, 0, nullptr);
  } else {
    LXY_TRY(scaff_positions(c, g->pop[g->cur].as<int32_t>(), P, g->n, g->inv.as<int32_t>()));
    // PMX hit lists + stack in shared memory when they fit (3 n ints)
    const size_t pmx_smem = 12 * (size_t)g->n;
    const bool in_smem = pmx_smem + 4096 <= (size_t)c->max_smem_optin && g->n < 32768;
    if (in_smem && pmx_smem > 48 * 1024) {
      static size_t configured[64] = {};
      std::lock_guard<std::mutex> cfg_lock(config_mutex());
      if (configured[c->cfg.device & 63] < pmx_smem) {
        LXY_CUDA(cudaFuncSetAttribute(k_ga_children<int32_t, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pmx_smem));
        configured[c->cfg.device & 63] = pmx_smem;
      }
    }
    k_ga_children<int32_t, false><<<P, 1024, in_smem ? pmx_smem : 0, st>>>(g->plan.as<lxy_ga_child>(), g->count.as<int32_t>(), g->n,
        g->pop[g->cur].as<int32_t>(), g->inv.as<int32_t>(), g->child.as<int32_t>(), g->scratch.as<int32_t>(), in_smem ? 1 : 0, 2,
        g->breed.as<int32_t>());
    k_ga_children<int32_t, false><<<3 * P, 256, 0, st>>>(g->plan.as<lxy_ga_child>(), g->count.as<int32_t>(), g->n,
        g->pop[g->cur].as<int32_t>(), g->inv.as<int32_t>(), g->child.as<int32_t>(), g->scratch.as<int32_t>(), 0, 1, nullptr);
  }
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(g->ev[1], st));
  // the child count is the one value the host needs per generation (launch shape of the scorer)
  LXY_CUDA(cudaMemcpyAsync(&g->n_children, g->count.p, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  LXY_CUDA(cudaStreamSynchronize(st));
  g->in_step = true;
  return LXY_OK;
}

int ga_end(lxy_ga* g) {
  lxy_ctx* c = g->ctx;
  if (!g->in_step) { set_error("GA: step_end without step_begin"); return LXY_ERR_STATE; }
  const int32_t P = g->prm.popsize, M = P + g->n_children;
  const int cur = g->cur, nxt = cur ^ 1;
  cudaStream_t st = c->stream;
  LXY_CUDA(cudaEventRecord(g->ev[2], st));
  k_ga_rank<<<(M + 255) / 256, 256, 0, st>>>(P, g->count.as<int32_t>(), g->score[cur].as<double>(),
                                            g->child_score.as<double>(), g->order.as<int32_t>());
  if (g->prm.kind == LXY_GA_PHASE)
    k_ga_gather<uint8_t><<<P, 256, 0, st>>>(P, g->n, g->order.as<int32_t>(), g->pop[cur].as<uint8_t>(), g->child.as<uint8_t>(),
        g->score[cur].as<double>(), g->child_score.as<double>(), g->x100[cur].as<int64_t>(), g->child_x100.as<int64_t>(),
        g->pop[nxt].as<uint8_t>(), g->score[nxt].as<double>(), g->x100[nxt].as<int64_t>());
  else
    k_ga_gather<int32_t><<<P, 256, 0, st>>>(P, g->n, g->order.as<int32_t>(), g->pop[cur].as<int32_t>(), g->child.as<int32_t>(),
        g->score[cur].as<double>(), g->child_score.as<double>(), g->x100[cur].as<int64_t>(), g->child_x100.as<int64_t>(),
        g->pop[nxt].as<int32_t>(), g->score[nxt].as<double>(), g->x100[nxt].as<int64_t>());
  LXY_CUDA(cudaGetLastError());
  LXY_CUDA(cudaEventRecord(g->ev[3], st));
  LXY_CUDA(cudaEventSynchronize(g->ev[3]));
  float a = 0, b = 0, d = 0, t = 0;
  cudaEventElapsedTime(&a, g->ev[0], g->ev[1]);
  cudaEventElapsedTime(&b, g->ev[1], g->ev[2]);
  cudaEventElapsedTime(&d, g->ev[2], g->ev[3]);
  cudaEventElapsedTime(&t, g->ev[0], g->ev[3]);
  g->stats.ms_operators = a; g->stats.ms_score = b; g->stats.ms_select = d; g->stats.ms_last = t;
  g->stats.last_children = g->n_children;
  g->stats.generations += 1;
  g->cur = nxt;
  g->gen += 1;
  g->in_step = false;
  return LXY_OK;
}

}  // namespace

extern "C" {

int lxy_ga_create(lxy_ctx* ctx, const lxy_ga_params* prm, lxy_ga** out) {
  if (!ctx || !prm || !out) { set_error("ga_create: null argument"); return LXY_ERR_ARG; }
  *out = nullptr;
  if (prm->kind != LXY_GA_SCAFF && prm->kind != LXY_GA_PHASE) { set_error("ga_create: bad kind"); return LXY_ERR_ARG; }
  if (prm->popsize < 2 || prm->popsize > (1 << 20)) { set_error("ga_create: popsize out of range"); return LXY_ERR_ARG; }
  if (prm->tour_k < 2 || prm->tour_k > 16 || !(prm->tour_p > 0.0)) { set_error("ga_create: tournament needs 2..16 contestants and p > 0"); return LXY_ERR_ARG; }
  if (prm->p_breed < 0 || prm->p_mutate < 0) { set_error("ga_create: negative probability"); return LXY_ERR_ARG; }
  std::lock_guard<std::mutex> lk(ctx->mu);
  if (cudaSetDevice(ctx->cfg.device) != cudaSuccess) { set_error("cudaSetDevice failed"); return LXY_ERR_CUDA; }
  int32_t n = 0;
  if (prm->kind == LXY_GA_SCAFF) {
    if (!ctx->scaff.ready || ctx->scaff.gprofile) { set_error("ga_create: set the scaffolding links first (lxy_ctx_set_links)"); return LXY_ERR_STATE; }
    n = ctx->scaff.n;
  } else {
    if (!ctx->phase.ready) { set_error("ga_create: set the phasing links first (lxy_ctx_set_phase_links)"); return LXY_ERR_STATE; }
    n = ctx->phase.n;
  }
  lxy_ga* g = new (std::nothrow) lxy_ga();
  if (!g) { set_error("out of memory"); return LXY_ERR_NOMEM; }
  g->ctx = ctx;
  g->prm = *prm;
  g->n = n;
  g->esz = prm->kind == LXY_GA_PHASE ? 1 : 4;
  const size_t P = (size_t)prm->popsize, row = (size_t)n * g->esz;
  int rc = LXY_OK;
  for (int b = 0; b < 2 && rc == LXY_OK; ++b) {
    rc = g->pop[b].reserve(P * row + 16);
    if (rc == LXY_OK) rc = g->score[b].reserve(8 * P);
    if (rc == LXY_OK) rc = g->x100[b].reserve(8 * P);
  }
  if (rc == LXY_OK) rc = g->child.reserve(3 * P * row + 16);
  if (rc == LXY_OK) rc = g->child_score.reserve(8 * (3 * P + 64));
  if (rc == LXY_OK) rc = g->child_x100.reserve(8 * (3 * P + 64));
  if (rc == LXY_OK) rc = g->plan.reserve(sizeof(lxy_ga_child) * 3 * P);
  if (rc == LXY_OK) rc = g->count.reserve(64);
  if (rc == LXY_OK) rc = g->order.reserve(4 * P);
  if (rc == LXY_OK) rc = g->breed.reserve(4 * P + 16);
  if (rc == LXY_OK && prm->kind == LXY_GA_SCAFF) {
    rc = g->inv.reserve(4 * P * (size_t)n);
    if (rc == LXY_OK) rc = g->scratch.reserve(4 * P * 3 * (size_t)n);  // hit lists + stack per breeding slot
  }
  for (auto& e : g->ev)
    if (rc == LXY_OK && cudaEventCreate(&e) != cudaSuccess) { set_error("cudaEventCreate failed"); rc = LXY_ERR_CUDA; }
  if (rc != LXY_OK) { lxy_ga_destroy(g); return rc; }
  *out = g;
  return LXY_OK;
}

void lxy_ga_destroy(lxy_ga* g) {
  if (!g) return;
  {
    std::lock_guard<std::mutex> lk(g->ctx->mu);
    cudaSetDevice(g->ctx->cfg.device);
    cudaStreamSynchronize(g->ctx->stream);
    DevBuf* bufs[] = {&g->pop[0], &g->pop[1], &g->score[0], &g->score[1], &g->x100[0], &g->x100[1], &g->child,
                      &g->child_score, &g->child_x100, &g->plan, &g->count, &g->order, &g->inv, &g->scratch, &g->breed};
    for (DevBuf* b : bufs) b->release();
    for (auto& e : g->ev) if (e) cudaEventDestroy(e);
  }
  delete g;
}

int lxy_ga_init(lxy_ga* g, const void* first, int32_t randomize) {
  CHECK_GA(g);
  lxy_ctx* c = g->ctx;
  const size_t row = (size_t)g->n * g->esz;
  void* dfirst = nullptr;
  if (first) {
    LXY_TRY(c->d_tmp0.reserve(row));
    LXY_CUDA(cudaMemcpyAsync(c->d_tmp0.p, first, row, cudaMemcpyHostToDevice, c->stream));
    dfirst = c->d_tmp0.p;
  }
  if (g->prm.kind == LXY_GA_PHASE)
    k_ga_randomize<uint8_t, true><<<g->prm.popsize, 256, 0, c->stream>>>(g->prm, g->n, (const uint8_t*)dfirst, randomize, g->pop[g->cur].as<uint8_t>());
  else
    k_ga_randomize<int32_t, false><<<g->prm.popsize, 256, 0, c->stream>>>(g->prm, g->n, (const int32_t*)dfirst, randomize, g->pop[g->cur].as<int32_t>());
  LXY_CUDA(cudaGetLastError());
  g->gen = 0;
  g->in_step = false;
  return ga_settle(g);
}

int lxy_ga_set_population(lxy_ga* g, const void* genes) {
  CHECK_GA(g);
  if (!genes) { set_error("ga_set_population: null genes"); return LXY_ERR_ARG; }
  lxy_ctx* c = g->ctx;
  const size_t bytes = (size_t)g->prm.popsize * g->n * g->esz;
  LXY_CUDA(cudaMemcpyAsync(g->pop[g->cur].p, genes, bytes, cudaMemcpyHostToDevice, c->stream));
  g->in_step = false;
  return ga_settle(g);
}

int lxy_ga_step_begin(lxy_ga* g, int32_t* n_children, void** children_dev, double** scores_dev, int64_t** x100_dev) {
  CHECK_GA(g);
  LXY_TRY(ga_begin(g));
  if (n_children) *n_children = g->n_children;
  if (children_dev) *children_dev = g->child.p;
  if (scores_dev) *scores_dev = g->child_score.as<double>();
  if (x100_dev) *x100_dev = g->child_x100.as<int64_t>();
  return LXY_OK;
}

int lxy_ga_score_children(lxy_ga* g, int32_t first, int32_t count) {
  CHECK_GA(g);
  if (!g->in_step) { set_error("ga_score_children: no generation in progress"); return LXY_ERR_STATE; }
  if (first < 0 || count < 0 || first + count > g->n_children) { set_error("ga_score_children: range outside the children"); return LXY_ERR_ARG; }
  return ga_score(g, g->child.p, first, count, g->child_score.as<double>(), g->child_x100.as<int64_t>());
}

int lxy_ga_step_end(lxy_ga* g) {
  CHECK_GA(g);
  return ga_end(g);
}

int lxy_ga_optimize(lxy_ga* g, int32_t generations) {
  CHECK_GA(g);
  for (int32_t it = 0; it < generations; ++it) {
    LXY_TRY(ga_begin(g));
    LXY_TRY(ga_score(g, g->child.p, 0, g->n_children, g->child_score.as<double>(), g->child_x100.as<int64_t>()));
    LXY_TRY(ga_end(g));
  }
  return LXY_OK;
}

int lxy_ga_set_sharding(lxy_ga* g, int32_t mode) {
  CHECK_GA(g);
  if (mode < -1 || mode > LXY_SHARD_LINKS) { set_error("ga_set_sharding: bad mode"); return LXY_ERR_ARG; }
  if (mode >= 0 && g->prm.kind != LXY_GA_SCAFF) { set_error("ga_set_sharding: only the scaffolding scorer shards"); return LXY_ERR_ARG; }
  if (mode >= 0 && g->ctx->world > 1 && !g->ctx->comm) { set_error("ga_set_sharding: the context has no communicator"); return LXY_ERR_STATE; }
  g->shard_mode = mode;
  return LXY_OK;
}

int lxy_ga_best(lxy_ga* g, void* gene, double* score) {
  CHECK_GA(g);
  if (!g->ready) { set_error("ga_best: population not initialised"); return LXY_ERR_STATE; }
  lxy_ctx* c = g->ctx;
  if (gene) LXY_CUDA(cudaMemcpyAsync(gene, g->pop[g->cur].p, (size_t)g->n * g->esz, cudaMemcpyDeviceToHost, c->stream));
  if (score) LXY_CUDA(cudaMemcpyAsync(score, g->score[g->cur].p, 8, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  return LXY_OK;
}

int lxy_ga_population(lxy_ga* g, void* genes, double* scores) {
  CHECK_GA(g);
  if (!g->ready) { set_error("ga_population: population not initialised"); return LXY_ERR_STATE; }
  lxy_ctx* c = g->ctx;
  const size_t P = (size_t)g->prm.popsize;
  if (genes) LXY_CUDA(cudaMemcpyAsync(genes, g->pop[g->cur].p, P * g->n * g->esz, cudaMemcpyDeviceToHost, c->stream));
  if (scores) LXY_CUDA(cudaMemcpyAsync(scores, g->score[g->cur].p, 8 * P, cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  return LXY_OK;
}

int lxy_ga_children(lxy_ga* g, void* genes, void* plan) {
  CHECK_GA(g);
  if (!g->in_step) { set_error("ga_children: no generation in progress"); return LXY_ERR_STATE; }
  lxy_ctx* c = g->ctx;
  const size_t C = (size_t)g->n_children;
  if (genes && C) LXY_CUDA(cudaMemcpyAsync(genes, g->child.p, C * g->n * g->esz, cudaMemcpyDeviceToHost, c->stream));
  if (plan && C) LXY_CUDA(cudaMemcpyAsync(plan, g->plan.p, C * sizeof(lxy_ga_child), cudaMemcpyDeviceToHost, c->stream));
  LXY_CUDA(cudaStreamSynchronize(c->stream));
  return LXY_OK;
}

int lxy_ga_get_stats(lxy_ga* g, lxy_ga_stats* out) {
  if (!g || !out) { set_error("ga_get_stats: null argument"); return LXY_ERR_ARG; }
  std::lock_guard<std::mutex> lk(g->ctx->mu);
  *out = g->stats;
  return LXY_OK;
}

}  // extern "C"
