// ga_oracle.cpp — CPU ORACLE for the GA operators around the fitness path.  TEST INFRASTRUCTURE ONLY
// (same rules as lxy_oracle.cpp: only tests/, smoke() and bench.py's CPU legs may load it).
//
// Restated from the reference, in its own sequential form and cost (file:line relative to
// /root/reference/):
//   permutation genome   src/lxy/scaff/genome.go   Crossover = PMX :47-91 with the O(n) scan pmxmap
//                        :37-44 per window element (O(n^2)); Splice :93-97; Switch :125-128;
//                        Invert :161-177; GAInvertMutator :184-209
//   bitstring genome     src/lxy/phase/genome.go   Crossover (two-point) :37-46; Switch :56-59;
//                        Invert (flip [p1, p2)) :95-109; GAInvertMutator :117-131
// The engine that calls them (go-galib) is absent from the reference tree; its generation loop, the
// tournament selector and the shift / switch mutators are restated from its call sites and the
// genome methods they can only act through and fixed in include/lxy_ga_plan.h,
// which this file shares with the CUDA engine so that both draw the same random decisions.
// PARITY: the genome methods are pinned by the reference source text only (its genome tests are
// empty: scaff/genome_test.go is 0 bytes, phase/genome_test.go:7-9 has an empty TestCrossover);
// the engine-level behaviour is "parity unpinned" (source absent).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>

#include "../include/lxy_ga_plan.h"

namespace {

// pmxmap (scaff/genome.go:37-44): first index outside [p1, p2] holding v.
bool pmxmap(const std::vector<int32_t>& g, int32_t v, int p1, int p2, int* at) {
  for (int i = 0; i < (int)g.size(); ++i)
    if (g[i] == v && (i < p1 || i > p2)) { *at = i; return true; }
  return false;
}

// Crossover (scaff/genome.go:47-91), line for line: windows exchanged, then the two stacks.
void pmx_ref(const int32_t* a, const int32_t* b, int n, int p1, int p2, int32_t* oa, int32_t* ob) {
  std::vector<int32_t> ca(a, a + n), cb(b, b + n);
  std::copy(b + p1, b + p2 + 1, ca.begin() + p1);
  std::copy(a + p1, a + p2 + 1, cb.begin() + p1);
  std::vector<int> amap, bmap;
  for (int i = p1; i <= p2; ++i) {
    int ma;
    if (pmxmap(ca, ca[i], p1, p2, &ma)) {
      amap.push_back(ma);
      if (!bmap.empty()) {
        const int i1 = amap.back(); amap.pop_back();
        const int i2 = bmap.back(); bmap.pop_back();
        std::swap(ca[i1], cb[i2]);
      }
    }
    int mb;
    if (pmxmap(cb, cb[i], p1, p2, &mb)) {
      bmap.push_back(mb);
      if (!amap.empty()) {
        const int i1 = amap.back(); amap.pop_back();
        const int i2 = bmap.back(); bmap.pop_back();
        std::swap(ca[i1], cb[i2]);
      }
    }
  }
  std::copy(ca.begin(), ca.end(), oa);
  std::copy(cb.begin(), cb.end(), ob);
}

template <class T>
void mutate_ref(int kind, const T* parent, int n, int p1, int p2, bool bits, T* out) {
  std::copy(parent, parent + n, out);
  if (kind == LXY_CHILD_SWITCH) {                 // Switch (scaff/genome.go:125-128, phase :56-59)
    std::swap(out[p1], out[p2]);
  } else if (kind == LXY_CHILD_SHIFT) {           // two Splice calls (genome.go:93-97): p2 -> p1, [p1,p2) right
    out[p1] = parent[p2];
    for (int i = p1; i < p2; ++i) out[i + 1] = parent[i];
  } else if (kind == LXY_CHILD_INVERT) {
    if (bits) {                                   // phase/genome.go:95-109: flip [p1, p2)
      for (int i = p1; i < p2; ++i) out[i] = out[i] ? 0 : 1;
    } else {                                      // scaff/genome.go:161-177: Switch(p1++, p2--) to the centre
      while (p1 < p2) { std::swap(out[p1], out[p2]); ++p1; --p2; }
    }
  }
}

}  // namespace

extern "C" {

void lxyo_pmx(const int32_t* a, const int32_t* b, int n, int p1, int p2, int32_t* oa, int32_t* ob) {
  pmx_ref(a, b, n, p1, p2, oa, ob);
}

void lxyo_mutate_perm(int kind, const int32_t* parent, int n, int p1, int p2, int32_t* out) {
  mutate_ref<int32_t>(kind, parent, n, p1, p2, false, out);
}
void lxyo_mutate_bits(int kind, const uint8_t* parent, int n, int p1, int p2, uint8_t* out) {
  mutate_ref<uint8_t>(kind, parent, n, p1, p2, true, out);
}

// The plan of one generation in go-galib's append order: for p = 0..popsize-1 the breed children
// (two), then the mutation child.  Returns the number of children; plan must hold 3*popsize entries.
int lxyo_ga_plan(uint64_t seed, uint64_t gen, int popsize, int len, double p_breed, double p_mutate,
                 double tour_p, int tour_k, uint32_t mutators, lxy_ga_child* plan) {
  int c = 0;
  for (int p = 0; p < popsize; ++p) {
    lxy_ga_child k[3];
    lxy_ga_plan_slot(seed, gen, p, popsize, len, p_breed, p_mutate, tour_p, tour_k, mutators, k);
    for (int j = 0; j < 3; ++j)
      if (k[j].kind != LXY_CHILD_NONE) plan[c++] = k[j];
  }
  return c;
}

// Children of one generation from a sorted population.  kind_bits = 0: permutations (int32 genes),
// 1: bitstrings (uint8 genes, two-point crossover phase/genome.go:37-46).
int lxyo_ga_children(uint64_t seed, uint64_t gen, int popsize, int len, double p_breed,
                     double p_mutate, double tour_p, int tour_k, uint32_t mutators, int kind_bits,
                     const void* pop, void* children) {
  std::vector<lxy_ga_child> plan(3 * (size_t)popsize);
  const int C = lxyo_ga_plan(seed, gen, popsize, len, p_breed, p_mutate, tour_p, tour_k, mutators, plan.data());
  for (int c = 0; c < C; ++c) {
    const lxy_ga_child& k = plan[c];
    if (!kind_bits) {
      const int32_t* P = (const int32_t*)pop;
      int32_t* out = (int32_t*)children + (size_t)c * len;
      if (k.kind == LXY_CHILD_BREED_A) {  // the pair is adjacent in the plan: A then B
        pmx_ref(P + (size_t)k.pa * len, P + (size_t)k.pb * len, len, k.p1, k.p2, out, out + len);
        ++c;
      } else {
        mutate_ref<int32_t>(k.kind, P + (size_t)k.pa * len, len, k.p1, k.p2, false, out);
      }
    } else {
      const uint8_t* P = (const uint8_t*)pop;
      uint8_t* out = (uint8_t*)children + (size_t)c * len;
      if (k.kind == LXY_CHILD_BREED_A) {
        const uint8_t *a = P + (size_t)k.pa * len, *b = P + (size_t)k.pb * len;
        std::copy(a, a + len, out);
        std::copy(b, b + len, out + len);
        std::copy(b + k.p1, b + k.p2 + 1, out + k.p1);
        std::copy(a + k.p1, a + k.p2 + 1, out + len + k.p1);
        ++c;
      } else {
        mutate_ref<uint8_t>(k.kind, P + (size_t)k.pa * len, len, k.p1, k.p2, true, out);
      }
    }
  }
  return C;
}

// Survivor selection: stable ascending sort of (population ++ children) by key, first popsize kept
// (go-galib sorts by Score() and truncates; ties keep append order here).  order_out[popsize].
void lxyo_ga_select(int popsize, int nchildren, const int64_t* keys_i64, const double* keys_f64,
                    int32_t* order_out) {
  const int M = popsize + nchildren;
  std::vector<int32_t> idx(M);
  std::iota(idx.begin(), idx.end(), 0);
  if (keys_i64)  // exact integers: larger x100 = lower fitness = better (fitness = -x100/100)
    std::stable_sort(idx.begin(), idx.end(), [&](int x, int y) { return keys_i64[x] > keys_i64[y]; });
  else
    std::stable_sort(idx.begin(), idx.end(), [&](int x, int y) { return keys_f64[x] < keys_f64[y]; });
  std::copy(idx.begin(), idx.begin() + popsize, order_out);
}


// ---- solution-quality metrics, the reference's loops as written ------------------------------------
// EvalScaffolding (src/lxy/scaff/scaff.go:329-384): O(n^3) walk over index triples.  scaff[i] and
// key[i] are contig ids here (the reference uses names; keyOrder maps a name to its index in key).
void lxyo_eval_scaffolding(const int32_t* scaff, const int32_t* key, int n, double* out2) {
  std::vector<int> keyOrder(n), keyOrderReverse(n);
  for (int i = 0; i < n; ++i) { keyOrder[key[i]] = i; keyOrderReverse[key[i]] = n - i; }
  double comparisons = 0, correct = 0, correctReverse = 0, neighbors = 0, neighborsCorrect = 0,
         neighborsCorrectReverse = 0;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j)
      for (int k = 0; k < n; ++k) {
        if (i < j && k > j) {
          const int v1 = scaff[i], v2 = scaff[j], v3 = scaff[k];
          comparisons += 1;
          if (keyOrder[v1] < keyOrder[v2] && keyOrder[v2] < keyOrder[v3]) correct += 1;
          if (keyOrderReverse[v1] < keyOrderReverse[v2] && keyOrderReverse[v2] < keyOrderReverse[v3]) correctReverse += 1;
          if (j == i + 1 && k == j + 1) {
            neighbors += 1;
            if (keyOrder[v1] < keyOrder[v2] && keyOrder[v2] < keyOrder[v3]) neighborsCorrect += 1;
            if (keyOrderReverse[v1] < keyOrderReverse[v2] && keyOrderReverse[v2] < keyOrderReverse[v3]) neighborsCorrectReverse += 1;
          }
        }
      }
  const double qscoreF = correct / comparisons, qscoreR = correctReverse / comparisons;
  if (qscoreF > qscoreR) { out2[0] = qscoreF; out2[1] = neighborsCorrect / neighbors; }
  else { out2[0] = qscoreR; out2[1] = neighborsCorrectReverse / neighbors; }
}

// EvalPhasing (src/lxy/phase/phase.go:211-263), including its counter quirk (:225-235).
void lxyo_eval_phasing(const uint8_t* phasing, const uint8_t* key, int n, double* out4) {
  double matches = 0, comparisons = 0, neighbor_matches = 0, neighbor_comparisons = 0;
  double second_neighbor_matches = 0, second_neighbor_comparisons = 0, third_neighbor_matches = 0,
         third_neighbor_comparisons = 0;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      comparisons += 1;
      if (i - j == 1 || j - i == 1) neighbor_comparisons += 1;
      if (i - j == 2 || j - i == 2) neighbor_comparisons += 1;
      if (i - j == 3 || j - i == 3) neighbor_comparisons += 1;
      const bool pmatch = (phasing[i] != 0) == (phasing[j] != 0), kmatch = (key[i] != 0) == (key[j] != 0);
      if ((pmatch && kmatch) || (!pmatch && !kmatch)) {
        matches += 1;
        if (i - j == 1 || j - i == 1) neighbor_matches += 1;
        if (i - j == 2 || j - i == 2) second_neighbor_matches += 1;
        if (i - j == 3 || j - i == 3) third_neighbor_matches += 1;
      }
    }
  out4[0] = matches / comparisons;
  out4[1] = neighbor_matches / neighbor_comparisons;
  out4[2] = (neighbor_matches + second_neighbor_matches) / (neighbor_comparisons + second_neighbor_comparisons);
  out4[3] = (neighbor_matches + second_neighbor_matches + third_neighbor_matches) /
            (neighbor_comparisons + second_neighbor_comparisons + third_neighbor_comparisons);
}

// EvalPhasingDev (src/lxy/phase/phase.go:266-325), the metric `lxy phase infer` logs (phase/cmd.go:152).
// The reference ranges over two maps keyed by variant-block name; `index[i]` stands for
// Atoi(Split(name_i, "_")[1]) (:279-280) and the n entries are the names present in BOTH maps (the
// others are skipped at :289-299 before anything is counted).  out3 = (all ordered pairs, pairs whose
// indices differ by <= 2, by <= 9), each matches / comparisons, i == j included.
void lxyo_eval_phasing_dev(const uint8_t* phasing, const uint8_t* key, const int32_t* index, int n, double* out3) {
  double matches = 0, comparisons = 0, matches300k = 0, matches1mbp = 0, comparisons300k = 0, comparisons1mbp = 0;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      long long inddiff = (long long)index[i] - (long long)index[j];
      if (inddiff < 0) inddiff = inddiff * -1;
      const bool is3k = inddiff <= 2, is1mbp = inddiff <= 9;
      const bool pmatch = (phasing[i] != 0) == (phasing[j] != 0), kmatch = (key[i] != 0) == (key[j] != 0);
      comparisons += 1;
      if (is3k) comparisons300k += 1;
      if (is1mbp) comparisons1mbp += 1;
      if ((pmatch && kmatch) || (!pmatch && !kmatch)) {
        matches += 1;
        if (is3k) matches300k += 1;
        if (is1mbp) matches1mbp += 1;
      }
    }
  out3[0] = matches / comparisons;
  out3[1] = matches300k / comparisons300k;
  out3[2] = matches1mbp / comparisons1mbp;
}

}  // extern "C"
