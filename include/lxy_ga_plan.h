/* lxy_ga_plan.h — the deterministic "plan" of one GA generation (who breeds, who mutates, where).
 *
 * The reference drives its scorers from github.com/thoj/go-galib (GAParallel.Optimize(1),
 * /root/reference/src/lxy/scaff/scaff.go:80-94,108; phase.go:48-63,78).  That engine's source is
 * absent from the reference tree and its RNG is time-seeded (scaff.go:70), so a reference run can
 * not be replayed.  This header DEFINES the replacement: every random decision of a generation is a
 * pure function of (seed, generation, slot, draw), so the device engine (lxy_b200/csrc/ga_kernels.cu)
 * and the CPU oracle (oracle/ga_oracle.cpp) make identical decisions and can be compared bit for
 * bit.  The shape of a generation follows go-galib as used by the reference:
 *
 *   for p in 0..popsize-1:                                   (the population is sorted, best first)
 *     if PBreed  > rand: a, b = tournament(), tournament();  p1 <= p2 random cut points
 *                        append a.Crossover(b, p1, p2)       (two children; PMX, scaff/genome.go:47-91;
 *                                                             two-point for bits, phase/genome.go:37-46)
 *     if PMutate > rand: append mutator(pop[p])              (one child; GAMultiMutator picks one of the
 *                                                             registered mutators uniformly, scaff.go:72-78)
 *   sort ascending by Score(), keep the first popsize       (lower is better, scaff.go:308)
 *
 * Tournament (NewGATournamentSelector(0.7, 5), scaff.go:82): draw `tour_k` individuals, order them by
 * score, walk from the best and accept each with probability `tour_p`; the last one is the
 * fallback.  Selection reads the population as it was at the start of the generation.
 *
 * Known deviation from the reference's trajectory (not from its scorer): phase/genome.go:95-109 `Invert`
 * flips bits WITHOUT calling Reset(), and Copy() carries `hasscore` (phase/genome.go:80), so from the
 * second generation on an invert-mutated bitstring child keeps its parent's memoised score in the
 * reference.  Here every child is scored (the scaffolding genome's Invert does Reset(),
 * scaff/genome.go:161-177, so only the phasing workflow differs); the stale memo looks like a reference
 * bug, and since the engine and its RNG cannot be replayed anyway nothing could be compared against it.
 *
 * Plain C, included by host code, CUDA code and the oracle alike. */
#ifndef LXY_GA_PLAN_H_
#define LXY_GA_PLAN_H_

#include <stdint.h>

#if defined(__CUDACC__)
#define LXY_HD __host__ __device__ __forceinline__
#else
#define LXY_HD static inline
#endif

/* mutators (bit mask in lxy_ga_params.mutators) */
#define LXY_MUT_SHIFT 1u  /* go-galib GAShiftMutator: the gene at p2 moves to p1, [p1, p2) shifts right */
#define LXY_MUT_SWITCH 2u /* go-galib GASwitchMutator: genome.Switch(p1, p2)                           */
#define LXY_MUT_INVERT 4u /* LXY's GAInvertMutator: scaff/genome.go:184-209 (segment reversal),        \
                             phase/genome.go:117-131 (bit flip of [p1, p2))                            */

/* child kinds */
#define LXY_CHILD_NONE 0
#define LXY_CHILD_BREED_A 1 /* first child of Crossover  (a's copy with b's window) */
#define LXY_CHILD_BREED_B 2 /* second child                                          */
#define LXY_CHILD_SHIFT 3
#define LXY_CHILD_SWITCH 4
#define LXY_CHILD_INVERT 5

typedef struct lxy_ga_child { /* one appended individual */
  int32_t kind;               /* LXY_CHILD_*                                        */
  int32_t pa, pb;             /* parent indices into the sorted population (pb: breed only) */
  int32_t p1, p2;             /* cut points, p1 <= p2                               */
  int32_t slot;               /* the population slot p whose draw appended this child */
} lxy_ga_child;

LXY_HD uint64_t lxy_ga_mix(uint64_t x) { /* splitmix64 finaliser */
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
/* the draw-th random word of slot `slot` in generation `gen` */
LXY_HD uint64_t lxy_ga_rand(uint64_t seed, uint64_t gen, uint64_t slot, uint64_t draw) {
  return lxy_ga_mix(lxy_ga_mix(lxy_ga_mix(seed ^ 0x4C58595F47415F31ull) + gen) + (slot << 20) + draw);
}
LXY_HD double lxy_ga_unit(uint64_t r) { return (double)(r >> 11) * (1.0 / 9007199254740992.0); }
LXY_HD int32_t lxy_ga_intn(uint64_t r, int32_t n) { /* [0, n) */
  return (int32_t)(((r >> 32) * (uint64_t)(uint32_t)n) >> 32);
}

/* Tournament over a population that is sorted best-first: rank == index, so "order the
 * contestants by score" is "order their indices"; equal scores keep population order. */
LXY_HD int32_t lxy_ga_tournament(uint64_t seed, uint64_t gen, uint64_t slot, uint64_t* draw,
                                 int32_t popsize, int32_t tour_k, double tour_p) {
  int32_t c[16];
  if (tour_k > 16) tour_k = 16;
  for (int32_t i = 0; i < tour_k; ++i) {
    int32_t v = lxy_ga_intn(lxy_ga_rand(seed, gen, slot, (*draw)++), popsize);
    int32_t j = i; /* insertion sort, ascending index = ascending score */
    while (j > 0 && c[j - 1] > v) { c[j] = c[j - 1]; --j; }
    c[j] = v;
  }
  for (int32_t i = 0; i + 1 < tour_k; ++i)
    if (tour_p > lxy_ga_unit(lxy_ga_rand(seed, gen, slot, (*draw)++))) return c[i];
  return c[tour_k - 1];
}

/* The (up to three) children slot p appends in generation gen; out[0..2]; kind NONE = not drawn. */
LXY_HD void lxy_ga_plan_slot(uint64_t seed, uint64_t gen, int32_t p, int32_t popsize, int32_t len,
                             double p_breed, double p_mutate, double tour_p, int32_t tour_k,
                             uint32_t mutators, lxy_ga_child out[3]) {
  uint64_t draw = 0;
  const uint64_t slot = (uint64_t)(uint32_t)p;
  out[0].kind = out[1].kind = out[2].kind = LXY_CHILD_NONE;
  if (p_breed > lxy_ga_unit(lxy_ga_rand(seed, gen, slot, draw++))) {
    const int32_t a = lxy_ga_tournament(seed, gen, slot, &draw, popsize, tour_k, tour_p);
    const int32_t b = lxy_ga_tournament(seed, gen, slot, &draw, popsize, tour_k, tour_p);
    int32_t p1 = lxy_ga_intn(lxy_ga_rand(seed, gen, slot, draw++), len);
    int32_t p2 = lxy_ga_intn(lxy_ga_rand(seed, gen, slot, draw++), len);
    if (p1 > p2) { int32_t t = p1; p1 = p2; p2 = t; }
    out[0].kind = LXY_CHILD_BREED_A; out[1].kind = LXY_CHILD_BREED_B;
    out[0].pa = out[1].pa = a; out[0].pb = out[1].pb = b;
    out[0].p1 = out[1].p1 = p1; out[0].p2 = out[1].p2 = p2;
  }
  out[0].slot = out[1].slot = out[2].slot = p;
  draw = 64; /* the mutation draws do not depend on whether the slot bred */
  if (p_mutate > lxy_ga_unit(lxy_ga_rand(seed, gen, slot, draw++))) {
    int32_t kinds[3], nk = 0;
    if (mutators & LXY_MUT_SHIFT) kinds[nk++] = LXY_CHILD_SHIFT;
    if (mutators & LXY_MUT_SWITCH) kinds[nk++] = LXY_CHILD_SWITCH;
    if (mutators & LXY_MUT_INVERT) kinds[nk++] = LXY_CHILD_INVERT;
    if (nk > 0) {
      const int32_t which = lxy_ga_intn(lxy_ga_rand(seed, gen, slot, draw++), nk);
      int32_t p1 = lxy_ga_intn(lxy_ga_rand(seed, gen, slot, draw++), len);
      int32_t p2 = lxy_ga_intn(lxy_ga_rand(seed, gen, slot, draw++), len);
      if (p1 > p2) { int32_t t = p1; p1 = p2; p2 = t; }
      out[2].kind = kinds[which];
      out[2].pa = p; out[2].pb = p; out[2].p1 = p1; out[2].p2 = p2;
    }
  }
}

#endif /* LXY_GA_PLAN_H_ */
