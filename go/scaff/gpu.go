// gpu.go — package scaff (drop into /src/lxy/scaff/ of cwbeitel/lxy): scaff.score and the generation
// loop on liblxyb200.  Keeps the reference's hook: NewOrderedIntGenome(ids, score) is unchanged.
package scaff

/*
#include <stdlib.h>
#include "lxy_b200.h"
*/
import "C"

import (
	"fmt"
	"math"
	"unsafe"

	util "sequtil"
)

// gpuScorer owns one lxy_ctx (one B200).  Calls may arrive from any goroutine: the library serialises
// them per context and never keeps a Go pointer after returning.
type gpuScorer struct {
	ctx *C.lxy_ctx
	n   int
}

func lastErr() error { return fmt.Errorf("liblxyb200: %s", C.GoString(C.lxy_last_error())) }

// newGPUScorer uploads the link table once (replaces the per-call map probes of Links.Get,
// src/sequtil/links.go:190-221).  A zero lxy_cfg is LXY_PATH_AUTO: the library picks the dense pair
// table or the record stream from the table's size; both return the same scores.
func newGPUScorer(links *util.Links, device int) (*gpuScorer, error) {
	cfg := C.lxy_cfg{device: C.int32_t(device)}
	s := &gpuScorer{n: links.Size()}
	if rc := C.lxy_ctx_create(&cfg, &s.ctx); rc != 0 {
		return nil, lastErr() // no CUDA device: there is no CPU fallback
	}
	a, b, w := links.Records() // one walk over the table (sequtil/links_gpu.go)
	var pa, pb *C.uint32_t
	var pw *C.double
	if len(a) > 0 {
		pa, pb, pw = (*C.uint32_t)(unsafe.Pointer(&a[0])), (*C.uint32_t)(unsafe.Pointer(&b[0])), (*C.double)(unsafe.Pointer(&w[0]))
	}
	if rc := C.lxy_ctx_set_links(s.ctx, C.int32_t(s.n), C.int64_t(len(a)), pa, pb, pw); rc != 0 {
		C.lxy_ctx_destroy(s.ctx)
		return nil, lastErr()
	}
	return s, nil
}

// newGPUScorerFromFile: LoadLinks on the device (links.go:427-491), the records never leave it.
// names[i] is Decode(i) (links.go:276-286).
func newGPUScorerFromFile(path string, device int) (*gpuScorer, []string, error) {
	cfg := C.lxy_cfg{device: C.int32_t(device)}
	s := &gpuScorer{}
	if rc := C.lxy_ctx_create(&cfg, &s.ctx); rc != 0 {
		return nil, nil, lastErr()
	}
	cpath := C.CString(path)
	defer C.free(unsafe.Pointer(cpath))
	var tbl *C.lxy_links
	var st C.lxy_load_stats
	if rc := C.lxy_ctx_load_links(s.ctx, cpath, 0, &tbl, &st); rc != 0 {
		C.lxy_ctx_destroy(s.ctx)
		return nil, nil, lastErr() // LoadLinks' error, or the line where it would panic
	}
	defer C.lxy_links_free(tbl)
	s.n = int(C.lxy_links_size(tbl))
	names := make([]string, s.n)
	for i := range names {
		names[i] = C.GoString(C.lxy_links_name(tbl, C.int32_t(i)))
	}
	return s, names, nil
}

func (s *gpuScorer) Close() { C.lxy_ctx_destroy(s.ctx) }

// rows16: the same rows as 16-bit genes (n <= 65,536): half the bytes over PCIe, lxy_scaff_score_batch_u16.
func (s *gpuScorer) rows16(pop []*GAOrderedIntGenome) []C.uint16_t {
	perm := make([]C.uint16_t, len(pop)*s.n)
	for p, g := range pop {
		for i, c := range g.Gene {
			perm[p*s.n+i] = C.uint16_t(c)
		}
	}
	return perm
}

func (s *gpuScorer) rows(pop []*GAOrderedIntGenome) []C.int32_t { // Go ints are 64-bit; the ABI takes int32 rows
	perm := make([]C.int32_t, len(pop)*s.n)
	for p, g := range pop {
		for i, c := range g.Gene {
			perm[p*s.n+i] = C.int32_t(c)
		}
	}
	return perm
}

// ScoreBatch fills score/hasscore of every genome that has no memo yet (genome.go:153-159): one
// device call per generation.
func (s *gpuScorer) ScoreBatch(pop []*GAOrderedIntGenome) error {
	var todo []*GAOrderedIntGenome
	for _, g := range pop {
		if !g.hasscore {
			todo = append(todo, g)
		}
	}
	if len(todo) == 0 {
		return nil
	}
	out := make([]C.double, len(todo))
	if s.n <= 65536 { // contig ids fit 16 bits: narrow while copying
		perm := s.rows16(todo)
		if rc := C.lxy_scaff_score_batch_u16(s.ctx, &perm[0], C.int32_t(len(todo)), &out[0], nil); rc != 0 {
			return lastErr()
		}
	} else {
		perm := s.rows(todo)
		if rc := C.lxy_scaff_score_batch(s.ctx, &perm[0], C.int32_t(len(todo)), &out[0], nil); rc != 0 {
			return lastErr()
		}
	}
	for p, g := range todo {
		g.score, g.hasscore = float64(out[p]), true
		scores++ // the reference's call counter (scaff.go:18,306)
	}
	return nil
}

var gpu *gpuScorer

// score keeps the reference's signature (scaff.go:246); a lone call is a batch of one (the library's
// HBM-bound single-ordering kernel).  Errors cannot be returned through this signature: +Inf is the
// worst fitness and can never become elite.
func score(g *GAOrderedIntGenome) float64 {
	perm := gpu.rows([]*GAOrderedIntGenome{g})
	var out C.double
	if rc := C.lxy_scaff_score_batch(gpu.ctx, &perm[0], 1, &out, nil); rc != 0 {
		return math.Inf(1)
	}
	scores++
	return float64(out)
}

// scaffoldOnDevice is Scaffold's loop (scaff.go:105-126) with go-galib's Optimize(1) / Best() replaced
// by the device-resident engine: selection, PMX, mutators, scoring, sort and truncation on the GPU.
func scaffoldOnDevice(s *gpuScorer, ids []int, iterations int, seed uint64, each func(ct int, best []int, score float64)) error {
	prm := C.lxy_ga_params{kind: C.LXY_GA_SCAFF, popsize: 40, p_breed: 0.2, p_mutate: 0.6, // scaff.go:80-94
		tour_p: 0.7, tour_k: 5, mutators: C.LXY_MUT_SHIFT | C.LXY_MUT_SWITCH | C.LXY_MUT_INVERT, seed: C.uint64_t(seed)}
	var ga *C.lxy_ga
	if rc := C.lxy_ga_create(s.ctx, &prm, &ga); rc != 0 {
		return lastErr()
	}
	defer C.lxy_ga_destroy(ga)
	first := make([]C.int32_t, s.n)
	for i, id := range ids {
		first[i] = C.int32_t(id)
	}
	if rc := C.lxy_ga_init(ga, unsafe.Pointer(&first[0]), 1); rc != 0 { // GARandomInitializer, gao.Init(40, genome)
		return lastErr()
	}
	gene := make([]C.int32_t, s.n)
	best := make([]int, s.n)
	var sc C.double
	for ct := 1; ct <= iterations; ct++ {
		if rc := C.lxy_ga_optimize(ga, 1); rc != 0 { // gao.Optimize(1)
			return lastErr()
		}
		C.lxy_ga_best(ga, unsafe.Pointer(&gene[0]), &sc) // gao.Best()
		for i, c := range gene {
			best[i] = int(c)
		}
		each(ct, best, float64(sc))
	}
	return nil
}
