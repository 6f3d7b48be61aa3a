// gpu.go — package phase (drop into /src/lxy/phase/ of cwbeitel/lxy): phase.score on liblxyb200.
// The hook stays: NewFixedBitstringGenome(init, score) (phase.go:59-61).
package phase

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

type gpuScorer struct {
	ctx *C.lxy_ctx
	n   int // links.Size() = genome length (phase.go:59); the gene index IS the link id (phase.go:185)
}

func lastErr() error { return fmt.Errorf("liblxyb200: %s", C.GoString(C.lxy_last_error())) }

func newGPUScorer(links *util.Links, device int) (*gpuScorer, error) {
	cfg := C.lxy_cfg{device: C.int32_t(device)}
	s := &gpuScorer{n: links.Size()}
	if rc := C.lxy_ctx_create(&cfg, &s.ctx); rc != 0 {
		return nil, lastErr()
	}
	a, b, w := links.Records()
	var pa, pb *C.uint32_t
	var pw *C.double
	if len(a) > 0 {
		pa, pb, pw = (*C.uint32_t)(unsafe.Pointer(&a[0])), (*C.uint32_t)(unsafe.Pointer(&b[0])), (*C.double)(unsafe.Pointer(&w[0]))
	}
	if rc := C.lxy_ctx_set_phase_links(s.ctx, C.int32_t(s.n), C.int64_t(len(a)), pa, pb, pw); rc != 0 {
		C.lxy_ctx_destroy(s.ctx)
		return nil, lastErr()
	}
	return s, nil
}

func (s *gpuScorer) Close() { C.lxy_ctx_destroy(s.ctx) }

// pack: []bool rows -> row-packed words, bit (i & 31) of word [p][i >> 5] = gene i: one eighth of the
// bytes of the byte-per-gene entry point over PCIe (the rows have to be copied into one buffer anyway).
func (s *gpuScorer) pack(pop []*GAFixedBitstringGenome) []C.uint32_t {
	W := (s.n + 31) / 32
	words := make([]C.uint32_t, len(pop)*W)
	for p, g := range pop {
		for i, bit := range g.Gene {
			if bit {
				words[p*W+i>>5] |= 1 << uint(i&31)
			}
		}
	}
	return words
}

// ScoreBatch fills score/hasscore of every genome without a memo (phase/genome.go:87-93).
func (s *gpuScorer) ScoreBatch(pop []*GAFixedBitstringGenome) error {
	var todo []*GAFixedBitstringGenome
	for _, g := range pop {
		if !g.hasscore {
			todo = append(todo, g)
		}
	}
	if len(todo) == 0 {
		return nil
	}
	words := s.pack(todo)
	out := make([]C.double, len(todo)) // the value phase.score returns (phase.go:206); out_i64 is the same number, exact
	if rc := C.lxy_phase_score_batch_bits(s.ctx, &words[0], C.int32_t(len(todo)), nil, &out[0]); rc != 0 {
		return lastErr()
	}
	for p, g := range todo {
		g.score, g.hasscore = float64(out[p]), true
		scores++
	}
	return nil
}

var gpu *gpuScorer

func score(g *GAFixedBitstringGenome) float64 { // phase.go:171 signature kept
	words := gpu.pack([]*GAFixedBitstringGenome{g})
	var out C.double
	if rc := C.lxy_phase_score_batch_bits(gpu.ctx, &words[0], 1, nil, &out); rc != 0 {
		return math.Inf(1)
	}
	scores++
	return float64(out)
}
