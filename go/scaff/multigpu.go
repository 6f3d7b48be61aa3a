// multigpu.go — package scaff: several B200s from the reference's single Go process (scaff.go:88-94 runs
// go-galib with 7 goroutines; here: one goroutine, locked to an OS thread, per device).
package scaff

/*
#include "lxy_b200.h"
*/
import "C"

import (
	"runtime"
	"sync"

	util "sequtil"
)

// gpuGang is one gpuScorer per device with an NCCL communicator across them (lxy_comm_init_all).
type gpuGang struct {
	s    []*gpuScorer
	mode C.int32_t // LXY_SHARD_POPULATION: links replicated; LXY_SHARD_LINKS: every device holds a share of the records
}

// newGang: population-sharded (the whole table on every device) when shardLinks is false; otherwise
// device r receives the records lxy_comm_link_owners assigns to it (whole tiles of the sorted store).
func newGang(links *util.Links, devices int, shardLinks bool) (*gpuGang, error) {
	g := &gpuGang{mode: C.LXY_SHARD_POPULATION}
	a, b, w := links.Records()
	owner := make([]C.int32_t, len(a))
	if shardLinks {
		g.mode = C.LXY_SHARD_LINKS
		if rc := C.lxy_comm_link_owners(C.int32_t(links.Size()), C.int64_t(len(a)), (*C.uint32_t)(&a[0]), (*C.uint32_t)(&b[0]),
			C.int32_t(devices), &owner[0]); rc != 0 {
			return nil, lastErr()
		}
	}
	ctxs := make([]*C.lxy_ctx, devices)
	for d := 0; d < devices; d++ {
		cfg := C.lxy_cfg{device: C.int32_t(d), path: C.LXY_PATH_STREAM}
		s := &gpuScorer{n: links.Size()}
		if rc := C.lxy_ctx_create(&cfg, &s.ctx); rc != 0 {
			return nil, lastErr()
		}
		var da, db []C.uint32_t
		var dw []C.double
		for i := range a {
			if !shardLinks || int(owner[i]) == d {
				da, db, dw = append(da, C.uint32_t(a[i])), append(db, C.uint32_t(b[i])), append(dw, C.double(w[i]))
			}
		}
		var pa, pb *C.uint32_t
		var pw *C.double
		if len(da) > 0 {
			pa, pb, pw = &da[0], &db[0], &dw[0]
		}
		if rc := C.lxy_ctx_set_links(s.ctx, C.int32_t(s.n), C.int64_t(len(da)), pa, pb, pw); rc != 0 {
			return nil, lastErr()
		}
		g.s, ctxs[d] = append(g.s, s), s.ctx
	}
	if rc := C.lxy_comm_init_all(&ctxs[0], C.int32_t(devices)); rc != 0 { // ncclCommInitRank inside ncclGroupStart/End
		return nil, lastErr()
	}
	return g, nil
}

// ScoreBatch scores every genome without a memo on all devices: every device receives the whole list
// (the library moves each row over PCIe once per box and the score vector over NVLink) and returns the
// whole score vector; device 0's copy is used.
func (g *gpuGang) ScoreBatch(pop []*GAOrderedIntGenome) error {
	var todo []*GAOrderedIntGenome
	for _, x := range pop {
		if !x.hasscore {
			todo = append(todo, x)
		}
	}
	if len(todo) == 0 {
		return nil
	}
	perm := g.s[0].rows(todo)
	outs := make([][]C.double, len(g.s))
	errs := make([]error, len(g.s))
	var wg sync.WaitGroup
	for d := range g.s {
		wg.Add(1)
		go func(d int) { // a sharded call returns when its collective has completed: one thread per device
			defer wg.Done()
			runtime.LockOSThread()
			defer runtime.UnlockOSThread()
			outs[d] = make([]C.double, len(todo))
			if rc := C.lxy_scaff_score_sharded(g.s[d].ctx, g.mode, &perm[0], C.int32_t(len(todo)), &outs[d][0], nil); rc != 0 {
				errs[d] = lastErr() // a bad row on ANY device is reported by every device
			}
		}(d)
	}
	wg.Wait()
	for _, e := range errs {
		if e != nil {
			return e
		}
	}
	for p, x := range todo {
		x.score, x.hasscore = float64(outs[0][p]), true
		scores++
	}
	return nil
}

func (g *gpuGang) Close() {
	for _, s := range g.s {
		C.lxy_ctx_destroy(s.ctx) // destroys the communicator of the context too
	}
}
