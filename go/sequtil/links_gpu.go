// links_gpu.go — additions to package sequtil (drop into /src/sequtil/ of cwbeitel/lxy) that hand the
// link table to liblxyb200 and replace the two SAM -> links producers by their device versions.
// Not compiled in the build image of lxy_b200 (no Go toolchain there); the same C entry points are
// exercised through ctypes by lxy_b200/_native.py and tests/.
//
//   CGO_CFLAGS="-I$REPO/include" CGO_LDFLAGS="-L$REPO/lxy_b200/_lib -llxyb200 -Wl,-rpath,$REPO/lxy_b200/_lib"
package sequtil

/*
#include <stdlib.h>
#include "lxy_b200.h"
*/
import "C"

import (
	"fmt"
	"unsafe"
)

func lastError() error { return fmt.Errorf("liblxyb200: %s", C.GoString(C.lxy_last_error())) }

// Records walks the table once (links.go:14-28: data[id1][id2] with id1 < id2, links.go:130-132) and
// returns it as flat records — what lxy_ctx_set_links takes.  O(pairs), not the O(n^2) Get loop.
func (l *Links) Records() (a, b []uint32, w []float64) {
	for id1, row := range l.data {
		for id2, v := range row {
			a, b, w = append(a, uint32(id1)), append(b, uint32(id2)), append(w, v)
		}
	}
	return
}

// ScaffoldLinksFromSamGPU has the signature of ScaffoldLinksFromSam (alignment.go:624), which prints open
// errors and carries on; this one returns them.  scaff/cmd.go:225 calls it instead.
func ScaffoldLinksFromSamGPU(samPath, outPath string) error {
	csam, cout := C.CString(samPath), C.CString(outPath)
	defer C.free(unsafe.Pointer(csam))
	defer C.free(unsafe.Pointer(cout))
	var st C.lxy_load_stats
	if rc := C.lxy_scaffold_links_from_sam(csam, cout, 0, &st); rc != 0 {
		return lastError() // LXY_ERR_ARG names the line where the reference would panic
	}
	return nil
}

// VarBlockBatch collects the (varPos1, varPos2) maps BlockLinksFromSam builds per read pair
// (alignment.go:520-521) for ONE chromosome, in the CSR form lxy_links_tabulate_varblocks takes.
type VarBlockBatch struct {
	Chrom      string
	off1, off2 []C.int64_t
	blk1, blk2 []C.int32_t
}

func NewVarBlockBatch(chrom string) *VarBlockBatch {
	return &VarBlockBatch{Chrom: chrom, off1: []C.int64_t{0}, off2: []C.int64_t{0}}
}

// Add appends one read pair: the two maps GetVariantBlocks returned (alignment.go:265-323).
func (v *VarBlockBatch) Add(varPos1, varPos2 map[int][]int) {
	for k, e := range varPos1 {
		v.blk1 = append(v.blk1, C.int32_t(k), C.int32_t(e[0]), C.int32_t(e[1]), C.int32_t(e[2]))
	}
	for k, e := range varPos2 {
		v.blk2 = append(v.blk2, C.int32_t(k), C.int32_t(e[0]), C.int32_t(e[1]), C.int32_t(e[2]))
	}
	v.off1 = append(v.off1, C.int64_t(len(v.blk1)/4))
	v.off2 = append(v.off2, C.int64_t(len(v.blk2)/4))
}

// gpuLinks is the library's own sequtil.Links (lxy_links): TabulateVariantBlockLinks accumulates into
// it on the device, Write emits the .links file LoadLinks reads (links.go:242-264).
type gpuLinks struct{ h *C.lxy_links }

func newGPULinks() (*gpuLinks, error) {
	g := &gpuLinks{}
	if rc := C.lxy_links_new(&g.h); rc != 0 {
		return nil, lastError()
	}
	return g, nil
}
func (g *gpuLinks) Close() { C.lxy_links_free(g.h) }
func (g *gpuLinks) Write(path string) error {
	cp := C.CString(path)
	defer C.free(unsafe.Pointer(cp))
	if rc := C.lxy_links_write(g.h, cp); rc != 0 {
		return lastError()
	}
	return nil
}

// TabulateVariantBlockLinks (links.go:372-420) for every read pair of the batch on GPU `device`
// (device < 0: the library's host loop); returns the reference's (ct, balance).
func (g *gpuLinks) TabulateVariantBlockLinks(v *VarBlockBatch, device int) (int, int, error) {
	cchrom := C.CString(v.Chrom)
	defer C.free(unsafe.Pointer(cchrom))
	var p1, p2 *C.int32_t
	if len(v.blk1) > 0 {
		p1 = &v.blk1[0]
	}
	if len(v.blk2) > 0 {
		p2 = &v.blk2[0]
	}
	var ct, bal C.int64_t
	if rc := C.lxy_links_tabulate_varblocks(g.h, cchrom, C.int64_t(len(v.off1)-1), &v.off1[0], p1, &v.off2[0], p2,
		C.int32_t(device), &ct, &bal); rc != 0 {
		return 0, 0, lastError()
	}
	return int(ct), int(bal), nil
}

// In BlockLinksFromSam (alignment.go:461-555) the call at :524
//     ct1, bal1 := links.TabulateVariantBlockLinks(hits[0].rname, varPos1, varPos2)
// becomes  batch[hits[0].rname].Add(varPos1, varPos2)  and, after the scan loop,
//     for _, b := range batch { ct, bal, err := g.TabulateVariantBlockLinks(b, 0); ... }
//     g.Write(outPath)
// The ids the library assigns differ from a Go run's (which are map-iteration random anyway); names
// and values are the same.
