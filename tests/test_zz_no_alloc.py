"""lxy_phase_score_batch_bits_noalloc_dev — the allocation-free variant of the packed phasing scorer
(no counterpart in the reference; scratch in a static device array of LXY_PHASE_NOALLOC_SCRATCH_BYTES).

Called through the C ABI with device pointers, twice in a row on different populations (the array
keeps the previous call's contents: the second call must not see them), and compared bit for bit
with lxy_phase_score_batch_bits_dev and with the oracle's exact integers (phase.go:171-207
restated).  Sizes beyond the static array are refused with LXY_ERR_RANGE.

The file sorts last on purpose.  It passes alone in a fresh process
(its own context, nothing scored before it) and after the other GPU tests.
"""
import numpy as np
import pytest

import _oracle as O
import _synth as S
import lxy_b200._native as N

pytestmark = pytest.mark.gpu

SCRATCH = 8 << 20     # LXY_PHASE_NOALLOC_SCRATCH_BYTES


def _device_arrays(torch, words, P):
    dw = torch.from_numpy(words.view(np.int32).copy()).cuda()
    oi = torch.full((P,), 12345, dtype=torch.int64, device="cuda")       # stale values must be overwritten
    of = torch.full((P,), -1.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()                                              # the library's stream is non-blocking
    return dw, oi, of


@pytest.mark.parametrize("n,P", [(2050, 200), (33, 7), (40_000, 64), (200_000, 130)])
def test_noalloc_variant_equals_the_allocating_entry_point_twice_in_a_row(n, P):
    torch = pytest.importorskip("torch")
    pa, pb, pw, truth = S.phase_links(n, 17, far_pairs=min(100, n))
    with N.Context(0) as c:
        c.set_phase_links(n, pa, pb, pw)
        for call, seed in enumerate((301, 302)):                          # two calls in a row, new genes each time
            genes = S.phase_population(n, P, seed, truth)
            _, exact = O.phase_score_links_batch(n, pa, pb, pw, genes, nthreads=4)
            words = N.pack_bits(genes)
            dw, oi, of = _device_arrays(torch, words, P)
            c.phase_score_batch_bits_noalloc_dev(dw.data_ptr(), P, oi.data_ptr(), of.data_ptr())
            c.sync()
            got_i, got_f = oi.cpu().numpy(), of.cpu().numpy()
            _, ri, rf = _device_arrays(torch, words, P)
            c.phase_score_batch_bits_dev(dw.data_ptr(), P, ri.data_ptr(), rf.data_ptr())
            c.sync()
            assert got_i.tolist() == ri.cpu().numpy().tolist() == exact.tolist(), call     # integers: bit-exact
            assert got_f.tolist() == rf.cpu().numpy().tolist() == (exact.astype(np.float64)).tolist()
        # either output may be left out, as in the allocating entry point
        dw, oi, of = _device_arrays(torch, words, P)
        c.phase_score_batch_bits_noalloc_dev(dw.data_ptr(), P, oi.data_ptr(), 0)
        c.sync()
        assert oi.cpu().numpy().tolist() == exact.tolist() and (of.cpu().numpy() == -1.0).all()


def test_noalloc_variant_after_a_smaller_population_sees_no_stale_scratch():
    """A big call then a small one then the big one again: the accumulator words of candidates the small
    call does not have still hold the big call's sums; every call clears what it uses."""
    torch = pytest.importorskip("torch")
    n = 5000
    pa, pb, pw, truth = S.phase_links(n, 19, far_pairs=100)
    with N.Context(0) as c:
        c.set_phase_links(n, pa, pb, pw)
        for P, seed in ((256, 311), (3, 312), (256, 313), (40, 314)):
            genes = S.phase_population(n, P, seed, truth)
            _, exact = O.phase_score_links_batch(n, pa, pb, pw, genes, nthreads=4)
            dw, oi, of = _device_arrays(torch, N.pack_bits(genes), P)
            c.phase_score_batch_bits_noalloc_dev(dw.data_ptr(), P, oi.data_ptr(), of.data_ptr())
            c.sync()
            assert oi.cpu().numpy().tolist() == exact.tolist(), P


def test_noalloc_variant_refuses_sizes_beyond_the_static_array():
    torch = pytest.importorskip("torch")
    n = 600_000
    pa, pb, pw, truth = S.phase_links(n, 23, far_pairs=100)
    with N.Context(0) as c:
        c.set_phase_links(n, pa, pb, pw)
        P_ok, P_big = 64, 128                                             # 600,000 * 64 / 8 = 4.8 MB; * 128 = 9.6 MB
        assert n * P_ok // 8 + 8 * P_ok <= SCRATCH < n * P_big // 8 + 8 * P_big
        genes = S.phase_population(n, P_big, 321, truth)
        words = N.pack_bits(genes)
        dw, oi, of = _device_arrays(torch, words, P_big)
        with pytest.raises(N.LxyError) as e:
            c.phase_score_batch_bits_noalloc_dev(dw.data_ptr(), P_big, oi.data_ptr(), of.data_ptr())
        assert e.value.status == -5 and "static" in str(e.value)
        torch.cuda.synchronize()
        assert (oi.cpu().numpy() == 12345).all()                          # outputs untouched
        # the same links, half the population: fits, and equals the allocating entry point
        _, exact = O.phase_score_links_batch(n, pa, pb, pw, genes[:P_ok], nthreads=4)
        dw, oi, of = _device_arrays(torch, np.ascontiguousarray(words[:P_ok]), P_ok)
        c.phase_score_batch_bits_noalloc_dev(dw.data_ptr(), P_ok, oi.data_ptr(), of.data_ptr())
        c.sync()
        assert oi.cpu().numpy().tolist() == exact.tolist()
    with N.Context(0) as c:                                               # links not set: a state error, not a crash
        with pytest.raises(N.LxyError):
            c.phase_score_batch_bits_noalloc_dev(dw.data_ptr(), P_ok, oi.data_ptr(), of.data_ptr())
