"""GPU parity of what round 2 added behind the C ABI: the LXY_PATH_AUTO dispatch (pair table vs
stream, duplicate records summed as Links.Add does, links.go:149-177), populations that are not a
multiple of 128 (whole groups through the batch kernel + the rest through streaming passes), the
row-packed phasing entry point, and the sharded scorers / GA over a communicator of ONE rank (the
collectives, pack / unpack kernels and error propagation all run; world > 1 is in
test_gpu_multi.py, which needs two devices).

Every case is compared with the oracle (scaff.go:246-310 / phase.go:171-207 restated) and with the
explicit LXY_PATH_STREAM context, bit for bit in the integers.
"""
import numpy as np
import pytest

import _oracle as O
import _synth as S
import lxy_b200._native as N

pytestmark = pytest.mark.gpu


def _dup_links(n, pairs, repeat, seed):
    """`pairs` distinct contig pairs, each repeated ~`repeat` times (one record per read pair)."""
    rng = np.random.default_rng(seed)
    a = rng.integers(0, n, pairs).astype(np.uint32)
    b = ((a + rng.integers(1, n, pairs)) % n).astype(np.uint32)
    idx = rng.integers(0, pairs, pairs * repeat)
    flip = rng.random(len(idx)) < 0.5
    ra, rb = a[idx], b[idx]
    return np.where(flip, rb, ra).astype(np.uint32), np.where(flip, ra, rb).astype(np.uint32)


def test_auto_path_picks_the_pair_table_for_dense_link_sets():
    n, P = 300, 64
    a, b = _dup_links(n, 20_000, 5, 1)                 # 100,000 records >= 256 n: dense
    perms = S.scaff_population(n, P, 101)
    _, ox100 = O.scaff_score_links_batch(n, a, b, np.ones(len(a)), perms, nthreads=4)
    with N.Context(0) as auto, N.Context(0, path=N.PATH_STREAM) as st:
        auto.set_links(n, a, b, None)
        st.set_links(n, a, b, None)
        path, recs = auto.path()
        assert path == N.PATH_PAIRS and recs < len(a) // 2          # pairs chosen, duplicates summed
        assert st.path() == (N.PATH_STREAM, len(a))
        ga, xa = auto.scaff_score_batch(perms, want_x100=True)
        gs, xs = st.scaff_score_batch(perms, want_x100=True)
        assert xa.tolist() == ox100.tolist() == xs.tolist() and ga.tolist() == gs.tolist()
        for i, j in [(0, 1), (int(a[0]), int(b[0])), (5, 5), (int(b[7]), int(a[7]))]:
            assert auto.links_get(i, j) == st.links_get(i, j)       # Links.Get on the summed store
        assert auto.links_get(-1, 2)[1] != 0


@pytest.mark.parametrize("weights", ["unit", "int"])
def test_auto_path_streams_sparse_link_sets_with_duplicates_summed(weights):
    n, P = 3000, 40
    a, b = _dup_links(n, 50_000, 4, 2)                 # 200,000 records < 256 n: the stream is cheaper
    w = None if weights == "unit" else np.random.default_rng(3).integers(-3, 9, len(a)).astype(np.float64)
    perms = S.scaff_population(n, P, 102)
    _, ox100 = O.scaff_score_links_batch(n, a, b, np.ones(len(a)) if w is None else w, perms, nthreads=4)
    ref = O.Links.from_records(n, a, b, np.ones(len(a)) if w is None else w).scaff_score_batch(perms, nthreads=4)
    with N.Context(0) as auto:
        auto.set_links(n, a, b, w)
        path, recs = auto.path()
        assert path == N.PATH_STREAM and recs <= len(a) // 2
        for rows in (perms, perms[:1], perms[:7], perms[:9]):        # batch, stream1, stream1, passes / batch
            got, x100 = auto.scaff_score_batch(rows, want_x100=True)
            assert x100.tolist() == ox100[:len(rows)].tolist()
            np.testing.assert_allclose(got, ref[:len(rows)], rtol=1e-9, atol=0)


def test_auto_path_keeps_raw_records_when_summing_does_not_halve_them():
    n = 2000
    a, b, _ = S.scaff_links(n, 100_000, 4)
    with N.Context(0) as auto:
        auto.set_links(n, a, b, None)
        path, recs = auto.path()
        assert path == N.PATH_STREAM and recs == len(a)
    real = np.random.default_rng(5).gamma(2.0, 1.5, len(a))
    with N.Context(0) as auto:
        auto.set_links(n, a, b, real)                                 # real weights: never summed
        assert auto.path() == (N.PATH_STREAM, len(a))


@pytest.mark.parametrize("P", [129, 133, 260, 300])
def test_population_not_a_multiple_of_128(P):
    """Whole groups of 128 go through k_scaff_score, the remainder through streaming passes when the
    cost model says a padded group would cost more (scaff_score_enqueue): same integers as the
    oracle and as the single-kernel route, flags of both parts survive, graph replay included."""
    n, L = 700, 3_000_000
    a, b, w = S.scaff_links(n, L, 6)
    perms = S.scaff_population(n, P, 106)
    _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=8)
    with N.Context(0, path=N.PATH_STREAM) as c, N.Context(0, path=N.PATH_STREAM, flags=N.FLAG_NO_STREAM1) as cb:
        c.set_links(n, a, b, None)
        cb.set_links(n, a, b, None)
        got, x100 = c.scaff_score_batch(perms, want_x100=True)
        rem = P % 128
        want_launches = 5 + 3 * ((rem + 7) // 8) if rem <= 40 else 5      # layout, check, score, head, finish (+ passes)
        assert c.timing()["launches"] == want_launches
        gb, xb = cb.scaff_score_batch(perms, want_x100=True)
        assert cb.timing()["launches"] == 5
        assert x100.tolist() == ox100.tolist() == xb.tolist() and got.tolist() == gb.tolist()
        for _ in range(3):
            g2, x2 = c.scaff_score_batch(perms, want_x100=True)
            assert x2.tolist() == ox100.tolist()
        for row in (3, P - 1):                                            # a bad row in either part
            bad = perms.copy()
            bad[row, 5] = bad[row, 6]
            with pytest.raises(N.LxyError) as e:
                c.scaff_score_batch(bad)
            assert e.value.status == -4


@pytest.mark.parametrize("n,P", [(11, 3), (2000, 40), (2050, 33), (300_000, 130), (1000, 1), (31, 64), (33, 65)])
def test_packed_phasing_entry_point_equals_the_byte_one(n, P):
    pa, pb, pw, truth = S.phase_links(n, 7, far_pairs=min(100, n))
    genes = S.phase_population(n, P, 107, truth)
    _, exact = O.phase_score_links_batch(n, pa, pb, pw, genes, nthreads=4)
    with N.Context(0) as c:
        c.set_phase_links(n, pa, pb, pw)
        oi, of = c.phase_score_batch(genes)
        words = N.pack_bits(genes)
        bi, bf = c.phase_score_batch_bits(words)
        assert oi.tolist() == exact.tolist() == bi.tolist() and of.tolist() == bf.tolist()
        if n % 32:                                                        # bits past n are ignored
            words[:, -1] |= np.uint32((0xFFFFFFFF << (n % 32)) & 0xFFFFFFFF)
        assert c.phase_score_batch_bits(words)[0].tolist() == exact.tolist()


@pytest.fixture()
def comm1():
    """Two contexts (stream / auto) with a communicator of ONE rank each."""
    cs = [N.Context(0, path=N.PATH_STREAM), N.Context(0)]
    for c in cs:
        c.comm_init_rank(1, 0, N.comm_unique_id())
        assert c.comm_info() == (1, 0)
    yield cs
    for c in cs:
        c.close()


@pytest.mark.parametrize("mode", [N.SHARD_POPULATION, N.SHARD_LINKS])
def test_sharded_scorer_world_one(comm1, mode):
    torch = pytest.importorskip("torch")
    n, L, P = 900, 300_000, 150
    a, b, w = S.scaff_links(n, L, 8)
    perms = S.scaff_population(n, P, 108)
    _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=8)
    wr = np.random.default_rng(9).gamma(2.0, 1.5, L)
    ref_real = O.Links.from_records(n, a, b, wr).scaff_score_batch(perms, nthreads=8)
    for c in comm1:
        c.set_links(n, a, b, None)
        plain, xp = c.scaff_score_batch(perms, want_x100=True)
        got, x100 = c.scaff_score_sharded(mode, perms, want_x100=True)
        assert x100.tolist() == ox100.tolist() == xp.tolist() and got.tolist() == plain.tolist()
        pd = torch.from_numpy(perms).cuda()
        od, xd = torch.zeros(P, dtype=torch.float64, device="cuda"), torch.zeros(P, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        c.scaff_score_sharded_dev(mode, pd.data_ptr(), P, od.data_ptr(), xd.data_ptr())
        c.sync()
        assert xd.cpu().numpy().tolist() == ox100.tolist() and od.cpu().numpy().tolist() == plain.tolist()
        bad = perms.copy()
        bad[P - 2, 1] = bad[P - 2, 2]
        with pytest.raises(N.LxyError) as e:
            c.scaff_score_sharded(mode, bad)
        assert e.value.status == -4
        c.set_links(n, a, b, wr)                               # real weights: fp64 partials, rank order
        got, x100 = c.scaff_score_sharded(mode, perms, want_x100=True)
        assert (x100 == N.INT64_MIN).all()
        np.testing.assert_allclose(got, ref_real, rtol=1e-9, atol=0)


@pytest.mark.parametrize("mode", [N.SHARD_POPULATION, N.SHARD_LINKS])
def test_sharded_ga_world_one_equals_the_local_loop(comm1, mode):
    n, L = 400, 120_000
    a, b, _ = S.scaff_links(n, L, 10)
    c = comm1[0]
    c.set_links(n, a, b, None)
    runs = []
    for shard in (-1, mode):
        with N.GA(c, N.GA_SCAFF, popsize=64, seed=11) as ga:
            ga.set_sharding(shard)
            ga.init()
            ga.optimize(6)
            runs.append((ga.population()[0].tolist(), ga.population()[1].tolist(), ga.stats()["scored"]))
    assert runs[0] == runs[1]


def test_total_length_boundary_of_the_g_profile():
    """Scaffold coordinates are 32-bit: a total length of 2^32 - 1 bp is accepted (and scored exactly),
    2^32 is rejected with LXY_ERR_RANGE; a context holds fewer than 2^32 records."""
    n = 8
    lens = np.full(n, (1 << 29), np.uint32)
    lens[-1] = (1 << 29) - 1                                   # total = 2^32 - 1
    rng = np.random.default_rng(12)
    L = 5000
    a = rng.integers(0, n, L).astype(np.uint32)
    b = ((a + rng.integers(1, n, L)) % n).astype(np.uint32)
    xa = (rng.random(L) * lens[a]).astype(np.uint32)
    xb = (rng.random(L) * lens[b]).astype(np.uint32)
    perms = S.scaff_population(n, 5, 112)
    oris = (rng.random((5, n)) < 0.5).astype(np.uint8)
    unit = 1 << 24
    with N.Context(0) as c:
        c.set_links_g(lens, unit, a, xa, b, xb, None)
        got, x100 = c.scaffg_score_batch(perms, oris, want_x100=True)
        _, ox = O.scaffg_score_links_batch(lens, unit, a, xa, b, xb, np.ones(L), perms, oris, nthreads=2)
        assert x100.tolist() == ox.tolist()
        lens[-1] += 1                                           # total = 2^32
        with pytest.raises(N.LxyError) as e:
            c.set_links_g(lens, unit, a, xa, b, xb, None)
        assert e.value.status == -5
        with pytest.raises(N.LxyError) as e:                    # the record-count limit is checked before any pointer is read
            N._check(N.lib().lxy_ctx_set_links(c._h, n, (1 << 32), None, None, None))
        assert e.value.status == -5


@pytest.mark.parametrize("P", [512, 1000, 1300])
def test_packed_phasing_from_page_locked_memory_uploads_in_slices(P):
    """Page-locked caller memory and P >= 512: the rows go up in slices on the side stream while the
    slices already there are scored (api.cu).  Same integers as the byte entry point and the oracle,
    call after call (the device buffer is reused)."""
    torch = pytest.importorskip("torch")
    n = 20_011
    pa, pb, pw, truth = S.phase_links(n, 9, far_pairs=50)
    genes = S.phase_population(n, P, 109, truth)
    _, exact = O.phase_score_links_batch(n, pa, pb, pw, genes, nthreads=8)
    words = torch.from_numpy(N.pack_bits(genes).view(np.int32)).pin_memory().numpy().view(np.uint32)
    with N.Context(0) as c:
        c.set_phase_links(n, pa, pb, pw)
        for _ in range(3):
            assert c.phase_score_batch_bits(words)[0].tolist() == exact.tolist()
        flipped = words.copy()                                  # pageable copy of a different population
        flipped[:, 0] ^= np.uint32(1)
        g2 = genes.copy()
        g2[:, 0] ^= 1
        _, exact2 = O.phase_score_links_batch(n, pa, pb, pw, g2, nthreads=8)
        assert c.phase_score_batch_bits(flipped)[0].tolist() == exact2.tolist()
        assert c.phase_score_batch_bits(words)[0].tolist() == exact.tolist()


def test_deferred_errors_are_reported_by_the_next_sync():
    """LXY_FLAG_DEFER_ERRORS: device-pointer scoring calls queue back to back; the on-device validation
    still runs in every call and its verdict comes with lxy_ctx_sync(), as one status for the calls since
    the last sync.  Host-buffer calls on the same context keep reporting at once."""
    torch = pytest.importorskip("torch")
    n, L, P = 600, 200_000, 40
    a, b, w = S.scaff_links(n, L, 12)
    perms = S.scaff_population(n, P, 112)
    _, ox100 = O.scaff_score_links_batch(n, a, b, w, perms, nthreads=4)
    bad = perms.copy()
    bad[7, 3] = bad[7, 4]
    with N.Context(0, path=N.PATH_STREAM, flags=N.FLAG_DEFER_ERRORS) as c:
        c.set_links(n, a, b, None)
        good_d, bad_d = torch.from_numpy(perms).cuda(), torch.from_numpy(bad).cuda()
        out = torch.zeros(P, dtype=torch.float64, device="cuda")
        x = torch.zeros(P, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        for _ in range(3):                                      # good calls: nothing to report
            c.scaff_score_batch_dev(good_d.data_ptr(), P, out.data_ptr(), x.data_ptr())
        c.sync()
        assert x.cpu().numpy().tolist() == ox100.tolist()
        c.scaff_score_batch_dev(good_d.data_ptr(), P, out.data_ptr(), x.data_ptr())
        c.scaff_score_batch_dev(bad_d.data_ptr(), P, out.data_ptr(), x.data_ptr())      # returns at once
        c.scaff_score_batch_dev(good_d.data_ptr(), P, out.data_ptr(), x.data_ptr())
        with pytest.raises(N.LxyError) as e:
            c.sync()
        assert e.value.status == -4 and "deferred" in str(e.value) and "3 scoring calls" in str(e.value)
        c.sync()                                                # reported once
        assert x.cpu().numpy().tolist() == ox100.tolist()       # the good call after the bad one is intact
        for _ in range(70):                                     # more unreported calls than slots
            c.scaff_score_batch_dev(good_d.data_ptr(), P, out.data_ptr(), x.data_ptr())
        c.sync()
        with pytest.raises(N.LxyError) as e:                    # host buffers: reported at once, as without the flag
            c.scaff_score_batch(bad)
        assert e.value.status == -4
        assert c.scaff_score_batch(perms, want_x100=True)[1].tolist() == ox100.tolist()


@pytest.mark.parametrize("path", [N.PATH_AUTO, N.PATH_STREAM])
def test_sixteen_bit_genes_entry_point(path):
    """lxy_scaff_score_batch_u16: the orderings as uint16 (half the PCIe bytes), same integers as the
    int32 entry point and the oracle on both scoring paths; a bad row is reported; n > 65,536 is refused."""
    n, P = 700, 150
    a, b = _dup_links(n, 40_000, 6, 13)                       # dense enough for the pair table under AUTO
    perms = S.scaff_population(n, P, 113)
    _, ox100 = O.scaff_score_links_batch(n, a, b, np.ones(len(a)), perms, nthreads=4)
    with N.Context(0, path=path) as c:
        c.set_links(n, a, b, None)
        s32, x32 = c.scaff_score_batch(perms, want_x100=True)
        for _ in range(3):                                     # repeated calls: the captured graph is replayed
            s16, x16 = c.scaff_score_batch_u16(perms.astype(np.uint16), want_x100=True)
            assert x16.tolist() == x32.tolist() == ox100.tolist() and s16.tolist() == s32.tolist()
        bad = perms.astype(np.uint16)
        bad[149, 3] = bad[149, 4]
        with pytest.raises(N.LxyError) as e:
            c.scaff_score_batch_u16(bad)
        assert e.value.status == -4
    big = 65_537
    ab = np.arange(big - 1, dtype=np.uint32)
    with N.Context(0, path=N.PATH_STREAM) as c:
        c.set_links(big, ab, ab + 1, None)
        with pytest.raises(N.LxyError) as e:
            c.scaff_score_batch_u16(np.zeros((1, big), np.uint16))
        assert e.value.status == -5
