"""Coordinate read -> probe assignment: RangeMap.getObjectsThatContainRangeInclusive
(nimblegen_heatseq/src/main/java/com/roche/heatseq/process/RangeMap.java:92-153) and ReadToProbeAssigner.assignProbesToReads
(…/PrimerReadExtensionAndPcrDuplicateIdentification.java:439-576).

The reference holds no asserting test for this path (RangeMap.main's go3/go4 only print, RangeMap.java:234-254), so the oracle
is pinned on those two inputs with hand-traced answers, on a brute-force containment scan where the reference's own
"no probe contains another" assumption holds (the two then agree by construction), and on a hand-traced nested case where
they differ (the quirks -- nested windows, a window starting exactly at the query start -- are part of the behaviour).  The CUDA path is compared with the oracle bit for bit."""
import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from bioinformatics_b200.batches import MappedRecordsHost, ProbeRangesHost
from oracle import oracle as O


def _ranges(windows, contig=None, reverse=None):
    n = len(windows)
    return ProbeRangesHost(contig if contig is not None else np.zeros(n), reverse if reverse is not None else np.zeros(n),
                           [w[0] for w in windows], [w[1] for w in windows])


def _records(queries, contig=None, negative=None, second=None, read_index=None, alt=None):
    n = len(queries)
    z = np.zeros(n)
    return MappedRecordsHost(contig if contig is not None else z, negative if negative is not None else z,
                             second if second is not None else z, [q[0] for q in queries], [q[1] for q in queries],
                             read_index if read_index is not None else np.arange(n), alt)


def test_oracle_rangemap_go3_go4():
    # go3 (RangeMap.java:234-242): only "d" [8712,8851] contains [8741,8787]
    out = O.assign_probes(_ranges([(2145, 2284), (6462, 6601), (8617, 8756), (8712, 8851)]), _records([(8741, 8787)]))
    assert out.probe_sets() == [(3,)]
    # go4 (:244-252): all four staggered windows contain [5,7]
    out = O.assign_probes(_ranges([(1, 10), (2, 11), (3, 12), (4, 13)]), _records([(5, 7)]))
    assert out.probe_sets() == [(0, 1, 2, 3)]
    # put() normalises (start, stop) and the query's order does not matter (:62-67, :93-94)
    out = O.assign_probes(_ranges([(10, 1), (11, 2)]), _records([(7, 5), (1, 1), (11, 11), (12, 12), (0, 0)]))
    assert out.probe_sets() == [(0, 1), (0,), (1,), (), ()]


def test_oracle_rangemap_nested_window_quirk():
    # A[1,100] contains B[10,20] and C[30,40] ("assumes that no probe completely overlaps another probe", RangeMap.java:125):
    #  [35,36]: the walk starts at C, stops at B (ends before the query), never reaches A            -> C only
    #  [15,16]: the search lands on B, the walk continues to A                                       -> A and B
    #  [50,60]: lands on C, which ends before the query: the walk stops at once, A is missed        -> nothing
    #  [5,6]:   lands on A                                                                            -> A
    out = O.assign_probes(_ranges([(1, 100), (10, 20), (30, 40)]), _records([(35, 36), (15, 16), (50, 60), (5, 6)]))
    assert out.probe_sets() == [(2,), (0, 1), (), (0,)]


def test_search_windows_from_probe_table():
    # getProbeSearchBoundaries (…Identification.java:287-300) over Probe.getStart/getStop (Probe.java:75-81)
    from bioinformatics_b200.batches import ProbeTableHost
    probes = ProbeTableHost([0, 1], ext_start=[100, 520], ext_stop=[120, 500], lig_start=[280, 330], lig_stop=[300, 310],
                            target_start=[121, 331], target_stop=[279, 499], ext=["ACGT", "ACGT"], lig=["ACGT", "ACGT"], target=["AC", "AC"])
    r = ProbeRangesHost.from_probe_table(probes, [0, 1])
    assert r.start.tolist() == [100, 310] and r.stop.tolist() == [300, 520] and r.reverse_strand.tolist() == [0, 1]
    r = ProbeRangesHost.from_probe_table(probes, [0, 1], 5, 7)
    assert r.start.tolist() == [105, 315] and r.stop.tolist() == [293, 513]


def _staggered_windows(rng, n, lo=1000):
    """Distinct starts, ends increasing with the starts: no window contains another."""
    starts = lo + np.cumsum(rng.integers(1, 120, n))
    ends = np.maximum.accumulate(starts + rng.integers(60, 400, n) + np.arange(n))  # strictly increasing
    ends = ends + np.arange(n)
    return starts.astype(np.int64), ends.astype(np.int64)


def test_oracle_equals_brute_force_without_containment():
    rng = np.random.default_rng(11)
    for trial in range(20):
        n = int(rng.integers(1, 60))
        starts, ends = _staggered_windows(rng, n)
        perm = rng.permutation(n)  # probe-file order is not coordinate order
        windows = [(int(starts[p]), int(ends[p])) for p in perm]
        qa = rng.integers(starts.min() - 50, ends.max() + 50, 300)
        qb = qa + rng.integers(0, 150, 300)
        out = O.assign_probes(_ranges(windows), _records(list(zip(qa.tolist(), qb.tolist()))), max_probes=16)
        assert not out.overflow.any()
        for k in range(300):
            if qa[k] in starts:
                continue  # a window starting exactly at the query start can be skipped by the search: next test
            brute = tuple(i for i, (s, e) in enumerate(windows) if s <= qa[k] and e >= qb[k])
            assert out.probe_sets()[k] == brute, (trial, k)


def test_oracle_rangemap_equal_start_landing_quirk():
    # hand trace of RangeMap.java:103-150 on 8 windows, query [1171,1306]: lo=1,hi=8 -> after=4 (1314 >= 1171, before 1242 > 1171)
    # -> hi=3 -> after=2: window[2].start == 1171 >= 1171 and window[1].start 1111 <= 1171: the search lands on window[1] and
    # walks left, so window[2] = [1171,1550], which also contains the query, is never looked at.
    windows = [(1016, 1126), (1111, 1309), (1171, 1550), (1242, 1551), (1314, 1552), (1399, 1653), (1403, 1654), (1461, 1791)]
    out = O.assign_probes(_ranges(windows), _records([(1171, 1306), (1172, 1306), (1170, 1306)]))
    assert out.probe_sets() == [(1,), (1, 2), (1,)]


def test_oracle_strand_and_contig_rules():
    # probe 0 forward on contig 0, probe 1 reverse on contig 0, probe 2 forward on contig 2; contig 1 has no probe
    ranges = ProbeRangesHost([0, 0, 2], [0, 1, 0], [100, 100, 100], [300, 300, 300])
    rec = MappedRecordsHost(contig=[0, 0, 0, 0, 1, 2, 5, -1], negative_strand=[0, 1, 0, 1, 0, 0, 0, 0], second_of_pair=[0, 0, 1, 1, 0, 1, 0, 0],
                            start=[150] * 8, end=[200] * 8, read_index=np.arange(8))
    out = O.assign_probes(ranges, rec)
    # read two carries the opposite strand of its probe (…Identification.java:452-455)
    assert out.probe_sets() == [(0,), (1,), (1,), (0,), (), (), (), ()]
    assert out.mate.tolist() == [-1, -1, -1, -1, A.ASSIGN_IGNORED, -1, A.ASSIGN_IGNORED, A.ASSIGN_IGNORED]


def test_oracle_pairing_rules():
    ranges = _ranges([(100, 300), (150, 350), (1000, 1200)])
    # name 7: two records -> pair, common = intersection; name 8: three -> first two pair, third alone;
    # name 9: four -> removed; name 10: one -> unpaired; name 11: five -> removed
    names = [7, 8, 9, 7, 8, 9, 10, 8, 9, 9, 11, 11, 11, 11, 11]
    q = [(160, 200), (110, 120), (160, 170), (120, 140), (160, 170), (160, 170), (160, 170), (160, 170), (160, 170), (160, 170)] + [(1000, 1001)] * 5
    out = O.assign_probes(ranges, _records(q, read_index=names))
    amb = A.ASSIGN_AMBIGUOUS_NAME
    assert out.mate.tolist() == [3, 4, amb, 0, 1, amb, -1, -1, amb, amb] + [amb] * 5
    assert out.probe_sets()[0] == (0, 1) and out.probe_sets()[3] == (0,)
    assert out.common_sets()[0] == (0,) and out.common_sets()[3] == (0,)
    assert out.common_sets()[1] == (0,) and out.common_sets()[4] == (0,)
    assert out.n_common[[2, 5, 6, 7, 8, 9]].tolist() == [0] * 6


def test_oracle_alternative_hits_and_overflow():
    ranges = ProbeRangesHost([0, 0, 1, 1], [0, 0, 0, 1], [100, 150, 100, 100], [300, 350, 300, 300])
    # record 0 (read one, forward): own hit -> probe 0; alt hits: contig 1 forward -> 2; contig 1 reverse -> 3; contig 9 -> nothing
    # record 1 (read two, forward): strand flipped for the record AND for its alternative hits (:555-562)
    alt = ([0, 3, 4], [1, 1, 9, 1], [0, 1, 0, 0], [150, 150, 150, 150], [200, 200, 200, 200])
    rec = MappedRecordsHost([0, 0], [0, 0], [0, 1], [110, 110], [120, 120], [5, 5], alt)
    out = O.assign_probes(ranges, rec)
    assert out.probe_sets() == [(0, 2, 3), (3,)]
    assert out.mate.tolist() == [1, 0] and out.common_sets() == [(3,), (3,)]
    tight = O.assign_probes(ranges, rec, max_probes=2)
    assert tight.overflow.tolist() == [1, 1]  # the mate of an overflowing record is tainted too
    assert tight.n_probes[0] == 2


def _random_problem(rng, n_probes, n_records, n_contigs):
    contig = rng.integers(0, n_contigs, n_probes)
    contig[contig == 3] = 4  # contig 3 holds no probe
    reverse = rng.integers(0, 2, n_probes)
    start = rng.integers(1, 200000, n_probes)
    length = rng.integers(80, 400, n_probes)
    # a fifth of the probes share a start with, or are nested in, another probe: the landing rules of the search matter
    dup = rng.random(n_probes) < 0.2
    src = rng.integers(0, n_probes, n_probes)
    start[dup] = start[src[dup]] + rng.integers(0, 3, int(dup.sum())) * 7
    contig[dup] = contig[src[dup]]
    reverse[dup] = reverse[src[dup]]
    stop = start + length
    swap = rng.random(n_probes) < 0.1
    start[swap], stop[swap] = stop[swap], start[swap]
    ranges = ProbeRangesHost(contig, reverse, start, stop)
    # records: most sit inside a probe window
    p = rng.integers(0, n_probes, n_records)
    lo = np.minimum(start, stop)[p] + rng.integers(-20, 60, n_records)
    hi = lo + rng.integers(20, 150, n_records)
    rc = contig[p].copy()
    stray = rng.random(n_records) < 0.05
    rc[stray] = rng.integers(-1, n_contigs + 2, int(stray.sum()))
    neg = reverse[p] ^ (rng.random(n_records) < 0.1)
    second = rng.integers(0, 2, n_records)
    neg = neg ^ second  # read two maps to the opposite strand
    names = rng.integers(0, max(1, n_records // 2), n_records)  # counts 0..6 per name
    n_alt = rng.integers(0, 3, n_records) * (rng.random(n_records) < 0.3)
    alt_off = np.concatenate([[0], np.cumsum(n_alt)])
    ta = int(alt_off[-1])
    ap = rng.integers(0, n_probes, ta)
    a_lo = np.minimum(start, stop)[ap] + rng.integers(-10, 40, ta)
    a_hi = a_lo + rng.integers(20, 100, ta)
    a_contig = contig[ap].copy()
    a_contig[rng.random(ta) < 0.1] = -1
    alt = (alt_off, a_contig, rng.integers(0, 2, ta), a_lo, a_hi)
    return ranges, MappedRecordsHost(rc, neg, second, lo, hi, names, alt)


def _same(a, b):
    assert (a.overflow == b.overflow).all()
    assert (a.mate == b.mate).all()
    ok = a.overflow == 0
    assert (a.n_probes[ok] == b.n_probes[ok]).all() and (a.n_common[ok] == b.n_common[ok]).all()
    cols = np.arange(a.max_probes)[None, :]
    live = ok[:, None] & (cols < a.n_probes[:, None])
    assert (a.probes[live] == b.probes[live]).all()
    live = ok[:, None] & (cols < a.n_common[:, None])
    assert (a.common[live] == b.common[live]).all()


def test_oracle_random_problem_is_self_consistent():
    rng = np.random.default_rng(3)
    ranges, rec = _random_problem(rng, 400, 5000, 8)
    out = O.assign_probes(ranges, rec, max_probes=8)
    paired = out.mate >= 0
    assert paired.any() and (out.mate == A.ASSIGN_AMBIGUOUS_NAME).any() and (out.mate == A.ASSIGN_IGNORED).any()
    k = np.nonzero(paired)[0]
    assert (out.mate[out.mate[k]] == k).all()  # mates point at each other
    assert (rec.read_index[out.mate[k]] == rec.read_index[k]).all()
    assert (out.n_common[k] == out.n_common[out.mate[k]]).all()
    assert (out.n_common <= out.n_probes).all()


@pytest.mark.gpu
def test_cuda_assign_hand_cases():
    from bioinformatics_b200._lib import Context
    with Context(0) as ctx:
        ctx.set_probe_ranges(_ranges([(2145, 2284), (6462, 6601), (8617, 8756), (8712, 8851)]))
        assert ctx.assign_probes(_records([(8741, 8787)])).probe_sets() == [(3,)]
        ctx.set_probe_ranges(_ranges([(1, 10), (2, 11), (3, 12), (4, 13)]))
        assert ctx.assign_probes(_records([(5, 7)])).probe_sets() == [(0, 1, 2, 3)]
        ctx.set_probe_ranges(_ranges([(1, 100), (10, 20), (30, 40)]))
        assert ctx.assign_probes(_records([(35, 36), (15, 16), (50, 60), (5, 6)])).probe_sets() == [(2,), (0, 1), (), (0,)]
        ranges = _ranges([(100, 300), (150, 350), (1000, 1200)])
        names = [7, 8, 9, 7, 8, 9, 10, 8, 9, 9, 11, 11, 11, 11, 11]
        q = [(160, 200), (110, 120), (160, 170), (120, 140)] + [(160, 170)] * 6 + [(1000, 1001)] * 5
        ctx.set_probe_ranges(ranges)
        rec = _records(q, read_index=names)
        _same(ctx.assign_probes(rec), O.assign_probes(ranges, rec))
        # no probes at all / no records at all
        ctx.set_probe_ranges(_ranges([]))
        out = ctx.assign_probes(_records([(1, 2), (3, 4)]))
        assert out.mate.tolist() == [A.ASSIGN_IGNORED] * 2 and out.n_probes.tolist() == [0, 0]
        assert ctx.assign_probes(_records([])).n == 0


@pytest.mark.gpu
@pytest.mark.parametrize("seed,n_probes,n_records,n_contigs,max_probes", [(1, 50, 3000, 3, 8), (2, 5000, 200000, 24, 8), (3, 2000, 50000, 6, 2),
                                                                          (4, 1, 1000, 1, 1), (5, 60000, 400000, 30, 4)])
def test_cuda_assign_matches_oracle(seed, n_probes, n_records, n_contigs, max_probes):
    from bioinformatics_b200._lib import Context
    rng = np.random.default_rng(seed)
    ranges, rec = _random_problem(rng, n_probes, n_records, n_contigs)
    want = O.assign_probes(ranges, rec, max_probes=max_probes)
    with Context(0) as ctx:
        ctx.set_probe_ranges(ranges)
        got = ctx.assign_probes(rec, max_probes=max_probes)
        _same(got, want)
        if max_probes == 2:
            assert want.overflow.any()
        # without alternative hits (alt_offsets NULL)
        rec2 = MappedRecordsHost(rec.contig, rec.negative_strand, rec.second_of_pair, rec.start, rec.end, rec.read_index)
        _same(ctx.assign_probes(rec2, max_probes=max_probes), O.assign_probes(ranges, rec2, max_probes=max_probes))


@pytest.mark.gpu
def test_cuda_assign_argument_errors():
    from bioinformatics_b200._lib import Context, HsqError
    with Context(0) as ctx:
        with pytest.raises(HsqError) as e:  # ranges not set
            ctx.assign_probes(_records([(1, 2)]))
        assert e.value.status == A.HSQ_ERR_INVALID_ARGUMENT
        with pytest.raises(HsqError):
            ctx.set_probe_ranges(ProbeRangesHost([-1], [0], [1], [2]))
        ctx.set_probe_ranges(_ranges([(1, 10)]))
        with pytest.raises(HsqError):
            ctx.assign_probes(_records([(1, 2)]), max_probes=0)
