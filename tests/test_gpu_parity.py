"""GPU parity: the CUDA path through the C ABI vs the CPU oracle on the same seeded inputs (bit-exact)."""
import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from bioinformatics_b200 import synth
from bioinformatics_b200._lib import Context, HsqError
from bioinformatics_b200.batches import PairBatchHost, ProbeTableHost, scorer
from oracle import oracle as O
from tests import helpers as H
from tests.test_oracle_kat import MD_KATS, T2_QRY, T2_REF, T3_QRY, T3_REF, T4_REF

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.close()


def test_align_reference_kats(ctx):
    refs = [k[0] for k in MD_KATS] + [T3_REF, T4_REF, T2_REF[::-1], "CATTCCG", "ACTGAAATTTGGGCCCAGTCTAAANNNNTTTTT"]
    qrys = [k[1] for k in MD_KATS] + [T3_QRY, T3_QRY, T2_QRY[::-1], "CACTCCG", "ATTTGGGC"]
    got = ctx.align_batch(refs, qrys)
    for k, kat in enumerate(MD_KATS):
        assert got.mds()[k] == kat[2]
    assert got.first_q[-1] == 6  # AlignmentTest.java:34-40
    assert got.op_strings()[-2].count("X") + got.op_strings()[-2].count("I") == 1  # CigarStringUtilTest.java:112-120
    H.assert_align_equal(got, O.align_batch(refs, qrys))


@pytest.mark.parametrize("flags", [0, 1, 2, 3])
@pytest.mark.parametrize("sc", [(1, -4, -1, -6), (2, -3, -2, -5), (1, -1, -1, -1), (5, -4, 0, -70), (1, -4, 1, 2)])
def test_align_random_vs_oracle(ctx, flags, sc):
    rng = np.random.default_rng(1000 + flags + 17 * abs(sc[3]))
    refs, qrys = H.align_cases(rng, 1200)
    s = scorer(*sc)
    got = ctx.align_batch(refs, qrys, s, flags)
    exp = O.align_batch(refs, qrys, s, flags, nthreads=8)
    H.assert_align_equal(got, exp)


def test_align_long(ctx):
    rng = np.random.default_rng(5)
    refs = [H.random_seq(rng, int(rng.integers(250, 320))) for _ in range(64)]
    qrys = [H.mutate(rng, r, 0.03, 0.02) for r in refs]
    got = ctx.align_batch(refs, qrys, cols_cap=800, md_cap=800)
    H.assert_align_equal(got, O.align_batch(refs, qrys, cols_cap=800, md_cap=800, nthreads=8))


def test_align_bad_character(ctx):
    with pytest.raises(HsqError) as e:
        ctx.align_batch(["ACGT", "ACXT"], ["ACGT", "ACGT"])
    assert e.value.status == A.HSQ_ERR_BAD_CHARACTER


def test_align_overflow_flag(ctx):
    with pytest.raises(HsqError) as e:
        ctx.align_batch(["ACGTACGTACGT"], ["ACGTACGAACGT"], cols_cap=4, md_cap=2)
    assert e.value.status == A.HSQ_ERR_OVERFLOW


@pytest.mark.parametrize("read_len,uid", [(150, 10), (100, 14), (60, 8)])
def test_extend_synthetic_vs_oracle(ctx, read_len, uid):
    probes = synth.make_probes(300, seed=11 + read_len)
    pairs = synth.make_pairs(probes, 6000, read_len=read_len, ext_uid_len=uid, seed=29 + read_len, sub_rate=0.02,
                             ins_rate=0.003, del_rate=0.003)
    ctx.set_probes(probes)
    ctx.set_scorer(scorer())
    got = ctx.extend_batch(pairs, cigar_cap=64, md_cap=96)
    exp = O.extend_batch(probes, pairs, cigar_cap=64, md_cap=96, nthreads=8)
    H.assert_extend_equal(got, exp)
    assert got.pair_ok.mean() > 0.5


def test_extend_adversarial_vs_oracle(ctx):
    """Random reads against random probes (most primers not found), short reads, IUPAC primers, custom scorer."""
    rng = np.random.default_rng(77)
    n_probes, n = 40, 1500
    ext = [H.random_seq(rng, int(rng.integers(5, 30)), "ACGTN" if k % 5 == 0 else "ACGT") for k in range(n_probes)]
    lig = [H.random_seq(rng, int(rng.integers(5, 30))) for _ in range(n_probes)]
    tgt = [H.random_seq(rng, int(rng.integers(1, 170))) for _ in range(n_probes)]
    z = np.arange(n_probes, dtype=np.int32) * 500
    probes = ProbeTableHost(np.arange(n_probes) % 2, z + 10, z + 30, z + 300, z + 320, z + 31, z + 299, ext, lig, tgt)
    pid = rng.integers(0, n_probes, n)
    r1, r2 = [], []
    for k in range(n):
        p = int(pid[k])
        kind = k % 4
        if kind == 0:
            a, b = H.random_seq(rng, int(rng.integers(0, 160))), H.random_seq(rng, int(rng.integers(0, 160)))
        else:
            mol = ext[p].replace("N", "A") + tgt[p] + lig[p]
            rc = mol[::-1].translate(str.maketrans("ACGT", "TGCA"))
            L = int(rng.integers(12, 170))
            a = (H.random_seq(rng, 6) + H.mutate(rng, mol, 0.04, 0.02) + H.random_seq(rng, 200))[:L]
            b = (H.mutate(rng, rc, 0.04, 0.02) + H.random_seq(rng, 200))[:L]
            if kind == 3:
                a = a.replace("A", "N", 1)
        r1.append(a)
        r2.append(b)
    pairs = PairBatchHost(pid, r1, r2)
    # the last two scorers exceed the 14-bit score range of the packed kernel and take the 32-bit kernel
    for sc in (scorer(), scorer(2, -3, -2, -4), scorer(1, -4, 1, 2), scorer(5, -4, 0, -70), scorer(40, -30, -20, -45)):
        ctx.set_probes(probes)
        ctx.set_scorer(sc)
        got = ctx.extend_batch(pairs, cigar_cap=1600, md_cap=2400)
        exp = O.extend_batch(probes, pairs, sc, cigar_cap=1600, md_cap=2400, nthreads=8)
        H.assert_extend_equal(got, exp)
    ctx.set_scorer(scorer())


def test_extend_forced_32bit_kernel():
    """hsq_config.force_dp32 routes hsq_extend_batch through the 32-bit capture kernel: same results."""
    probes = synth.make_probes(100, seed=5)
    pairs = synth.make_pairs(probes, 3000, read_len=150, ext_uid_len=10, seed=6, sub_rate=0.02, ins_rate=0.003, del_rate=0.003)
    with Context(0, config={"force_dp32": 1}) as c32:
        c32.set_probes(probes)
        got = c32.extend_batch(pairs, cigar_cap=64, md_cap=96)
    H.assert_extend_equal(got, O.extend_batch(probes, pairs, cigar_cap=64, md_cap=96, nthreads=8))


@pytest.mark.parametrize("chunks", [2, 3, 7])
def test_extend_chunked_pipeline(ctx, chunks):
    """hsq_config.extend_chunks forces the H2D | kernels | D2H pipeline of hsq_extend_batch on a small batch: same results."""
    ctx.configure(extend_chunks=chunks, uniform_chunks=chunks == 3)
    probes = synth.make_probes(80, seed=15)
    pairs = synth.make_pairs(probes, 5001, read_len=150, ext_uid_len=10, seed=16, sub_rate=0.02, ins_rate=0.003, del_rate=0.003)
    ctx.set_probes(probes)
    ctx.set_scorer(scorer())
    got = ctx.extend_batch(pairs, cigar_cap=64, md_cap=96)
    H.assert_extend_equal(got, O.extend_batch(probes, pairs, cigar_cap=64, md_cap=96, nthreads=8))
    with pytest.raises(HsqError) as e:  # a too small slot in some chunk is reported, after all chunks ran
        ctx.extend_batch(pairs, cigar_cap=3, md_cap=96)
    assert e.value.status == A.HSQ_ERR_OVERFLOW
    ctx.configure()


def test_extend_non_acgt_capture_target(ctx):
    """IUPAC / N codes inside the capture target (the DP reference) take the kernels' slow-column path."""
    rng = np.random.default_rng(123)
    n_probes, n = 24, 2000
    ext = [H.random_seq(rng, 20) for _ in range(n_probes)]
    lig = [H.random_seq(rng, 21) for _ in range(n_probes)]
    tgt = []
    for k in range(n_probes):
        t = list(H.random_seq(rng, 120))
        for _ in range(k % 4):
            t[int(rng.integers(0, 120))] = str(rng.choice(list("NRYKMU")))
        tgt.append("".join(t))
    z = np.arange(n_probes, dtype=np.int32) * 500
    probes = ProbeTableHost(np.arange(n_probes) % 2, z + 10, z + 30, z + 300, z + 320, z + 31, z + 299, ext, lig, tgt)
    pid = rng.integers(0, n_probes, n)
    comp = str.maketrans("ACGTNRYKMU", "TGCANYRMKA")
    r1, r2 = [], []
    for k in range(n):
        p = int(pid[k])
        clean = "".join(c if c in "ACGT" else "A" for c in tgt[p])
        mol = ext[p] + clean + lig[p]
        rc = mol[::-1].translate(comp)
        a = (H.random_seq(rng, 10) + H.mutate(rng, mol, 0.02, 0.005) + H.random_seq(rng, 200))[:150]
        b = (H.mutate(rng, rc, 0.02, 0.005) + H.random_seq(rng, 200))[:150]
        r1.append(a)
        r2.append(b)
    pairs = PairBatchHost(pid, r1, r2)
    ctx.set_probes(probes)
    ctx.set_scorer(scorer())
    got = ctx.extend_batch(pairs, cigar_cap=200, md_cap=300)
    exp = O.extend_batch(probes, pairs, cigar_cap=200, md_cap=300, nthreads=8)
    H.assert_extend_equal(got, exp)
    assert got.pair_ok.mean() > 0.5


def test_dedup_vs_oracle(ctx):
    probes = synth.make_probes(50, seed=3)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 20000, read_len=150, ext_uid_len=6, lig_uid_len=2, seed=4))
    # force quality-sum ties so the read-index tie-break is exercised
    pairs.q1.data[:] = ord("I")
    pairs.q2.data[:] = ord("I")
    pairs.q1.data[::7] = ord("5")
    for (u, v) in ((6, 2), (6, 0), (0, 0), (10, 4)):
        ib, gs, st = ctx.dedup_batch(pairs, u, v, probes.n)
        eb, egs, est = O.dedup_batch(pairs, u, v, probes.n)
        assert (ib == eb).all() and (gs == egs).all()
        assert (st.unique_pairs == est.unique_pairs).all() and (st.duplicate_pairs == est.duplicate_pairs).all()
        assert (st.max_pairs_per_uid == est.max_pairs_per_uid).all()


def _dedup_equal(got, exp):
    (ib, gs, st), (eb, egs, est) = got, exp
    assert (ib == eb).all() and (gs == egs).all()
    assert (st.unique_pairs == est.unique_pairs).all() and (st.duplicate_pairs == est.duplicate_pairs).all()
    assert (st.max_pairs_per_uid == est.max_pairs_per_uid).all()


def test_dedup_segment_and_radix_paths(ctx):
    """The per-probe segment path (counting sort by probe + one block sorting and reducing each probe) and the radix-sort path give
    the oracle's answer: skewed panels, a probe larger than a block takes (whole batch falls back), probes without pairs, a sparse
    200 k-probe table, repeated read indexes (ties go to the earlier pair)."""
    from bioinformatics_b200._lib import Context
    rng = np.random.default_rng(5)
    probes = synth.make_probes(60, seed=23)
    base, _ = synth.drop_n_pairs(synth.make_pairs(probes, 30000, read_len=100, ext_uid_len=5, lig_uid_len=1, seed=24))
    base.q1.data[::5] = ord("5")
    cases = []
    skew = base.subset(np.arange(base.n))
    skew.probe_id[:] = np.minimum(rng.geometric(0.15, size=base.n) - 1, 59)  # probe 0 holds ~4500 pairs, the tail almost none
    cases.append(("skewed", skew, 60, 5, 1))
    hot = base.subset(np.arange(base.n))
    hot.probe_id[:] = np.where(rng.random(base.n) < 0.5, 7, hot.probe_id)  # ~15 k pairs on one probe: more than a block sorts
    cases.append(("hot probe", hot, 60, 5, 1))
    sparse = base.subset(np.arange(5000))
    sparse.probe_id[:] = rng.integers(0, 200000, size=5000)
    cases.append(("sparse", sparse, 200000, 5, 1))
    ties = base.subset(np.arange(8000))
    ties.read_index[:] = rng.integers(0, 50, size=8000)
    ties.q1.data[:] = ord("I")
    ties.q2.data[:] = ord("I")
    cases.append(("ties", ties, 60, 3, 0))
    cases.append(("no uid", base.head(3000), 60, 0, 0))
    cases.append(("one pair", base.head(1), 60, 5, 1))
    # oversized probes are split into UID-hash buckets; what a bucket cannot hold either falls back to the radix path:
    nouid_hot = base.subset(np.arange(6000))
    nouid_hot.probe_id[:] = 3
    cases.append(("no uid, one probe above a block's table", nouid_hot, 60, 0, 0))
    one_group = base.subset(np.arange(6000))
    one_group.probe_id[:] = 9
    one_group.r1.data.reshape(6000, -1)[:, :5] = np.frombuffer(b"ACGTA", dtype=np.uint8)
    one_group.r2.data.reshape(6000, -1)[:, :1] = ord("C")
    cases.append(("one UID group above a block's table", one_group, 60, 5, 1))
    small_panel = base.subset(np.arange(base.n))
    small_panel.probe_id[:] = rng.integers(0, 4, size=base.n)  # every probe oversized (7.5 k pairs each): all of them split
    cases.append(("every probe split", small_panel, 60, 5, 1))
    expected = {name: O.dedup_batch(pb, u, v, npr) for name, pb, npr, u, v in cases}
    for name, pb, npr, u, v in cases:
        _dedup_equal(ctx.dedup_batch(pb, u, v, npr), expected[name])
    with Context(0, config={"force_dedup_radix": 1}) as radix:
        for name, pb, npr, u, v in cases:
            _dedup_equal(radix.dedup_batch(pb, u, v, npr), expected[name])


def test_dedup_sharded_entry_points(ctx):
    """hsq_dedup_keys -> (device records) -> hsq_dedup_reduce_keys equals hsq_dedup_batch; records equal the stand-in."""
    import torch
    from bioinformatics_b200 import sharded
    probes = synth.make_probes(50, seed=3)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 20000, read_len=150, ext_uid_len=6, lig_uid_len=2, seed=4))
    pairs.q1.data[::7] = ord("5")
    be = sharded.DeviceBackend(ctx)
    keys = be.keys(pairs, 6, 2, probes.n, 1000)
    exp_keys = O.dedup_keys(pairs, 6, 2, 1000)
    assert (keys.cpu().numpy().reshape(-1) == exp_keys.view(np.uint8)).all()
    perm = torch.randperm(keys.shape[0], device=keys.device)  # the owner sees records in arbitrary order
    best, gsize, stats = be.reduce(keys[perm], 8, probes.n)
    eb, egs, est = O.dedup_batch(pairs, 6, 2, probes.n)
    p = perm.cpu().numpy()
    assert (best.cpu().numpy() == eb[p]).all() and (gsize.cpu().numpy() == egs[p]).all()
    assert (stats.unique_pairs == est.unique_pairs).all() and (stats.max_pairs_per_uid == est.max_pairs_per_uid).all()


def test_dedup_config3_shape(ctx):
    """BASELINE.json configs[2] shape at test size: 14 bp UID, 10 k probes."""
    probes = synth.make_probes(10000, seed=13)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 150000, read_len=150, ext_uid_len=14, seed=14))
    ib, gs, st = ctx.dedup_batch(pairs, 14, 0, probes.n)
    eb, egs, est = O.dedup_batch(pairs, 14, 0, probes.n)
    assert (ib == eb).all() and (gs == egs).all() and (st.unique_pairs == est.unique_pairs).all()
    assert (st.duplicate_pairs == est.duplicate_pairs).all() and (st.max_pairs_per_uid == est.max_pairs_per_uid).all()


@pytest.mark.parametrize("read_len", [200, 300])
def test_extend_long_reads(ctx, read_len):
    """configs[4] lengths: the packed kernels' score range still holds at 2 x 300 bp."""
    probes = synth.make_probes(100, seed=17, target_len=(read_len - 50, read_len + 10))
    pairs = synth.make_pairs(probes, 3000, read_len=read_len, ext_uid_len=10, seed=18, sub_rate=0.02, ins_rate=0.003, del_rate=0.003)
    ctx.set_probes(probes)
    ctx.set_scorer(scorer())
    got = ctx.extend_batch(pairs, cigar_cap=64, md_cap=128)
    H.assert_extend_equal(got, O.extend_batch(probes, pairs, cigar_cap=64, md_cap=128, nthreads=8))
    assert got.pair_ok.mean() > 0.5


def test_dedup_int16_wraparound(ctx):
    """2 x 450 bp at Q41+ overflows a Java short: best-pair choice must follow the wrapped sum."""
    rng = np.random.default_rng(9)
    n, L = 64, 450
    r = ["ACGTAC" + H.random_seq(rng, L - 6) for _ in range(n)]
    q1 = ["".join(chr(33 + int(x)) for x in rng.integers(30, 60, L)) for _ in range(n)]
    q2 = ["".join(chr(33 + int(x)) for x in rng.integers(30, 60, L)) for _ in range(n)]
    pairs = PairBatchHost(np.zeros(n, dtype=np.int32), r, r, q1, q2, np.arange(n)[::-1].copy())
    ib, gs, st = ctx.dedup_batch(pairs, 6, 0, 1)
    eb, egs, est = O.dedup_batch(pairs, 6, 0, 1)
    assert (ib == eb).all() and ib.sum() == 1 and (gs == egs).all()


def test_dedup_rejects_non_acgt_uid(ctx):
    pairs = PairBatchHost([0], ["ACNTACGTAC"], ["ACGTACGTAC"], ["IIIIIIIIII"], ["IIIIIIIIII"], [0])
    with pytest.raises(HsqError) as e:
        ctx.dedup_batch(pairs, 4, 0, 1)
    assert e.value.status == A.HSQ_ERR_UNSUPPORTED


def test_trim_vs_oracle(ctx):
    rng = np.random.default_rng(21)
    reads = [H.random_seq(rng, int(rng.integers(30, 160)), "ACGTN") for _ in range(5000)]
    quals = ["".join(chr(33 + int(x)) for x in rng.integers(2, 42, len(r))) for r in reads]
    for (first, last, tp) in ((5, 14, True), (22, 120, True), (0, 1000, True), (29, 10, False), (10, 40, False)):
        gs, gq = ctx.trim_batch(reads, quals, first, last, tp)
        es, eq = O.trim_batch(reads, quals, first, last, tp)
        assert (gs.offsets == es.offsets).all() and (gs.data == es.data).all() and (gq.data == eq.data).all()
    # FastqReadTrimmerTest.java:35-40
    s, q = ctx.trim_batch(["ACGTACGTACGTACGTACGT"], ["ABCDEFGHIJKLMNOPQRST"], 5, 14, True)
    assert s.strings() == ["CGTACGTACG"] and q.strings() == ["FGHIJKLMNO"]
    with pytest.raises(HsqError) as e:
        ctx.trim_batch(["ACGT"], ["IIII"], 4, 10, True)
    assert e.value.status == A.HSQ_ERR_OUT_OF_RANGE
    s, q = ctx.trim_batch([], [], 1, 5, True)
    assert s.n == 0
    # the chunked copy-in | kernel | copy-out pipeline (hsq_config.trim_chunks) on a small batch: same bytes
    for chunks in (2, 3, 8):
        ctx.configure(trim_chunks=chunks)
        gs, gq = ctx.trim_batch(reads, quals, 7, 99, True)
        es, eq = O.trim_batch(reads, quals, 7, 99, True)
        assert (gs.offsets == es.offsets).all() and (gs.data == es.data).all() and (gq.data == eq.data).all()
        gs, gq = ctx.trim_batch(reads[:3], quals[:3], 7, 99, True)  # fewer reads than chunks
        assert gs.strings() == [r[7:100] for r in reads[:3]]
    ctx.configure()


def test_resolve_multi_probe_vs_oracle(ctx):
    rng = np.random.default_rng(31)
    probes = synth.make_probes(30, seed=8)
    pairs = synth.make_pairs(probes, 400, read_len=120, ext_uid_len=8, seed=9, with_quals=False)
    offs, cand = [0], []
    for k in range(pairs.n):
        c = {int(pairs.probe_id[k])} if k % 3 else set()
        while len(c) < 2 + k % 3:
            c.add(int(rng.integers(0, probes.n)))
        c = list(c)
        if k % 5 == 0:  # duplicate candidate -> exact tie -> read removed from all probes
            c = [c[0], c[0]]
        cand += c
        offs.append(len(cand))
    ctx.set_probes(probes)
    gb, gm, ge = ctx.resolve_multi_probe(pairs.r1, pairs.r2, offs, cand, 8, 0)
    eb, em, ee = O.resolve_multi_probe(probes, pairs.r1, pairs.r2, offs, cand, 8, 0)
    assert (gm == em).all() and (ge == ee).all() and (gb == eb).all()
    assert (gb[::5] == -1).all() and (gb >= 0).sum() > 100
