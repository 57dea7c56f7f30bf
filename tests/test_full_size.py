"""BASELINE.json configurations at their FULL sizes on the GPU, checked through a vectorised numpy statement of the rule
(exact, not sampled) and through size-independent properties:
  configs[0]  trim, 100 k read pairs 2 x 150 bp                       (FastqReadTrimmer.java:309-328)
  configs[2]  dedup, 10 M read pairs, 14 bp UIDs, 10 k probes          (FilterByUid.java:134-203,281-297)
  configs[3]  probe-count scale of the 8-GPU panel: 200 k probes       (one GPU's share of the pairs)
configs[1] (10 M read ends through the DP kernels) is asserted at full size inside bench.py (chunked == resident digest and
GPU == oracle on the CPU sample); the C++ oracle cannot finish these sizes in seconds, hence numpy here."""
import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from bioinformatics_b200.batches import PairBatchHost

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


def _dedup_problem(n, n_probes, uid_len, read_len, seed):
    """Random reads whose first uid_len bases of read one come from ~n/4 molecules (PCR families), Phred 2..41."""
    rng = np.random.default_rng(seed)
    n_mol = max(1, n // 4)
    mol_probe = rng.integers(0, n_probes, n_mol).astype(np.int32)
    mol_uid = rng.integers(0, 4, (n_mol, uid_len), dtype=np.uint8)
    mol = rng.integers(0, n_mol, n)
    r1 = ACGT[rng.integers(0, 4, (n, read_len), dtype=np.uint8)]
    r1[:, :uid_len] = ACGT[mol_uid[mol]]
    r2 = ACGT[rng.integers(0, 4, (n, read_len), dtype=np.uint8)]
    q1 = rng.integers(35, 75, (n, read_len), dtype=np.uint8)
    q2 = rng.integers(35, 75, (n, read_len), dtype=np.uint8)
    q1[rng.random(n) < 0.5] = ord("I")  # many equal quality sums: the read-index tie-break decides
    q2[:] = ord("I")
    read_index = rng.permutation(n).astype(np.int32)
    pairs = PairBatchHost(mol_probe[mol], A.SeqBatch.from_fixed(r1), A.SeqBatch.from_fixed(r2), A.SeqBatch.from_fixed(q1),
                          A.SeqBatch.from_fixed(q2), read_index)
    return pairs, r1, q1, q2


def _numpy_dedup(pairs, r1, q1, q2, uid_len, n_probes):
    """FilterByUid.reduceReadsByProbeAndUid + findBestData, vectorised: group by (probe, UID), best = highest
    (short)((short)sum(q1-33) + (short)sum(q2-33)), ties to the smaller read index."""
    n = pairs.n
    code = np.zeros(256, dtype=np.uint64)
    code[ACGT] = np.arange(4, dtype=np.uint64)
    uid = np.zeros(n, dtype=np.uint64)
    for k in range(uid_len):
        uid = (uid << np.uint64(2)) | code[r1[:, k]]
    key = (pairs.probe_id.astype(np.uint64) << np.uint64(2 * uid_len)) | uid
    s1 = (q1.sum(axis=1, dtype=np.int64) - 33 * q1.shape[1]).astype(np.int16)
    s2 = (q2.sum(axis=1, dtype=np.int64) - 33 * q2.shape[1]).astype(np.int16)
    score = (s1.astype(np.int32) + s2.astype(np.int32)).astype(np.int16)
    order = np.lexsort((pairs.read_index, -score.astype(np.int32), key))
    ks = key[order]
    head = np.ones(n, dtype=bool)
    head[1:] = ks[1:] != ks[:-1]
    is_best = np.zeros(n, dtype=np.uint8)
    is_best[order[head]] = 1
    starts = np.nonzero(head)[0]
    sizes = np.diff(np.append(starts, n))
    group_size = np.zeros(n, dtype=np.int32)
    group_size[order] = np.repeat(sizes, sizes)
    probe_of_group = pairs.probe_id[order[starts]]
    unique = np.bincount(probe_of_group, minlength=n_probes)
    maxsz = np.zeros(n_probes, dtype=np.int64)
    np.maximum.at(maxsz, probe_of_group, sizes)
    total = np.bincount(pairs.probe_id, minlength=n_probes)
    return is_best, group_size, unique, total - unique, maxsz


def test_numpy_statement_matches_oracle():
    """The vectorised statement used at full size agrees with the C++ oracle where the oracle is fast enough."""
    from oracle import oracle as O
    pairs, r1, q1, q2 = _dedup_problem(30000, 50, 6, 40, seed=5)
    want = O.dedup_batch(pairs, 6, 0, 50)
    best, gs, unique, dup, mx = _numpy_dedup(pairs, r1, q1, q2, 6, 50)
    assert (best == want[0]).all() and (gs == want[1]).all()
    assert (unique == want[2].unique_pairs).all() and (dup == want[2].duplicate_pairs).all() and (mx == want[2].max_pairs_per_uid).all()


@pytest.mark.gpu
@pytest.mark.parametrize("n,n_probes", [(10_000_000, 10_000), (2_000_000, 200_000)], ids=["config3_10M_pairs_10k_probes", "config4_200k_probes"])
def test_dedup_full_size(n, n_probes):
    from bioinformatics_b200._lib import Context
    uid_len, read_len = 14, 150
    pairs, r1, q1, q2 = _dedup_problem(n, n_probes, uid_len, read_len, seed=n_probes)
    want_best, want_gs, want_unique, want_dup, want_max = _numpy_dedup(pairs, r1, q1, q2, uid_len, n_probes)
    with Context(0) as ctx:
        best, gs, st = ctx.dedup_batch(pairs, uid_len, 0, n_probes)
        assert (best == want_best).all() and (gs == want_gs).all()
        assert (st.unique_pairs == want_unique).all() and (st.duplicate_pairs == want_dup).all() and (st.max_pairs_per_uid == want_max).all()
        # properties that hold at any size
        assert int(best.sum()) == int(st.unique_pairs.sum()) and int(st.unique_pairs.sum() + st.duplicate_pairs.sum()) == pairs.n
        assert (gs[best == 1] >= 1).all() and int((1.0 / gs).sum().round()) == int(best.sum())  # every group counted once
        # idempotence: the survivors are all unique
        keep = np.nonzero(best)[0][:2_000_000]
        best2, gs2, _ = ctx.dedup_batch(pairs.subset(keep), uid_len, 0, n_probes)
        assert best2.all() and (gs2 == 1).all()


@pytest.mark.gpu
def test_trim_full_size_config1():
    """configs[0]: 100 k read pairs 2 x 150 bp with the offsets HSQUtils trim derives (UID + primer at 5', 3' clip)."""
    from bioinformatics_b200._lib import Context
    rng = np.random.default_rng(1)
    n, L = 100_000, 150
    first, last = 10 + 22, L - 23
    with Context(0) as ctx:
        for mate in range(2):
            reads = ACGT[rng.integers(0, 4, (n, L), dtype=np.uint8)]
            quals = rng.integers(35, 75, (n, L), dtype=np.uint8)
            seq, qual = ctx.trim_batch(A.SeqBatch.from_fixed(reads), A.SeqBatch.from_fixed(quals), first, last, True)
            keep = last - first + 1
            assert (seq.lengths() == keep).all() and (qual.lengths() == keep).all()
            assert (seq.data.reshape(n, keep) == reads[:, first:last + 1]).all()
            assert (qual.data.reshape(n, keep) == quals[:, first:last + 1]).all()
            # trimming the result with a window that covers it is the identity
            seq2, _ = ctx.trim_batch(seq, qual, 0, keep - 1, True)
            assert (seq2.data == seq.data).all()


@pytest.mark.gpu
def test_ingest_full_size():
    """f2 at a size where every scan spans many tiles and the string sorts run over > 2^19 records: 400 k read pairs, both FASTQ texts indexed,
    ~730 k SAM lines indexed and joined, all equal to the CPU oracle; then size-independent properties of the join."""
    from bioinformatics_b200 import synth
    from bioinformatics_b200._lib import Context
    from bioinformatics_b200.batches import NameJoinHost
    from oracle import oracle as O
    case = synth.make_ingest_case(400000, read_len=100, seed=41, trim=(10, 89, 0, 89))
    sam_text = synth.sam_text_of(case, seed=42, xa_rate=0.1)
    with Context(0) as ctx:
        for text in (case["fq1_text"], case["fq2_text"]):
            g, e = ctx.index_fastq(text), O.index_fastq(text)
            assert g.n == e.n == 400000 and all((g.head()[f] == e.head()[f]).all() for f in g.FIELDS)
        gs, es = ctx.index_sam(sam_text), O.index_sam(sam_text)
        assert gs.n == es.n == len(case["sam_names"]) and all((getattr(gs, c) == getattr(es, c)).all() for c in gs.columns())
        join = NameJoinHost(case["sam_names"], gs.flag, gs.seq_len, case["fq1_headers"], case["fq2_headers"], case["fq_seq_len"], case["fq_seq_len"],
                            sam_qual_len=gs.qual_len, r1_first=10, r1_stop=89, r2_first=0, r2_stop=89)
        g, e = ctx.join_names(join), O.join_names(join)
        assert g.n_emitted == e.n_emitted and g.total_matching_pairs == e.total_matching_pairs
        assert all((a == b).all() for a, b in zip(g.emitted(), e.emitted())) and (g.status == e.status).all()
        rec, fq, _ = g.emitted()
        # every returned record carries the FASTQ ordinal of its own name; the iterator's order is the order of the names
        ids = np.array(case["ids"])
        names = np.array(case["sam_names"])
        assert (names[rec] == ids[fq]).all()
        assert (names[rec][1:] >= names[rec][:-1]).all()
        # idempotence: joining only the returned records returns all of them, in the same order
        again = NameJoinHost(list(names[rec]), gs.flag[rec], gs.seq_len[rec], case["fq1_headers"], case["fq2_headers"], check_trim=False)
        g2 = ctx.join_names(again)
        assert g2.n_emitted == len(rec) and (g2.emitted()[0] == np.arange(len(rec))).all() and (g2.emitted()[1] == fq).all()


@pytest.mark.gpu
def test_band_equals_full_matrices_at_scale():
    """configs[4] at bench scale (2 M read pairs, 2 x 150 bp): with the opt-in band every column of hsq_extend_out and every
    CIGAR / MD string equals the full-matrix run's, and the two counters add up (banded = read ends that reached DP#2)."""
    from bioinformatics_b200 import synth
    from bioinformatics_b200._lib import Context
    from bioinformatics_b200.batches import ExtendResult
    n = 2_000_000
    probes = synth.make_probes(1000)
    pairs = synth.make_pairs(probes, n, read_len=150, ext_uid_len=10, with_quals=False)
    with Context(0) as ctx:
        ctx.set_probes(probes)
        full = None
        for band in (0, 12, 40):
            ctx.configure(band=band)
            out = ExtendResult(n, 32, 64)
            ctx.extend_stage(pairs)
            ctx.extend_run(32, 64)
            ctx.extend_fetch(out)
            banded, redone = ctx.band_stats()
            if band == 0:
                full = out
                assert (banded, redone) == (0, 0)
                continue
            for f in ExtendResult.FIELDS_I32 + ("ok", "pair_ok", "overflow"):
                assert np.array_equal(getattr(out, f), getattr(full, f)), f
            assert out.cigars() == full.cigars() and out.mds() == full.mds() and out.cells == full.cells
            reached = int(((full.primer_end >= 0) & (full.primer_end != A.HSQ_PRIMER_NOT_FOUND)).sum())
            assert 0.95 * reached <= banded <= 2 * n and 0 <= redone < 0.2 * banded
