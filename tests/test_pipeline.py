"""hsq_dedup_extend_batch: the body of one probe task (reduce by UID, then extend the unique pairs -- and the duplicates when
they are kept or marked) for a batch spanning probes
(nimblegen_heatseq/src/main/java/com/roche/heatseq/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:1172-1216).
Expected values: the oracle's dedup, then the oracle's extend on exactly the pairs the reference would extend."""
import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from bioinformatics_b200 import synth
from bioinformatics_b200.batches import scorer
from oracle import oracle as O
from tests import helpers as H


def _expected(probes, pairs, u, v, extend_duplicates, caps):
    best, gs, stats = O.dedup_batch(pairs, u, v, probes.n)
    idx = np.arange(pairs.n) if extend_duplicates else np.nonzero(best)[0]
    ext = O.extend_batch(probes, pairs.subset(idx), cigar_cap=caps[0], md_cap=caps[1], nthreads=8)
    unable_u = np.zeros(probes.n, dtype=np.int32)
    unable_d = np.zeros(probes.n, dtype=np.int32)
    failed = idx[ext.pair_ok == 0]
    np.add.at(unable_u, pairs.probe_id[failed[best[failed] == 1]], 1)
    np.add.at(unable_d, pairs.probe_id[failed[best[failed] == 0]], 1)
    return best, gs, stats, idx, ext, unable_u, unable_d


def test_ragged_take_matches_python_loop():
    rng = np.random.default_rng(1)
    strings = ["".join(rng.choice(list("ACGT"), int(rng.integers(0, 12)))) for _ in range(50)]
    sb = A.SeqBatch.from_strings(strings)
    idx = rng.integers(0, 50, 80)
    assert sb.take(idx).strings() == [strings[k] for k in idx]
    assert sb.take([]).strings() == []


@pytest.mark.gpu
@pytest.mark.parametrize("extend_duplicates", [False, True])
@pytest.mark.parametrize("n_pairs,read_len,u,v", [(8000, 150, 8, 0), (3000, 100, 6, 2)])
def test_dedup_extend_vs_oracle(extend_duplicates, n_pairs, read_len, u, v):
    from bioinformatics_b200._lib import Context
    probes = synth.make_probes(120, seed=5 + read_len)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, n_pairs, read_len=read_len, ext_uid_len=u, lig_uid_len=v, seed=9 + read_len,
                                                   sub_rate=0.02, ins_rate=0.003, del_rate=0.003))
    # a tenth of the pairs get a random read one: their primer is not found, so "unable to extend" is exercised
    rng = np.random.default_rng(2)
    for k in rng.choice(pairs.n, pairs.n // 10, replace=False):
        a, b = pairs.r1.offsets[k] + u, pairs.r1.offsets[k + 1]
        pairs.r1.data[a:b] = rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), b - a)
    caps = (640, 960)  # the randomised reads give gap-rich alignments
    best, gs, stats, idx, ext, unable_u, unable_d = _expected(probes, pairs, u, v, extend_duplicates, caps)
    with Context(0) as ctx:
        ctx.set_probes(probes)
        ctx.set_scorer(scorer())
        got = ctx.dedup_extend_batch(pairs, u, v, extend_duplicates, cigar_cap=caps[0], md_cap=caps[1])
        assert (got.is_best == best).all() and (got.group_size == gs).all()
        assert (got.stats.unique_pairs == stats.unique_pairs).all() and (got.stats.duplicate_pairs == stats.duplicate_pairs).all()
        assert (got.stats.max_pairs_per_uid == stats.max_pairs_per_uid).all()
        assert got.n_extended == len(idx) and (got.extended_pair[:got.n_extended] == idx).all()
        H.assert_extend_equal(got.extend.head(got.n_extended), ext)
        assert (got.unique_unable_to_extend == unable_u).all() and (got.duplicate_unable_to_extend == unable_d).all()
        assert unable_u.sum() > 0 and (extend_duplicates == (unable_d.sum() > 0))
        # the context is reusable for the plain calls afterwards
        plain = ctx.extend_batch(pairs.subset(idx[:500]), cigar_cap=caps[0], md_cap=caps[1])
        assert (plain.pair_ok == ext.pair_ok[:500]).all()


@pytest.mark.gpu
def test_extend_selected_vs_oracle():
    from bioinformatics_b200._lib import Context, HsqError
    probes = synth.make_probes(80, seed=21)
    pairs = synth.make_pairs(probes, 4000, read_len=150, ext_uid_len=8, seed=22, sub_rate=0.02, ins_rate=0.003, del_rate=0.003)
    rng = np.random.default_rng(23)
    keep = (rng.random(pairs.n) < 0.3).astype(np.uint8)
    idx = np.nonzero(keep)[0]
    want = O.extend_batch(probes, pairs.subset(idx), cigar_cap=64, md_cap=96, nthreads=8)
    with Context(0) as ctx:
        ctx.set_probes(probes)
        ctx.set_scorer(scorer())
        m, lst, got = ctx.extend_selected(pairs, keep, cigar_cap=64, md_cap=96)
        assert m == len(idx) and (lst[:m] == idx).all()
        H.assert_extend_equal(got.head(m), want)
        m, lst, got = ctx.extend_selected(pairs, np.zeros(pairs.n, dtype=np.uint8), cigar_cap=64, md_cap=96)
        assert m == 0
        m, lst, got = ctx.extend_selected(pairs, np.ones(pairs.n, dtype=np.uint8), cigar_cap=64, md_cap=96)
        H.assert_extend_equal(got, O.extend_batch(probes, pairs, cigar_cap=64, md_cap=96, nthreads=8))
        bad = pairs.subset(np.arange(10))
        bad.probe_id[3] = 999
        with pytest.raises(HsqError):
            ctx.extend_selected(bad, np.ones(10, dtype=np.uint8))


@pytest.mark.gpu
def test_dedup_extend_empty_and_all_duplicates():
    from bioinformatics_b200._lib import Context
    probes = synth.make_probes(10, seed=1)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 200, read_len=120, ext_uid_len=6, seed=3))
    with Context(0) as ctx:
        ctx.set_probes(probes)
        ctx.set_scorer(scorer())
        got = ctx.dedup_extend_batch(pairs.head(0), 6, 0)
        assert got.n_extended == 0
        one = pairs.subset(np.zeros(64, dtype=np.int64))  # 64 copies of one pair: one group
        one.read_index[:] = np.arange(64)
        got = ctx.dedup_extend_batch(one, 6, 0)
        assert got.n_extended == 1 and got.is_best.sum() == 1 and got.extended_pair[0] == 0
        assert got.stats.unique_pairs.sum() == 1 and got.stats.duplicate_pairs.sum() == 63
