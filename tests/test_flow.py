"""bioinformatics_b200.flow: mapped records -> probe assignment -> tie-break -> N filter -> dedup + extension -> record fields,
i.e. the data path of PrimerReadExtensionAndPcrDuplicateIdentification.java:302-402 (assignment), :704-765 (tie-break),
:805-830 (pair table, N filter), :1172-1216 (probe task) and ExtendReadsToPrimer.java:131-181 (records), chained over the C ABI.
The same flow function runs over an adapter of the CPU oracle; the GPU run must reproduce every array of it."""
import types

import numpy as np
import pytest

from bioinformatics_b200 import flow, synth
from bioinformatics_b200.batches import scorer
from oracle import oracle as O
from tests import helpers as H

U, V = 8, 0


class OracleBackend:
    """The calls flow.py makes, answered by the CPU oracle."""

    def __init__(self, probes, ranges):
        self.probes, self.ranges = probes, ranges

    def assign_probes(self, records, max_probes=4):
        return O.assign_probes(self.ranges, records, max_probes)

    def resolve_multi_probe(self, r1, r2, cand_offsets, cand_probe, u, v):
        return O.resolve_multi_probe(self.probes, r1, r2, cand_offsets, cand_probe, u, v)

    def flag_n_pairs(self, pairs):
        # readString.toLowerCase().contains("n") on either mate (…Identification.java:815-821)
        return np.array([1 if "n" in (pairs.r1.string(k) + pairs.r2.string(k)).lower() else 0 for k in range(pairs.n)], dtype=np.uint8)

    def dedup_extend_batch(self, pairs, u, v, extend_duplicates=False, cigar_cap=32, md_cap=48):
        best, gs, stats = O.dedup_batch(pairs, u, v, self.probes.n)
        idx = np.arange(pairs.n) if extend_duplicates else np.nonzero(best)[0]
        ext = O.extend_batch(self.probes, pairs.subset(idx), cigar_cap=cigar_cap, md_cap=md_cap, nthreads=8)
        unable_u = np.zeros(self.probes.n, dtype=np.int32)
        unable_d = np.zeros(self.probes.n, dtype=np.int32)
        failed = idx[ext.pair_ok == 0]
        np.add.at(unable_u, pairs.probe_id[failed[best[failed] == 1]], 1)
        np.add.at(unable_d, pairs.probe_id[failed[best[failed] == 0]], 1)
        lst = np.zeros(pairs.n, dtype=np.int32)
        lst[:len(idx)] = idx
        return types.SimpleNamespace(is_best=best, group_size=gs, stats=stats, unique_unable_to_extend=unable_u,
                                     duplicate_unable_to_extend=unable_d, n_extended=len(idx), extended_pair=lst, extend=ext)

    def extend_records(self, pairs, ext, duplicate=None):
        return O.extend_records(self.probes, pairs, duplicate=duplicate)


def _problem(n_probes=60, n_pairs=3000, seed=3):
    probes = synth.make_probes(n_probes, seed=seed)
    fastq = synth.make_pairs(probes, n_pairs, read_len=150, ext_uid_len=U, lig_uid_len=V, seed=seed + 1, sub_rate=0.01, n_rate=0.01)
    ranges, records = synth.make_mapped_records(probes, fastq, seed=seed + 2)
    return probes, fastq, ranges, records


def test_flow_on_oracle_backend_recovers_the_true_probes():
    probes, fastq, ranges, records = _problem(n_probes=40, n_pairs=1200)
    be = OracleBackend(probes, ranges)
    pairs, info = flow.pair_table(be, records, fastq, U, V)
    # every pair handed to the probe tasks carries the probe its reads were made from (the tie-break picks the true probe among
    # the two candidates of a pair with an alternative hit), and none of them holds an N
    assert (pairs.probe_id == fastq.probe_id[pairs.read_index]).all()
    assert info["multi_probe_pairs"] > 0 and info["with_n"] > 0 and info["unresolved"] == 0
    assert not be.flag_n_pairs(pairs).any()
    a = info["assignment"]
    assert (a.mate == -1).sum() > 0 and (a.mate == -2).sum() == 12 and (a.mate == -3).sum() > 0
    assert len(np.unique(pairs.read_index)) == pairs.n
    out = flow.dedup_flow(be, records, fastq, U, V, cigar_cap=64, md_cap=96)
    m = out["dedup"].n_extended
    assert m == out["dedup"].is_best.sum() and out["records"].n == m
    out2 = flow.dedup_flow(be, records, fastq, U, V, mark_duplicates=True, cigar_cap=64, md_cap=96)
    assert out2["records"].n == pairs.n
    dup_flag = (out2["records"].flag[0::2] & 0x400) != 0
    assert (dup_flag == (out2["dedup"].is_best == 0)).all()


def test_flow_keeps_the_range_map_landing_rule():
    """With several windows per (contig, strand) a record that starts exactly on its window's first base can miss that window
    (RangeMap.java:108-131: the bisection may land on the entry before it); the pair then keeps only the probes both mates
    reported.  The flow does not repair this: the pair is handed to the alternative-hit probe or to none, as in the reference."""
    probes, fastq, ranges, records = _problem(n_probes=200, n_pairs=4000, seed=11)
    pairs, info = flow.pair_table(OracleBackend(probes, ranges), records, fastq, U, V)
    wrong = np.nonzero(pairs.probe_id != fastq.probe_id[pairs.read_index])[0]
    assert 0 < len(wrong) < pairs.n // 50 and info["without_probe"] > 0
    for j in wrong:
        rows = np.nonzero(records.read_index == pairs.read_index[j])[0]
        true_probe = fastq.probe_id[pairs.read_index[j]]
        assert (records.start[rows] == ranges.start[true_probe]).any()


@pytest.mark.gpu
@pytest.mark.parametrize("mark_duplicates", [False, True])
def test_flow_gpu_matches_oracle_backend(mark_duplicates):
    from bioinformatics_b200._lib import Context
    probes, fastq, ranges, records = _problem(n_probes=200, n_pairs=20000, seed=11)
    # caps: a pair that lost its true probe to the RangeMap landing rule (see the CPU test below) aligns against a foreign target
    want = flow.dedup_flow(OracleBackend(probes, ranges), records, fastq, U, V, mark_duplicates, cigar_cap=640, md_cap=960)
    with Context(0) as ctx:
        ctx.set_probes(probes)
        ctx.set_scorer(scorer())
        ctx.set_probe_ranges(ranges)
        got = flow.dedup_flow(ctx, records, fastq, U, V, mark_duplicates, cigar_cap=640, md_cap=960)
    for k in ("pairs_seen", "multi_probe_pairs", "unresolved", "without_probe", "with_n"):
        assert got["info"][k] == want["info"][k], k
    ga, wa = got["info"]["assignment"], want["info"]["assignment"]
    for f in ("n_probes", "mate", "n_common", "overflow"):
        assert (getattr(ga, f) == getattr(wa, f)).all(), f
    assert ga.probe_sets() == wa.probe_sets() and ga.common_sets() == wa.common_sets()  # slots past the counts are not defined
    gp, wp = got["pairs"], want["pairs"]
    assert (gp.probe_id == wp.probe_id).all() and (gp.read_index == wp.read_index).all()
    assert (gp.r1.data == wp.r1.data).all() and (gp.r2.data == wp.r2.data).all()
    gd, wd = got["dedup"], want["dedup"]
    assert (gd.is_best == wd.is_best).all() and (gd.group_size == wd.group_size).all()
    for f in ("unique_pairs", "duplicate_pairs", "max_pairs_per_uid"):
        assert (getattr(gd.stats, f) == getattr(wd.stats, f)).all(), f
    assert (gd.unique_unable_to_extend == wd.unique_unable_to_extend).all()
    assert (gd.duplicate_unable_to_extend == wd.duplicate_unable_to_extend).all()
    assert gd.n_extended == wd.n_extended and (gd.extended_pair[:gd.n_extended] == wd.extended_pair[:wd.n_extended]).all()
    H.assert_extend_equal(got["extend"], want["extend"])
    gr, wr = got["records"], want["records"]
    assert (gr.flag == wr.flag).all() and (gr.mate_pos == wr.mate_pos).all() and (gr.offsets == wr.offsets).all()
    total = int(wr.offsets[-1])
    assert (gr.seq[:total] == wr.seq[:total]).all() and (gr.qual[:total] == wr.qual[:total]).all()
    assert want["info"]["multi_probe_pairs"] > 100 and gd.n_extended > 1000


@pytest.mark.gpu
def test_flag_n_pairs_gpu():
    from bioinformatics_b200._abi import SeqBatch
    from bioinformatics_b200._lib import Context
    from bioinformatics_b200.batches import PairBatchHost
    rng = np.random.default_rng(4)
    r1, r2 = [], []
    for k in range(3000):
        a = H.random_seq(rng, int(rng.integers(0, 200)), "ACGT")
        b = H.random_seq(rng, int(rng.integers(0, 200)), "ACGT")
        kind = k % 6
        if kind == 1 and a:
            j = int(rng.integers(0, len(a)))
            a = a[:j] + "N" + a[j + 1:]
        if kind == 2 and b:
            b = b[:-1] + "n"  # last byte of the read, lower case
        if kind == 3 and a:
            a = "N" + a[1:]
        r1.append(a)
        r2.append(b)
    pairs = PairBatchHost(np.zeros(3000, dtype=np.int32), SeqBatch.from_strings(r1), SeqBatch.from_strings(r2))
    want = np.array([1 if "n" in (x + y).lower() else 0 for x, y in zip(r1, r2)], dtype=np.uint8)
    with Context(0) as ctx:
        assert (ctx.flag_n_pairs(pairs) == want).all()
        assert len(ctx.flag_n_pairs(pairs.head(0))) == 0
    assert want.sum() > 900
