"""Hardening of the oracle where the reference pins nothing (A3-A7, A9): tests/pyref.py is a second statement of
ExtendReadsToPrimer / FilterByUid / the multi-probe tie-break / the alignment getters, written in plain Python from the Java
sources.  Here the C++ oracle (what the CUDA path is tested against) must agree with it on seeded cases, on the reference's orphan
fixtures (real reads, probes with N; tests/golden/reference_fixtures/) and on the quadruples of ExtendReadsToPrimerTest.java:23-28.
The GPU tests at the end run the same fixtures through the CUDA path."""
import os

import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from bioinformatics_b200 import synth
from bioinformatics_b200.batches import PairBatchHost, ProbeTableHost, scorer
from oracle import oracle as O
from tests import helpers as H
from tests import pyref as P

FIX = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_fixtures")
EXT_FIELDS = ("ok", "primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen")


def probe_dicts(probes):
    ext, lig, tgt = probes.ext.strings(), probes.lig.strings(), probes.target.strings()
    return [dict(reverse=bool(probes.reverse_strand[k]), ext=ext[k], lig=lig[k], target=tgt[k], ext_start=int(probes.ext_start[k]),
                 ext_stop=int(probes.ext_stop[k]), lig_start=int(probes.lig_start[k]), lig_stop=int(probes.lig_stop[k])) for k in range(probes.n)]


def assert_oracle_equals_pyref_extend(probes, pairs, sc=None, cigar_cap=700, md_cap=1000):
    exp = O.extend_batch(probes, pairs, sc, cigar_cap=cigar_cap, md_cap=md_cap, nthreads=8)
    pd = probe_dicts(probes)
    psc = None if sc is None else P.Scorer(sc.match, sc.mismatch, sc.gap_extend, sc.gap_start)
    r1, r2 = pairs.r1.strings(), pairs.r2.strings()
    cig, md = exp.cigars(), exp.mds()
    for k in range(pairs.n):
        mates, pair_ok = P.extend_read_pair(pd[int(pairs.probe_id[k])], r1[k], r2[k], psc)
        assert bool(exp.pair_ok[k]) == pair_ok, k
        for mate in range(2):
            m, e = mates[mate], 2 * k + mate
            got = dict(ok=bool(exp.ok[e]), primer_end=None if exp.primer_end[e] == A.HSQ_PRIMER_NOT_FOUND else int(exp.primer_end[e]),
                       score=int(exp.score[e]), pos=int(exp.pos[e]), read_start=int(exp.read_start[e]), read_stop=int(exp.read_stop[e]),
                       nm=int(exp.nm[e]), tlen=int(exp.tlen[e]), cigar=cig[e], md=md[e])
            want = {f: m[f] for f in got}
            # a mate that does not extend has no record in the reference: only the stage it reached is compared for it
            if m["ok"]:
                assert got == want, (k, mate, got, want)
            else:
                assert not got["ok"] and got["primer_end"] == want["primer_end"] and got["score"] == want["score"], (k, mate, got, want)
    return exp


def test_extend_oracle_equals_python_statement_short_reads():
    """10 000 read ends: short amplicons (fast in pure Python), both strands, substitutions and indels, short mates, N / IUPAC / lower case."""
    rng = np.random.default_rng(11)
    probes = synth.make_probes(40, seed=71, target_len=(24, 44), primer_len=(8, 13))
    base = synth.make_pairs(probes, 5000, read_len=52, ext_uid_len=4, lig_uid_len=2, seed=72, sub_rate=0.04, ins_rate=0.01, del_rate=0.01, n_rate=0.01)
    r1, r2 = base.r1.strings(), base.r2.strings()
    for k in range(0, 5000, 9):
        r1[k] = r1[k][:int(rng.integers(0, 52))]
    for k in range(3, 5000, 11):
        r2[k] = r2[k][:int(rng.integers(1, 52))].lower()
    for k in range(5, 5000, 13):
        s = list(r2[k])
        s[int(rng.integers(0, len(s)))] = str(rng.choice(list("RYKMSWN")))
        r2[k] = "".join(s)
    pairs = PairBatchHost(base.probe_id, r1, r2)
    exp = assert_oracle_equals_pyref_extend(probes, pairs)
    assert 0.3 < exp.pair_ok.mean() < 0.999  # both outcomes are exercised


def test_extend_oracle_equals_python_statement_full_size_and_scorers():
    """2 x 150 bp reads on HEAT-Seq sized probes, default and non-default scorers."""
    probes = synth.make_probes(12, seed=73)
    pairs = synth.make_pairs(probes, 60, read_len=150, ext_uid_len=10, seed=74, sub_rate=0.03, ins_rate=0.006, del_rate=0.006)
    assert_oracle_equals_pyref_extend(probes, pairs)
    assert_oracle_equals_pyref_extend(probes, pairs.head(20), scorer(2, -3, -2, -4))


def quadruples():
    read_one = ["GGAGGTGGAGGCGGCAGTGGCGGTGGTGGTGAGGGAGGGGGTGGCCCCTGAGCGTCATCTGCCCCC", "CGTCCCTCGCAAGTCAGGGGATCCGAGGTACCTGCAGCAGCAGCAGCAGCGCCGAGAGGCTGCGGC",
                "TCCATTTCTTAGAGGGAATGGTATATATACTTTATATAATATATAGTAATAATAGAATAAAAATAA"]
    read_two = ["TCGCTGCCCAGCACCGCCGTCTGGTTGGCCGGCAGCCCCGCCTGCAGGATGGGCCGGTGCGGGGAGCGCTCTGTGG",
                "CCAGCCCGGCCCGACCCGACCGCACCCGGCGCCTGCCCTCGCTCGGCGTCCCCGGCCAGCCATGGGCCCTTGGAGC", "CAAGATGATGGGAGGCCTAGGATTTCATGAAGTCCTCAATAATGTAAGTAAACCTGAAAATCAAACCACAATAATT"]
    primer_one = ["GTGGGGGCGGCAGTGGCGG", "ATGCGTCCCTCGCAAGTCAGGG", "CTTAGAGGGAATGGTATA"]
    primer_two = ["GCGGTGCTGGGCAGCGACGTG", "GGTCGGGTCGGGCCGGGC", "CTAGGCCTCCCATCATCTTGGTCCCCA"]
    return read_one, read_two, primer_one, primer_two


def test_primer_quadruples_of_the_reference_test():
    """The inputs of ExtendReadsToPrimerTest.java:23-28 (its assertions are commented out, :45-56): oracle == Python statement."""
    read_one, read_two, primer_one, primer_two = quadruples()
    for k in range(3):
        for primer, read in ((primer_one[k], read_one[k]), (primer_two[k], P.complement(P.reverse(read_two[k])))):
            want = P.primer_end_index_in_read(primer, read)
            got, _ = O.primer_end(primer, read)
            assert got == want, (k, primer, got, want)


def fixture_pairs(n=200):
    def reads(path):
        lines = open(os.path.join(FIX, path)).read().split("\n")
        return [lines[i + 1] for i in range(0, 4 * n, 4)], [lines[i + 3] for i in range(0, 4 * n, 4)]
    s1, q1 = reads("reverse_test_one.head200.fastq")
    s2, q2 = reads("reverse_test_two.head200.fastq")
    return s1, q1, s2, q2


def fixture_probes(path, limit=None):
    rows = [ln.split("\t") for ln in open(os.path.join(FIX, path)).read().split("\n") if ln.strip()]
    col = {name: k for k, name in enumerate(rows[0])}
    rows = rows[1:limit + 1 if limit else None]
    g = lambda name: [r[col[name]] for r in rows]
    gi = lambda name: [int(x) for x in g(name)]
    return ProbeTableHost([1 if s == "-" else 0 for s in g("probe_strand")], gi("ext_start"), gi("ext_stop"), gi("lig_start"), gi("lig_stop"),
                          gi("target_start"), gi("target_stop"), g("ext_sequence"), g("lig_sequence"), g("target_sequence"))


def orphan_cases():
    """(probe table, pair batch) built from the reference's orphan fixtures."""
    s1, q1, s2, q2 = fixture_pairs()
    out = []
    for path in ("forward_test_probe.txt", "reverse_test_probe.txt"):
        probes = fixture_probes(path)
        out.append((path, probes, PairBatchHost(np.zeros(len(s1), dtype=np.int32), s1, s2, q1, q2, np.arange(len(s1), dtype=np.int32))))
    probes = fixture_probes("probes.head40.txt")  # primers and targets with N: reads synthesised from them
    pairs = synth.make_pairs(probes, 400, read_len=150, ext_uid_len=8, seed=75, sub_rate=0.01, ins_rate=0.002, del_rate=0.002)
    out.append(("probes.head40.txt", probes, pairs))
    return out


def test_orphan_fixtures_oracle_equals_python_statement():
    """reverse_test_{one,two}.fastq against the + and the - probe of their amplicon, and probes.txt's N-laden probes."""
    for name, probes, pairs in orphan_cases():
        sub = pairs.head(40 if name != "probes.head40.txt" else 30)
        exp = assert_oracle_equals_pyref_extend(probes, sub)
        if name == "forward_test_probe.txt":
            assert exp.pair_ok.mean() > 0.5, "the real reads extend on their own probe"


def test_dedup_oracle_equals_python_statement():
    """FilterByUid per probe (a dict of UID strings per probe, findBestData's loop) == the oracle's batch statement, incl. the Java short
    wrap of the quality sum and repeated read indexes."""
    rng = np.random.default_rng(12)
    probes = synth.make_probes(30, seed=76)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 12000, read_len=60, ext_uid_len=5, lig_uid_len=2, seed=77))
    pairs.q1.data[::4] = ord("#")
    eb, egs, est = O.dedup_batch(pairs, 5, 2, probes.n)
    r1, r2, q1, q2 = pairs.r1.strings(), pairs.r2.strings(), pairs.q1.strings(), pairs.q2.strings()
    for p in range(probes.n):
        idx = np.nonzero(pairs.probe_id == p)[0]
        best, size, uniq, dup, mx = P.reduce_by_uid([(int(pairs.read_index[k]), r1[k], r2[k], q1[k], q2[k]) for k in idx], 5, 2)
        assert np.array_equal(eb[idx], best) and np.array_equal(egs[idx], size), p
        assert (est.unique_pairs[p], est.duplicate_pairs[p], est.max_pairs_per_uid[p]) == (uniq, dup, mx)
    # quality sums beyond the short range: 2 x 450 bases of '~' (93 each) wrap
    long_q = PairBatchHost([0, 0, 0], ["ACGTA" + "A" * 445] * 3, ["CC" + "A" * 448] * 3, ["~" * 450, "I" * 450, "5" * 450], ["~" * 450, "I" * 450, "5" * 450],
                           [5, 6, 7])
    eb, _, _ = O.dedup_batch(long_q, 5, 2, 1)
    best, _, _, _, _ = P.reduce_by_uid([(5 + k, long_q.r1.string(k), long_q.r2.string(k), long_q.q1.string(k), long_q.q2.string(k)) for k in range(3)], 5, 2)
    assert eb.tolist() == best and P.java_short(2 * 93 * 450) != 2 * 93 * 450


def test_tie_break_oracle_equals_python_statement():
    """The multi-probe tie-break (…Identification.java:704-765): matches, edit distance and the verdict per read."""
    rng = np.random.default_rng(13)
    probes = synth.make_probes(24, seed=78, target_len=(30, 50), primer_len=(9, 13))
    pairs = synth.make_pairs(probes, 60, read_len=70, ext_uid_len=4, lig_uid_len=3, seed=79, sub_rate=0.03, ins_rate=0.01, del_rate=0.01)
    cand_offsets, cand_probe = [0], []
    for k in range(pairs.n):
        c = {int(pairs.probe_id[k])} | {int(x) for x in rng.integers(0, probes.n, int(rng.integers(1, 3)))}
        if k % 10 == 0:  # an exact tie: the same probe twice under two rows would need equal sequences; use the probe itself twice
            c = [int(pairs.probe_id[k])] * 2
        cand_probe += list(c)
        cand_offsets.append(len(cand_probe))
    best, m, e = O.resolve_multi_probe(probes, pairs.r1, pairs.r2, cand_offsets, cand_probe, 4, 3)
    ext, lig, tgt = probes.ext.strings(), probes.lig.strings(), probes.target.strings()
    r1, r2 = pairs.r1.strings(), pairs.r2.strings()
    for k in range(pairs.n):
        cs = cand_probe[cand_offsets[k]:cand_offsets[k + 1]]
        bk, ms, es = P.tie_break(r1[k], r2[k], [(ext[c], tgt[c], lig[c]) for c in cs], 4, 3)
        assert ms == m[cand_offsets[k]:cand_offsets[k + 1]].tolist() and es == e[cand_offsets[k]:cand_offsets[k + 1]].tolist(), k
        assert int(best[k]) == (cs[bk] if bk >= 0 else -1), k


def test_alignment_getters_oracle_equals_python_statement():
    """The operator seam (NeedlemanWunschGlobalAlignment + AlignmentPair getters) on random, tie-heavy, IUPAC and empty inputs, under the
    four terminal-gap flag combinations."""
    rng = np.random.default_rng(14)
    refs, qrys = H.align_cases(rng, 400, max_len=48)
    for flags in range(4):
        got = O.align_batch(refs, qrys, flags=flags, cols_cap=200, md_cap=300)
        ra, qa, ops, mds, rmds = got.ref_alns(), got.qry_alns(), got.op_strings(), got.mds(), got.rmds()
        sc = P.Scorer(1, -4, -1, -6, bool(flags & 1), bool(flags & 2))
        for k in range(0, 400, 2 if flags else 1):
            score, a1, a2 = P.needleman_wunsch(P.codes(refs[k]), P.codes(qrys[k]), sc)
            ap = P.AlignmentPair(a1, a2)
            assert (score, a1, a2) == (int(got.score[k]), ra[k], qa[k]), (k, flags)
            assert ap.expanded_cigar() == ops[k] and ap.md() == mds[k] and ap.reverse_md() == rmds[k], (k, flags)
            assert (ap.first_q, ap.last_q, ap.first_r, ap.last_r) == (got.first_q[k], got.last_q[k], got.first_r[k], got.last_r[k]), (k, flags)
            assert (ap.n_matches(), ap.n_mismatches(), ap.n_non_terminal_gaps(), ap.edit_distance()) == \
                (got.matches[k], got.mismatches[k], got.non_terminal_gaps[k], got.edit_distance[k]), (k, flags)


# ---- the same fixtures through the CUDA path -----------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_gpu_orphan_fixtures_equal_oracle():
    from bioinformatics_b200._lib import Context
    with Context(0) as ctx:
        for name, probes, pairs in orphan_cases():
            ctx.set_probes(probes)
            got = ctx.extend_batch(pairs, cigar_cap=256, md_cap=400)
            H.assert_extend_equal(got, O.extend_batch(probes, pairs, cigar_cap=256, md_cap=400, nthreads=8))
            clean, _ = synth.drop_n_pairs(pairs) if name == "probes.head40.txt" else (pairs, None)
            keep = np.array(["N" not in (a + b).upper() for a, b in zip(clean.r1.strings(), clean.r2.strings())])
            clean = clean.subset(np.nonzero(keep)[0])
            b, g, st = ctx.dedup_batch(clean, 8, 0, probes.n)
            eb, eg, est = O.dedup_batch(clean, 8, 0, probes.n)
            assert np.array_equal(b, eb) and np.array_equal(g, eg) and np.array_equal(st.unique_pairs, est.unique_pairs)


@pytest.mark.gpu
def test_gpu_primer_quadruples():
    """getPrimerEndIndexInRead of the reference test's quadruples through hsq_extend_batch: a + probe with ext = primerOne and
    lig = reverse(primerTwo), read two = reverse(readTwo) makes mate 2 align primerTwo to revcomp(readTwo) (ExtendReadsToPrimer.java:159-161)."""
    from bioinformatics_b200._lib import Context
    read_one, read_two, primer_one, primer_two = quadruples()
    probes = ProbeTableHost([0, 0, 0], [100] * 3, [120] * 3, [300] * 3, [320] * 3, [121] * 3, [299] * 3, primer_one, [p[::-1] for p in primer_two],
                            ["ACGTACGTACGTACGTACGTAAGGCCTTAAGG"] * 3)
    pairs = PairBatchHost([0, 1, 2], read_one, [r[::-1] for r in read_two])
    with Context(0) as ctx:
        ctx.set_probes(probes)
        got = ctx.extend_batch(pairs, cigar_cap=128, md_cap=200)
    for k in range(3):
        for mate, (primer, read) in enumerate(((primer_one[k], read_one[k]), (primer_two[k], P.complement(P.reverse(read_two[k]))))):
            want = P.primer_end_index_in_read(primer, read)
            g = int(got.primer_end[2 * k + mate])
            assert (None if g == A.HSQ_PRIMER_NOT_FOUND else g) == want, (k, mate)
