"""Candidate stage of FastqFindAndTrimTool: trimReadBasedOnQuality
(bioinformatics_common/src/main/java/com/roche/sequencing/bioinformatics/common/fastqtool/FastqFindAndTrimTool.java:1188-1219), the N
test (:920) and SimpleMapper (…/common/mapping/SimpleMapper.java:90-217).  The oracle is pinned on the reference's SimpleMapperTest
(bioinformatics_common/src/test/java/…/mapping/SimpleMapperTest.java:35-63) and cross-checked with a second, dictionary-based
statement of the same rules; the GPU path must equal the oracle array by array."""
import types

import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from bioinformatics_b200 import flow
from oracle import oracle as O
from tests import helpers as H


def test_simple_mapper_reference_kats():
    # mappingOneTest: no candidate; mappingTwoTest: one candidate (default mapper: k = 5, 50 references per k-mer)
    r = O.find_candidates(["AACCGGTT", "ATATATAAT", "CGATCGATT"], ["ATACGA"], ["IIIIII"], max_refs_per_kmer=50)
    assert r.candidate_lists() == [[]]
    r = O.find_candidates(["TGAAGGGAGGATGGGC", "ATATATAAT", "CGATCGATT"], ["TGAAGGGAGGATGGGC"], ["I" * 16], max_refs_per_kmer=50)
    assert r.candidate_lists() == [[(0, 11)]]  # 16 - 5 query positions, the last 5-mer is never looked up (:211)


def test_quality_trim_by_hand():
    # Phred+33: '#' = 2, '5' = 20, 'I' = 40
    reads = ["ACGTACGTAC", "ACGTACGTAC", "ACGTACGTAC", "ACGT", ""]
    quals = ["##5IIIII##", "##########", "IIIIIIIIII", "#II#", ""]
    r = O.find_candidates(["ACGTACGTAC"], reads, quals, leading_quality=20, trailing_quality=30)
    assert list(r.lead_trimmed) == [2, 10, 0, 1, 0]
    assert list(r.trail_trimmed) == [2, 0, 0, 1, 0]  # the trailing walk never crosses the leading trim (:1207)
    r = O.find_candidates(["ACGTACGTAC"], ["ACGTNCGTAC", "NCGTACGTAn"], ["IIIIIIIIII", "#IIIIIIII#"], leading_quality=10, trailing_quality=10)
    assert list(r.has_n) == [1, 0] and list(r.lead_trimmed) == [0, 1] and list(r.trail_trimmed) == [0, 1]  # N trimmed away: kept


def test_kmer_cap_limit_and_order_dependence_by_hand():
    # three sequences share AAAAA...: with a cap of 2 that 5-mer leaves the index and only the private 5-mers count
    seqs = ["AAAAAAAC", "AAAAAAAG", "AAAAAAAT"]
    r = O.find_candidates(seqs, ["AAAAAAACC"], ["I" * 9], max_refs_per_kmer=2)
    assert r.candidate_lists() == [[]]  # AAAAA is excluded; AAAAC sits at the last start of sequence 0 and is never indexed (:93)
    r = O.find_candidates(seqs, ["AAAAAAACC"], ["I" * 9], max_refs_per_kmer=3)
    assert r.candidate_lists() == [[(0, 3), (1, 3), (2, 3)]]
    # 14 identical sequences tie: the limit of 10 admits the whole tie group (:172 needs hits < lastAddedHitSize to stop)
    r = O.find_candidates(["ACGTTGCATG"] * 14, ["ACGTTGCATG"], ["I" * 10], candidate_limit=10, max_candidates=16)
    assert r.n_candidates[0] == 14 and not r.overflow[0]
    r = O.find_candidates(["ACGTTGCATG"] * 14, ["ACGTTGCATG"], ["I" * 10], candidate_limit=10, max_candidates=8)
    assert r.n_candidates[0] == 8 and r.overflow[0]
    # a second, lower group after more than `limit` candidates is cut
    r = O.find_candidates(["ACGTTGCATG"] * 12 + ["ACGTTGCCCC"], ["ACGTTGCATG"], ["I" * 10], candidate_limit=10)
    assert r.n_candidates[0] == 12 and set(r.candidates[0, :12]) == set(range(12))
    # equal tallies, one entry passes its reference ratio and one fails: the list depends on which the HashMap yields first
    short, long_ = "ACGTTGCAG", "ACGTTGCA" + "C" * 30  # 4 indexed 5-mers of each occur in the query
    query = "ACGTTGCA" + "GGGGGGGGGGGGGGGGGGGGGGGG"
    r = O.find_candidates([short, long_], [query], ["I" * len(query)])
    assert r.order_dependent[0] == 1 and r.candidate_lists() == [[(0, 4)]]
    r = O.find_candidates([long_, short], [query], ["I" * len(query)])
    assert r.order_dependent[0] == 1 and r.candidate_lists() == [[]]


def _second_statement(seqs, read, qual, k, cap, lead_q, trail_q, limit, ratio):
    """Plain dictionaries, written from the rules rather than from the oracle's code."""
    lead = 0
    while lead < len(read) and ord(qual[lead]) - 33 < lead_q:
        lead += 1
    trail = 0
    while trail < len(read) - lead and ord(qual[len(read) - 1 - trail]) - 33 < trail_q:
        trail += 1
    t = read[lead:len(read) - trail].upper()
    if "N" in t:
        return lead, trail, 1, []
    holders = {}
    for row, s in enumerate(seqs):
        for i in range(len(s) - k):
            holders.setdefault(s[i:i + k].upper(), set()).add(row)
    holders = {m: rows for m, rows in holders.items() if cap <= 0 or len(rows) <= cap}
    tally = {}
    for i in range(len(t) - k):
        for row in holders.get(t[i:i + k], ()):
            tally[row] = tally.get(row, 0) + 1
    out = []
    for row, hits in sorted(tally.items(), key=lambda e: (-e[1], e[0])):
        if hits < ratio * (len(t) - k) and hits < ratio * (len(seqs[row]) - k):
            break
        if len(out) > limit and hits < out[-1][1]:
            break
        out.append((row, hits))
    return lead, trail, 0, out


def _problem(rng, n_seqs, n_reads, seq_len=(18, 40), read_len=(0, 160), related=0.8):
    seqs = [H.random_seq(rng, int(rng.integers(seq_len[0], seq_len[1] + 1))) for _ in range(n_seqs)]
    for j in range(1, n_seqs, 7):  # near copies: shared k-mers and tied tallies
        s = list(seqs[j - 1])
        s[int(rng.integers(0, len(s)))] = "ACGT"[int(rng.integers(0, 4))]
        seqs[j] = "".join(s)
    reads, quals = [], []
    for k in range(n_reads):
        L = int(rng.integers(read_len[0], read_len[1] + 1))
        body = H.random_seq(rng, L)
        if rng.random() < related and L > 45:
            s = H.mutate(rng, seqs[int(rng.integers(0, n_seqs))], 0.03, 0.01)
            at = int(rng.integers(0, L - len(s))) if L > len(s) else 0
            body = (body[:at] + s + body[at + len(s):])[:L]
        if k % 17 == 3 and L:
            at = int(rng.integers(0, L))
            body = body[:at] + "nN"[k % 2] + body[at + 1:]
        if k % 11 == 5:
            body = body.lower()
        q = np.clip(np.round(rng.normal(32, 8, L)), 2, 41).astype(int)
        cut = int(rng.integers(0, 6))
        q[:cut] = 2
        q[L - int(rng.integers(0, 6)):] = 3
        reads.append(body)
        quals.append("".join(chr(33 + v) for v in q))
    return seqs, reads, quals


@pytest.mark.parametrize("k,cap,limit,ratio", [(5, 15000, 10, 0.5), (5, 3, 2, 0.5), (3, 40, 10, 0.25), (8, 50, 0, 0.7)])
def test_oracle_equals_second_statement(k, cap, limit, ratio):
    rng = np.random.default_rng(100 + k)
    seqs, reads, quals = _problem(rng, 60, 300)
    r = O.find_candidates(seqs, reads, quals, kmer=k, max_refs_per_kmer=cap, leading_quality=10, trailing_quality=15, candidate_limit=limit,
                          min_hit_ratio=ratio, max_candidates=64)
    lists = r.candidate_lists()
    assert not r.overflow.any()
    for j in range(len(reads)):
        lead, trail, has_n, want = _second_statement(seqs, reads[j], quals[j], k, cap, 10, 15, limit, ratio)
        assert (r.lead_trimmed[j], r.trail_trimmed[j], r.has_n[j]) == (lead, trail, has_n), j
        assert lists[j] == want, j
    assert sum(len(x) for x in lists) > 50 and r.has_n.sum() > 5


def test_oracle_rejects_what_the_reference_throws_on():
    with pytest.raises(O.OracleError) as e:
        O.find_candidates(["ACGTNACGT"], ["ACGTACGT"], ["IIIIIIII"])
    assert e.value.status == A.HSQ_ERR_BAD_CHARACTER
    with pytest.raises(O.OracleError) as e:
        O.find_candidates(["ACGT"], ["ACGTACGT"], ["IIIIIIII"])
    assert e.value.status == A.HSQ_ERR_OUT_OF_RANGE
    with pytest.raises(O.OracleError) as e:
        O.find_candidates(["ACGTACGT"], ["ACGTRCGT"], ["IIIIIIII"])  # NucleotideCodeSequence(query) throws on R
    assert e.value.status == A.HSQ_ERR_BAD_CHARACTER


FIND_FIELDS = ("lead_trimmed", "trail_trimmed", "has_n", "n_candidates", "overflow", "order_dependent")


def _assert_find_equal(got, want):
    for f in FIND_FIELDS:
        a, b = getattr(got, f), getattr(want, f)
        bad = np.nonzero(a != b)[0]
        assert len(bad) == 0, "field %s differs at %s: got %s expected %s" % (f, bad[:5], a[bad[:5]], b[bad[:5]])
    assert got.candidate_lists() == want.candidate_lists()


@pytest.mark.gpu
@pytest.mark.parametrize("n_seqs,k,cap,limit,ratio,slots", [(300, 5, 15000, 10, 0.5, 32), (300, 5, 20, 2, 0.5, 4), (40, 3, 30, 10, 0.25, 48),
                                                           (2000, 8, 50, 0, 0.7, 16), (1, 5, 0, 10, 0.5, 2), (30000, 6, 15000, 10, 0.5, 32),
                                                           (60000, 7, 15000, 10, 0.5, 32)])
def test_find_candidates_gpu_vs_oracle(n_seqs, k, cap, limit, ratio, slots):
    from bioinformatics_b200._lib import Context
    rng = np.random.default_rng(n_seqs + k)
    seqs, reads, quals = _problem(rng, n_seqs, 3000 if n_seqs < 10000 else 600)
    reads += ["", "A", "ACGTA", "ACGTAC", "nnnnn"]
    quals += ["", "I", "IIIII", "IIIIII", "IIIII"]
    want = O.find_candidates(seqs, reads, quals, kmer=k, max_refs_per_kmer=cap, leading_quality=10, trailing_quality=15, candidate_limit=limit,
                             min_hit_ratio=ratio, max_candidates=slots)
    with Context(0) as ctx:
        ctx.set_find_sequences(seqs, k, cap)
        got = ctx.find_candidates(reads, quals, 10, 15, limit, ratio, max_candidates=slots)
        _assert_find_equal(got, want)
        # a window of a larger batch (offsets not starting at 0) and an empty batch
        rb, qb = A.SeqBatch.from_strings(reads), A.SeqBatch.from_strings(quals)
        win = A.SeqBatch(rb.data, rb.offsets[100:301])
        got = ctx.find_candidates(win, A.SeqBatch(qb.data, qb.offsets[100:301]), 10, 15, limit, ratio, max_candidates=slots)
        assert got.candidate_lists() == want.candidate_lists()[100:300]
        assert ctx.find_candidates([], [], 10, 15, limit, ratio).n == 0
    assert want.n_candidates.sum() > 50


@pytest.mark.gpu
def test_find_errors_gpu():
    from bioinformatics_b200._lib import Context, HsqError
    with Context(0) as ctx:
        with pytest.raises(HsqError) as e:
            ctx.find_candidates(["ACGT"], ["IIII"])
        assert e.value.status == A.HSQ_ERR_INVALID_ARGUMENT  # no index yet
        for seqs, status in ((["ACGTNACGT"], A.HSQ_ERR_BAD_CHARACTER), (["ACGT"], A.HSQ_ERR_OUT_OF_RANGE)):
            with pytest.raises(HsqError) as e:
                ctx.set_find_sequences(seqs)
            assert e.value.status == status
        with pytest.raises(HsqError) as e:
            ctx.set_find_sequences(["ACGTACGT"], kmer=11)
        assert e.value.status == A.HSQ_ERR_UNSUPPORTED
        ctx.set_find_sequences(["ACGTACGTAA"])
        with pytest.raises(HsqError) as e:
            ctx.find_candidates(["ACGTRCGTAA"], ["I" * 10])
        assert e.value.status == A.HSQ_ERR_BAD_CHARACTER
        assert ctx.find_candidates(["ACGTNCGTAA"], ["I" * 10]).has_n[0] == 1  # an N is a skip, not an error
        with pytest.raises(HsqError) as e:  # new IupacNucleotideCodeSequence(readString) throws before the N test is reached
            ctx.find_candidates(["ACGTN*GTAA"], ["I" * 10])
        assert e.value.status == A.HSQ_ERR_BAD_CHARACTER


@pytest.mark.gpu
def test_find_candidates_long_and_mixed_length_reads_gpu():
    from bioinformatics_b200._lib import Context
    rng = np.random.default_rng(5)
    seqs = [H.random_seq(rng, int(rng.integers(20, 60))) for _ in range(500)]
    reads, quals = [], []
    for k in range(400):
        L = int(rng.choice([30, 150, 300, 1200, 3000]))
        body = list(H.random_seq(rng, L))
        for _ in range(L // 60):  # several sequences to find planted per read: many tally groups, long candidate walks
            s_ = seqs[int(rng.integers(0, 500))]
            at = int(rng.integers(0, max(1, L - len(s_))))
            body[at:at + len(s_)] = list(s_[:max(0, L - at)])
        reads.append("".join(body)[:L])
        quals.append("I" * L)
    for limit, ratio in ((10, 0.5), (3, 0.2), (40, 0.05)):
        want = O.find_candidates(seqs, reads, quals, candidate_limit=limit, min_hit_ratio=ratio, max_candidates=64)
        with Context(0) as ctx:
            ctx.set_find_sequences(seqs, 5, 15000)
            _assert_find_equal(ctx.find_candidates(reads, quals, 0, 0, limit, ratio, max_candidates=64), want)
        assert want.n_candidates.max() > 3


class _OracleFind:
    def __init__(self, seqs, k, cap):
        self.seqs, self.k, self.cap = seqs, k, cap

    def find_candidates(self, reads, quals, lead_q, trail_q, max_candidates=16):
        return O.find_candidates(self.seqs, reads, quals, kmer=self.k, max_refs_per_kmer=self.cap, leading_quality=lead_q, trailing_quality=trail_q,
                                 max_candidates=max_candidates)

    def align_batch(self, refs, qrys, flags=0, cols_cap=640, md_cap=640):
        return O.align_batch(refs, qrys, flags=flags, cols_cap=cols_cap, md_cap=md_cap, nthreads=8)


def _find_flow_problem():
    rng = np.random.default_rng(77)
    seqs, reads, quals = _problem(rng, 120, 1500, read_len=(40, 150), related=0.9)
    return A.SeqBatch.from_strings(seqs), A.SeqBatch.from_strings(reads), A.SeqBatch.from_strings(quals)


def test_find_flow_on_oracle():
    seqs, reads, quals = _find_flow_problem()
    out = flow.find_flow(_OracleFind(seqs, 5, 15000), seqs, reads, quals, 10, 15, allowed_mismatches=2, allowed_insertion_gaps=1,
                         allowed_insertion_bases=1, allowed_deletion_gaps=1, allowed_deletion_bases=1, max_candidates=32)
    n = out["n_matches"]
    assert (n == 1).sum() > 200 and (n == 0).sum() > 50 and (n > 1).sum() > 5  # single, none, multiple matches all occur
    # an accepted alignment really is within the limits, on the trimmed read
    k = int(np.nonzero(out["accepted"])[0][0])
    assert (out["alignments"].find_counts[k] <= [2, 1, 1, 1, 1]).all()


@pytest.mark.gpu
def test_find_flow_gpu_matches_oracle():
    from bioinformatics_b200._lib import Context
    seqs, reads, quals = _find_flow_problem()
    kw = dict(allowed_mismatches=2, allowed_insertion_gaps=1, allowed_insertion_bases=1, allowed_deletion_gaps=1, allowed_deletion_bases=1,
              max_candidates=32)
    want = flow.find_flow(_OracleFind(seqs, 5, 15000), seqs, reads, quals, 10, 15, **kw)
    with Context(0) as ctx:
        ctx.set_find_sequences(seqs, 5, 15000)
        got = flow.find_flow(ctx, seqs, reads, quals, 10, 15, **kw)
    _assert_find_equal(got["candidates"], want["candidates"])
    assert (got["item_read"] == want["item_read"]).all() and (got["item_seq"] == want["item_seq"]).all()
    H.assert_align_equal(got["alignments"], want["alignments"])
    assert (got["accepted"] == want["accepted"]).all() and (got["n_matches"] == want["n_matches"]).all()


@pytest.mark.gpu
def test_find_matches_fused_equals_the_two_call_flow_on_the_oracle():
    from bioinformatics_b200._lib import Context, HsqError
    seqs, reads, quals = _find_flow_problem()
    limits = (2, 1, 1, 1, 1)
    K = 32
    want = flow.find_flow(_OracleFind(seqs, 5, 15000), seqs, reads, quals, 10, 15, *limits, max_candidates=K)
    with Context(0) as ctx:
        ctx.set_find_sequences(seqs, 5, 15000)
        got = ctx.find_matches(reads, quals, 10, 15, limits, max_candidates=K)
        _assert_find_equal(got.cand, want["candidates"])
        assert (got.n_matches == want["n_matches"]).all()
        mask = np.arange(K)[None, :] < got.cand.n_candidates[:, None]
        assert (got.counts[mask] == want["alignments"].find_counts).all()
        assert (got.accepted[mask].astype(bool) == want["accepted"]).all()
        assert not got.accepted[~mask].any() and got.cells == want["alignments"].cells
        with pytest.raises(HsqError) as e:  # too few slots: an error, never a silently shorter candidate list
            ctx.find_matches(reads, quals, 10, 15, limits, max_candidates=2)
        assert e.value.status == A.HSQ_ERR_OVERFLOW
        assert ctx.find_matches([], [], 10, 15, limits).n == 0
