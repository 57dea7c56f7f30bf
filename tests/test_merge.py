"""mergeSequences: the reference's six known-answer tests
(nimblegen_heatseq/src/test/java/com/roche/heatseq/process/PrimerReadExtensionAndPcrDuplicateIdentificationTest.java:30-93)
pin the oracle; the CUDA path is checked against the KATs and, on random cases, against the oracle."""
import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from oracle import oracle as O

# (negative, upstream, downstream, upstream quality, downstream quality, upstream start, downstream start, merged, merged quality)
KATS = [
    (0, "AACCGGTT", "TTGGAACC", "HHHHHHHH", "HHHHHHHH", 10, 16, "AACCGGTTGGAACC", "HHHHHHHHHHHHHH"),      # standardOneTest :30-39
    (0, "AACCGGTT", "TTGGAACC", "HHHHHHHH", "HHHHHHHH", 10, 20, "AACCGGTTNNTTGGAACC", "HHHHHHHH!!HHHHHHHH"),  # gapOneTest :41-50
    (0, "AACCGGTC", "TTGGAACC", "HHHHHHHI", "HHHHHHHH", 10, 16, "AACCGGTCGGAACC", "HHHHHHHIHHHHHH"),      # disagreementOneTest :52-61
    (0, "AACCGGTC", "TTGGAACC", "HHHHHHHH", "HHHHHHHH", 1, 9, "AACCGGTCTTGGAACC", "HHHHHHHHHHHHHHHH"),    # noOverlapOneTest :63-72
    (0, "AACCGGTC", "AACCGGTC", "HHHHHHHH", "HHHHHHHH", 1, 1, "AACCGGTC", "HHHHHHHH"),                    # onlyOverlapOneTest :74-83
    (0, "TAACCGGTCC", "AACCGGTC", "HHHHHHHHHH", "HHHHHHHH", 25, 26, "TAACCGGTCC", "HHHHHHHHHH"),          # longerUpstreamOneTest :85-93
]


def _cols(cases):
    return [list(c) for c in zip(*cases)]


def _random_cases(rng, n):
    cases = []
    for k in range(n):
        ul, dl = int(rng.integers(0, 60)), int(rng.integers(0, 60))
        alpha = list("ACGT") if k % 4 else list("ACGTNRYacgtnU-")
        up = "".join(rng.choice(alpha, ul))
        down = "".join(rng.choice(alpha, dl))
        uq = "".join(chr(33 + int(x)) for x in rng.integers(0, 42, ul))
        dq = "".join(chr(33 + int(x)) for x in rng.integers(0, 42, dl))
        us = int(rng.integers(1, 50))
        ds = us + int(rng.integers(0, 90))  # overlap, abutting, gap, contained
        cases.append((int(rng.integers(0, 2)), up, down, uq, dq, us, ds))
    return cases


def test_oracle_reference_kats():
    neg, up, down, uq, dq, us, ds, exp_s, exp_q = _cols(KATS)
    s, q = O.merge_pairs(neg, up, down, uq, dq, us, ds)
    assert s.strings() == exp_s and q.strings() == exp_q


def test_oracle_negative_strand_and_errors():
    s, q = O.merge_pairs([1], ["AACCGGTT"], ["TTGGAACC"], ["ABCDEFGH"], ["HHHHHHHH"], [10], [20])
    assert s.strings() == ["GGTTCCAANNAACCGGTT"] and q.strings() == ["HHHHHHHH!!HGFEDCBA"]
    with pytest.raises(O.OracleError) as e:  # downstream starts before upstream inside an overlap: Java substring throws
        O.merge_pairs([0], ["AACCGGTT"], ["TTGGAACC"], ["HHHHHHHH"], ["HHHHHHHH"], [10], [8])
    assert e.value.status == A.HSQ_ERR_OUT_OF_RANGE


@pytest.mark.gpu
def test_cuda_merge_kats_and_random():
    from bioinformatics_b200._lib import Context, HsqError
    rng = np.random.default_rng(5)
    with Context(0) as ctx:
        neg, up, down, uq, dq, us, ds, exp_s, exp_q = _cols(KATS)
        s, q = ctx.merge_pairs(neg, up, down, uq, dq, us, ds)
        assert s.strings() == exp_s and q.strings() == exp_q
        cases = _random_cases(rng, 4000)
        cols = _cols(cases)
        gs, gq = ctx.merge_pairs(*cols)
        es, eq = O.merge_pairs(*cols)
        assert (gs.offsets == es.offsets).all() and (gs.data == es.data).all() and (gq.data == eq.data).all()
        with pytest.raises(HsqError) as e:
            ctx.merge_pairs([0], ["AACCGGTT"], ["TTGGAACC"], ["HHHHHHHH"], ["HHHHHHHH"], [10], [8])
        assert e.value.status == A.HSQ_ERR_OUT_OF_RANGE
        with pytest.raises(HsqError) as e:
            ctx.merge_pairs([0], ["AACXGGTT"], ["TTGGAACC"], ["HHHHHHHH"], ["HHHHHHHH"], [10], [12])
        assert e.value.status == A.HSQ_ERR_BAD_CHARACTER
        with pytest.raises(HsqError) as e:  # explicit small capacity: overflow, offsets still say what is needed
            ctx.merge_pairs([0], ["AACCGGTT"], ["TTGGAACC"], ["HHHHHHHH"], ["HHHHHHHH"], [10], [16], out_cap=4)
        assert e.value.status == A.HSQ_ERR_OVERFLOW
        s, q = ctx.merge_pairs([], [], [], [], [], [], [])
        assert s.n == 0
