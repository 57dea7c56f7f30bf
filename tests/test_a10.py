"""SURVEY 8(a) row A10 (optional, dead in the CLI): getVariableLengthUid (HSQ/utils/SAMRecordUtil.java:223-261, quirk :238-250 kept) and the
mergePairs composition -- mergeSequences, then the merged read aligned to the capture target under the default scorer, CIGAR and MD of that
alignment on the merged record (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:1274-1339, DP at :1306).
CPU: C++ oracle == the independent Python statement (tests/pyref.py).  GPU: CUDA == oracle."""
import numpy as np
import pytest

from bioinformatics_b200.batches import scorer
from oracle import oracle as O
from tests import helpers as H
from tests import pyref as P


def uid_cases(rng, n):
    """Reads = UID (0..14 bases) ++ primer copy (exact, mutated, or unrelated) ++ tail; primers 12..26 bases."""
    reads, primers = [], []
    for k in range(n):
        primer = H.random_seq(rng, int(rng.integers(12, 27)))
        uid = H.random_seq(rng, int(rng.integers(0, 15)))
        kind = k % 5
        body = primer if kind == 0 else H.mutate(rng, primer, 0.06, 0.03) if kind in (1, 2) else H.mutate(rng, primer, 0.25, 0.1) if kind == 3 \
            else H.random_seq(rng, len(primer))
        tail = H.random_seq(rng, int(rng.integers(0, 30)))
        read = uid + body + tail
        if k % 17 == 0:
            read = read.lower()
        if k % 23 == 0 and len(read) > 3:
            read = read[:2] + "N" + read[3:]
        reads.append(read)
        primers.append(primer)
    return reads, primers


def test_variable_length_uid_oracle_equals_python_statement():
    rng = np.random.default_rng(21)
    reads, primers = uid_cases(rng, 600)
    seen = set()
    for sc_args in ((1, -4, -1, -6), (2, -3, -2, -4)):
        for r, p in zip(reads, primers):
            want = P.variable_length_uid(r, p, P.Scorer(*sc_args))
            got = O.variable_length_uid(r, p, scorer(*sc_args))
            assert (None if want is None else len(want)) == got, (r, p, want, got)
            seen.add(got is None)
    assert seen == {True, False}
    # hand case: a 5-base UID in front of an exact 20-base primer: columns = 5 D + 20 '=' + 3 D -> edit 8 >= 20 // 4: the reference's own
    # cutoff (with its never-decremented deletion count) rejects it; a 2-base UID and no tail passes
    primer = "ACGTTGCAAGGCTTAACCGG"
    assert O.variable_length_uid("TTTTT" + primer + "AAA", primer) is None
    assert O.variable_length_uid("TT" + primer, primer) == 2


def merge_cases(rng, n):
    out = []
    for k in range(n):
        target = H.random_seq(rng, int(rng.integers(40, 90)))
        a = int(rng.integers(20, len(target)))
        b = int(rng.integers(0, a + 5))  # overlap, abutting or a gap of a few bases
        up = H.mutate(rng, target[:a], 0.03, 0.01)
        down = H.mutate(rng, target[b:], 0.03, 0.01)
        uq = "".join(chr(33 + int(x)) for x in rng.integers(2, 42, len(up)))
        dq = "".join(chr(33 + int(x)) for x in rng.integers(2, 42, len(down)))
        out.append((int(rng.integers(0, 2)), up, down, uq, dq, 100, 100 + min(b, len(up) + 4), target))
    return out


def test_merge_pairs_composition_oracle_equals_python_statement():
    """mergePairs: the merged sequence (oracle mergeSequences, pinned on the reference's 6 KATs in tests/test_merge.py) against the capture
    target: standard CIGAR and MD from the oracle's alignment getters == the Python statement's."""
    rng = np.random.default_rng(22)
    cases = merge_cases(rng, 120)
    neg, up, down, uq, dq, us, ds, target = [list(c) for c in zip(*cases)]
    merged, _ = O.merge_pairs(neg, up, down, uq, dq, us, ds)
    got = O.align_batch(target, merged.strings(), cols_cap=400, md_cap=600)
    ops, mds = got.op_strings(), got.mds()
    for k, m in enumerate(merged.strings()):
        score, ra, qa = P.needleman_wunsch(P.codes(target[k]), P.codes(m))
        ap = P.AlignmentPair(ra, qa)
        assert ap.expanded_cigar() == ops[k] and ap.md() == mds[k] and score == int(got.score[k]), k
        assert ap.standard_cigar() == P.summarize(ops[k].replace("X", "M").replace("=", "M"))


@pytest.mark.gpu
def test_gpu_variable_length_uid_equals_oracle():
    from bioinformatics_b200._lib import Context
    rng = np.random.default_rng(23)
    reads, primers = uid_cases(rng, 5000)
    reads += ["", "ACGT", "ACGTACGTACGTACGTACGT"]
    primers += ["ACGTACGTACGT", "", "ACGTACGTACGTACGTACGT"]
    with Context(0) as ctx:
        for sc_args, flags in (((1, -4, -1, -6), 0), ((2, -3, -2, -4), 0), ((1, -4, -1, -6), 3)):
            got = ctx.variable_length_uid(reads, primers, scorer(*sc_args), flags)
            want = [O.variable_length_uid(r, p, scorer(*sc_args), flags) for r, p in zip(reads, primers)]
            assert got.tolist() == [-1 if w is None else w for w in want]
        assert len(ctx.variable_length_uid([], [])) == 0


@pytest.mark.gpu
def test_gpu_merge_pairs_composition_equals_oracle():
    """hsq_merge_pairs -> hsq_align_batch(capture target, merged read) on the device == the same composition over the oracle."""
    from bioinformatics_b200._lib import Context
    rng = np.random.default_rng(24)
    cases = merge_cases(rng, 1500)
    neg, up, down, uq, dq, us, ds, target = [list(c) for c in zip(*cases)]
    with Context(0) as ctx:
        gm, gq = ctx.merge_pairs(neg, up, down, uq, dq, us, ds)
        em, eq = O.merge_pairs(neg, up, down, uq, dq, us, ds)
        assert gm.strings() == em.strings() and gq.strings() == eq.strings()
        got = ctx.align_batch(target, gm, cols_cap=400, md_cap=600)
    H.assert_align_equal(got, O.align_batch(target, em, cols_cap=400, md_cap=600))
