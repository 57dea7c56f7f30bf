"""The opt-in band of the capture-target DP (hsq_config.band; BASELINE.json configs[4]).

The reference fills full matrices (NeedlemanWunschGlobalAlignment.java:72-78).  With a band the packed capture kernel computes
a staircase around the main diagonal, accepts a read end only if its traceback stays strictly inside the band, and redoes every
other read end with full matrices: the results checked here are those of the oracle's full matrices, for every band."""
import numpy as np
import pytest

from bioinformatics_b200 import synth
from bioinformatics_b200._lib import Context
from bioinformatics_b200.batches import PairBatchHost, scorer
from oracle import oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu


def _with_long_indels(pairs, rng, every=7):
    """every `every`-th pair gets a deletion or an insertion of 5..40 bases somewhere behind its primer: its path leaves a narrow band"""
    r1, r2 = pairs.r1.strings(), pairs.r2.strings()
    for k in range(0, pairs.n, every):
        for reads in (r1, r2):
            s = reads[k]
            at = int(rng.integers(45, max(46, len(s) - 20)))
            d = int(rng.integers(5, 41))
            if rng.random() < 0.5:
                s = s[:at] + s[at + d:]
            else:
                s = s[:at] + H.random_seq(rng, d) + s[at:]
            reads[k] = s
    return PairBatchHost(pairs.probe_id, r1, r2)


@pytest.mark.parametrize("band", [1, 8, 16, 32, 64])
def test_band_results_equal_full_matrices(band):
    rng = np.random.default_rng(500 + band)
    probes = synth.make_probes(200, seed=41)
    pairs = synth.make_pairs(probes, 8000, read_len=150, ext_uid_len=10, seed=43, sub_rate=0.02, ins_rate=0.004, del_rate=0.004,
                             with_quals=False)
    pairs = _with_long_indels(pairs, rng)
    exp = O.extend_batch(probes, pairs, cigar_cap=96, md_cap=160, nthreads=8)
    with Context(0, config={"band": band}) as ctx:
        ctx.set_probes(probes)
        got = ctx.extend_batch(pairs, cigar_cap=96, md_cap=160)
        banded, redone = ctx.band_stats()
    H.assert_extend_equal(got, exp)
    reached = int((got.primer_end >= 0).sum())
    assert 0 < banded <= 2 * pairs.n and banded >= reached * 0.9  # every read end that reached DP#2 went through the band first
    assert 0 < redone <= banded                                    # the long indels (and the unaligned read ends) were redone in full
    if 16 <= band <= 32:
        assert redone < banded * 0.6                               # ... and a good part of the read ends was not
    if band == 64:
        assert redone > 0.9 * banded                               # 2 * 64 + 16 rows and columns: hardly a matrix of a 150-base read is that large


def test_band_off_and_unavailable():
    """band = 0 reports no banded read ends; a scorer whose scores leave no room for the value of an uncomputed cell, and the 32-bit
    kernels, ignore the band."""
    probes = synth.make_probes(60, seed=45)
    pairs = synth.make_pairs(probes, 1500, read_len=120, ext_uid_len=8, seed=46, sub_rate=0.02, ins_rate=0.003, del_rate=0.003,
                             with_quals=False)
    exp = O.extend_batch(probes, pairs, cigar_cap=64, md_cap=96, nthreads=8)
    with Context(0) as ctx:
        ctx.set_probes(probes)
        H.assert_extend_equal(ctx.extend_batch(pairs, cigar_cap=64, md_cap=96), exp)
        assert ctx.band_stats() == (0, 0)
        ctx.configure(band=16, force_dp32=1)
        H.assert_extend_equal(ctx.extend_batch(pairs, cigar_cap=64, md_cap=96), exp)
        assert ctx.band_stats() == (0, 0)
        ctx.configure(band=16)
        sc = scorer(9, -9, -9, -18)  # fits the packed kernel's 14 bits, but not twice
        ctx.set_scorer(sc)
        H.assert_extend_equal(ctx.extend_batch(pairs, cigar_cap=64, md_cap=96), O.extend_batch(probes, pairs, sc, cigar_cap=64, md_cap=96, nthreads=8))


@pytest.mark.parametrize("chunks", [1, 3])
def test_band_through_the_chunked_and_packed_calls(chunks):
    probes = synth.make_probes(80, seed=47)
    pairs = synth.make_pairs(probes, 6001, read_len=150, ext_uid_len=10, seed=48, sub_rate=0.02, ins_rate=0.003, del_rate=0.003,
                             with_quals=False)
    exp = O.extend_batch(probes, pairs, cigar_cap=64, md_cap=96, nthreads=8)
    with Context(0, config={"band": 12, "extend_chunks": chunks}) as ctx:
        ctx.set_probes(probes)
        H.assert_extend_equal(ctx.extend_batch(pairs, cigar_cap=64, md_cap=96), exp)
        banded, redone = ctx.band_stats()
        assert banded > 0
        from tests.test_packed import assert_pooled_equal, packed_pairs_of
        assert_pooled_equal(ctx.extend_packed(packed_pairs_of(pairs, 2)), exp)
        assert ctx.band_stats() == (banded, redone)


@pytest.mark.parametrize("band", [2, 5, 16, 40])
def test_band_on_adversarial_shapes(band):
    """Random reads against random probes, capture targets of 1 .. 170 bases against read remainders of 0 .. 160 (targets much shorter
    and much longer than the read, so the free terminal gaps run along the last column or the last row for most of the matrix), IUPAC
    primers, N in reads, three scorers the band accepts and one it must ignore: every result equals the oracle's full matrices."""
    from bioinformatics_b200.batches import ProbeTableHost
    rng = np.random.default_rng(900 + band)
    n_probes, n = 40, 1500
    ext = [H.random_seq(rng, int(rng.integers(5, 30)), "ACGTN" if k % 5 == 0 else "ACGT") for k in range(n_probes)]
    lig = [H.random_seq(rng, int(rng.integers(5, 30))) for _ in range(n_probes)]
    tgt = [H.random_seq(rng, int(rng.integers(1, 170)), "ACGTR" if k % 7 == 0 else "ACGT") for k in range(n_probes)]
    z = np.arange(n_probes, dtype=np.int32) * 500
    probes = ProbeTableHost(np.arange(n_probes) % 2, z + 10, z + 30, z + 300, z + 320, z + 31, z + 299, ext, lig, tgt)
    pid = rng.integers(0, n_probes, n)
    r1, r2 = [], []
    for k in range(n):
        p = int(pid[k])
        mol = ext[p].replace("N", "A") + tgt[p].replace("R", "G") + lig[p]
        rc = mol[::-1].translate(str.maketrans("ACGT", "TGCA"))
        L = int(rng.integers(12, 190))
        if k % 4 == 0:
            a, b = H.random_seq(rng, int(rng.integers(0, 160))), H.random_seq(rng, int(rng.integers(0, 160)))
        else:
            shift = H.random_seq(rng, int(rng.integers(0, 12))) if k % 3 == 0 else ""  # reads that start off the expected diagonal
            a = (H.random_seq(rng, 6) + H.mutate(rng, mol[:len(ext[p])], 0.02, 0.0) + shift + H.mutate(rng, mol[len(ext[p]):], 0.04, 0.03) + H.random_seq(rng, 200))[:L]
            b = (H.mutate(rng, rc, 0.04, 0.03) + H.random_seq(rng, 200))[:L]
            if k % 4 == 3:
                a = a.replace("A", "N", 1)
        r1.append(a)
        r2.append(b)
    pairs = PairBatchHost(pid, r1, r2)
    with Context(0, config={"band": band}) as ctx:
        ctx.set_probes(probes)
        for sc in (scorer(), scorer(2, -3, -2, -4), scorer(5, -4, 0, -7), scorer(1, -4, 1, 2)):
            ctx.set_scorer(sc)
            got = ctx.extend_batch(pairs, cigar_cap=1600, md_cap=2400)
            exp = O.extend_batch(probes, pairs, sc, cigar_cap=1600, md_cap=2400, nthreads=8)
            H.assert_extend_equal(got, exp)
            banded, redone = ctx.band_stats()
            if sc.gap_extend > 0:
                assert banded == 0  # gapExtend > 0: not a scorer the acceptance bound holds for, full matrices
            else:
                assert banded > 0 and redone > 0
