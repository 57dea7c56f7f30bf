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
    if band >= 16:
        assert redone < banded * 0.6                               # ... and a good part of the read ends was not


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
