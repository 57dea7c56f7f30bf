"""A plain-Python model of the opt-in band (hsq_config.band, csrc/hsq_dp2.cuh band_cols + csrc/hsq_extend.cuh capture2_body<true>), runnable
without a GPU: the reference's pointer-driven fill (NeedlemanWunschGlobalAlignment.java:261-376, free terminal gaps at both ends as
ExtendReadsToPrimer.java:211 uses it) restricted to the staircase, cells that are not computed unavailable as predecessors, and the three
acceptance tests.  On adversarial shapes every read end the model accepts has the full matrix's score and path; the same set without
test 0 contains read ends that do not (tiny capture targets), which is how that test came about.  The CUDA kernel is checked against the
oracle on the same kind of data in tests/test_band.py."""
import numpy as np

from tests import helpers as H
from tests import pyref as P

K = 16


def fill(reference, query, sc, band=0):
    nq, nr = len(query), len(reference)
    n_strips = (nq + K - 1) // K

    def cols(s):
        if band == 0:
            return 1, nr
        lo = min(nr, max(1, s * K + 1 - band))
        hi = nr if s == n_strips - 1 else min(nr, s * K + K + band)
        return lo, hi
    S = [[None] * (nr + 1) for _ in range(nq + 1)]
    src = [[None] * (nr + 1) for _ in range(nq + 1)]
    for c in range(nr + 1):
        S[0][c] = 0
        src[0][c] = (0, c - 1) if c else None
    for r in range(nq + 1):
        S[r][0] = 0
        src[r][0] = (r - 1, 0) if r else None
    for r in range(1, nq + 1):
        lo, hi = cols((r - 1) // K)
        for c in range(lo, hi + 1):
            cands = []
            if S[r][c - 1] is not None and (c - 1 >= lo or c - 1 == 0):
                h = S[r][c - 1]
                if r != nq:  # horizontal moves are free in the last row
                    h += sc.gap_extend if (src[r][c - 1] is not None and src[r][c - 1][0] == r) else sc.gap_start
                cands.append((h, 2, (r, c - 1)))
            if S[r - 1][c] is not None:
                v = S[r - 1][c]
                if c != nr:  # vertical moves are free in the last column
                    v += sc.gap_extend if (src[r - 1][c] is not None and src[r - 1][c][1] == c) else sc.gap_start
                cands.append((v, 1, (r - 1, c)))
            if S[r - 1][c - 1] is not None:
                cands.append((S[r - 1][c - 1] + (sc.match if P.matches(query[r - 1], reference[c - 1]) else sc.mismatch), 0, (r - 1, c - 1)))
            if cands:
                best = max(cands, key=lambda t: (t[0], t[1]))  # ties: horizontal > vertical > diagonal
                S[r][c], src[r][c] = best[0], best[2]
    return S, src


def path(src, nq, nr):
    out, r, c = [], nq, nr
    while src[r][c] is not None:
        out.append((r, c))
        r, c = src[r][c]
    return out


def accepted(reference, query, sc, band, test0=True):
    nq, nr = len(query), len(reference)
    if test0 and min(nq, nr) < 2 * band + K:
        return None
    S, src = fill(reference, query, sc, band)
    s = S[nq][nr]
    if s is None or s <= 0 or s <= sc.match * min(nq, nr) + sc.gap_start + (band - 1) * sc.gap_extend:
        return None
    p = path(src, nq, nr)
    if not all(abs(c - r) < band or r == nq or c == nr for r, c in p if r > 0 and c > 0):
        return None
    return s, p


def cases(seed, n):
    rng = np.random.default_rng(seed)
    out = []
    for k in range(n):
        target = H.random_seq(rng, int(rng.integers(1, 150)))
        shift = H.random_seq(rng, int(rng.integers(0, 10))) if k % 3 == 0 else ""
        read = (shift + H.mutate(rng, target, 0.04, 0.03) + H.random_seq(rng, 60))[:int(rng.integers(1, 160))]
        if k % 5 == 0:
            read = H.random_seq(rng, len(read))
        out.append((target, read))
    return out


def test_model_accepts_only_what_the_full_matrix_gives():
    sc = P.Scorer()
    n_acc = 0
    for band in (3, 9):
        for ref, qry in cases(100 + band, 120):
            got = accepted(ref, qry, sc, band)
            if got is None:
                continue
            n_acc += 1
            S, src = fill(ref, qry, sc, 0)
            assert got[0] == S[len(qry)][len(ref)] and got[1] == path(src, len(qry), len(ref))
            assert got[0] == P.needleman_wunsch(ref, qry, sc)[0]  # and the full model is the reference statement of tests/pyref.py
    assert n_acc > 10


def test_model_without_test0_is_wrong_on_tiny_targets():
    """Why test 0 exists: a three-base target matches anywhere along a read, the band only looks near the start."""
    sc = P.Scorer()
    rng = np.random.default_rng(7)
    wrong = 0
    for _ in range(300):
        target = H.random_seq(rng, 3)
        read = H.random_seq(rng, 60) + target  # the only good placement is at the far end of the read
        got = accepted(target, read, sc, 2, test0=False)
        if got is None:
            continue
        S, src = fill(target, read, sc, 0)
        wrong += got[0] != S[len(read)][len(target)] or got[1] != path(src, len(read), len(target))
        assert accepted(target, read, sc, 2, test0=True) is None
    assert wrong > 0
