"""Shared helpers for the parity tests: random/adversarial inputs and field-by-field comparison."""
import numpy as np

ALIGN_FIELDS = ("score", "n_cols", "first_q", "last_q", "first_r", "last_r", "matches", "mismatches", "non_terminal_gaps",
                "edit_distance", "md_len", "rmd_len", "find_counts")
EXTEND_FIELDS = ("ok", "primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen", "cigar_len", "md_len",
                 "overflow", "pair_ok")


def random_seq(rng, n, alphabet="ACGT"):
    return "".join(rng.choice(list(alphabet), n)) if n else ""


def mutate(rng, s, sub=0.05, indel=0.03):
    out = []
    for ch in s:
        r = rng.random()
        if r < indel:
            continue
        if r < 2 * indel:
            out.append(rng.choice(list("ACGT")))
        if rng.random() < sub:
            out.append(rng.choice(list("ACGT")))
        else:
            out.append(ch)
    return "".join(out)


def align_cases(rng, n, max_len=90):
    """Random related / unrelated / tie-heavy / empty / IUPAC pairs."""
    refs, qrys = [], []
    for k in range(n):
        kind = k % 8
        if kind == 0:
            r = random_seq(rng, int(rng.integers(0, max_len)))
            q = mutate(rng, r)
        elif kind == 1:
            r = random_seq(rng, int(rng.integers(0, max_len)))
            q = random_seq(rng, int(rng.integers(0, max_len)))
        elif kind == 2:  # homopolymer / low complexity: many ties
            r = random_seq(rng, int(rng.integers(1, max_len)), "AT")
            q = random_seq(rng, int(rng.integers(1, max_len)), "AT")
        elif kind == 3:  # query is a window of the reference (free end gaps matter)
            r = random_seq(rng, int(rng.integers(20, max_len)))
            a = int(rng.integers(0, len(r) // 2))
            q = mutate(rng, r[a:a + int(rng.integers(1, len(r) - a + 1))], 0.02, 0.01)
        elif kind == 4:  # IUPAC + N + gap characters + U + lower case
            r = random_seq(rng, int(rng.integers(1, 40)), "ACGTNRYSWKMBDHVU-acgtn")
            q = random_seq(rng, int(rng.integers(1, 40)), "ACGTNRYSWKMBDHVU-acgtn")
        elif kind == 5:  # lengths around the 16-row strip boundaries
            L = int(rng.choice([1, 15, 16, 17, 31, 32, 33, 47, 48, 49]))
            q = random_seq(rng, L)
            r = mutate(rng, q, 0.1, 0.05)
        elif kind == 6:
            r = random_seq(rng, int(rng.integers(0, 4)))
            q = random_seq(rng, int(rng.integers(0, 4)))
        else:
            r = "A" * int(rng.integers(1, 50))
            q = "A" * int(rng.integers(1, 50))
        refs.append(r)
        qrys.append(q)
    return refs, qrys


def assert_align_equal(got, exp):
    for f in ALIGN_FIELDS:
        a, b = getattr(got, f), getattr(exp, f)
        bad = np.nonzero(a != b)[0]
        assert len(bad) == 0, "field %s differs at %s: got %s expected %s" % (f, bad[:5], a[bad[:5]], b[bad[:5]])
    assert got.ref_alns() == exp.ref_alns()
    assert got.qry_alns() == exp.qry_alns()
    assert got.op_strings() == exp.op_strings()
    assert got.mds() == exp.mds()
    assert got.rmds() == exp.rmds()
    assert got.cells == exp.cells


def assert_extend_equal(got, exp):
    for f in EXTEND_FIELDS:
        a, b = getattr(got, f), getattr(exp, f)
        bad = np.nonzero(a != b)[0]
        assert len(bad) == 0, "field %s differs at %s: got %s expected %s" % (f, bad[:5], a[bad[:5]], b[bad[:5]])
    assert got.cigars() == exp.cigars()
    assert got.mds() == exp.mds()
    assert got.cells == exp.cells
