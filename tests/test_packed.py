"""N1 (north_star subsystem 1): packed read batches and pooled CIGAR / MD output.

CPU part: the host packing helper of the library against a plain-Python statement of the layout in include/hsq_b200.h.
GPU part: hsq_unpack_reads round trip, and hsq_extend_packed == the oracle's extendReadsToPrimers on the same reads."""
import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from bioinformatics_b200 import synth
from bioinformatics_b200._lib import HsqError, pack_reads, quality_sums
from bioinformatics_b200.batches import PackedPairsHost, PooledExtendResult
from tests import helpers as H

CODE2 = {"A": 0, "C": 1, "T": 2, "G": 3}
IUPAC4 = {"A": 1, "C": 2, "G": 4, "T": 8, "R": 5, "Y": 10, "S": 6, "W": 9, "K": 12, "M": 3, "B": 14, "D": 13, "H": 11, "V": 7, "N": 15,
          "_": 0, "-": 0, ".": 0}


def py_pack(strings, bits):
    """The layout of hsq_packed_reads, restated: words per read, base i at bits [bits*(i % per), +bits) of word i // per."""
    per = 32 // bits
    words, lens, exc = [], [], []
    for k, s in enumerate(strings):
        lens.append(len(s))
        w = [0] * ((len(s) * bits + 31) // 32)
        for i, ch in enumerate(s):
            if bits == 2:
                v = CODE2.get(ch.upper())
                if v is None:
                    exc.append((k, i, ord(ch)))
                    v = 0
            else:
                v = IUPAC4[ch.upper()]
            w[i // per] |= v << (bits * (i % per))
        words += w
    return np.array(words, dtype=np.uint32), np.array(lens, dtype=np.int32), exc


def random_reads(rng, n, alphabet="ACGT", lo=0, hi=70):
    return ["".join(rng.choice(list(alphabet), int(rng.integers(lo, hi)))) for _ in range(n)]


@pytest.mark.parametrize("bits", [2, 4])
def test_pack_reads_layout(bits):
    rng = np.random.default_rng(bits)
    strings = random_reads(rng, 3000) + ["", "A", "ACGTACGTACGTACGT", "ACGTACGTACGTACGTA", "N" * 33, "acgtn"]
    strings += random_reads(rng, 6000, "ACGTNacgtRYSWKMBDHV-", 0, 40)  # enough reads for the multi-threaded path
    p = pack_reads(strings, bits, threads=3)
    words, lens, exc = py_pack(strings, bits)
    assert np.array_equal(p.len, lens)
    assert np.array_equal(p.data.view(np.uint32), words)
    if bits == 2:
        assert list(zip(p.exc_read.tolist(), p.exc_pos.tolist(), p.exc_char.tolist())) == exc
    else:
        assert len(p.exc_read) == 0


def test_pack_reads_errors():
    with pytest.raises(HsqError) as e:
        pack_reads(["ACGT", "AC!T"], 2)
    assert e.value.status == A.HSQ_ERR_BAD_CHARACTER
    with pytest.raises(HsqError) as e:
        pack_reads(["ACGU"], 4)
    assert e.value.status == A.HSQ_ERR_UNSUPPORTED
    assert pack_reads([], 2).n == 0


def test_quality_sums_wrap():
    """(short) sum(q - 33): Java short arithmetic wraps (BamFileUtil.java:74-85)."""
    q = A.SeqBatch.from_strings(["I" * 10, "", "~" * 400, "!" * 7])
    assert quality_sums(q).tolist() == [400, 0, ((93 * 400 + 32768) % 65536) - 32768, 0]


# ---- GPU --------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def ctx():
    from bioinformatics_b200._lib import Context
    with Context(0) as c:
        yield c


def packed_pairs_of(pairs, bits=2, alloc=None):
    return PackedPairsHost(pairs.probe_id, pack_reads(pairs.r1, bits, alloc), pack_reads(pairs.r2, bits, alloc),
                           None if pairs.q1 is None else pairs.q1.data, None if pairs.q2 is None else pairs.q2.data, pairs.read_index)


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [2, 4])
def test_unpack_round_trip(ctx, bits):
    rng = np.random.default_rng(10 + bits)
    strings = random_reads(rng, 5000, "ACGTN" if bits == 2 else "ACGTNRYSWKMBDHV", 0, 200) + ["", "A" * 16, "C" * 17, "N" * 5]
    got = ctx.unpack_reads(pack_reads(strings, bits))
    assert got.strings() == strings


def assert_pooled_equal(got, exp):
    for f in ("ok", "primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen", "cigar_len", "md_len", "pair_ok"):
        a, b = getattr(got, f), getattr(exp, f)
        bad = np.nonzero(a != b)[0]
        assert len(bad) == 0, "field %s differs at %s: got %s expected %s" % (f, bad[:5], a[bad[:5]], b[bad[:5]])
    assert got.cigars() == exp.cigars()
    assert got.mds() == exp.mds()
    assert got.cells == exp.cells
    assert got.pool_bytes == int(exp.cigar_len.sum() + exp.md_len.sum())
    assert np.array_equal(got.str_off, np.concatenate([[0], np.cumsum(exp.cigar_len + exp.md_len)[:-1]]))


@pytest.mark.gpu
@pytest.mark.parametrize("bits,chunks,pinned", [(2, 0, False), (4, 0, False), (2, 3, True), (2, 7, False)])
def test_extend_packed_equals_oracle(ctx, bits, chunks, pinned):
    """Packed reads (2 bit with the N exceptions, 4 bit) through the chunked pipeline, pool in pinned (written by the device) and in
    pageable memory: every field, string and the pool layout equal the oracle's extendReadsToPrimers."""
    from oracle import oracle as O
    probes = synth.make_probes(90, seed=41)
    pairs = synth.make_pairs(probes, 6001, read_len=150, ext_uid_len=10, seed=42, sub_rate=0.02, ins_rate=0.004, del_rate=0.004)
    ctx.configure(extend_chunks=chunks)
    ctx.set_probes(probes)
    exp = O.extend_batch(probes, pairs, cigar_cap=64, md_cap=96, nthreads=8)
    pp = packed_pairs_of(pairs, bits)
    out = PooledExtendResult(pairs.n, 2 * pairs.n * 40, alloc=ctx.pinned_array if pinned else None)
    got = ctx.extend_packed(pp, out=out)
    assert_pooled_equal(got, exp)
    # a pool that is too small: the call says how much it needs
    with pytest.raises(HsqError) as e:
        ctx.extend_packed(pp, pool_cap=1000)
    assert e.value.status == A.HSQ_ERR_OVERFLOW
    # only the columns asked for are delivered
    few = ctx.extend_packed(pp, fields=("pos", "pair_ok"))
    assert np.array_equal(few.pos, exp.pos) and np.array_equal(few.pair_ok, exp.pair_ok) and few.score is None
    ctx.configure()


@pytest.mark.gpu
def test_extend_packed_edge_cases(ctx):
    """Empty batch, ragged read lengths (short mates that never reach the capture DP), IUPAC / lower-case reads through the exception list."""
    from oracle import oracle as O
    from bioinformatics_b200.batches import PairBatchHost
    rng = np.random.default_rng(8)
    probes = synth.make_probes(40, seed=43)
    ctx.set_probes(probes)
    empty = PackedPairsHost(np.zeros(0, dtype=np.int32), pack_reads([], 2), pack_reads([], 2))
    got = ctx.extend_packed(empty)
    assert got.cells == 0 and got.pool_bytes == 0
    base = synth.make_pairs(probes, 900, read_len=150, ext_uid_len=10, seed=44, sub_rate=0.03, ins_rate=0.01, del_rate=0.01)
    r1, r2 = base.r1.strings(), base.r2.strings()
    for k in range(0, 900, 3):  # ragged: cut mates to random lengths, sprinkle IUPAC codes and lower case
        r1[k] = r1[k][:int(rng.integers(0, 150))]
        r2[k] = r2[k][:int(rng.integers(5, 150))].lower()
    for k in range(1, 900, 7):
        s = list(r1[k])
        s[int(rng.integers(0, len(s)))] = str(rng.choice(list("NRYKM")))
        r1[k] = "".join(s)
    pairs = PairBatchHost(base.probe_id, r1, r2)
    exp = O.extend_batch(probes, pairs, cigar_cap=64, md_cap=96, nthreads=8)
    assert_pooled_equal(ctx.extend_packed(packed_pairs_of(pairs, 2)), exp)
    assert_pooled_equal(ctx.extend_packed(packed_pairs_of(pairs, 4)), exp)


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [2, 4])
def test_dedup_packed_equals_oracle(ctx, bits):
    """hsq_dedup_packed (quality bytes on the device, and host-made quality sums) == FilterByUid restatement; a UID with an N is refused
    like hsq_dedup_batch refuses it."""
    from oracle import oracle as O
    probes = synth.make_probes(50, seed=51)
    raw = synth.make_pairs(probes, 20000, read_len=120, ext_uid_len=7, lig_uid_len=3, seed=52)
    pairs, _ = synth.drop_n_pairs(raw)
    pairs.q1.data[::3] = ord("#")
    exp = O.dedup_batch(pairs, 7, 3, probes.n)
    pp = packed_pairs_of(pairs, bits)
    for variant in ("bytes", "sums"):
        if variant == "sums":
            pp = PackedPairsHost(pairs.probe_id, pp.r1, pp.r2, None, None, pairs.read_index, quality_sums(pairs.q1), quality_sums(pairs.q2))
        b, g, st = ctx.dedup_packed(pp, 7, 3, probes.n)
        assert np.array_equal(b, exp[0]) and np.array_equal(g, exp[1])
        for f in ("unique_pairs", "duplicate_pairs", "max_pairs_per_uid"):
            assert np.array_equal(getattr(st, f), getattr(exp[2], f))
    # N inside a UID (first pair): refused, not grouped
    r1 = pairs.r1.strings()
    r1[0] = "N" + r1[0][1:]
    from bioinformatics_b200.batches import PairBatchHost
    bad = PairBatchHost(pairs.probe_id, r1, pairs.r2, pairs.q1, pairs.q2, pairs.read_index)
    with pytest.raises(HsqError) as e:
        ctx.dedup_packed(packed_pairs_of(bad, bits), 7, 3, probes.n)
    assert e.value.status == A.HSQ_ERR_UNSUPPORTED


@pytest.mark.gpu
@pytest.mark.parametrize("mark,sums", [(False, False), (True, True), (False, True)])
def test_dedup_extend_packed_equals_ascii_call(ctx, mark, sums):
    """hsq_dedup_extend_packed (packed reads in, pooled strings out; quality bytes or host-made sums) returns what hsq_dedup_extend_batch
    returns for the same pairs, and that equals oracle dedup -> oracle extend (tests/test_pipeline.py pins the ASCII call)."""
    from oracle import oracle as O
    probes = synth.make_probes(70, seed=61)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 9000, read_len=150, ext_uid_len=8, lig_uid_len=2, seed=62, sub_rate=0.02, ins_rate=0.004,
                                                   del_rate=0.004))
    ctx.set_probes(probes)
    ref = ctx.dedup_extend_batch(pairs, 8, 2, extend_duplicates=mark, cigar_cap=64, md_cap=96)
    pp = packed_pairs_of(pairs, 2)
    if sums:
        pp = PackedPairsHost(pairs.probe_id, pp.r1, pp.r2, None, None, pairs.read_index, quality_sums(pairs.q1), quality_sums(pairs.q2))
    got = ctx.dedup_extend_packed(pp, 8, 2, extend_duplicates=mark)
    m = ref.n_extended
    assert got.n_extended == m and np.array_equal(got.is_best, ref.is_best) and np.array_equal(got.group_size, ref.group_size)
    assert np.array_equal(got.extended_pair[:m], ref.extended_pair[:m])
    for f in ("unique_pairs", "duplicate_pairs", "max_pairs_per_uid"):
        assert np.array_equal(getattr(got.stats, f), getattr(ref.stats, f))
    assert np.array_equal(got.unique_unable_to_extend, ref.unique_unable_to_extend)
    assert np.array_equal(got.duplicate_unable_to_extend, ref.duplicate_unable_to_extend)
    g, r = got.pooled, ref.extend.head(m)
    for f in ("ok", "primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen", "cigar_len", "md_len"):
        assert np.array_equal(getattr(g, f)[:2 * m], getattr(r, f)), f
    assert np.array_equal(g.pair_ok[:m], r.pair_ok)
    assert g.cigars()[:2 * m] == r.cigars() and g.mds()[:2 * m] == r.mds()
    assert g.cells == ref.extend.cells
    # and the oracle, end to end
    eb, _, _ = O.dedup_batch(pairs, 8, 2, probes.n)
    idx = np.arange(pairs.n) if mark else np.nonzero(eb)[0]
    ext = O.extend_batch(probes, pairs.subset(idx), cigar_cap=64, md_cap=96, nthreads=8)
    assert np.array_equal(got.is_best, eb) and np.array_equal(g.pos[:2 * m], ext.pos) and g.cigars()[:2 * m] == ext.cigars()


# ---- keys first: hsq_uid_keys (host) and hsq_dedup_extend_packed_keys ------------------------------------------------------------------
def test_uid_keys_host_helper():
    """hsq_uid_keys against the definition: R1[0:u] ++ R2[0:v], 2 bit per base in the order A C G T, first base most significant."""
    from bioinformatics_b200._lib import uid_keys
    rng = np.random.default_rng(3)
    r1 = [H.random_seq(rng, int(rng.integers(12, 40))) for _ in range(500)]
    r2 = [H.random_seq(rng, int(rng.integers(12, 40))) for _ in range(500)]
    order = {"A": 0, "C": 1, "G": 2, "T": 3}
    for bits in (2, 4):
        p1, p2 = pack_reads(A.SeqBatch.from_strings(r1), bits), pack_reads(A.SeqBatch.from_strings(r2), bits)
        for u, v in ((10, 0), (7, 5), (0, 12), (0, 0)):
            want = [int("".join(str(order[c]) for c in (a[:u] + b[:v])) or "0", 4) for a, b in zip(r1, r2)]
            assert uid_keys(p1, p2, u, v).tolist() == want
    bad = list(r1)
    bad[17] = bad[17][:3] + "N" + bad[17][4:]
    for bits in (2, 4):
        pb = pack_reads(A.SeqBatch.from_strings(bad), bits)
        with pytest.raises(HsqError) as e:
            uid_keys(pb, pack_reads(A.SeqBatch.from_strings(r2), bits), 10, 0)
        assert e.value.status == A.HSQ_ERR_UNSUPPORTED
        assert len(uid_keys(pb, pack_reads(A.SeqBatch.from_strings(r2), bits), 3, 0)) == 500  # the N lies behind a 3-base UID
    with pytest.raises(HsqError) as e:
        uid_keys(pack_reads(A.SeqBatch.from_strings(r1), 2), pack_reads(A.SeqBatch.from_strings(r2), 2), 41, 0)
    assert e.value.status in (A.HSQ_ERR_UNSUPPORTED, A.HSQ_ERR_OUT_OF_RANGE)


@pytest.mark.gpu
@pytest.mark.parametrize("chunks,mark", [(0, False), (3, False), (7, False), (3, True)])
def test_dedup_extend_packed_keys_equals_the_sequential_call(ctx, chunks, mark):
    """The keys-first probe task (reduction from 20 bytes per pair, reads in chunks under the kernels) against hsq_dedup_extend_packed and the
    oracle: flags, group sizes, counters, slot numbering, every pooled column and string."""
    from bioinformatics_b200._lib import uid_keys
    from bioinformatics_b200.batches import DedupExtendResult
    from oracle import oracle as O
    probes = synth.make_probes(120, seed=51)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 9001, read_len=150, ext_uid_len=10, seed=52, sub_rate=0.02, ins_rate=0.004, del_rate=0.004,
                                                   mean_family=3.0))
    ctx.configure(extend_chunks=chunks)
    ctx.set_probes(probes)
    pk = PackedPairsHost(pairs.probe_id, pack_reads(pairs.r1, 2), pack_reads(pairs.r2, 2), None, None, pairs.read_index, quality_sums(pairs.q1),
                         quality_sums(pairs.q2))
    keys = uid_keys(pk.r1, pk.r2, 10, 0)
    a = ctx.dedup_extend_packed(pk, 10, 0, extend_duplicates=mark)
    b = ctx.dedup_extend_packed_keys(pk, keys, 10, 0, extend_duplicates=mark)
    assert a.n_extended == b.n_extended and a.n_extended > 0
    m = a.n_extended
    for f in ("is_best", "group_size", "unique_unable_to_extend", "duplicate_unable_to_extend"):
        assert np.array_equal(getattr(a, f), getattr(b, f)), f
    assert np.array_equal(a.extended_pair[:m], b.extended_pair[:m])
    for f in ("unique_pairs", "duplicate_pairs", "max_pairs_per_uid"):
        assert np.array_equal(getattr(a.stats, f), getattr(b.stats, f)), f
    for f in ("primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen", "cigar_len", "md_len", "ok"):
        assert np.array_equal(getattr(a.pooled, f)[:2 * m], getattr(b.pooled, f)[:2 * m]), f
    assert np.array_equal(a.pooled.pair_ok[:m], b.pooled.pair_ok[:m])
    assert a.pooled.cigars()[:2 * m] == b.pooled.cigars()[:2 * m] and a.pooled.mds()[:2 * m] == b.pooled.mds()[:2 * m]
    assert a.pooled.cells == b.pooled.cells and a.pooled.pool_bytes == b.pooled.pool_bytes
    # and against the oracle: the reduction, then the extension of the kept pairs
    best, gs, st = O.dedup_batch(pairs, 10, 0, probes.n)
    assert np.array_equal(b.is_best, best) and np.array_equal(b.group_size, gs)
    kept = np.arange(pairs.n) if mark else np.nonzero(best)[0]
    assert np.array_equal(b.extended_pair[:m], kept)
    exp = O.extend_batch(probes, pairs.subset(kept), cigar_cap=64, md_cap=96, nthreads=8)
    assert np.array_equal(b.pooled.pos[:2 * m], exp.pos) and b.pooled.cigars()[:2 * m] == exp.cigars() and b.pooled.mds()[:2 * m] == exp.mds()
    ctx.configure()


@pytest.mark.gpu
def test_dedup_extend_packed_keys_refusals_and_empty_batch(ctx):
    """Missing key inputs are refused before anything runs; an empty batch is a no-op; pool overflow is reported with the size needed."""
    from bioinformatics_b200._lib import uid_keys
    from bioinformatics_b200.batches import DedupExtendResult
    probes = synth.make_probes(30, seed=61)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 700, read_len=150, ext_uid_len=10, seed=62))
    ctx.set_probes(probes)
    pk = PackedPairsHost(pairs.probe_id, pack_reads(pairs.r1, 2), pack_reads(pairs.r2, 2), None, None, pairs.read_index, quality_sums(pairs.q1),
                         quality_sums(pairs.q2))
    keys = uid_keys(pk.r1, pk.r2, 10, 0)
    no_sums = PackedPairsHost(pairs.probe_id, pk.r1, pk.r2, pairs.q1.data, pairs.q2.data, pairs.read_index)
    with pytest.raises(HsqError) as e:
        ctx.dedup_extend_packed_keys(no_sums, keys, 10, 0)
    assert e.value.status == A.HSQ_ERR_INVALID_ARGUMENT
    with pytest.raises(HsqError) as e:  # more UID bases than fit a key
        ctx.dedup_extend_packed_keys(pk, keys, 20, 20)
    assert e.value.status == A.HSQ_ERR_UNSUPPORTED
    with pytest.raises(HsqError) as e:  # a pool that cannot hold the strings
        ctx.dedup_extend_packed_keys(pk, keys, 10, 0, pool_cap=64)
    assert e.value.status == A.HSQ_ERR_OVERFLOW
    ok = ctx.dedup_extend_packed_keys(pk, keys, 10, 0)  # and the context is usable afterwards
    ref = ctx.dedup_extend_packed(pk, 10, 0)
    assert ok.n_extended == ref.n_extended and np.array_equal(ok.pooled.pos[:2 * ok.n_extended], ref.pooled.pos[:2 * ref.n_extended])
    empty = PackedPairsHost(np.zeros(0, dtype=np.int32), pack_reads([], 2), pack_reads([], 2), None, None, np.zeros(0, dtype=np.int32),
                            np.zeros(0, dtype=np.int16), np.zeros(0, dtype=np.int16))
    out = ctx.dedup_extend_packed_keys(empty, np.zeros(0, dtype=np.uint64), 10, 0)
    assert out.n_extended == 0 and out.pooled.pool_bytes == 0
