"""hsq_index_bam: record index of the level-0 BGZF BAMs the reference writes between its stages
(HSQ/process/FastqAndBamFileMerger.java:131-133,386-396; read back at ...Identification.java:439-474).
CPU part: the oracle's sequential parser against the test writer (two independent statements of the BAM layout).
GPU part: the library against the oracle, including records that span BGZF blocks, several stored blocks per BGZF block, a
record whose qualities look like a record header (exact walk), and the refusals."""
import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from bioinformatics_b200._lib import HsqError
from oracle import oracle as O
from tests import bamio, helpers as H


@pytest.fixture(scope="module")
def ctx():
    from bioinformatics_b200._lib import Context
    with Context(0) as c:
        yield c

REFS = [("chr1", 5_000_000), ("chr2", 3_000_000), ("chrUn_gl000220", 161_802)]
TEXT = b"@HD\tVN:1.4\tSO:unsorted\n@SQ\tSN:chr1\tLN:5000000\n@SQ\tSN:chr2\tLN:3000000\n@SQ\tSN:chrUn_gl000220\tLN:161802\n"


def make_records(n, seed, long_every=0):
    rng = np.random.default_rng(seed)
    recs = []
    for k in range(n):
        L = int(rng.integers(0, 160)) if not (long_every and k % long_every == 0) else int(rng.integers(70_000, 90_000))
        mapped = rng.random() < 0.9 and L > 0
        cigar = []
        if mapped:
            left = L
            while left > 0:
                op = "MIDS=X"[int(rng.integers(0, 6))]
                ln = int(rng.integers(1, left + 1)) if op != "D" else int(rng.integers(1, 20))
                cigar.append((op, ln))
                if op != "D":
                    left -= ln
        tags = b"NMC" + bytes([int(rng.integers(0, 9))]) + (b"XAZ" + ("chr2,+%d,50M,1;" % rng.integers(1, 10**6)).encode() + b"\0" if k % 5 == 0 else b"")
        recs.append(dict(name="%d" % k if k % 3 else "M01:%d:000-A:1:%d" % (k, rng.integers(1, 10**5)), flag=int(rng.integers(0, 4096)),
                         ref_id=int(rng.integers(0, len(REFS))) if mapped else -1, pos=int(rng.integers(0, 10**6)) if mapped else -1,
                         mapq=int(rng.integers(0, 61)), cigar=cigar, seq=H.random_seq(rng, L, "ACGTN"), qual=rng.integers(0, 42, L).astype(np.uint8).tobytes(),
                         next_ref_id=int(rng.integers(-1, len(REFS))), next_pos=int(rng.integers(-1, 10**6)), tlen=int(rng.integers(-500, 500)), tags=tags))
    return recs


def check_against_writer(idx, stream, offsets, recs):
    assert idx["n"] == len(recs) and idx["stream"] == stream
    assert idx["record_off"].tolist() == offsets
    for k in (0, len(recs) // 2, len(recs) - 1):
        r = recs[k]
        assert stream[idx["qname_off"][k]:idx["qname_off"][k] + idx["qname_len"][k]].decode() == r["name"]
        assert idx["seq_len"][k] == len(r["seq"]) and idx["cigar_len"][k] == len(r["cigar"])
        assert stream[idx["qual_off"][k]:idx["qual_off"][k] + idx["qual_len"][k]] == r["qual"]
        assert stream[idx["tags_off"][k]:idx["tags_off"][k] + idx["tags_len"][k]] == r["tags"]
        assert idx["pos"][k] == r["pos"] + 1 and idx["flag"][k] == r["flag"] and idx["refid"][k] == r["ref_id"]
        assert idx["ref_len"][k] == sum(n for op, n in r["cigar"] if op in "MDN=X")


def test_oracle_parser_equals_the_writer():
    recs = make_records(3000, 1)
    stream, offsets = bamio.bam_stream(TEXT, REFS, recs)
    for data in (bamio.bgzf(stream), bamio.bgzf(stream, level=6), bamio.bgzf(stream, block=4000, pieces=3)):
        idx = O.bam_index(data)
        check_against_writer(idx, stream, offsets, recs)
        assert idx["n_ref"] == 3 and idx["text_bytes"] == len(TEXT) and idx["first_record"] == offsets[0]


def assert_index_equal(got, exp):
    assert (got.n, got.n_ref, got.text_bytes, got.first_record, got.payload_bytes) == (exp["n"], exp["n_ref"], exp["text_bytes"], exp["first_record"],
                                                                                       len(exp["stream"]))
    cols = got.columns()
    for k, v in cols.items():
        bad = np.nonzero(v != exp[k])[0]
        assert len(bad) == 0, "column %s differs at %s: got %s expected %s" % (k, bad[:5], v[bad[:5]], exp[k][bad[:5]])


@pytest.mark.gpu
@pytest.mark.parametrize("n,block,pieces,long_every", [(20000, 65280, None, 0), (5000, 65280, 2, 0), (3000, 777, 1, 0), (400, 65280, None, 50), (0, 65280, None, 0)])
def test_index_bam_equals_oracle(ctx, n, block, pieces, long_every):
    recs = make_records(n, 7 + n, long_every)
    stream, _ = bamio.bam_stream(TEXT, REFS, recs)
    data = bamio.bgzf(stream, block=block, pieces=pieces)
    got = ctx.index_bam(data, want_payload=True)
    exp = O.bam_index(data)
    assert_index_equal(got, exp)
    assert got.exact_walk == 0
    assert got.payload[:got.payload_bytes].tobytes() == exp["stream"]


@pytest.mark.gpu
def test_index_bam_record_header_inside_a_record(ctx):
    """qualities that spell a plausible record header: the parallel test sees one record too many, the exact walk sorts it out"""
    import struct
    recs = make_records(500, 3)
    fake = struct.pack("<iiiBBHHHiiii", 33, 0, 5, 1, 0, 0, 0, 0, 0, -1, -1, 0) + b"\0"
    victim = recs[250]
    victim["seq"] = "A" * 120
    victim["cigar"] = [("M", 120)]
    victim["qual"] = bytes([30] * 20) + fake + bytes([30] * (100 - len(fake)))
    stream, _ = bamio.bam_stream(TEXT, REFS, recs)
    data = bamio.bgzf(stream)
    got = ctx.index_bam(data)
    assert got.exact_walk == 1
    assert_index_equal(got, O.bam_index(data))


@pytest.mark.gpu
def test_index_bam_refusals(ctx):
    recs = make_records(200, 5)
    stream, _ = bamio.bam_stream(TEXT, REFS, recs)
    with pytest.raises(HsqError) as e:  # a deflate-compressed block: no inflate in this library
        ctx.index_bam(bamio.bgzf(stream, level=6))
    assert e.value.status == A.HSQ_ERR_UNSUPPORTED
    data = bamio.bgzf(stream)
    with pytest.raises(HsqError) as e:
        ctx.index_bam(data, cap=10)
    assert e.value.status == A.HSQ_ERR_OVERFLOW
    with pytest.raises(HsqError) as e:  # the last record cut short
        ctx.index_bam(bamio.bgzf(stream[:-7]))
    assert e.value.status == A.HSQ_ERR_MALFORMED
    with pytest.raises(HsqError) as e:
        ctx.index_bam(data[:1000])
    assert e.value.status == A.HSQ_ERR_MALFORMED
    with pytest.raises(HsqError) as e:
        ctx.index_bam(bamio.bgzf(b"SAM\1" + stream[4:]))
    assert e.value.status == A.HSQ_ERR_MALFORMED
