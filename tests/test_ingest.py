"""FASTQ ingest and the FASTQ <-> SAM name join (HSQ/process/FastqAndBamFileMerger.java:65-94,107-158,213-345,
386-396; HSQ/utils/FastqSorter.java:90-124,291-339; HSQ/utils/BamSorter.java:322-361;
COMMON/utils/IlluminaFastQReadNameUtil.java:40-58).

The reference's tests hold nothing for FastqAndBamFileMerger (parity unpinned by the reference).  The oracle is the sequential
two-cursor loop restated line by line; it is pinned here on hand traces of that loop and cross-checked against a second,
dictionary-based statement of the join; the CUDA path is then compared with the oracle array by array."""
import re
from collections import defaultdict

import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from bioinformatics_b200 import synth
from bioinformatics_b200.batches import NameJoinHost
from oracle import oracle as O


def join_of(case, **kw):
    f1, s1, f2, s2 = case["trim"]
    args = dict(fq1_seq_len=case["fq_seq_len"], fq2_seq_len=case["fq_seq_len"], r1_first=f1, r1_stop=s1, r2_first=f2, r2_stop=s2,
                three_prime=case["three_prime"])
    args.update(kw)
    return NameJoinHost(case["sam_names"], case["sam_flags"], case["sam_seq_len"], case["fq1_headers"], case["fq2_headers"], **args)


def second_statement(join, by_index=False):
    """The join without cursors: FASTQ positions of every unique id in a dictionary, a per-id count of consumed positions."""
    uid = lambda h: re.split("[ /]", h, maxsplit=1)[0]
    h1, h2, names = join.fq1_headers.strings(), join.fq2_headers.strings(), join.sam_names.strings()
    nf = min(len(h1), len(h2))
    flags = join.sam_flags
    if by_index:
        o1, o2 = list(range(len(h1))), list(range(len(h2)))
        sam_order = sorted(range(len(names)), key=lambda r: (int(names[r]), 0 if flags[r] & 0x40 else 1))
        key_of = lambda r: int(names[r])
        positions = {k: [k] for k in range(nf)}
    else:
        o1 = sorted(range(len(h1)), key=lambda k: (h1[k] + " _ %d" % k).encode())
        o2 = sorted(range(len(h2)), key=lambda k: (h2[k] + " _ %d" % k).encode())
        sam_order = sorted(range(len(names)), key=lambda r: (names[r].encode(), 0 if flags[r] & 0x40 else 1))
        key_of = lambda r: uid(names[r])
        positions = defaultdict(list)
        for q in range(nf):
            positions[uid(h1[o1[q]])].append(q)
    used = defaultdict(int)
    status = np.full(len(names), 2, dtype=np.uint8)
    sam_record, fastq_record, name_len = [], [], []
    pairs = 0
    for r in sam_order:
        if flags[r] & 0x100:
            status[r] = 0
            continue
        k = key_of(r)
        pos = positions.get(k, [])
        c = used[k]
        first = bool(flags[r] & 0x40)
        if not first:
            used[k] += 1
        if c < len(pos):
            q = pos[c]
            status[r] = 1
            sam_record.append(r)
            fastq_record.append(o1[q] if first else o2[q])
            name_len.append(len(uid(h1[o1[q]])))
            pairs += 0 if first else 1
    return sam_record, fastq_record, name_len, status, pairs


def assert_same(got, exp):
    assert got.n_emitted == exp.n_emitted and got.total_matching_pairs == exp.total_matching_pairs
    for a, b in zip(got.emitted(), exp.emitted()):
        assert (a == b).all()
    assert (got.status == exp.status).all()


KAT_BY_NAME = dict(
    sam_names=["r3", "r1", "r1", "r2", "r2", "zz", "r1"], sam_flags=[0x81, 0x41, 0x81, 0x141, 0x41, 0x41, 0x81], sam_seq_len=[5] * 7,
    fq1=["r2 1:N", "r1 1:N", "r3 1:N"], fq2=["r2 2:N", "r1 2:N", "r3 2:N"],
    # hand trace of MergedSamIterator: FASTQ sorted = r1 (file ordinal 1), r2 (0), r3 (2); SAM sorted = r1/1 [1], r1/2 [2], r1/2 [6],
    # r2 not primary [3], r2/1 [4], r3/2 [0], zz [5].  [1] takes FASTQ one of r1; [2] takes FASTQ two of r1 and moves the FASTQ
    # cursor to r2, so [6] compares r1 < r2 and is skipped; [3] is dropped as not primary; [4] takes r2; the cursor moves to r3 for
    # [0], which is a read two: matched, cursor past the end, loop over, [5] is never looked at.
    sam_record=[1, 2, 4, 0], fastq_record=[1, 1, 0, 2], name_len=[2, 2, 2, 2], status=[1, 1, 1, 0, 1, 2, 2], pairs=2)

KAT_BY_INDEX = dict(
    sam_names=["2", "0", "0", "10", "-1", "1"], sam_flags=[0x81, 0x41, 0x81, 0x41, 0x41, 0x41], sam_seq_len=[5] * 6,
    fq1=["a:1 1:N", "b:22/1", "c"], fq2=["a:1 2:N", "b:22/2", "c"],
    # sorted by Integer.parseInt: -1 [4], 0/1 [1], 0/2 [2], 1 [5], 2 [0], 10 [3]; FASTQ position k has key k.  -1 < 0: skipped; [1] and
    # [2] take position 0; [5] takes position 1 (FASTQ one); [0] is a read two at position 2: matched, FASTQ exhausted; [3] unreached.
    sam_record=[1, 2, 5, 0], fastq_record=[0, 0, 1, 2], name_len=[3, 3, 4, 1], status=[1, 1, 1, 2, 2, 1], pairs=2)


def kat_join(k, by_index):
    return NameJoinHost(k["sam_names"], k["sam_flags"], k["sam_seq_len"], k["fq1"], k["fq2"], [9, 9, 9], [9, 9, 9], by_index=by_index, check_trim=False)


def check_kat(res, k):
    s, f, n = res.emitted()
    assert list(s) == k["sam_record"] and list(f) == k["fastq_record"] and list(n) == k["name_len"]
    assert list(res.status) == k["status"] and res.total_matching_pairs == k["pairs"]


def error_cases():
    """(join, expected status) for every exception of the path."""
    base = dict(sam_names=["r1", "r1", "r2", "r2"], sam_flags=[0x41, 0x81, 0x41, 0x81], sam_seq_len=[8, 8, 8, 8])
    fq1, fq2 = ["r1 1", "r2 1"], ["r1 2", "r2 2"]
    cases = []
    # FASTQ two holds another read at the second position: IllegalStateException :270-273
    cases.append((NameJoinHost(fq1_headers=fq1, fq2_headers=["r1 2", "rX 2"], fq1_seq_len=[10, 10], fq2_seq_len=[10, 10], check_trim=False, **base),
                  A.HSQ_ERR_NAME_MISMATCH))
    # ... but a mismatch behind the last position the loop fetches is never seen
    cases.append((NameJoinHost(["r1", "r1"], [0x41, 0x81], [8, 8], fq1, ["r1 2", "rX 2"], [10, 10], [10, 10], check_trim=False), A.HSQ_OK))
    # SAM read longer than the trimmed FASTQ read: UnableToMergeFastqAndBamFilesException :366-371 (10 - 3 = 7 < 8)
    cases.append((NameJoinHost(fq1_headers=fq1, fq2_headers=fq2, fq1_seq_len=[10, 10], fq2_seq_len=[10, 10], r1_first=3, **base), A.HSQ_ERR_NOT_TRIMMED))
    # same lengths with the check switched off by --trimmingSkipped
    cases.append((NameJoinHost(fq1_headers=fq1, fq2_headers=fq2, fq1_seq_len=[10, 10], fq2_seq_len=[10, 10], r1_first=3, trimming_skipped=True, **base),
                  A.HSQ_OK))
    # three prime trimming: kept [0, 6] = 7 bases < 8
    cases.append((NameJoinHost(fq1_headers=fq1, fq2_headers=fq2, fq1_seq_len=[10, 10], fq2_seq_len=[10, 10], three_prime=True, r1_stop=9, r2_stop=6, **base),
                  A.HSQ_ERR_NOT_TRIMMED))
    # qualities longer than the trimmed read although the bases fit
    cases.append((NameJoinHost(fq1_headers=fq1, fq2_headers=fq2, fq1_seq_len=[10, 10], fq2_seq_len=[10, 10], sam_qual_len=[8, 8, 11, 8], **base),
                  A.HSQ_ERR_NOT_TRIMMED))
    # first base to keep outside the FASTQ read: FastqReadTrimmer.trim throws (FastqReadTrimmer.java:310-312)
    cases.append((NameJoinHost(fq1_headers=fq1, fq2_headers=fq2, fq1_seq_len=[10, 10], fq2_seq_len=[10, 10], r2_first=10, **base), A.HSQ_ERR_OUT_OF_RANGE))
    # the trim failure of r1 comes before the name mismatch at r2 ...
    cases.append((NameJoinHost(fq1_headers=fq1, fq2_headers=["r1 2", "rX 2"], fq1_seq_len=[10, 10], fq2_seq_len=[10, 10], r1_first=3, **base),
                  A.HSQ_ERR_NOT_TRIMMED))
    # ... and the other way round
    cases.append((NameJoinHost(fq1_headers=fq1, fq2_headers=["rX 2", "r2 2"], fq1_seq_len=[10, 10], fq2_seq_len=[10, 8], r2_first=1, **base),
                  A.HSQ_ERR_NAME_MISMATCH))
    # nothing matches: IllegalStateException :143-148 (only read ones match -> getTotalMatches() is still 0)
    cases.append((NameJoinHost(["q1", "q2"], [0x41, 0x81], [8, 8], fq1, fq2, [10, 10], [10, 10], check_trim=False), A.HSQ_ERR_NO_MATCHES))
    cases.append((NameJoinHost(["r1", "r2"], [0x41, 0x41], [8, 8], fq1, fq2, [10, 10], [10, 10], check_trim=False), A.HSQ_ERR_NO_MATCHES))
    cases.append((NameJoinHost(["r1"], [0x181], [8], fq1, fq2, [10, 10], [10, 10], check_trim=False), A.HSQ_ERR_NO_MATCHES))
    cases.append((NameJoinHost(["r1"], [0x81], [8], [], [], [], [], check_trim=False), A.HSQ_ERR_NO_MATCHES))
    # by index with a name that is not an integer: NumberFormatException
    cases.append((NameJoinHost(["0", "x1"], [0x81, 0x81], [8, 8], fq1, fq2, [10, 10], [10, 10], by_index=True, check_trim=False), A.HSQ_ERR_UNSUPPORTED))
    cases.append((NameJoinHost(["0", "99999999999"], [0x81, 0x81], [8, 8], fq1, fq2, [10, 10], [10, 10], by_index=True, check_trim=False),
                  A.HSQ_ERR_UNSUPPORTED))
    # unequal FASTQ files: the loop stops with the shorter one
    cases.append((NameJoinHost(fq1_headers=fq1, fq2_headers=["r1 2"], fq1_seq_len=[10, 10], fq2_seq_len=[10], check_trim=False, **base), A.HSQ_OK))
    return cases


def run_status(fn, join):
    try:
        return A.HSQ_OK, fn(join)
    except Exception as e:  # OracleError / HsqError both carry .status
        return e.status, None


FASTQ_BAD = [
    (b"@r1\nACGT\n+\nIII\n", A.HSQ_ERR_MALFORMED),           # lengths differ
    (b"r1\nACGT\n+\nIIII\n", A.HSQ_ERR_MALFORMED),           # no '@'
    (b"@r1\nACGT\n-\nIIII\n", A.HSQ_ERR_MALFORMED),          # no '+'
    (b"@r1\nACGT\n+\nIIII\n@r2\nAC\n+\n", A.HSQ_ERR_MALFORMED),  # too short
    (b"@r1\nACGT\n+\nIIII\n\n", A.HSQ_ERR_MALFORMED),        # blank line at the end
    (b"@r1\n\n+\n\n", A.HSQ_ERR_MALFORMED),                  # empty read
    (b"@r1\nACGT\n+\nIIII\n   \nAC\n+\nII\n", A.HSQ_ERR_MALFORMED),  # blank header
]
FASTQ_GOOD = [b"", b"@r1 x/1\nACGT\n+\nIIII\n", b"@r1\nACGT\n+r1\nIIII", b"@r1\r\nACGT\r\n+\r\nIIII\r\n@r2\r\nA\r\n+\r\n#\r\n", b"@a\nA\n+\n@\n@b\n@\n+\n+\n",
              b"@r1\r\nACGT\r\n+\r\nIIII\r"]


def split_statement(text):
    """FASTQ records by str.splitlines-style reading (second statement of the index for well-formed text)."""
    lines, pos = [], 0
    while pos < len(text):
        e = text.find(b"\n", pos)
        e = len(text) if e < 0 else e
        stop = e - 1 if e > pos and text[e - 1:e] == b"\r" else e
        lines.append((pos, stop - pos))
        pos = e + 1
    recs = [lines[k:k + 4] for k in range(0, len(lines), 4)]
    return dict(header_off=[r[0][0] + 1 for r in recs], header_len=[r[0][1] - 1 for r in recs], seq_off=[r[1][0] for r in recs],
                seq_len=[r[1][1] for r in recs], qual_off=[r[3][0] for r in recs], qual_len=[r[3][1] for r in recs])


def sam_split_statement(text):
    """SAM columns by bytes.split: the second statement of the field index for well-formed text."""
    cols = defaultdict(list)
    pos, header = 0, 0
    ref_ops = set("MDN=X")
    for raw in text.split(b"\n"):
        line = raw[:-1] if raw.endswith(b"\r") else raw
        start = pos
        pos += len(raw) + 1
        if not line:
            continue
        if line.startswith(b"@"):
            header += 1
            continue
        f = line.split(b"\t")
        offs = [start]
        for x in f[:-1]:
            offs.append(offs[-1] + len(x) + 1)
        for name, c in (("qname", 0), ("rname", 2), ("cigar", 5), ("seq", 9), ("qual", 10)):
            cols[name + "_off"].append(offs[c])
            cols[name + "_len"].append(len(f[c]))
        cols["tags_off"].append(offs[11] if len(f) > 11 else start + len(line))
        cols["tags_len"].append(start + len(line) - cols["tags_off"][-1])
        xa = [(offs[i] + 5, len(f[i]) - 5) for i in range(11, len(f)) if f[i].startswith(b"XA:Z:")]
        xo, xl = xa[0] if xa and xa[0][1] > 0 else (0, 0)
        cols["xa_off"].append(xo)
        cols["xa_len"].append(xl)
        for name, c in (("flag", 1), ("pos", 3), ("mapq", 4), ("pnext", 7), ("tlen", 8)):
            cols[name].append(int(f[c]))
        cig = f[5].decode()
        cols["ref_len"].append(0 if cig == "*" else sum(int(n) for n, op in re.findall(r"(\d+)([A-Z=])", cig) if op in ref_ops))
    return header, cols


SAM_SMALL = (b"@HD\tVN:1.0\n@SQ\tSN:chr1\tLN:100\nr1\t99\tchr1\t10\t60\t5M1I3M2D4M\t=\t50\t60\tACGTACGTACGTA\tIIIIIIIIIIIII\tNM:i:1\tXA:Z:chr2,+5,13M,1;\tMD:Z:3\n"
             b"r2\t4\t*\t0\t0\t*\t*\t0\t0\t*\t*\nr3\t-1\tc\t+7\t0\t3=1X2N\t=\t1\t-9\tAC\t##\tXA:Z:\tXA:Z:q\nr4\t0\tc\t7\t0\t2H3M\t=\t1\t0\tAC\t##\t")
SAM_BAD = [b"r1\t99\tchr1\t10\t60\t5M\t=\t50\t60\tACGTA\n",                        # ten columns
           b"r1\t99\tchr1\t1x\t60\t5M\t=\t50\t60\tACGTA\tIIIII\n",                  # POS is not a number
           b"r1\t99\tchr1\t10\t60\t5M3\t=\t50\t60\tACGTA\tIIIII\n",                 # CIGAR ends inside a count
           b"r1\t99\tchr1\t10\t60\t5Q\t=\t50\t60\tACGTA\tIIIII\n",                  # unknown operator
           b"r1\t99\tchr1\t10\t60\tM\t=\t50\t60\tACGTA\tIIIII\n",                   # operator without count
           b"r1\t99\tchr1\t10\t60\t5M\t=\t50\t60\tACGTA\tIIIII\n@CO\tlate header\n",  # header line after an alignment line
           b"@HD\tVN:1.0\n\nr1\t99\tchr1\t10\t60\t5M\t=\t50\t60\tACGTA\tIIIII\n",      # empty line
           b"r1\t99999999999\tchr1\t10\t60\t5M\t=\t50\t60\tACGTA\tIIIII\n"]          # FLAG outside int


def check_sam_index(ix, text):
    header, cols = sam_split_statement(text)
    assert ix.n_header_lines == header and ix.n == len(cols["flag"])
    for c in ix.columns():
        assert list(getattr(ix, c)) == cols[c], c


def test_oracle_sam_index():
    check_sam_index(O.index_sam(SAM_SMALL), SAM_SMALL)
    ix = O.index_sam(SAM_SMALL)
    assert list(ix.ref_len) == [14, 0, 6, 3] and list(ix.flag) == [99, 4, -1, 0] and list(ix.xa_len) == [14, 0, 0, 0] and list(ix.tags_len) == [33, 0, 12, 0]
    for text in SAM_BAD:
        with pytest.raises(O.OracleError) as e:
            O.index_sam(text)
        assert e.value.status == A.HSQ_ERR_MALFORMED
    assert O.index_sam(b"").n == 0 and O.index_sam(b"@HD\tVN:1.0\n").n_header_lines == 1
    case = synth.make_ingest_case(800, read_len=40, seed=4, trim=(3, 30, 0, 30))
    for kw in ({}, dict(crlf=True), dict(final_newline=False)):
        text = synth.sam_text_of(case, **kw)
        check_sam_index(O.index_sam(text), text)


SAM_RECORDS_KAT = (b"@HD\tVN:1.0\n"
                   b"7\t99\tchr1\t10\t60\t5M1I3M2D4M\t=\t50\t60\tACGTACGTACGTA\tIIIIIIIIIIIII\tNM:i:1\t"
                   b"XA:Z:chr2,+5,13M,1;chrX,-77,3M2D3M,0;chr1,x,3M,1;chr1,+9,3M;chr2,-4,2S3M,0,;;\n"
                   b"8\t77\t*\t0\t0\t*\t*\t0\t0\tAC\t##\n"
                   b"9\t147\tchrU\t20\t0\t4M\t=\t1\t0\tACGT\t####\n")
# by hand: line 0 (flag 99: paired, first, forward) chr1 -> 0, end = 10 + (5+3+2+4) - 1 = 23; XA keeps chr2,+5,13M (stop 17), chrX,-77,3M2D3M
# (contig -1, reverse, stop 77+8-1 = 84) and chr2,-4,2S3M,0, (String.split drops the trailing empty field; stop 4+3-1 = 6); "chr1,x,..." is not a
# number, "chr1,+9,3M" has three fields.  Line 1 (flag 77) is unmapped: dropped.  Line 2 (flag 147: reverse, second) sits on a contig without
# probes: -1, its name is never parsed.
SAM_RECORDS_KAT_EXPECTED = dict(sam_line=[0, 2], contig=[0, -1], negative_strand=[0, 1], second_of_pair=[0, 1], start=[10, 20], end=[23, 23],
                                read_index=[7, 0], alt_offsets=[0, 3, 3], alt_contig=[1, -1, 1], alt_negative=[0, 1, 1], alt_start=[5, 77, 4],
                                alt_stop=[17, 84, 6])
SAM_RECORDS_BAD = [(b"7\t0\tchr1\t10\t60\t5M\t=\t50\t60\tACGTA\tIIIII\n", A.HSQ_ERR_UNSUPPORTED),            # not paired
                   (b"r7\t99\tchr1\t10\t60\t5M\t=\t50\t60\tACGTA\tIIIII\n", A.HSQ_ERR_BAD_CHARACTER),        # name is not a number
                   (b"7\t99\tchr1\t10\t60\t5M\t=\t50\t60\tACGTA\tIIIII\tXA:Z:chr1,+5,3Q,0;\n", A.HSQ_ERR_MALFORMED)]  # CIGAR of a hit


def check_sam_records(r, exp):
    n, m = len(exp["contig"]), len(exp["alt_contig"])
    assert r.n == n and r.n_alt == m
    for c in r.COLUMNS:
        assert list(getattr(r, c)[:n]) == list(exp[c]), c
    assert list(r.alt_offsets[:n + 1]) == list(exp["alt_offsets"])
    for c in r.ALT_COLUMNS:
        assert list(getattr(r, c)[:m]) == list(exp[c]), c


def sam_records_statement(text, contig_names):
    """Second statement with str.split and re (String.split semantics: trailing empty strings dropped)."""
    def jsplit(s, sep):
        f = s.split(sep)
        while f and f[-1] == "":
            f.pop()
        return f
    exp = defaultdict(list)
    exp["alt_offsets"].append(0)
    cid = {c: i for i, c in enumerate(contig_names)}
    k = -1
    for raw in text.decode().split("\n"):
        line = raw.rstrip("\r")
        if not line or line.startswith("@"):
            continue
        k += 1
        f = line.split("\t")
        fl = int(f[1])
        if fl & 0xC:
            continue
        ref = sum(int(n) for n, op in re.findall(r"(\d+)([A-Z=])", f[5]) if op in "MDN=X")
        contig = cid.get(f[2], -1)
        for name, v in (("sam_line", k), ("contig", contig), ("negative_strand", 1 if fl & 0x10 else 0), ("second_of_pair", 1 if fl & 0x80 else 0),
                        ("start", int(f[3])), ("end", int(f[3]) + ref - 1), ("read_index", int(f[0]) if contig >= 0 else 0)):
            exp[name].append(v)
        xa = [t[5:] for t in f[11:] if t.startswith("XA:Z:")]
        for entry in (jsplit(xa[0], ";") if xa else []):
            g = jsplit(entry, ",")
            if len(g) != 4 or not re.fullmatch(r"[+-]?\d+", g[1]):
                continue
            v = int(g[1])
            rl = sum(int(n) for n, op in re.findall(r"(\d+)([A-Z=])", g[2]) if op in "MDN=X")
            exp["alt_contig"].append(cid.get(g[0], -1)); exp["alt_negative"].append(1 if v < 0 else 0)
            exp["alt_start"].append(abs(v)); exp["alt_stop"].append(abs(v) + rl - 1)
        exp["alt_offsets"].append(len(exp["alt_contig"]))
    return exp


CONTIGS = ["chr%d" % c for c in range(1, 23)]


def test_oracle_sam_mapped_records():
    check_sam_records(O.sam_mapped_records(SAM_RECORDS_KAT, ["chr1", "chr2"]), SAM_RECORDS_KAT_EXPECTED)
    check_sam_records(O.sam_mapped_records(SAM_RECORDS_KAT, ["chr1", "chr2"]), sam_records_statement(SAM_RECORDS_KAT, ["chr1", "chr2"]))
    for text, want in SAM_RECORDS_BAD:
        with pytest.raises(O.OracleError) as e:
            O.sam_mapped_records(text, ["chr1"])
        assert e.value.status == want
    case = synth.make_ingest_case(1500, read_len=40, seed=6, trim=(3, 30, 0, 30))
    text = synth.sam_text_of(case, xa_rate=0.4, index_names=True)
    r = O.sam_mapped_records(text, CONTIGS)
    check_sam_records(r, sam_records_statement(text, CONTIGS))
    assert r.n_alt > 1000 and (r.alt_contig[:r.n_alt] == -1).sum() > 20


# ---------------------------------------------------------------------------------------------------------------- CPU: the oracle
def test_oracle_hand_traces():
    check_kat(O.join_names(kat_join(KAT_BY_NAME, False)), KAT_BY_NAME)
    check_kat(O.join_names(kat_join(KAT_BY_INDEX, True)), KAT_BY_INDEX)


def test_oracle_unique_id():
    """IlluminaFastQReadNameUtil.getUniqueIdForReadHeader on the two sample headers of its source (:30-33), through the join."""
    j = NameJoinHost(["MS5_15454:1:1110:12527:26507#26", "M01077:35:000000000-A3J96:1:1102:13646:7860"], [0x81, 0x81], [5, 5],
                     ["M01077:35:000000000-A3J96:1:1102:13646:7860 1:N:0:1", "MS5_15454:1:1110:12527:26507#26/1"],
                     ["M01077:35:000000000-A3J96:1:1102:13646:7860 2:N:0:1", "MS5_15454:1:1110:12527:26507#26/2"], check_trim=False)
    r = O.join_names(j)
    s, f, n = r.emitted()
    assert list(s) == [1, 0] and list(f) == [0, 1] and list(n) == [43, 31]


@pytest.mark.parametrize("old_style", [False, True])
def test_oracle_equals_second_statement(old_style):
    case = synth.make_ingest_case(3000, read_len=40, seed=11 + old_style, old_style=old_style, trim=(3, 30, 0, 30))
    j = join_of(case)
    exp = second_statement(j)
    got = O.join_names(j)
    s, f, n = got.emitted()
    assert list(s) == exp[0] and list(f) == exp[1] and list(n) == exp[2] and (got.status == exp[3]).all() and got.total_matching_pairs == exp[4]
    assert got.n_emitted > 4000 and (got.status == 0).sum() > 10 and (got.status == 2).sum() > 10


def by_index_case(case, seed=3):
    """The second configuration: the records of the first join renamed to the FASTQ ordinal they took (:319-321)."""
    first = O.join_names(join_of(case))
    s, f, _ = first.emitted()
    rng = np.random.default_rng(seed)
    names = [str(int(v)) for v in f] + ["-5", str(case["n_pairs"] + 7), str(int(f[0]))]
    flags = np.concatenate([case["sam_flags"][s], [0x41, 0x81, 0x81]]).astype(np.int32)
    order = rng.permutation(len(names))
    return NameJoinHost([names[i] for i in order], flags[order], np.full(len(names), 30, dtype=np.int32), case["fq1_headers"], case["fq2_headers"],
                        by_index=True, check_trim=False)


def test_oracle_by_index_equals_second_statement():
    case = synth.make_ingest_case(2000, read_len=40, seed=5, trim=(3, 30, 0, 30))
    j = by_index_case(case)
    exp = second_statement(j, by_index=True)
    got = O.join_names(j)
    s, f, n = got.emitted()
    assert list(s) == exp[0] and list(f) == exp[1] and list(n) == exp[2] and (got.status == exp[3]).all() and got.total_matching_pairs == exp[4]


def test_oracle_errors():
    for j, want in error_cases():
        st, _ = run_status(O.join_names, j)
        assert st == want


def test_oracle_fastq_index():
    for text in FASTQ_GOOD:
        ix = O.index_fastq(text)
        exp = split_statement(text)
        assert ix.n == len(exp["header_off"])
        for f in ix.FIELDS:
            assert list(ix.head()[f]) == exp[f], f
    for text, want in FASTQ_BAD:
        with pytest.raises(O.OracleError) as e:
            O.index_fastq(text)
        assert e.value.status == want
    case = synth.make_ingest_case(500, read_len=30, seed=2, final_newline=False)
    ix = O.index_fastq(case["fq1_text"])
    t = np.frombuffer(case["fq1_text"], dtype=np.uint8)
    assert ix.n == 500 and (ix.seq_len[:500] == 30).all()
    for k in (0, 17, 499):
        assert t[ix.header_off[k]:ix.header_off[k] + ix.header_len[k]].tobytes().decode() == case["fq1_headers"][k]
        assert (t[ix.seq_off[k]:ix.seq_off[k] + 30] == case["bases"][0][k]).all() and (t[ix.qual_off[k]:ix.qual_off[k] + 30] == case["quals"][0][k]).all()


# ---------------------------------------------------------------------------------------------------------------- GPU: parity
@pytest.fixture(scope="module")
def ctx():
    from bioinformatics_b200._lib import Context
    with Context(0) as c:
        yield c


@pytest.mark.gpu
def test_gpu_hand_traces(ctx):
    check_kat(ctx.join_names(kat_join(KAT_BY_NAME, False)), KAT_BY_NAME)
    check_kat(ctx.join_names(kat_join(KAT_BY_INDEX, True)), KAT_BY_INDEX)


@pytest.mark.gpu
@pytest.mark.parametrize("n_pairs,old_style,three_prime,seed", [(1, False, False, 1), (2, True, False, 2), (700, False, True, 3), (700, True, False, 4),
                                                                  (60000, False, False, 5), (60000, True, True, 6)])
def test_gpu_join_by_name_equals_oracle(ctx, n_pairs, old_style, three_prime, seed):
    case = synth.make_ingest_case(n_pairs, read_len=50, seed=seed, old_style=old_style, trim=(4, 40, 2, 44), three_prime=three_prime,
                                  mapped_frac=1.0 if n_pairs < 10 else 0.9)
    j = join_of(case)
    assert_same(ctx.join_names(j), O.join_names(j))


@pytest.mark.gpu
def test_gpu_join_repeated_fastq_names_and_heavy_repeats(ctx):
    """FASTQ files that hold a name more than once, SAM names with many records each, '/1' '/2' names kept by the mapper."""
    rng = np.random.default_rng(9)
    ids = ["n%02d" % k for k in range(40)]
    h = [ids[int(k)] for k in rng.integers(0, 40, size=300)]
    fq1 = [x + " 1:N:0:1" for x in h]
    fq2 = [x + " 2:N:0:1" for x in h]
    picks = rng.integers(0, 44, size=2500)
    names = [("n%02d" % k) + ("" if k % 3 else ("/1" if k % 2 else "/2")) for k in picks]
    flags = rng.choice([0x41, 0x81, 0x141, 0x181, 0x841, 0x1], size=2500).astype(np.int32)
    j = NameJoinHost(names, flags, np.full(2500, 9, dtype=np.int32), fq1, fq2, check_trim=False)
    exp = O.join_names(j)
    assert exp.n_emitted > 200
    assert_same(ctx.join_names(j), exp)


@pytest.mark.gpu
def test_gpu_join_by_index_equals_oracle(ctx):
    for n, seed in ((300, 1), (40000, 2)):
        case = synth.make_ingest_case(n, read_len=40, seed=seed, trim=(3, 30, 0, 30))
        j = by_index_case(case)
        assert_same(ctx.join_names(j), O.join_names(j))


@pytest.mark.gpu
def test_gpu_join_errors(ctx):
    from bioinformatics_b200._lib import HsqError
    for j, want in error_cases():
        st, got = run_status(ctx.join_names, j)
        assert st == want, ctx.lib.hsq_last_error(ctx.h)
        if want == A.HSQ_OK:
            assert_same(got, O.join_names(j))
    # ids that do not ascend along the sorted FASTQ side ("a/1" sorts after "a-b/1", the id "a" before "a-b"): refused
    j = NameJoinHost(["a", "a-b"], [0x81, 0x81], [5, 5], ["a/1", "a-b/1"], ["a/2", "a-b/2"], check_trim=False)
    with pytest.raises(HsqError) as e:
        ctx.join_names(j)
    assert e.value.status == A.HSQ_ERR_UNSUPPORTED


@pytest.mark.gpu
def test_gpu_fastq_index_equals_oracle(ctx):
    from bioinformatics_b200._lib import HsqError
    for text in FASTQ_GOOD:
        got, exp = ctx.index_fastq(text), O.index_fastq(text)
        assert got.n == exp.n
        for f in got.FIELDS:
            assert (got.head()[f] == exp.head()[f]).all(), f
    for text, want in FASTQ_BAD:
        with pytest.raises(HsqError) as e:
            ctx.index_fastq(text)
        assert e.value.status == want
    with pytest.raises(HsqError) as e:
        ctx.index_fastq(b"@r1\rACGT\n+\nIIII\n")
    assert e.value.status == A.HSQ_ERR_UNSUPPORTED
    for n, kw in ((1, {}), (4097, dict(crlf=True)), (30000, dict(final_newline=False)), (30000, dict(old_style=True))):
        case = synth.make_ingest_case(n, read_len=75, seed=n % 7, **kw)
        for text in (case["fq1_text"], case["fq2_text"]):
            got, exp = ctx.index_fastq(text), O.index_fastq(text)
            assert got.n == exp.n == n
            for f in got.FIELDS:
                assert (got.head()[f] == exp.head()[f]).all(), f
    with pytest.raises(HsqError) as e:
        ctx.index_fastq(case["fq1_text"], cap=10)
    assert e.value.status == A.HSQ_ERR_OVERFLOW


@pytest.mark.gpu
def test_gpu_sam_index_equals_oracle(ctx):
    from bioinformatics_b200._lib import HsqError
    check_sam_index(ctx.index_sam(SAM_SMALL), SAM_SMALL)
    for text in SAM_BAD:
        with pytest.raises(HsqError) as e:
            ctx.index_sam(text)
        assert e.value.status == A.HSQ_ERR_MALFORMED
    assert ctx.index_sam(b"").n == 0 and ctx.index_sam(b"@HD\tVN:1.0\n").n_header_lines == 1
    for n, kw in ((800, {}), (800, dict(crlf=True)), (25000, dict(final_newline=False, xa_rate=0.3))):
        case = synth.make_ingest_case(n, read_len=70, seed=n % 5, trim=(3, 60, 0, 60))
        text = synth.sam_text_of(case, **kw)
        got, exp = ctx.index_sam(text), O.index_sam(text)
        assert got.n == exp.n == len(case["sam_names"]) and got.n_header_lines == exp.n_header_lines == 24
        for c in got.columns():
            assert (getattr(got, c) == getattr(exp, c)).all(), c
    with pytest.raises(HsqError) as e:
        ctx.index_sam(text, cap=10)
    assert e.value.status == A.HSQ_ERR_OVERFLOW


@pytest.mark.gpu
def test_gpu_sam_mapped_records_equals_oracle(ctx):
    from bioinformatics_b200._lib import HsqError
    check_sam_records(ctx.sam_mapped_records(SAM_RECORDS_KAT, ["chr1", "chr2"]), SAM_RECORDS_KAT_EXPECTED)
    for text, want in SAM_RECORDS_BAD:
        with pytest.raises(HsqError) as e:
            ctx.sam_mapped_records(text, ["chr1"])
        assert e.value.status == want
    assert ctx.sam_mapped_records(b"", CONTIGS).n == 0 and ctx.sam_mapped_records(b"@HD\tVN:1.0\n", CONTIGS).n == 0
    for n, kw in ((600, dict(crlf=True)), (30000, dict(xa_rate=0.3))):
        case = synth.make_ingest_case(n, read_len=70, seed=n % 5, trim=(3, 60, 0, 60))
        text = synth.sam_text_of(case, index_names=True, **kw)
        got, exp = ctx.sam_mapped_records(text, CONTIGS), O.sam_mapped_records(text, CONTIGS)
        assert got.n == exp.n and got.n_alt == exp.n_alt and got.n_alt > 20
        for c in got.COLUMNS:
            assert (getattr(got, c)[:got.n] == getattr(exp, c)[:exp.n]).all(), c
        assert (got.alt_offsets[:got.n + 1] == exp.alt_offsets[:exp.n + 1]).all()
        for c in got.ALT_COLUMNS:
            assert (getattr(got, c)[:got.n_alt] == getattr(exp, c)[:exp.n_alt]).all(), c
    with pytest.raises(HsqError) as e:
        ctx.sam_mapped_records(text, CONTIGS, cap=100)
    assert e.value.status == A.HSQ_ERR_OVERFLOW
    with pytest.raises(HsqError) as e:
        ctx.sam_mapped_records(text, CONTIGS, alt_cap=3)
    assert e.value.status == A.HSQ_ERR_OVERFLOW
    # straight into the read -> probe assignment: the records built on the device give the assignment the oracle's records give
    from bioinformatics_b200.batches import ProbeRangesHost
    rng = np.random.default_rng(3)
    starts = rng.integers(1, 40_000_000, size=4000)
    ranges = ProbeRangesHost(rng.integers(0, 22, size=4000), rng.integers(0, 2, size=4000), starts, starts + 400_000)
    ctx.set_probe_ranges(ranges)
    a, b = ctx.assign_probes(got.records(), 8), O.assign_probes(ranges, exp.records(), 8)
    assert (a.overflow == b.overflow).all() and a.overflow.sum() < 0.02 * len(a.overflow)
    fits = a.overflow == 0  # a record that overlaps more windows than slots lists an unspecified subset
    for f in ("n_probes", "mate", "n_common"):
        assert (getattr(a, f)[fits] == getattr(b, f)[fits]).all(), f
    ps_a, ps_b, cs_a, cs_b = a.probe_sets(), b.probe_sets(), a.common_sets(), b.common_sets()
    assert all(ps_a[k] == ps_b[k] and cs_a[k] == cs_b[k] for k in np.nonzero(fits)[0])
    assert (a.n_probes > 0).sum() > 1000


class OraclePort:
    """The oracle behind the calls flow.ingest_flow makes (gather_rows: numpy fancy indexing)."""
    index_fastq = staticmethod(O.index_fastq)
    trim_fastq = staticmethod(O.trim_fastq)
    index_sam = staticmethod(O.index_sam)
    join_names = staticmethod(O.join_names)

    @staticmethod
    def gather_rows(text, off, length):
        t = np.frombuffer(text, dtype=np.uint8)
        off, length = np.asarray(off, dtype=np.int64), np.asarray(length, dtype=np.int64)
        offsets = np.zeros(len(off) + 1, dtype=np.int64)
        np.cumsum(length, out=offsets[1:])
        src = np.repeat(off - offsets[:-1], length) + np.arange(offsets[-1], dtype=np.int64)
        return A.SeqBatch(t[src], offsets)


def test_flow_over_oracle_against_generator():
    """flow.ingest_flow over the oracle: every returned record carries the FASTQ ordinal of its own name and the raw rows."""
    from bioinformatics_b200 import flow
    case = synth.make_ingest_case(3000, read_len=60, seed=22, trim=(5, 50, 0, 50), old_style=True)
    ts = synth.sam_text_of(case, seed=9, crlf=True)
    got = flow.ingest_flow(OraclePort, case["fq1_text"], case["fq2_text"], ts, trim=case["trim"])
    ids, first, fq = case["ids"], got["first_of_pair"], got["read_index"]
    assert len(fq) > 4000 and all(case["sam_names"][int(r)] == ids[int(k)] for r, k in zip(got["sam_line"], fq))
    assert (got["seq_one"].data.reshape(-1, 60) == case["bases"][0][fq[first]]).all()
    assert (got["qual_two"].data.reshape(-1, 60) == case["quals"][1][fq[~first]]).all()
    assert (got["mapped_read_length"] == np.where(first, 55, 60)).all()


@pytest.mark.gpu
def test_gpu_files_to_merged_records(ctx):
    """FASTQ texts + SAM text -> the merged records' read index, ML tag, raw bases and raw qualities: flow.ingest_flow over the
    C ABI against the same flow over the oracle, and against what the generator put into the files."""
    from bioinformatics_b200 import flow
    from bioinformatics_b200._lib import HsqError
    case = synth.make_ingest_case(20000, read_len=60, seed=21, trim=(5, 50, 0, 50))
    t1, t2, ts = case["fq1_text"], case["fq2_text"], synth.sam_text_of(case, seed=8)
    got = flow.ingest_flow(ctx, t1, t2, ts, trim=case["trim"])
    exp = flow.ingest_flow(OraclePort, t1, t2, ts, trim=case["trim"])
    assert set(got) == set(exp) and len(got["sam_line"]) > 30000
    for k in got:
        if isinstance(got[k], A.SeqBatch):
            assert (got[k].data == exp[k].data).all() and (got[k].offsets == exp[k].offsets).all(), k
        else:
            assert np.array_equal(got[k], exp[k]), k
    # against the generator: record -> its own name's FASTQ ordinal, raw (untrimmed) rows of the right file
    ids, first, fq = case["ids"], got["first_of_pair"], got["read_index"]
    assert all(case["sam_names"][int(r)] == ids[int(k)] for r, k in zip(got["sam_line"][:3000], fq[:3000]))
    assert (got["seq_one"].data.reshape(-1, 60) == case["bases"][0][fq[first]]).all()
    assert (got["qual_two"].data.reshape(-1, 60) == case["quals"][1][fq[~first]]).all()
    assert (got["mapped_read_length"] == np.where(first, 55, 60)).all()
    with pytest.raises(HsqError) as e:
        ctx.gather_rows(t1, [len(t1) - 2], [5])
    assert e.value.status == A.HSQ_ERR_OUT_OF_RANGE
    with pytest.raises(HsqError) as e:  # the BAM was made from untrimmed reads but the caller says trimmed
        flow.ingest_flow(ctx, t1, t2, ts, trim=(9, 50, 0, 50))
    assert e.value.status == A.HSQ_ERR_NOT_TRIMMED


def write_merged(sam_text, merged):
    """The writer of createBamFileWithDataFromRawFastqFiles, as text: header kept, the returned records in the iterator's order, QNAME
    replaced by the FASTQ ordinal (storeFastqInfoInRecord, FastqAndBamFileMerger.java:386-396; SEQ / QUAL already equal the raw reads here)."""
    lines = sam_text.decode().split("\n")
    header = [x for x in lines if x.startswith("@")]
    body = [x for x in lines if x and not x.startswith("@")]
    out = list(header)
    for line, idx in zip(merged["sam_line"], merged["read_index"]):
        f = body[int(line)].split("\t")
        f[0] = str(int(idx))
        out.append("\t".join(f))
    return ("\n".join(out) + "\n").encode()


def _files_problem(repeat_names):
    from tests.test_flow import U, V
    probes = synth.make_probes(50, seed=4)
    fastq = synth.make_pairs(probes, 2500, read_len=150, ext_uid_len=U, lig_uid_len=V, seed=5, sub_rate=0.01, n_rate=0.01)
    ranges, records = synth.make_mapped_records(probes, fastq, seed=6, repeat_names=repeat_names)
    return probes, fastq, ranges, records, synth.dedup_input_files(probes, fastq, records)


def _same_flow_outputs(a, b):
    from tests import helpers as H
    assert (a["merged"]["sam_line"] == b["merged"]["sam_line"]).all() and (a["merged"]["read_index"] == b["merged"]["read_index"]).all()
    for f in ("contig", "negative_strand", "second_of_pair", "start", "end", "read_index", "alt_offsets", "alt_contig", "alt_negative", "alt_start", "alt_stop"):
        assert (getattr(a["mapped_records"], f) == getattr(b["mapped_records"], f)).all(), f
    assert (a["pairs"].probe_id == b["pairs"].probe_id).all() and (a["pairs"].read_index == b["pairs"].read_index).all()
    assert (a["dedup"].is_best == b["dedup"].is_best).all() and (a["dedup"].group_size == b["dedup"].group_size).all()
    for f in H.EXTEND_FIELDS:
        assert (getattr(a["extend"], f) == getattr(b["extend"], f)).all(), f
    assert a["extend"].cigars() == b["extend"].cigars() and a["extend"].mds() == b["extend"].mds()
    ra, rb = a["records"], b["records"]
    assert (ra.flag == rb.flag).all() and (ra.mate_pos == rb.mate_pos).all() and (ra.offsets == rb.offsets).all()
    assert (ra.seq[:ra.offsets[-1]] == rb.seq[:rb.offsets[-1]]).all() and (ra.qual[:ra.offsets[-1]] == rb.qual[:rb.offsets[-1]]).all()


def test_dedup_from_files_on_oracle_equals_dedup_from_tables():
    """Files -> merge -> mapped records -> dedup gives what dedup_flow gives on the tables the files were written from."""
    from bioinformatics_b200 import flow
    from tests.test_flow import OracleBackend, U, V
    probes, fastq, ranges, records, (fq1, fq2, sam, contigs) = _files_problem(repeat_names=0)

    class Backend(OracleBackend, OraclePort):
        sam_mapped_records = staticmethod(O.sam_mapped_records)

    be = Backend(probes, ranges)
    out = flow.dedup_from_files(be, fq1, fq2, sam, contigs, U, V, write_merged)
    direct = flow.dedup_flow(be, records, fastq, U, V)
    assert out["mapped_records"].n == records.n and len(out["merged"]["sam_line"]) == records.n
    key = lambda o: sorted(zip(o["pairs"].read_index.tolist(), o["pairs"].probe_id.tolist(), o["dedup"].is_best.tolist()))
    assert key(out) == key(direct) and out["records"].n == direct["records"].n > 300
    pos = lambda o: sorted(zip(o["extended_pairs"].read_index.tolist(), o["extend"].pos[0::2].tolist(), o["extend"].cigars()[0::2]))
    assert pos(out) == pos(direct)


@pytest.mark.gpu
def test_gpu_dedup_from_files_equals_oracle(ctx):
    """The whole dedup data path from the three input files on the device against the same flow over the oracle (names that sit on
    four records included: the merge keeps three of them, FastqAndBamFileMerger.java:304-309)."""
    from bioinformatics_b200 import flow
    from tests.test_flow import OracleBackend, U, V
    probes, fastq, ranges, records, (fq1, fq2, sam, contigs) = _files_problem(repeat_names=3)

    class Backend(OracleBackend, OraclePort):
        sam_mapped_records = staticmethod(O.sam_mapped_records)

    ctx.set_probes(probes)
    ctx.set_probe_ranges(ranges)
    got = flow.dedup_from_files(ctx, fq1, fq2, sam, contigs, U, V, write_merged)
    exp = flow.dedup_from_files(Backend(probes, ranges), fq1, fq2, sam, contigs, U, V, write_merged)
    _same_flow_outputs(got, exp)
    assert len(got["merged"]["sam_line"]) == records.n - 3 and got["records"].n > 300


# -------------------------------------------------------------------------------------------------- `HSQutils trim`, file to file
def trim_statement(text, first, last, three_prime):
    """FastqReadTrimmer.trimReads with str.splitlines (second statement)."""
    lines = text.decode().splitlines()
    out = []
    for k in range(0, len(lines), 4):
        h, s_, p, q = lines[k:k + 4]
        lb = min(last, len(s_) - 1) if three_prime else len(s_) - 1
        out += [h, s_[first:lb + 1], p, q[first:lb + 1]]
    return ("\n".join(out) + "\n").encode() if out else b""


TRIM_TEXT = b"@r1 x\nACGTACGTACGTACGTACGT\n+r1 x\nABCDEFGHIJKLMNOPQRST\n@r2\r\nACGTACGTAC\r\n+\r\n0123456789"
# FastqReadTrimmerTest.java:35-40: a 20 base read, first base to keep 5, last 14 -> 10 bases; the second read is shorter than the stop
TRIM_EXPECTED = b"@r1 x\nCGTACGTACG\n+r1 x\nFGHIJKLMNO\n@r2\nCGTAC\n+\n56789\n"
TRIM_BAD = [((b"@r\nACGT\n+\nIIII\n", 4, 9, True), A.HSQ_ERR_OUT_OF_RANGE),      # first base to keep behind the read (:310-312)
            ((b"@r\nACGT\n+\nIIII\n", -1, 9, True), A.HSQ_ERR_INVALID_ARGUMENT),  # :264-266
            ((b"@r\nACGT\n+\nIIII\n", 3, 3, True), A.HSQ_ERR_INVALID_ARGUMENT),   # :268-270
            ((b"@r\nACGT\n+\nIII\n", 1, 3, True), A.HSQ_ERR_MALFORMED)]


def test_oracle_trim_fastq():
    t, n = O.trim_fastq(TRIM_TEXT, 5, 14, True)
    assert n == 2 and t.tobytes() == TRIM_EXPECTED
    assert O.trim_fastq(TRIM_TEXT, 5, 14, False)[0].tobytes() == b"@r1 x\nCGTACGTACGTACGT\n+r1 x\nFGHIJKLMNOPQRST\n@r2\nCGTAC\n+\n56789\n"
    for args, want in TRIM_BAD:
        with pytest.raises(O.OracleError) as e:
            O.trim_fastq(*args)
        assert e.value.status == want
    case = synth.make_ingest_case(700, read_len=60, seed=8)
    for first, last, tp in ((12, 47, True), (0, 200, True), (7, 8, False)):
        assert O.trim_fastq(case["fq2_text"], first, last, tp)[0].tobytes() == trim_statement(case["fq2_text"], first, last, tp)


def test_trim_flow_over_oracle():
    """`HSQutils trim` file to file: the offsets of getProbeTrimmingInformation, then every output record is its input record cut."""
    from bioinformatics_b200 import flow
    probes = synth.make_probes(30, seed=2)
    fastq = synth.make_pairs(probes, 400, read_len=150, ext_uid_len=10, seed=3)
    fq1, fq2, _, _ = synth.dedup_input_files(probes, fastq, synth.make_mapped_records(probes, fastq, seed=4)[1])
    t1, t2, info = flow.trim_flow(OraclePort, probes, fq1, fq2, 10, 0)
    ext, lig, tgt = probes.ext.lengths(), probes.lig.lengths(), probes.target.lengths()
    assert info == (10 + ext.max(), tgt.min() + lig.min(), lig.max(), tgt.min() + ext.min())
    i1 = O.index_fastq(t1.tobytes())
    assert i1.n == 400 and (i1.seq_len[:400] == min(info[1], 149) + 1 - info[0]).all()
    assert t2.tobytes() == trim_statement(fq2, info[2], info[3], True)


@pytest.mark.gpu
def test_gpu_trim_fastq_equals_oracle(ctx):
    from bioinformatics_b200._lib import HsqError
    from bioinformatics_b200 import flow
    t, n = ctx.trim_fastq(TRIM_TEXT, 5, 14, True)
    assert n == 2 and t.tobytes() == TRIM_EXPECTED
    for args, want in TRIM_BAD:
        with pytest.raises(HsqError) as e:
            ctx.trim_fastq(*args)
        assert e.value.status == want
    assert ctx.trim_fastq(b"", 1, 5)[1] == 0
    for n_pairs, kw in ((1, {}), (3000, dict(crlf=True)), (40000, dict(final_newline=False, old_style=True))):
        case = synth.make_ingest_case(n_pairs, read_len=101, seed=n_pairs % 9, **kw)
        for first, last, tp in ((12, 88, True), (0, 300, True), (30, 31, False)):
            got, exp = ctx.trim_fastq(case["fq1_text"], first, last, tp), O.trim_fastq(case["fq1_text"], first, last, tp)
            assert got[1] == exp[1] == n_pairs and got[0].tobytes() == exp[0].tobytes()
    with pytest.raises(HsqError) as e:
        ctx.trim_fastq(case["fq1_text"], 12, 88, True, out=np.zeros(1000, dtype=np.uint8))
    assert e.value.status == A.HSQ_ERR_OVERFLOW
    probes = synth.make_probes(30, seed=2)
    fastq = synth.make_pairs(probes, 2000, read_len=150, ext_uid_len=10, seed=3)
    fq1, fq2, _, _ = synth.dedup_input_files(probes, fastq, synth.make_mapped_records(probes, fastq, seed=4)[1])
    g, o = flow.trim_flow(ctx, probes, fq1, fq2, 10, 0), flow.trim_flow(OraclePort, probes, fq1, fq2, 10, 0)
    assert g[0].tobytes() == o[0].tobytes() and g[1].tobytes() == o[1].tobytes() and g[2] == o[2]


# ------------------------------------------------------------------------------------- the final coordinate sort (f4, output side)
def coord_problem(n, seed, n_names=None):
    rng = np.random.default_rng(seed)
    ref = rng.integers(-1, 4, size=n)
    pos = rng.integers(1, 60, size=n)
    flags = rng.choice([99, 147, 83, 163, 1123, 1171, 65, 129, 81, 161], size=n)
    mapq = rng.choice([0, 60, 37], size=n)
    mate_ref = rng.integers(-1, 3, size=n)
    mate_pos = rng.integers(0, 5, size=n)
    tlen = rng.integers(-3, 4, size=n)
    names = ["M0:%d:%s" % (int(k), "x" * int(k % 5)) for k in rng.integers(0, n_names or max(2, n // 3), size=n)]
    return ref, pos, flags, mapq, mate_ref, mate_pos, tlen, names


def coord_statement(ref, pos, flags, mapq, mate_ref, mate_pos, tlen, names):
    """sorted() with the comparator as a key tuple (Python's sort is stable)."""
    key = lambda k: ((1, 0) if ref[k] == -1 else (0, int(ref[k]), int(pos[k])), 1 if flags[k] & 0x10 else 0, names[k].encode(), int(flags[k]), int(mapq[k]),
                     int(mate_ref[k]), int(mate_pos[k]), int(tlen[k]))
    return sorted(range(len(ref)), key=key)


# by hand: reference 0 before reference 2, unplaced (-1) last; inside reference 0: start 5 before 9; at (0, 5) the forward record [3] before the
# reverse ones; the two reverse records at (0, 5) by name "a" < "b"; at (2, 7) same strand and name: flags 99 before 163
COORD_KAT = ([2, -1, 0, 0, 0, 2, 0], [7, 3, 5, 5, 9, 7, 5], [163, 77, 83, 99, 99, 99, 83], [60] * 7, [0] * 7, [1] * 7, [0] * 7, ["q", "u", "b", "z", "c", "q", "a"])
COORD_KAT_ORDER = [3, 6, 2, 4, 5, 0, 1]


def test_oracle_coordinate_order():
    assert list(O.coordinate_order(*COORD_KAT)) == COORD_KAT_ORDER
    for n, seed in ((1, 1), (50, 2), (3000, 3)):
        p = coord_problem(n, seed)
        assert list(O.coordinate_order(*p)) == coord_statement(*p)


@pytest.mark.gpu
def test_gpu_coordinate_order_equals_oracle(ctx):
    assert list(ctx.coordinate_order(*COORD_KAT)) == COORD_KAT_ORDER
    for n, seed, names in ((1, 1, None), (2, 2, None), (700, 3, 5), (50000, 4, None), (50000, 5, 40)):
        p = coord_problem(n, seed, names)
        assert (ctx.coordinate_order(*p) == O.coordinate_order(*p)).all()
    # the records dedup_flow writes, ordered for the output BAM: positions from the extension, names = read indexes as strings
    rng = np.random.default_rng(9)
    n = 20000
    p = (rng.integers(0, 22, size=n), rng.integers(1, 10 ** 8, size=n), rng.choice([99, 147, 83, 163], size=n), np.full(n, 60), rng.integers(0, 22, size=n),
         rng.integers(1, 10 ** 8, size=n), rng.integers(-400, 400, size=n), [str(int(k)) for k in rng.integers(0, 10 ** 7, size=n)])
    got = ctx.coordinate_order(*p)
    assert (got == O.coordinate_order(*p)).all()
    key = p[0][got].astype(np.int64) * 10 ** 9 + p[1][got]
    assert (np.diff(key) >= 0).all()
