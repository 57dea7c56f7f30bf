"""TEST INFRASTRUCTURE ONLY: ctypes binding of the CPU parity oracle (oracle/libhsq_oracle.so).

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from bioinformatics_b200 import _abi as A
from bioinformatics_b200.batches import (SamRecordsResult, SamIndexResult, FastqIndexResult, NameJoinResult, AlignResult, AssignResult, ExtendResult, FindResult, RecordResult, PairBatchHost, ProbeStatsHost, ProbeTableHost,
                                         merge_args, scorer)

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libhsq_oracle.so")
_lib = None


def build(force=False):
    if force or not os.path.exists(_SO):
        subprocess.check_call(["make", "-C", _HERE] + (["-B"] if force else []), stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        try:
            _lib = C.CDLL(_SO)
        except OSError:
            build(force=True)
            _lib = C.CDLL(_SO)
        _lib.hsqo_last_error.restype = C.c_char_p
        for name in ("hsqo_align_batch", "hsqo_trim_batch", "hsqo_dedup_batch", "hsqo_extend_batch",
                     "hsqo_resolve_multi_probe", "hsqo_primer_end", "hsqo_index_of_first_match_in_reference",
                     "hsqo_cigar_edit_distance", "hsqo_variable_length_uid", "hsqo_md_trimmed_with_full_cigar", "hsqo_merge_pairs", "hsqo_assign_probes",
                     "hsqo_extend_records", "hsqo_find_candidates", "hsqo_index_fastq", "hsqo_trim_fastq", "hsqo_index_sam", "hsqo_coordinate_order", "hsqo_sam_mapped_records", "hsqo_join_names"):
            getattr(_lib, name).restype = C.c_int32
    return _lib


class OracleError(RuntimeError):
    def __init__(self, status):
        super().__init__("oracle status %d: %s" % (status, lib().hsqo_last_error().decode()))
        self.status = status


def _check(st):
    if st != 0:
        raise OracleError(st)


def align_batch(refs, qrys, sc=None, flags=0, cols_cap=640, md_cap=640, nthreads=1):
    refs = refs if isinstance(refs, A.SeqBatch) else A.SeqBatch.from_strings(refs)
    qrys = qrys if isinstance(qrys, A.SeqBatch) else A.SeqBatch.from_strings(qrys)
    sc = sc or scorer()
    out = AlignResult(refs.n, cols_cap, md_cap)
    rc, qc, oc = refs.c(), qrys.c(), out.c()
    _check(lib().hsqo_align_batch(C.byref(rc), C.byref(qc), C.byref(sc), C.c_uint32(flags), C.byref(oc),
                                  C.c_int32(nthreads)))
    return out.finish()


def trim_batch(reads, quals, first, last, three_prime=True):
    reads = reads if isinstance(reads, A.SeqBatch) else A.SeqBatch.from_strings(reads)
    quals = quals if isinstance(quals, A.SeqBatch) else A.SeqBatch.from_strings(quals)
    out_seq = np.zeros(max(1, len(reads.data)), dtype=np.uint8)
    out_qual = np.zeros(max(1, len(reads.data)), dtype=np.uint8)
    out_off = np.zeros(reads.n + 1, dtype=np.int64)
    rc = reads.c()
    _check(lib().hsqo_trim_batch(C.byref(rc), A.ptr(quals.data, A.u8p), C.c_int32(first), C.c_int32(last),
                                 C.c_int32(1 if three_prime else 0), A.ptr(out_seq, A.u8p), A.ptr(out_qual, A.u8p),
                                 A.ptr(out_off, A.i64p)))
    total = int(out_off[-1])
    return A.SeqBatch(out_seq[:total], out_off), A.SeqBatch(out_qual[:total], out_off)


def merge_pairs(negative, up, down, upq, downq, up_start, down_start):
    mb, keep = merge_args(negative, up, down, upq, downq, up_start, down_start)
    off = np.zeros(mb.n + 1, dtype=np.int64)
    cap = 16
    for _ in range(2):
        seq = np.zeros(max(1, cap), dtype=np.uint8)
        qual = np.zeros(max(1, cap), dtype=np.uint8)
        st = lib().hsqo_merge_pairs(C.byref(mb), A.ptr(seq, A.u8p), A.ptr(qual, A.u8p), A.ptr(off, A.i64p), C.c_int64(cap))
        if st == A.HSQ_ERR_OVERFLOW:
            cap = int(off[-1])
            continue
        break
    _check(st)
    total = int(off[-1])
    return A.SeqBatch(seq[:total], off), A.SeqBatch(qual[:total], off.copy())


def extend_records(probes, pairs, sc=None, duplicate=None):
    """extendReadPair + createRecord + setSAMRecordsAsPair on the CPU: FLAG, PNEXT, read string, qualities per record."""
    sc = sc or scorer()
    cap = int(len(pairs.r1.data) + len(pairs.r2.data))
    dup = None if duplicate is None else np.ascontiguousarray(duplicate, dtype=np.uint8)
    out = RecordResult(pairs.n, cap)
    tc, pc, oc = probes.c(), pairs.c(), out.c()
    _check(lib().hsqo_extend_records(C.byref(tc), C.byref(sc), C.byref(pc), A.ptr(dup, A.u8p), C.byref(oc), C.c_int64(cap)))
    return out


def assign_probes(ranges, records, max_probes=4):
    """ReadToProbeAssigner.assignProbesToReads + RangeMap on the CPU (hsq_oracle_assign.hpp)."""
    out = AssignResult(records.n, max_probes)
    pc, rc, oc = ranges.c(), records.c(), out.c()
    _check(lib().hsqo_assign_probes(C.byref(pc), C.byref(rc), C.byref(oc)))
    return out


def index_fastq(text, cap=None):
    """FastqReader record index of FASTQ text on the CPU (hsq_oracle_ingest.hpp)."""
    text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
    if cap is None:
        cap = len(text) // 8 + 1
    out = FastqIndexResult(cap)
    oc = out.c()
    _check(lib().hsqo_index_fastq(A.ptr(text, A.u8p), C.c_int64(len(text)), C.byref(oc)))
    out.n = int(oc.n)
    return out


def trim_fastq(text, first, last, three_prime=True):
    """FastqReadTrimmer.trimReads on FASTQ text on the CPU (hsq_oracle_ingest.hpp)."""
    text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
    out = np.zeros(len(text) + 2, dtype=np.uint8)
    nb, nrec = C.c_int64(0), C.c_int64(0)
    _check(lib().hsqo_trim_fastq(A.ptr(text, A.u8p), C.c_int64(len(text)), C.c_int32(first), C.c_int32(last), C.c_int32(1 if three_prime else 0),
                                 A.ptr(out, A.u8p), C.c_int64(len(out)), C.byref(nb), C.byref(nrec)))
    return out[:nb.value], int(nrec.value)


def index_sam(text, cap=None):
    """Field index of SAM text on the CPU (hsq_oracle_ingest.hpp)."""
    text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
    if cap is None:
        cap = len(text) // 22 + 1
    out = SamIndexResult(cap)
    oc = out.c()
    _check(lib().hsqo_index_sam(A.ptr(text, A.u8p), C.c_int64(len(text)), C.byref(oc)))
    return out.take(oc)


def sam_mapped_records(text, contig_names, cap=None, alt_cap=None):
    """SAM text -> mapped records + alternative hits on the CPU (hsq_oracle_ingest.hpp)."""
    text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
    names = contig_names if isinstance(contig_names, A.SeqBatch) else A.SeqBatch.from_strings(contig_names)
    out = SamRecordsResult(len(text) // 22 + 1 if cap is None else cap, len(text) // 8 + 1 if alt_cap is None else alt_cap)
    nc, oc = names.c(), out.c()
    _check(lib().hsqo_sam_mapped_records(A.ptr(text, A.u8p), C.c_int64(len(text)), C.byref(nc), C.byref(oc)))
    return out.take(oc)


def coordinate_order(ref_index, pos, flags, mapq, mate_ref_index, mate_pos, tlen, names):
    """Order of the final coordinate sort on the CPU: std::stable_sort with htsjdk's SAMRecordCoordinateComparator restated."""
    from bioinformatics_b200._lib import _coord_args
    keep = _coord_args(ref_index, pos, flags, mapq, mate_ref_index, mate_pos, tlen, names)
    order = np.zeros(len(keep[0][0]), dtype=np.int32)
    _check(lib().hsqo_coordinate_order(C.byref(keep[2]), A.ptr(order, A.i32p)))
    return order


def join_names(join):
    """The two sorts and the merge loop of FastqAndBamFileMerger on the CPU, sequential as in the reference."""
    out = NameJoinResult(join.n)
    jc, oc = join.c(), out.c()
    _check(lib().hsqo_join_names(C.byref(jc), C.byref(oc)))
    return out.take(oc)


def find_candidates(seqs, reads, quals, kmer=5, max_refs_per_kmer=15000, leading_quality=0, trailing_quality=0, candidate_limit=10,
                    min_hit_ratio=0.5, max_candidates=16):
    """trimReadBasedOnQuality + the N test + SimpleMapper.getBestCandidateReferences per read (hsq_oracle_find.hpp)."""
    seqs = seqs if isinstance(seqs, A.SeqBatch) else A.SeqBatch.from_strings(seqs)
    reads = reads if isinstance(reads, A.SeqBatch) else A.SeqBatch.from_strings(reads)
    quals = quals if isinstance(quals, A.SeqBatch) else A.SeqBatch.from_strings(quals)
    out = FindResult(reads.n, max_candidates)
    sc, rc, oc = seqs.c(), reads.c(), out.c()
    _check(lib().hsqo_find_candidates(C.byref(sc), C.c_int32(kmer), C.c_int32(max_refs_per_kmer), C.byref(rc), A.ptr(quals.data, A.u8p),
                                      C.c_int32(leading_quality), C.c_int32(trailing_quality), C.c_int32(candidate_limit),
                                      C.c_double(min_hit_ratio), C.byref(oc)))
    return out


def dedup_batch(pairs, ext_uid_len, lig_uid_len, n_probes):
    is_best = np.zeros(pairs.n, dtype=np.uint8)
    group_size = np.zeros(pairs.n, dtype=np.int32)
    stats = ProbeStatsHost(n_probes)
    pc, sc = pairs.c(), stats.c()
    _check(lib().hsqo_dedup_batch(C.byref(pc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len), A.ptr(is_best, A.u8p),
                                  A.ptr(group_size, A.i32p), C.byref(sc)))
    return is_best, group_size, stats


def extend_batch(probes, pairs, sc=None, cigar_cap=32, md_cap=48, nthreads=1):
    sc = sc or scorer()
    out = ExtendResult(pairs.n, cigar_cap, md_cap)
    tc, pc, oc = probes.c(), pairs.c(), out.c()
    _check(lib().hsqo_extend_batch(C.byref(tc), C.byref(sc), C.byref(pc), C.byref(oc), C.c_int32(nthreads)))
    return out.finish()


def resolve_multi_probe(probes, r1, r2, cand_offsets, cand_probe, ext_uid_len, lig_uid_len):
    r1 = r1 if isinstance(r1, A.SeqBatch) else A.SeqBatch.from_strings(r1)
    r2 = r2 if isinstance(r2, A.SeqBatch) else A.SeqBatch.from_strings(r2)
    cand_offsets = np.ascontiguousarray(cand_offsets, dtype=np.int64)
    cand_probe = np.ascontiguousarray(cand_probe, dtype=np.int32)
    best = np.zeros(r1.n, dtype=np.int32)
    m = np.zeros(len(cand_probe), dtype=np.int32)
    e = np.zeros(len(cand_probe), dtype=np.int32)
    cc = A.Candidates(r1.n, r1.c(), r2.c(), A.ptr(cand_offsets, A.i64p), A.ptr(cand_probe, A.i32p))
    tc = probes.c()
    _check(lib().hsqo_resolve_multi_probe(C.byref(tc), C.byref(cc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                          A.ptr(best, A.i32p), A.ptr(m, A.i32p), A.ptr(e, A.i32p)))
    return best, m, e


def primer_end(primer, read):
    w = C.c_double(0)
    r = lib().hsqo_primer_end(primer.encode(), read.encode(), C.byref(w))
    return (None if r == A.HSQ_PRIMER_NOT_FOUND else r), w.value


def variable_length_uid(read, primer, sc=None, flags=0):
    """SAMRecordUtil.getVariableLengthUid: length of the UID prefix of `read`, None where the reference returns null."""
    sc = sc or scorer()
    r = lib().hsqo_variable_length_uid(read.encode(), primer.encode(), C.byref(sc), C.c_uint32(flags))
    return None if r < 0 else r


def index_of_first_match_in_reference(ref, qry):
    return lib().hsqo_index_of_first_match_in_reference(ref.encode(), qry.encode())


def cigar_edit_distance(ref, qry):
    return lib().hsqo_cigar_edit_distance(ref.encode(), qry.encode())


def md_trimmed_with_full_cigar(ref, qry, reverse=False):
    buf = C.create_string_buffer(4096)
    lib().hsqo_md_trimmed_with_full_cigar(ref.encode(), qry.encode(), C.c_int32(1 if reverse else 0), buf, 4096)
    return buf.value.decode()


# ---- CPU stand-ins for the two device calls of the sharded dedup (tests/test_sharded_gloo.py only) -------------------
KEY_DTYPE = np.dtype([("uid", "<u8"), ("rank", "<u8"), ("probe", "<i4"), ("origin", "<u4")])  # = hsq_dedup_key, 24 bytes


def dedup_keys(pairs, ext_uid_len, lig_uid_len, origin_base=0):
    """Key records as hsq_dedup_keys writes them (FilterByUid.java:100-101,136-138; findBestData :281-297;
    BamFileUtil.java:74-85 quality score; ReadPair.java:101-103 short arithmetic)."""
    code = {ord("A"): 0, ord("C"): 1, ord("G"): 2, ord("T"): 3}
    out = np.zeros(pairs.n, dtype=KEY_DTYPE)
    for k in range(pairs.n):
        a0, a1 = int(pairs.r1.offsets[k]), int(pairs.r1.offsets[k + 1])
        b0, b1 = int(pairs.r2.offsets[k]), int(pairs.r2.offsets[k + 1])
        uid = 0
        for ch in list(pairs.r1.data[a0:a0 + ext_uid_len]) + list(pairs.r2.data[b0:b0 + lig_uid_len]):
            uid = (uid << 2) | code[int(ch)]
        total = int(pairs.q1.data[a0:a1].astype(np.int64).sum() - 33 * (a1 - a0) +
                    pairs.q2.data[b0:b1].astype(np.int64).sum() - 33 * (b1 - b0))
        qsum = ((total + 32768) % 65536) - 32768  # Java (short)
        rank = (((32767 - qsum) & 0xFFFF) << 32) | ((int(pairs.read_index[k]) ^ 0x80000000) & 0xFFFFFFFF)
        out[k] = (uid, rank, int(pairs.probe_id[k]), origin_base + k)
    return out


def dedup_reduce_keys(keys, n_probes):
    """Best pair per (probe, UID) group over key records: (is_best, group_size) per record + per-probe counters."""
    n = len(keys)
    is_best = np.zeros(n, dtype=np.uint8)
    gsize = np.zeros(n, dtype=np.int32)
    stats = ProbeStatsHost(n_probes)
    if n == 0:
        return is_best, gsize, stats
    # last key: the record's origin (the pair's place in the whole batch) -- the single-process path breaks exact ties of
    # (quality sum, read index) by the earlier pair, and with globally unique origins the sharded answer is the same
    order = np.lexsort((keys["origin"], keys["rank"], keys["uid"], keys["probe"]))
    p, u = keys["probe"][order], keys["uid"][order]
    head = np.ones(n, dtype=bool)
    head[1:] = (p[1:] != p[:-1]) | (u[1:] != u[:-1])
    starts = np.nonzero(head)[0]
    sizes = np.diff(np.append(starts, n))
    is_best[order[starts]] = 1
    gsize[order] = np.repeat(sizes, sizes)
    for s, sz in zip(starts, sizes):
        pr = int(p[s])
        stats.unique_pairs[pr] += 1
        stats.duplicate_pairs[pr] += sz - 1
        stats.max_pairs_per_uid[pr] = max(stats.max_pairs_per_uid[pr], sz)
    return is_best, gsize, stats


def bam_index(data):
    """CPU statement of hsq_index_bam: what a sequential reader of the BAM (htsjdk BAMFileReader / BAMRecordCodec: SAM/BAM format
    specification 4.1-4.2) sees, record by record.  Inflates whatever level the blocks were written at.  Returns a dict with the
    uncompressed stream, header sizes and, per record, the columns of include/hsq_b200.h's hsq_bam_index."""
    import struct
    import zlib
    pos, stream = 0, bytearray()
    data = bytes(data)
    while pos < len(data):
        if data[pos:pos + 4] != b"\x1f\x8b\x08\x04":
            raise ValueError("not a BGZF block at %d" % pos)
        xlen = struct.unpack_from("<H", data, pos + 10)[0]
        x, bsize = pos + 12, None
        while x < pos + 12 + xlen:
            si1, si2, slen = data[x], data[x + 1], struct.unpack_from("<H", data, x + 2)[0]
            if si1 == 66 and si2 == 67 and slen == 2:
                bsize = struct.unpack_from("<H", data, x + 4)[0]
            x += 4 + slen
        if bsize is None:
            raise ValueError("no BC field")
        cdata = data[pos + 12 + xlen:pos + bsize + 1 - 8]
        crc, isize = struct.unpack_from("<II", data, pos + bsize + 1 - 8)
        raw = zlib.decompress(cdata, -15) if isize else b""
        if len(raw) != isize or (zlib.crc32(raw) & 0xFFFFFFFF) != crc:
            raise ValueError("BGZF block at %d: size or CRC mismatch" % pos)
        stream += raw
        pos += bsize + 1
    stream = bytes(stream)
    if stream[:4] != b"BAM\1":
        raise ValueError("not a BAM stream")
    l_text = struct.unpack_from("<i", stream, 4)[0]
    p = 8 + l_text
    n_ref = struct.unpack_from("<i", stream, p)[0]
    p += 4
    for _ in range(n_ref):
        l_name = struct.unpack_from("<i", stream, p)[0]
        p += 4 + l_name + 4
    first = p
    names = ("record", "qname", "cigar", "seq", "qual", "tags")
    cols = {k + "_off": [] for k in names}
    cols.update({k + "_len": [] for k in names})
    for k in ("refid", "pos", "flag", "mapq", "next_refid", "pnext", "tlen", "ref_len"):
        cols[k] = []
    while p < len(stream):
        bs = struct.unpack_from("<i", stream, p)[0]
        ref_id, bpos, l_name, mapq, _bin, n_cig, flag, l_seq, nref, npos, tlen = struct.unpack_from("<iiBBHHHiiii", stream, p + 4)
        name = p + 36
        cigar = name + l_name
        seq = cigar + 4 * n_cig
        qual = seq + (l_seq + 1) // 2
        tags = qual + l_seq
        end = p + 4 + bs
        if tags > end or end > len(stream) or stream[cigar - 1] != 0:
            raise ValueError("broken record at %d" % p)
        ref_len = 0
        for c in range(n_cig):
            op = struct.unpack_from("<I", stream, cigar + 4 * c)[0]
            if (op & 15) in (0, 2, 3, 7, 8):
                ref_len += op >> 4
        for key, o, ln in (("record", p, bs + 4), ("qname", name, l_name - 1), ("cigar", cigar, n_cig), ("seq", seq, l_seq), ("qual", qual, l_seq),
                           ("tags", tags, end - tags)):
            cols[key + "_off"].append(o)
            cols[key + "_len"].append(ln)
        for key, v in (("refid", ref_id), ("pos", bpos + 1), ("flag", flag), ("mapq", mapq), ("next_refid", nref), ("pnext", npos + 1), ("tlen", tlen),
                       ("ref_len", ref_len)):
            cols[key].append(v)
        p = end
    out = {k: np.asarray(v, dtype=np.int64 if k.endswith("_off") else np.int32) for k, v in cols.items()}
    out.update(stream=stream, n=len(cols["pos"]), n_ref=n_ref, text_bytes=l_text, first_record=first)
    return out
