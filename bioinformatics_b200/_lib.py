"""Loads the CUDA C-ABI library (bioinformatics_b200/libhsq_b200.so) and wraps a context.

There is no CPU fallback anywhere in this package: if the library is missing or no B200-class GPU is
present, creating a context raises.
"""
import ctypes as C
import os

import numpy as np

from . import _abi as A
from .batches import (PackedReadsHost, PackedPairsHost, PooledExtendResult, SamRecordsResult, SamIndexResult, FastqIndexResult, NameJoinHost, NameJoinResult, AlignResult, AssignResult, FindResult, FindMatchResult, DedupExtendResult, ExtendResult, RecordResult, MappedRecordsHost, PairBatchHost, ProbeRangesHost, ProbeStatsHost,
                      ProbeTableHost, merge_args, scorer)

_PKG = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_PKG, "libhsq_b200.so")
_lib = None

EXPORTS = ("hsq_ctx_create", "hsq_ctx_destroy", "hsq_last_error", "hsq_version", "hsq_pinned_alloc", "hsq_pinned_free",
           "hsq_set_scorer", "hsq_set_probes", "hsq_last_timing", "hsq_align_batch", "hsq_trim_batch", "hsq_dedup_batch",
           "hsq_extend_batch", "hsq_resolve_multi_probe", "hsq_extend_stage", "hsq_extend_run", "hsq_extend_fetch",
           "hsq_timer_start", "hsq_timer_stop", "hsq_int32_peak", "hsq_dedup_keys", "hsq_dedup_reduce_keys", "hsq_merge_pairs",
           "hsq_set_probe_ranges", "hsq_assign_probes", "hsq_dedup_extend_batch", "hsq_extend_records", "hsq_extend_selected", "hsq_flag_n_pairs",
           "hsq_set_find_sequences", "hsq_find_candidates", "hsq_find_matches", "hsq_index_fastq", "hsq_trim_fastq", "hsq_index_sam", "hsq_sam_mapped_records", "hsq_join_names", "hsq_gather_rows", "hsq_coordinate_order",
           "hsq_ctx_create_ex", "hsq_ctx_configure", "hsq_packed_bytes", "hsq_pack_reads", "hsq_quality_sums", "hsq_unpack_reads",
           "hsq_extend_packed", "hsq_dedup_packed", "hsq_comm_unique_id", "hsq_comm_init_rank", "hsq_comm_init_all", "hsq_comm_destroy",
           "hsq_dedup_sharded", "hsq_dedup_extend_sharded", "hsq_dedup_extend_multi_gpu", "hsq_last_spans", "hsq_dedup_extend_packed", "hsq_variable_length_uid", "hsq_align_scores_noalloc", "hsq_band_stats", "hsq_dedup_batch_report", "hsq_pair_reasons", "hsq_index_bam", "hsq_uid_keys", "hsq_dedup_extend_packed_keys")
_NOT_STATUS = {"hsq_last_error": C.c_char_p, "hsq_version": C.c_char_p, "hsq_ctx_destroy": None, "hsq_packed_bytes": C.c_int64}


def _coord_args(ref_index, pos, flags, mapq, mate_ref_index, mate_pos, tlen, names):
    cols = [np.ascontiguousarray(x, dtype=np.int32) for x in (ref_index, pos, flags, mapq, mate_ref_index, mate_pos, tlen)]
    names = names if isinstance(names, A.SeqBatch) else A.SeqBatch.from_strings(names)
    return cols, names, A.CoordSortIn(len(cols[0]), *[A.ptr(c, A.i32p) for c in cols], names.c())


class HsqError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("hsq status %d: %s" % (status, message))
        self.status = status


def load():
    """dlopen the CUDA library.  Raises if it has not been built (python -m bioinformatics_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise HsqError(A.HSQ_ERR_NO_DEVICE, "CUDA extension %s is missing; run __graft_entry__.build()" % SO_PATH)
        lib = C.CDLL(SO_PATH)
        lib.hsq_last_error.restype = C.c_char_p
        lib.hsq_last_error.argtypes = [C.c_void_p]
        lib.hsq_version.restype = C.c_char_p
        lib.hsq_ctx_destroy.restype = None
        lib.hsq_ctx_destroy.argtypes = [C.c_void_p]
        for name in EXPORTS:
            getattr(lib, name).restype = _NOT_STATUS.get(name, C.c_int32)
        _lib = lib
    return _lib


def pack_reads(reads, bits=2, alloc=None, threads=0):
    """hsq_pack_reads (host helper of the library, no device): SeqBatch -> PackedReadsHost.  alloc(nbytes) -> uint8 array (e.g. pinned)."""
    lib = load()
    reads = reads if isinstance(reads, A.SeqBatch) else A.SeqBatch.from_strings(reads)
    alloc = alloc or (lambda nbytes: np.zeros(nbytes, dtype=np.uint8))
    nbytes = int(lib.hsq_packed_bytes(A.ptr(reads.offsets, A.i64p), C.c_int64(reads.n), C.c_int32(bits)))
    if nbytes < 0:
        raise HsqError(A.HSQ_ERR_INVALID_ARGUMENT, "hsq_packed_bytes: bad arguments")
    data = alloc(max(nbytes, 4))[:max(nbytes, 4)]
    lens = alloc(4 * max(reads.n, 1)).view(np.int32)[:reads.n]
    n_exc = C.c_int64(0)
    rc_ = reads.c()
    cap = 0
    exc = (None, None, None)
    while True:
        st = lib.hsq_pack_reads(C.byref(rc_), C.c_int32(bits), A.ptr(data, A.u8p), A.ptr(lens, A.i32p), C.c_int64(cap), C.byref(n_exc),
                                A.ptr(exc[0], A.u32p), A.ptr(exc[1], A.u32p), A.ptr(exc[2], A.u8p), C.c_int32(threads))
        if st == A.HSQ_ERR_OVERFLOW and n_exc.value > cap:  # size the exception arrays and pack again
            cap = int(n_exc.value)
            exc = (alloc(4 * cap).view(np.uint32)[:cap], alloc(4 * cap).view(np.uint32)[:cap], alloc(cap)[:cap])
            continue
        if st != 0:
            raise HsqError(st, lib.hsq_last_error(None).decode())
        break
    k = int(n_exc.value)
    return PackedReadsHost(bits, data[:nbytes], lens, *(None if e is None else e[:k] for e in exc))


def comm_unique_id():
    """hsq_comm_unique_id: the 128-byte id a communicator is built from (rank 0 makes it, the host program distributes it)."""
    lib = load()
    buf = (C.c_uint8 * A.HSQ_COMM_ID_BYTES)()
    st = lib.hsq_comm_unique_id(buf)
    if st != 0:
        raise HsqError(st, lib.hsq_last_error(None).decode())
    return bytes(buf)


def dedup_extend_multi_gpu(contexts, pair_batches, ext_uid_len, lig_uid_len, extend_duplicates=False, cigar_cap=32, md_cap=48):
    """hsq_dedup_extend_multi_gpu: one process, one context per GPU, pair_batches[k] belongs to contexts[k]."""
    lib = load()
    n = len(contexts)
    outs = [DedupExtendResult(pb.n, ctx.n_probes, cigar_cap, md_cap) for ctx, pb in zip(contexts, pair_batches)]
    handles = (C.c_void_p * n)(*[ctx.h for ctx in contexts])
    pcs = (A.PairBatch * n)(*[pb.c() for pb in pair_batches])
    ocs = (A.DedupExtendOut * n)(*[o.c() for o in outs])
    st = lib.hsq_dedup_extend_multi_gpu(handles, C.c_int32(n), pcs, C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                        C.c_int32(1 if extend_duplicates else 0), ocs)
    for o, oc in zip(outs, ocs):
        o.n_extended = int(oc.n_extended)
        o.extend.cells = int(o._ext_c.cells)
    if st != 0:
        msgs = [lib.hsq_last_error(ctx.h).decode() for ctx in contexts]
        raise HsqError(st, " | ".join(m for m in msgs if m))
    return outs


def quality_sums(quals, threads=0):
    """hsq_quality_sums (host helper): (short) sum(q - 33) per read of a quality SeqBatch."""
    lib = load()
    out = np.zeros(quals.n, dtype=np.int16)
    st = lib.hsq_quality_sums(A.ptr(quals.data, A.u8p), A.ptr(quals.offsets, A.i64p), C.c_int64(quals.n), A.ptr(out, A.i16p), C.c_int32(threads))
    if st != 0:
        raise HsqError(st, lib.hsq_last_error(None).decode())
    return out


def uid_keys(r1, r2, ext_uid_len, lig_uid_len):
    """hsq_uid_keys (host helper): uint64 UID code per pair from two PackedReadsHost."""
    lib = load()
    out = np.zeros(r1.n, dtype=np.uint64)
    c1, c2 = r1.c(), r2.c()
    st = lib.hsq_uid_keys(C.byref(c1), C.byref(c2), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len), out.ctypes.data_as(C.POINTER(C.c_uint64)))
    if st != 0:
        raise HsqError(st, "hsq_uid_keys failed (a UID holds a base other than A/C/G/T, or a read is shorter than its UID)")
    return out


class Context:
    """One hsq_ctx (one GPU)."""

    def __init__(self, device=0, scorer_=None, config=None):
        """config: dict of hsq_config fields (force_dp32, force_dedup_radix, extend_chunks, uniform_chunks, band)."""
        self.lib = load()
        h = C.c_void_p()
        if config:
            cfg = A.Config(**config)
            st = self.lib.hsq_ctx_create_ex(C.c_int32(device), C.byref(cfg), C.byref(h))
        else:
            st = self.lib.hsq_ctx_create(C.c_int32(device), C.byref(h))
        if st != 0:
            raise HsqError(st, self.lib.hsq_last_error(None).decode())
        self.h = h
        self.device = device
        self.n_probes = 0
        self._pinned = []
        self.set_scorer(scorer_ or scorer())

    def close(self):
        if getattr(self, "h", None):
            for p in self._pinned:
                self.lib.hsq_pinned_free(self.h, p)
            self._pinned = []
            self.lib.hsq_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, st):
        if st != 0:
            raise HsqError(st, self.lib.hsq_last_error(self.h).decode())

    def configure(self, **config):
        cfg = A.Config(**config)
        self._check(self.lib.hsq_ctx_configure(self.h, C.byref(cfg)))

    def pack_reads(self, reads, bits=2, pinned=False, threads=0):
        return pack_reads(reads, bits, alloc=self.pinned_array if pinned else None, threads=threads)

    def unpack_reads(self, packed):
        """hsq_unpack_reads: packed words -> upper-case ASCII on the device (round-trip check of the packed layout)."""
        total = int(packed.len.sum())
        data = np.zeros(max(total, 1), dtype=np.uint8)
        off = np.zeros(packed.n + 1, dtype=np.int64)
        pc = packed.c()
        self._check(self.lib.hsq_unpack_reads(self.h, C.byref(pc), A.ptr(data, A.u8p), A.ptr(off, A.i64p)))
        return A.SeqBatch(data[:total], off)

    def extend_packed(self, pairs, out=None, pool_cap=None, fields=None, cigar_cap=0, md_cap=0):
        """hsq_extend_packed: packed reads in, integer columns + CIGAR / MD pool out."""
        out = out or PooledExtendResult(pairs.n, pool_cap if pool_cap is not None else 2 * pairs.n * 48, fields, cigar_cap=cigar_cap, md_cap=md_cap)
        pc, oc = pairs.c(), out.c()
        st = self.lib.hsq_extend_packed(self.h, C.byref(pc), C.byref(oc))
        out.finish()
        self._check(st)
        return out

    def last_timing(self):
        ms, n = C.c_float(0), C.c_int32(0)
        self.lib.hsq_last_timing(self.h, C.byref(ms), C.byref(n))
        return ms.value, n.value

    def band_stats(self):
        """hsq_band_stats: (read ends aligned inside the band, of them redone with full matrices) of the last extend call."""
        a, b = C.c_int64(0), C.c_int64(0)
        self.lib.hsq_band_stats(self.h, C.byref(a), C.byref(b))
        return int(a.value), int(b.value)

    def last_spans(self):
        """hsq_last_spans: kernel-only device times (ms) of the last call's kernel groups; None where the call has no such span."""
        ms = (C.c_float * 4)()
        self.lib.hsq_last_spans(self.h, ms)
        return [None if v < 0 else float(v) for v in ms]

    def pinned_array(self, nbytes):
        """uint8 numpy array backed by pinned host memory owned by this context."""
        p = C.c_void_p()
        self._check(self.lib.hsq_pinned_alloc(self.h, C.c_size_t(nbytes), C.byref(p)))
        buf = (C.c_uint8 * max(1, nbytes)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=np.uint8, count=nbytes)
        self._pinned.append(p)  # freed in close()
        return arr

    def set_scorer(self, sc):
        self.scorer = sc
        self._check(self.lib.hsq_set_scorer(self.h, C.byref(sc)))

    def set_probes(self, probes: ProbeTableHost):
        t = probes.c()
        self._check(self.lib.hsq_set_probes(self.h, C.byref(t)))
        self.n_probes = probes.n

    def align_batch(self, refs, qrys, sc=None, flags=0, cols_cap=640, md_cap=640):
        refs = refs if isinstance(refs, A.SeqBatch) else A.SeqBatch.from_strings(refs)
        qrys = qrys if isinstance(qrys, A.SeqBatch) else A.SeqBatch.from_strings(qrys)
        sc = sc or scorer()
        out = AlignResult(refs.n, cols_cap, md_cap)
        rc, qc, oc = refs.c(), qrys.c(), out.c()
        st = self.lib.hsq_align_batch(self.h, C.byref(rc), C.byref(qc), C.byref(sc), C.c_uint32(flags), C.byref(oc))
        out.finish()
        self._check(st)
        return out

    def align_scores_noalloc(self, d_ref_data, d_ref_offsets, ref_bytes, d_qry_data, d_qry_offsets, qry_bytes, n, max_ref_len, max_qry_len,
                             d_score, d_n_cols, d_matches, d_mismatches, d_status, sc=None, flags=0):
        """hsq_align_scores_noalloc: every d_* argument is a DEVICE address (int); the call allocates nothing."""
        sc = sc or scorer()
        v = C.c_void_p
        self._check(self.lib.hsq_align_scores_noalloc(self.h, v(d_ref_data), v(d_ref_offsets), C.c_int64(ref_bytes), v(d_qry_data), v(d_qry_offsets),
                                                      C.c_int64(qry_bytes), C.c_int64(n), C.c_int32(max_ref_len), C.c_int32(max_qry_len), C.byref(sc),
                                                      C.c_uint32(flags), v(d_score), v(d_n_cols), v(d_matches), v(d_mismatches), v(d_status)))

    def variable_length_uid(self, reads, primers, sc=None, flags=0):
        """hsq_variable_length_uid: per item the UID prefix length of the read, -1 where the reference returns null."""
        reads = reads if isinstance(reads, A.SeqBatch) else A.SeqBatch.from_strings(reads)
        primers = primers if isinstance(primers, A.SeqBatch) else A.SeqBatch.from_strings(primers)
        sc = sc or scorer()
        out = np.zeros(reads.n, dtype=np.int32)
        rc, pc = reads.c(), primers.c()
        self._check(self.lib.hsq_variable_length_uid(self.h, C.byref(rc), C.byref(pc), C.byref(sc), C.c_uint32(flags), A.ptr(out, A.i32p)))
        return out

    def trim_batch(self, reads, quals, first, last, three_prime=True, out=None):
        """out: optional (out_seq, out_qual, out_offsets) arrays to write into (e.g. pinned), each large enough."""
        reads = reads if isinstance(reads, A.SeqBatch) else A.SeqBatch.from_strings(reads)
        quals = quals if isinstance(quals, A.SeqBatch) else A.SeqBatch.from_strings(quals)
        if out is not None:
            out_seq, out_qual, out_off = out
        else:
            out_seq = np.zeros(max(1, len(reads.data)), dtype=np.uint8)
            out_qual = np.zeros(max(1, len(reads.data)), dtype=np.uint8)
            out_off = np.zeros(reads.n + 1, dtype=np.int64)
        rc = reads.c()
        self._check(self.lib.hsq_trim_batch(self.h, C.byref(rc), A.ptr(quals.data, A.u8p), C.c_int32(first), C.c_int32(last),
                                            C.c_int32(1 if three_prime else 0), A.ptr(out_seq, A.u8p),
                                            A.ptr(out_qual, A.u8p), A.ptr(out_off, A.i64p)))
        total = int(out_off[-1])
        return A.SeqBatch(out_seq[:total], out_off), A.SeqBatch(out_qual[:total], out_off)

    def dedup_batch(self, pairs: PairBatchHost, ext_uid_len, lig_uid_len, n_probes=None):
        n_probes = n_probes or self.n_probes
        is_best = np.zeros(pairs.n, dtype=np.uint8)
        group_size = np.zeros(pairs.n, dtype=np.int32)
        stats = ProbeStatsHost(n_probes)
        pc, sc = pairs.c(), stats.c()
        self._check(self.lib.hsq_dedup_batch(self.h, C.byref(pc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                             A.ptr(is_best, A.u8p), A.ptr(group_size, A.i32p), C.byref(sc)))
        return is_best, group_size, stats

    def index_bam(self, data, cap=None, want_payload=False):
        """hsq_index_bam: level-0 BGZF BAM bytes -> BamIndexResult (offsets into the uncompressed stream, kept on the device)."""
        from .batches import BamIndexResult
        data = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8) if isinstance(data, (bytes, bytearray)) else data)
        cap = cap if cap is not None else max(1, len(data) // 37)
        res = BamIndexResult(cap, len(data) if want_payload else 0)
        rc_ = res.c()
        st = self.lib.hsq_index_bam(self.h, A.ptr(data, A.u8p), C.c_int64(len(data)), C.byref(rc_))
        res.finish()
        self._check(st)
        return res

    def dedup_batch_report(self, pairs: PairBatchHost, ext_uid_len, lig_uid_len, n_probes=None):
        """hsq_dedup_batch_report: hsq_dedup_batch + the per-probe UID composition tables [n_probes, uid_len, 4] (A, C, G, T):
        returns (is_best, group_size, stats, unique_by_position, weighted_by_position)."""
        n_probes = n_probes or self.n_probes
        L = ext_uid_len + lig_uid_len
        is_best = np.zeros(pairs.n, dtype=np.uint8)
        group_size = np.zeros(pairs.n, dtype=np.int32)
        stats = ProbeStatsHost(n_probes)
        unique = np.zeros((n_probes, L, 4), dtype=np.int32)
        weighted = np.zeros((n_probes, L, 4), dtype=np.int32)
        rep = A.UidReport(0, A.ptr(unique, A.i32p), A.ptr(weighted, A.i32p))
        pc, sc = pairs.c(), stats.c()
        self._check(self.lib.hsq_dedup_batch_report(self.h, C.byref(pc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                                    A.ptr(is_best, A.u8p), A.ptr(group_size, A.i32p), C.byref(sc), C.byref(rep)))
        assert rep.uid_len == L
        return is_best, group_size, stats, unique, weighted

    def pair_reasons(self, result, extend_duplicates=False):
        """hsq_pair_reasons: one HSQ_PAIR_* code per pair of a DedupExtendResult (host helper)."""
        reason = np.zeros(result.n, dtype=np.uint8)
        oc = result.c()
        oc.n_extended = result.n_extended
        self._check(self.lib.hsq_pair_reasons(C.byref(oc), C.c_int64(result.n), C.c_int32(1 if extend_duplicates else 0), A.ptr(reason, A.u8p)))
        return reason

    def dedup_extend_batch(self, pairs: PairBatchHost, ext_uid_len, lig_uid_len, extend_duplicates=False, cigar_cap=32, md_cap=48,
                           n_probes=None, out=None):
        """hsq_dedup_extend_batch: one probe task body (reduce by UID, extend the unique pairs) for the whole batch."""
        out = out or DedupExtendResult(pairs.n, n_probes or self.n_probes, cigar_cap, md_cap)
        pc, oc = pairs.c(), out.c()
        st = self.lib.hsq_dedup_extend_batch(self.h, C.byref(pc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                             C.c_int32(1 if extend_duplicates else 0), C.byref(oc))
        out.finish()
        self._check(st)
        return out

    def flag_n_pairs(self, pairs: PairBatchHost):
        """hsq_flag_n_pairs: 1 where either read of the pair holds an N (…Identification.java:815-821)."""
        out = np.zeros(pairs.n, dtype=np.uint8)
        pc = pairs.c()
        self._check(self.lib.hsq_flag_n_pairs(self.h, C.byref(pc), A.ptr(out, A.u8p)))
        return out

    def extend_selected(self, pairs: PairBatchHost, keep, cigar_cap=32, md_cap=48, out=None, extended_pair=None):
        """hsq_extend_selected: extend only the pairs with keep[k] != 0; returns (n_extended, extended_pair, ExtendResult)."""
        out = out or ExtendResult(pairs.n, cigar_cap, md_cap)
        keep = np.ascontiguousarray(keep, dtype=np.uint8)
        lst = extended_pair if extended_pair is not None else np.zeros(pairs.n, dtype=np.int32)
        m = C.c_int64(0)
        pc, oc = pairs.c(), out.c()
        st = self.lib.hsq_extend_selected(self.h, C.byref(pc), A.ptr(keep, A.u8p), C.byref(m), A.ptr(lst, A.i32p), C.byref(oc))
        out.finish()
        self._check(st)
        return int(m.value), lst, out

    def extend_records(self, pairs: PairBatchHost, ext: ExtendResult, duplicate=None, out_cap=None, out=None):
        """hsq_extend_records: FLAG / PNEXT / read string / qualities of the records of the pairs `ext` extended.
        out: optional RecordResult to write into (e.g. with pinned arrays)."""
        cap = out.cap if out is not None else out_cap if out_cap is not None else int(len(pairs.r1.data) + len(pairs.r2.data))
        dup = None if duplicate is None else np.ascontiguousarray(duplicate, dtype=np.uint8)
        out = out if out is not None else RecordResult(pairs.n, cap)
        pc, ec, oc = pairs.c(), ext.c(), out.c()
        self._check(self.lib.hsq_extend_records(self.h, C.byref(pc), C.byref(ec), A.ptr(dup, A.u8p), C.byref(oc), C.c_int64(cap)))
        return out

    def merge_pairs(self, negative, up, down, upq, downq, up_start, down_start, out_cap=None):
        """hsq_merge_pairs: returns (merged sequences, merged qualities) as SeqBatch; sizes itself with a second call."""
        mb, keep = merge_args(negative, up, down, upq, downq, up_start, down_start)
        n = mb.n
        cap = out_cap if out_cap is not None else int(len(keep[0].data) + len(keep[1].data)) + 16
        off = np.zeros(n + 1, dtype=np.int64)
        for _ in range(2):
            seq = np.zeros(max(1, cap), dtype=np.uint8)
            qual = np.zeros(max(1, cap), dtype=np.uint8)
            st = self.lib.hsq_merge_pairs(self.h, C.byref(mb), A.ptr(seq, A.u8p), A.ptr(qual, A.u8p), A.ptr(off, A.i64p), C.c_int64(cap))
            if st == A.HSQ_ERR_OVERFLOW and out_cap is None:
                cap = int(off[-1])
                continue
            break
        self._check(st)
        total = int(off[-1])
        return A.SeqBatch(seq[:total], off), A.SeqBatch(qual[:total], off.copy())

    def set_probe_ranges(self, ranges: ProbeRangesHost):
        """hsq_set_probe_ranges: the per (contig, strand) RangeMaps, device resident."""
        rc = ranges.c()
        self._check(self.lib.hsq_set_probe_ranges(self.h, C.byref(rc)))

    def assign_probes(self, records: MappedRecordsHost, max_probes=4):
        """hsq_assign_probes: ReadToProbeAssigner.assignProbesToReads for a batch of mapped records."""
        out = AssignResult(records.n, max_probes)
        rc, oc = records.c(), out.c()
        self._check(self.lib.hsq_assign_probes(self.h, C.byref(rc), C.byref(oc)))
        return out

    def index_fastq(self, text, cap=None, out=None):
        """hsq_index_fastq: record index of FASTQ text (htsjdk FastqReader grammar).  out: a FastqIndexResult to fill (e.g. pinned)."""
        text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
        if cap is None:
            cap = len(text) // 8 + 1  # a record is at least "@x\nA\n+\nI\n"
        out = out or FastqIndexResult(cap)
        oc = out.c()
        self._check(self.lib.hsq_index_fastq(self.h, A.ptr(text, A.u8p), C.c_int64(len(text)), C.byref(oc)))
        out.n = int(oc.n)
        return out

    def trim_fastq(self, text, first, last, three_prime=True, out=None):
        """hsq_trim_fastq: FastqReadTrimmer.trimReads, FASTQ text in, trimmed FASTQ text out (uint8 array) and the record count."""
        text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
        out = np.zeros(len(text) + 2, dtype=np.uint8) if out is None else out  # a record only shrinks; a missing last line feed is added
        nb, nrec = C.c_int64(0), C.c_int64(0)
        self._check(self.lib.hsq_trim_fastq(self.h, A.ptr(text, A.u8p), C.c_int64(len(text)), C.c_int32(first), C.c_int32(last),
                                            C.c_int32(1 if three_prime else 0), A.ptr(out, A.u8p), C.c_int64(len(out)), C.byref(nb), C.byref(nrec)))
        return out[:nb.value], int(nrec.value)

    def index_sam(self, text, cap=None, out=None):
        """hsq_index_sam: field index of SAM text (the columns the dedup path reads off a SAMRecord).  out: a SamIndexResult to fill."""
        text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
        if cap is None:
            cap = len(text) // 22 + 1  # an alignment line is at least 11 one-byte columns, 10 tabs and a line feed
        out = out or SamIndexResult(cap)
        oc = out.c()
        self._check(self.lib.hsq_index_sam(self.h, A.ptr(text, A.u8p), C.c_int64(len(text)), C.byref(oc)))
        return out.take(oc)

    def sam_mapped_records(self, text, contig_names, cap=None, alt_cap=None):
        """hsq_sam_mapped_records: SAM text -> the columns hsq_assign_probes takes (records().n may be smaller than the line count)."""
        text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
        names = contig_names if isinstance(contig_names, A.SeqBatch) else A.SeqBatch.from_strings(contig_names)
        if cap is None:
            cap = len(text) // 22 + 1
        if alt_cap is None:
            alt_cap = len(text) // 8 + 1  # an XA entry is at least "c,1,*,0;"
        out = SamRecordsResult(cap, alt_cap)
        nc, oc = names.c(), out.c()
        self._check(self.lib.hsq_sam_mapped_records(self.h, A.ptr(text, A.u8p), C.c_int64(len(text)), C.byref(nc), C.byref(oc)))
        return out.take(oc)

    def coordinate_order(self, ref_index, pos, flags, mapq, mate_ref_index, mate_pos, tlen, names):
        """hsq_coordinate_order: order of the final coordinate sort (SAMRecordCoordinateComparator)."""
        keep = _coord_args(ref_index, pos, flags, mapq, mate_ref_index, mate_pos, tlen, names)
        order = np.zeros(len(keep[0][0]), dtype=np.int32)
        self._check(self.lib.hsq_coordinate_order(self.h, C.byref(keep[2]), A.ptr(order, A.i32p)))
        return order

    def join_names(self, join: NameJoinHost):
        """hsq_join_names: both sorts and the merge loop of FastqAndBamFileMerger's MergedSamIterator."""
        out = NameJoinResult(join.n)
        jc, oc = join.c(), out.c()
        self._check(self.lib.hsq_join_names(self.h, C.byref(jc), C.byref(oc)))
        return out.take(oc)

    def gather_rows(self, text, off, length, out_cap=None):
        """hsq_gather_rows: rows text[off[e] : off[e] + length[e]] as a ragged batch."""
        text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int64)
        length = np.ascontiguousarray(length, dtype=np.int32)
        n = len(off)
        if out_cap is None:
            out_cap = int(length.astype(np.int64).sum())
        data = np.zeros(max(out_cap, 1), dtype=np.uint8)
        offsets = np.zeros(n + 1, dtype=np.int64)
        self._check(self.lib.hsq_gather_rows(self.h, A.ptr(text, A.u8p), C.c_int64(len(text)), A.ptr(off, A.i64p), A.ptr(length, A.i32p),
                                             C.c_int64(n), A.ptr(data, A.u8p), A.ptr(offsets, A.i64p), C.c_int64(out_cap)))
        return A.SeqBatch(data[:offsets[-1]], offsets)

    def set_find_sequences(self, seqs, kmer=5, max_refs_per_kmer=15000):
        """hsq_set_find_sequences: SimpleMapper index of the sequences to find (FastqFindAndTrimTool.java:231-237)."""
        seqs = seqs if isinstance(seqs, A.SeqBatch) else A.SeqBatch.from_strings(seqs)
        sc = seqs.c()
        self._check(self.lib.hsq_set_find_sequences(self.h, C.byref(sc), C.c_int32(kmer), C.c_int32(max_refs_per_kmer)))

    def find_candidates(self, reads, quals, leading_quality=0, trailing_quality=0, candidate_limit=10, min_hit_ratio=0.5, max_candidates=16):
        """hsq_find_candidates: quality trim, N test and SimpleMapper.getBestCandidateReferences per read."""
        reads = reads if isinstance(reads, A.SeqBatch) else A.SeqBatch.from_strings(reads)
        quals = quals if isinstance(quals, A.SeqBatch) else A.SeqBatch.from_strings(quals)
        assert (quals.offsets == reads.offsets).all(), "qualities must match the reads' lengths"
        out = FindResult(reads.n, max_candidates)
        rc, oc = reads.c(), out.c()
        self._check(self.lib.hsq_find_candidates(self.h, C.byref(rc), A.ptr(quals.data, A.u8p), C.c_int32(leading_quality),
                                                 C.c_int32(trailing_quality), C.c_int32(candidate_limit), C.c_double(min_hit_ratio), C.byref(oc)))
        return out

    def find_matches(self, reads, quals, leading_quality=0, trailing_quality=0, limits=(0, 0, 0, 0, 0), candidate_limit=10, min_hit_ratio=0.5,
                     max_candidates=16):
        """hsq_find_matches: candidates, one alignment per candidate and the five-limit test, all on the device.
        limits = (mismatches, insertion gaps, insertion bases, deletion gaps, deletion bases)."""
        reads = reads if isinstance(reads, A.SeqBatch) else A.SeqBatch.from_strings(reads)
        quals = quals if isinstance(quals, A.SeqBatch) else A.SeqBatch.from_strings(quals)
        assert (quals.offsets == reads.offsets).all(), "qualities must match the reads' lengths"
        out = FindMatchResult(reads.n, max_candidates)
        rc, oc, lc = reads.c(), out.c(), A.FindLimits(*[int(v) for v in limits])
        st = self.lib.hsq_find_matches(self.h, C.byref(rc), A.ptr(quals.data, A.u8p), C.c_int32(leading_quality), C.c_int32(trailing_quality),
                                       C.c_int32(candidate_limit), C.c_double(min_hit_ratio), C.byref(lc), C.byref(oc))
        out.finish()
        self._check(st)
        return out

    def dedup_packed(self, pairs: PackedPairsHost, ext_uid_len, lig_uid_len, n_probes=None):
        """hsq_dedup_packed: hsq_dedup_batch for packed reads (quality bytes, or host-made quality sums)."""
        n_probes = n_probes or self.n_probes
        is_best = np.zeros(pairs.n, dtype=np.uint8)
        group_size = np.zeros(pairs.n, dtype=np.int32)
        stats = ProbeStatsHost(n_probes)
        pc, sc = pairs.c(), stats.c()
        self._check(self.lib.hsq_dedup_packed(self.h, C.byref(pc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                              A.ptr(is_best, A.u8p), A.ptr(group_size, A.i32p), C.byref(sc)))
        return is_best, group_size, stats

    def dedup_extend_packed(self, pairs: PackedPairsHost, ext_uid_len, lig_uid_len, extend_duplicates=False, sharded=False, out=None, pool_cap=None,
                            n_probes=None):
        """hsq_dedup_extend_packed: the probe task for a packed batch, results pooled (out.pooled); sharded: collective over the communicator."""
        if out is None:
            out = DedupExtendResult(pairs.n, n_probes or self.n_probes, pooled=PooledExtendResult(pairs.n, pool_cap if pool_cap is not None else 2 * pairs.n * 48))
        pc, oc = pairs.c(), out.c()
        st = self.lib.hsq_dedup_extend_packed(self.h, C.byref(pc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                              C.c_int32(1 if extend_duplicates else 0), C.c_int32(1 if sharded else 0), C.byref(oc))
        out.finish()
        self._check(st)
        return out

    def dedup_extend_packed_keys(self, pairs: PackedPairsHost, uid, ext_uid_len, lig_uid_len, extend_duplicates=False, sharded=False, out=None,
                                 pool_cap=None, n_probes=None):
        """hsq_dedup_extend_packed_keys: the same probe task, reduction from (uid, quality sums, probe, read index) first, reads in chunks."""
        if out is None:
            out = DedupExtendResult(pairs.n, n_probes or self.n_probes, pooled=PooledExtendResult(pairs.n, pool_cap if pool_cap is not None else 2 * pairs.n * 48))
        uid = np.ascontiguousarray(uid, dtype=np.uint64)
        pc, oc = pairs.c(), out.c()
        st = self.lib.hsq_dedup_extend_packed_keys(self.h, C.byref(pc), uid.ctypes.data_as(C.POINTER(C.c_uint64)), C.c_int32(ext_uid_len),
                                                   C.c_int32(lig_uid_len), C.c_int32(1 if extend_duplicates else 0), C.c_int32(1 if sharded else 0),
                                                   C.byref(oc))
        out.finish()
        self._check(st)
        return out

    # ---- multi-GPU inside the library (NCCL) ----
    def comm_init_rank(self, id_bytes, rank, world):
        """hsq_comm_init_rank: join the communicator whose id rank 0 made with comm_unique_id()."""
        buf = (C.c_uint8 * A.HSQ_COMM_ID_BYTES).from_buffer_copy(bytes(id_bytes))
        self._check(self.lib.hsq_comm_init_rank(self.h, buf, C.c_int32(rank), C.c_int32(world)))

    def comm_init_torch(self, group=None):
        """One process per GPU under torch.distributed: rank 0 makes the id, broadcast, every rank joins."""
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        box = [comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0, group=group)
        self.comm_init_rank(box[0], rank, world)

    def dedup_sharded(self, pairs: PairBatchHost, ext_uid_len, lig_uid_len, n_probes=None, out=None):
        """hsq_dedup_sharded (collective): dedup over the union of all ranks' pairs; answers for this rank's pairs."""
        n_probes = n_probes or self.n_probes
        is_best, group_size = out if out is not None else (np.zeros(pairs.n, dtype=np.uint8), np.zeros(pairs.n, dtype=np.int32))
        stats = ProbeStatsHost(n_probes)
        pc, sc = pairs.c(), stats.c()
        self._check(self.lib.hsq_dedup_sharded(self.h, C.byref(pc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                               A.ptr(is_best, A.u8p), A.ptr(group_size, A.i32p), C.byref(sc)))
        return is_best, group_size, stats

    def dedup_extend_sharded(self, pairs: PairBatchHost, ext_uid_len, lig_uid_len, extend_duplicates=False, cigar_cap=32, md_cap=48,
                             n_probes=None, out=None):
        """hsq_dedup_extend_sharded (collective)."""
        out = out or DedupExtendResult(pairs.n, n_probes or self.n_probes, cigar_cap, md_cap)
        pc, oc = pairs.c(), out.c()
        st = self.lib.hsq_dedup_extend_sharded(self.h, C.byref(pc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                               C.c_int32(1 if extend_duplicates else 0), C.byref(oc))
        out.finish()
        self._check(st)
        return out

    def dedup_keys(self, pairs: PairBatchHost, ext_uid_len, lig_uid_len, n_probes, origin_base, d_keys_ptr):
        """hsq_dedup_keys: key records of this rank's pairs into DEVICE memory at d_keys_ptr (24 B per pair)."""
        pc = pairs.c()
        self._check(self.lib.hsq_dedup_keys(self.h, C.byref(pc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                            C.c_int64(n_probes), C.c_uint32(origin_base), C.c_void_p(d_keys_ptr)))

    def dedup_reduce_keys(self, d_keys_ptr, n, uid_len, d_is_best_ptr, d_group_size_ptr, n_probes):
        """hsq_dedup_reduce_keys over n DEVICE key records; answers go to DEVICE memory; returns the per-probe counters."""
        stats = ProbeStatsHost(n_probes)
        sc = stats.c()
        self._check(self.lib.hsq_dedup_reduce_keys(self.h, C.c_void_p(d_keys_ptr), C.c_int64(n), C.c_int32(uid_len),
                                                   C.c_void_p(d_is_best_ptr), C.c_void_p(d_group_size_ptr), C.byref(sc)))
        return stats

    def extend_batch(self, pairs: PairBatchHost, cigar_cap=32, md_cap=48, out=None):
        out = out or ExtendResult(pairs.n, cigar_cap, md_cap)
        pc, oc = pairs.c(), out.c()
        st = self.lib.hsq_extend_batch(self.h, C.byref(pc), C.byref(oc))
        out.finish()
        self._check(st)
        return out

    def extend_stage(self, pairs: PairBatchHost):
        pc = pairs.c()
        self._check(self.lib.hsq_extend_stage(self.h, C.byref(pc)))

    def extend_run(self, cigar_cap=32, md_cap=48):
        self._check(self.lib.hsq_extend_run(self.h, C.c_int32(cigar_cap), C.c_int32(md_cap)))

    def extend_fetch(self, out: ExtendResult):
        oc = out.c()
        st = self.lib.hsq_extend_fetch(self.h, C.byref(oc))
        out.finish()
        self._check(st)
        return out

    def timer_start(self):
        self._check(self.lib.hsq_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float(0)
        self._check(self.lib.hsq_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def int32_peak(self, seconds=1.0):
        v = C.c_double(0)
        self._check(self.lib.hsq_int32_peak(self.h, C.c_double(seconds), C.byref(v)))
        return v.value

    def resolve_multi_probe(self, r1, r2, cand_offsets, cand_probe, ext_uid_len, lig_uid_len):
        r1 = r1 if isinstance(r1, A.SeqBatch) else A.SeqBatch.from_strings(r1)
        r2 = r2 if isinstance(r2, A.SeqBatch) else A.SeqBatch.from_strings(r2)
        cand_offsets = np.ascontiguousarray(cand_offsets, dtype=np.int64)
        cand_probe = np.ascontiguousarray(cand_probe, dtype=np.int32)
        best = np.zeros(r1.n, dtype=np.int32)
        m = np.zeros(max(1, len(cand_probe)), dtype=np.int32)
        e = np.zeros(max(1, len(cand_probe)), dtype=np.int32)
        cc = A.Candidates(r1.n, r1.c(), r2.c(), A.ptr(cand_offsets, A.i64p), A.ptr(cand_probe, A.i32p))
        self._check(self.lib.hsq_resolve_multi_probe(self.h, C.byref(cc), C.c_int32(ext_uid_len), C.c_int32(lig_uid_len),
                                                     A.ptr(best, A.i32p), A.ptr(m, A.i32p), A.ptr(e, A.i32p)))
        return best, m[:len(cand_probe)], e[:len(cand_probe)]
