"""Host-side batch containers (numpy) for the C ABI of include/hsq_b200.h.

They mirror the data the reference hands to its per-probe task
(HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:836-846): a parsed probe file
(COMMON/utils/probeinfo/Probe.java:53-69) and raw read pairs with qualities and FASTQ ordinals.
No compute happens here.
"""
import ctypes as C

import numpy as np

from . import _abi as A


class ProbeTableHost:
    def __init__(self, reverse_strand, ext_start, ext_stop, lig_start, lig_stop, target_start, target_stop, ext, lig,
                 target, probe_ids=None, chromosomes=None):
        self.reverse_strand = np.ascontiguousarray(reverse_strand, dtype=np.uint8)
        self.ext_start = np.ascontiguousarray(ext_start, dtype=np.int32)
        self.ext_stop = np.ascontiguousarray(ext_stop, dtype=np.int32)
        self.lig_start = np.ascontiguousarray(lig_start, dtype=np.int32)
        self.lig_stop = np.ascontiguousarray(lig_stop, dtype=np.int32)
        self.target_start = np.ascontiguousarray(target_start, dtype=np.int32)
        self.target_stop = np.ascontiguousarray(target_stop, dtype=np.int32)
        self.ext = ext if isinstance(ext, A.SeqBatch) else A.SeqBatch.from_strings(ext)
        self.lig = lig if isinstance(lig, A.SeqBatch) else A.SeqBatch.from_strings(lig)
        self.target = target if isinstance(target, A.SeqBatch) else A.SeqBatch.from_strings(target)
        self.n = len(self.reverse_strand)
        self.probe_ids = probe_ids
        self.chromosomes = chromosomes

    def c(self):
        return A.ProbeTable(self.n, A.ptr(self.reverse_strand, A.u8p), A.ptr(self.ext_start, A.i32p),
                            A.ptr(self.ext_stop, A.i32p), A.ptr(self.lig_start, A.i32p), A.ptr(self.lig_stop, A.i32p),
                            A.ptr(self.target_start, A.i32p), A.ptr(self.target_stop, A.i32p), self.ext.c(),
                            self.lig.c(), self.target.c())


class PairBatchHost:
    def __init__(self, probe_id, r1, r2, q1=None, q2=None, read_index=None):
        self.probe_id = np.ascontiguousarray(probe_id, dtype=np.int32)
        self.n = len(self.probe_id)
        self.read_index = None if read_index is None else np.ascontiguousarray(read_index, dtype=np.int32)
        self.r1 = r1 if isinstance(r1, A.SeqBatch) else A.SeqBatch.from_strings(r1)
        self.r2 = r2 if isinstance(r2, A.SeqBatch) else A.SeqBatch.from_strings(r2)
        self.q1 = None if q1 is None else (q1 if isinstance(q1, A.SeqBatch) else A.SeqBatch.from_strings(q1))
        self.q2 = None if q2 is None else (q2 if isinstance(q2, A.SeqBatch) else A.SeqBatch.from_strings(q2))

    def c(self):
        return A.PairBatch(self.n, A.ptr(self.probe_id, A.i32p), A.ptr(self.read_index, A.i32p), self.r1.c(),
                           self.r2.c(), A.ptr(None if self.q1 is None else self.q1.data, A.u8p),
                           A.ptr(None if self.q2 is None else self.q2.data, A.u8p))

    def head(self, k):
        """First k pairs (views, no copies)."""
        k = min(k, self.n)

        def cut(sb):
            if sb is None:
                return None
            return A.SeqBatch(sb.data[:sb.offsets[k]], sb.offsets[:k + 1])

        return PairBatchHost(self.probe_id[:k], cut(self.r1), cut(self.r2), cut(self.q1), cut(self.q2),
                             None if self.read_index is None else self.read_index[:k])

    def subset(self, idx):
        idx = np.asarray(idx, dtype=np.int64)

        def take(sb):
            return None if sb is None else sb.take(idx)

        return PairBatchHost(self.probe_id[idx], take(self.r1), take(self.r2), take(self.q1), take(self.q2),
                             None if self.read_index is None else self.read_index[idx])


class ExtendResult:
    """Owns the output arrays of hsq_extend_batch for n pairs."""
    FIELDS_I32 = ("primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen", "cigar_len", "md_len")

    def __init__(self, n, cigar_cap=32, md_cap=48):
        self.n, self.cigar_cap, self.md_cap = n, cigar_cap, md_cap
        for f in self.FIELDS_I32:
            setattr(self, f, np.zeros(2 * n, dtype=np.int32))
        self.ok = np.zeros(2 * n, dtype=np.uint8)
        self.overflow = np.zeros(2 * n, dtype=np.uint8)
        self.pair_ok = np.zeros(n, dtype=np.uint8)
        self.cigar = np.zeros(2 * n * cigar_cap, dtype=np.uint8)
        self.md = np.zeros(2 * n * md_cap, dtype=np.uint8)
        self.cells = 0
        self._c = None

    def c(self):
        p = A.ptr
        self._c = A.ExtendOut(self.cigar_cap, self.md_cap, p(self.ok, A.u8p), p(self.primer_end, A.i32p),
                              p(self.score, A.i32p), p(self.pos, A.i32p), p(self.read_start, A.i32p),
                              p(self.read_stop, A.i32p), p(self.nm, A.i32p), p(self.tlen, A.i32p),
                              p(self.cigar_len, A.i32p), p(self.cigar, A.u8p), p(self.md_len, A.i32p), p(self.md, A.u8p),
                              p(self.overflow, A.u8p), p(self.pair_ok, A.u8p), 0)
        return self._c

    def finish(self):
        self.cells = int(self._c.cells)
        return self

    def head(self, m):
        """The first m pair slots as a result of their own (views)."""
        r = ExtendResult(0, self.cigar_cap, self.md_cap)
        r.n = m
        for f in self.FIELDS_I32 + ("ok", "overflow"):
            setattr(r, f, getattr(self, f)[:2 * m])
        r.pair_ok = self.pair_ok[:m]
        r.cigar = self.cigar[:2 * m * self.cigar_cap]
        r.md = self.md[:2 * m * self.md_cap]
        r.cells = self.cells
        return r

    def cigars(self):
        return A.slot_strings(self.cigar, self.cigar_len, self.cigar_cap)

    def mds(self):
        return A.slot_strings(self.md, self.md_len, self.md_cap)


class PackedReadsHost:
    """hsq_packed_reads: 2 / 4 bit words, lengths and (2 bit) the exception list.  Built by `_lib.pack_reads`."""

    def __init__(self, bits, data, lens, exc_read=None, exc_pos=None, exc_char=None):
        self.bits = bits
        self.data = data
        self.len = np.ascontiguousarray(lens, dtype=np.int32) if not isinstance(lens, np.ndarray) or lens.dtype != np.int32 else lens
        self.n = len(self.len)
        self.exc_read = np.zeros(0, dtype=np.uint32) if exc_read is None else exc_read
        self.exc_pos = np.zeros(0, dtype=np.uint32) if exc_pos is None else exc_pos
        self.exc_char = np.zeros(0, dtype=np.uint8) if exc_char is None else exc_char
        self.n_bases = int(self.len.sum(dtype=np.int64))  # the totals travel with the batch: the library need not add up the lengths

    @property
    def nbytes(self):
        return self.data.nbytes + self.len.nbytes + self.exc_read.nbytes + self.exc_pos.nbytes + self.exc_char.nbytes

    def c(self):
        ne = len(self.exc_read)
        return A.PackedReads(self.bits, self.n, A.ptr(self.data, A.u8p), A.ptr(self.len, A.i32p), ne,
                             A.ptr(self.exc_read if ne else None, A.u32p), A.ptr(self.exc_pos if ne else None, A.u32p),
                             A.ptr(self.exc_char if ne else None, A.u8p), self.data.nbytes, self.n_bases)


class PackedPairsHost:
    """hsq_packed_pairs.  q1 / q2: quality bytes back to back in read order (numpy uint8) or None; qsum1 / qsum2: int16 or None."""

    def __init__(self, probe_id, r1, r2, q1=None, q2=None, read_index=None, qsum1=None, qsum2=None):
        self.probe_id = probe_id if isinstance(probe_id, np.ndarray) and probe_id.dtype == np.int32 else np.ascontiguousarray(probe_id, dtype=np.int32)
        self.n = len(self.probe_id)
        self.r1, self.r2, self.q1, self.q2, self.qsum1, self.qsum2 = r1, r2, q1, q2, qsum1, qsum2
        self.read_index = None if read_index is None else np.ascontiguousarray(read_index, dtype=np.int32)

    def c(self):
        return A.PackedPairs(self.n, A.ptr(self.probe_id, A.i32p), A.ptr(self.read_index, A.i32p), self.r1.c(), self.r2.c(),
                             A.ptr(self.q1, A.u8p), A.ptr(self.q2, A.u8p), A.ptr(self.qsum1, A.i16p), A.ptr(self.qsum2, A.i16p))


class PooledExtendResult:
    """hsq_extend_pooled_out for n pairs: integer columns (those named in `fields`; the others are not fetched) + the CIGAR / MD pool."""
    FIELDS_I32 = ExtendResult.FIELDS_I32

    def __init__(self, n, pool_cap, fields=None, alloc=None, cigar_cap=0, md_cap=0):
        alloc = alloc or (lambda nbytes: np.zeros(nbytes, dtype=np.uint8))
        self.n, self.pool_cap, self.cigar_cap, self.md_cap = n, pool_cap, cigar_cap, md_cap
        want = set(fields) if fields is not None else set(self.FIELDS_I32) | {"ok", "pair_ok", "str_off"}
        for f in self.FIELDS_I32:
            setattr(self, f, alloc(8 * n).view(np.int32) if f in want else None)
        self.ok = alloc(2 * n) if "ok" in want else None
        self.pair_ok = alloc(n) if "pair_ok" in want else None
        self.str_off = alloc(16 * n).view(np.int64) if "str_off" in want else None
        self.pool = alloc(max(pool_cap, 1))
        self.pool_bytes = 0
        self.cells = 0
        self._c = None

    def nbytes(self):
        """Bytes that cross PCIe towards the host per call (columns fetched + pool bytes used)."""
        cols = [getattr(self, f) for f in self.FIELDS_I32 + ("ok", "pair_ok", "str_off")]
        return sum(a.nbytes for a in cols if a is not None) + int(self.pool_bytes)

    def c(self):
        p = A.ptr
        self._c = A.ExtendPooledOut(self.cigar_cap, self.md_cap, p(self.ok, A.u8p), p(self.primer_end, A.i32p), p(self.score, A.i32p),
                                    p(self.pos, A.i32p), p(self.read_start, A.i32p), p(self.read_stop, A.i32p), p(self.nm, A.i32p),
                                    p(self.tlen, A.i32p), p(self.cigar_len, A.i32p), p(self.md_len, A.i32p), p(self.str_off, A.i64p),
                                    p(self.pool, A.u8p), self.pool_cap, 0, p(self.pair_ok, A.u8p), 0)
        return self._c

    def finish(self):
        self.cells = int(self._c.cells)
        self.pool_bytes = int(self._c.pool_bytes)
        return self

    def _strings(self, which):
        off = self.str_off + (self.cigar_len if which == "md" else 0)
        ln = self.md_len if which == "md" else self.cigar_len
        buf = self.pool.tobytes()
        return [buf[o:o + l].decode("ascii") for o, l in zip(off.tolist(), ln.tolist())]

    def cigars(self):
        return self._strings("cigar")

    def mds(self):
        return self._strings("md")


class AlignResult:
    FIELDS_I32 = ("score", "n_cols", "first_q", "last_q", "first_r", "last_r", "matches", "mismatches",
                  "non_terminal_gaps", "edit_distance", "md_len", "rmd_len")

    def __init__(self, n, cols_cap=640, md_cap=640):
        self.n, self.cols_cap, self.md_cap = n, cols_cap, md_cap
        for f in self.FIELDS_I32:
            setattr(self, f, np.zeros(n, dtype=np.int32))
        self.ref_aln = np.zeros(n * cols_cap, dtype=np.uint8)
        self.qry_aln = np.zeros(n * cols_cap, dtype=np.uint8)
        self.ops = np.zeros(n * cols_cap, dtype=np.uint8)
        self.md = np.zeros(n * md_cap, dtype=np.uint8)
        self.rmd = np.zeros(n * md_cap, dtype=np.uint8)
        self.overflow = np.zeros(n, dtype=np.uint8)
        # FastqFindAndTrimTool's filter counters over the query-trimmed alignment:
        # mismatches, insertion gaps, insertion bases, deletion gaps, deletion bases
        self.find_counts = np.zeros((n, 5), dtype=np.int32)
        self.cells = 0
        self._c = None

    def c(self):
        p = A.ptr
        self._c = A.AlignOut(self.cols_cap, self.md_cap, p(self.score, A.i32p), p(self.n_cols, A.i32p),
                             p(self.ref_aln, A.u8p), p(self.qry_aln, A.u8p), p(self.ops, A.u8p), p(self.first_q, A.i32p),
                             p(self.last_q, A.i32p), p(self.first_r, A.i32p), p(self.last_r, A.i32p),
                             p(self.matches, A.i32p), p(self.mismatches, A.i32p), p(self.non_terminal_gaps, A.i32p),
                             p(self.edit_distance, A.i32p), p(self.md_len, A.i32p), p(self.md, A.u8p),
                             p(self.rmd_len, A.i32p), p(self.rmd, A.u8p), p(self.overflow, A.u8p), 0, p(self.find_counts, A.i32p))
        return self._c

    def finish(self):
        self.cells = int(self._c.cells)
        return self

    def ref_alns(self):
        return A.slot_strings(self.ref_aln, self.n_cols, self.cols_cap)

    def qry_alns(self):
        return A.slot_strings(self.qry_aln, self.n_cols, self.cols_cap)

    def op_strings(self):
        return A.slot_strings(self.ops, self.n_cols, self.cols_cap)

    def mds(self):
        return A.slot_strings(self.md, self.md_len, self.md_cap)

    def rmds(self):
        return A.slot_strings(self.rmd, self.rmd_len, self.md_cap)


class ProbeStatsHost:
    def __init__(self, n_probes):
        self.n_probes = n_probes
        self.unique_pairs = np.zeros(n_probes, dtype=np.int32)
        self.duplicate_pairs = np.zeros(n_probes, dtype=np.int32)
        self.max_pairs_per_uid = np.zeros(n_probes, dtype=np.int32)

    def c(self):
        return A.ProbeStats(self.n_probes, A.ptr(self.unique_pairs, A.i32p), A.ptr(self.duplicate_pairs, A.i32p),
                            A.ptr(self.max_pairs_per_uid, A.i32p))


class DedupExtendResult:
    """hsq_dedup_extend_out with host arrays for n pairs and n_probes probes."""

    def __init__(self, n, n_probes, cigar_cap=32, md_cap=48, pooled=None):
        """pooled: a PooledExtendResult for hsq_dedup_extend_packed (the slot result `extend` is then not allocated)."""
        self.n = n
        self.pooled = pooled
        self.is_best = np.zeros(n, dtype=np.uint8)
        self.group_size = np.zeros(n, dtype=np.int32)
        self.stats = ProbeStatsHost(n_probes)
        self.unique_unable_to_extend = np.zeros(n_probes, dtype=np.int32)
        self.duplicate_unable_to_extend = np.zeros(n_probes, dtype=np.int32)
        self.extended_pair = np.zeros(n, dtype=np.int32)
        self.extend = ExtendResult(0 if pooled is not None else n, cigar_cap, md_cap)
        self.n_extended = 0
        self._c = None

    def c(self):
        self._stats_c = self.stats.c()
        self._ext_c = self.extend.c()
        self._pool_c = None if self.pooled is None else self.pooled.c()
        self._c = A.DedupExtendOut(A.ptr(self.is_best, A.u8p), A.ptr(self.group_size, A.i32p), C.pointer(self._stats_c),
                                   A.ptr(self.unique_unable_to_extend, A.i32p), A.ptr(self.duplicate_unable_to_extend, A.i32p), 0,
                                   A.ptr(self.extended_pair, A.i32p), C.pointer(self._ext_c),
                                   None if self._pool_c is None else C.cast(C.pointer(self._pool_c), C.c_void_p))
        return self._c

    def finish(self):
        self.n_extended = int(self._c.n_extended)
        self.extend.cells = int(self._ext_c.cells)
        if self.pooled is not None:
            self.pooled.finish()
        return self


class RecordResult:
    """hsq_record_out with host arrays: FLAG, PNEXT, read string and base qualities of the two records of each pair."""

    def __init__(self, n, cap):
        self.n, self.cap = n, cap
        self.flag = np.zeros(2 * n, dtype=np.uint16)
        self.mate_pos = np.zeros(2 * n, dtype=np.int32)
        self.offsets = np.zeros(2 * n + 1, dtype=np.int64)
        self.seq = np.zeros(max(1, cap), dtype=np.uint8)
        self.qual = np.zeros(max(1, cap), dtype=np.uint8)

    def c(self):
        return A.RecordOut(C.cast(self.flag.ctypes.data, C.POINTER(C.c_uint16)), A.ptr(self.mate_pos, A.i32p), A.ptr(self.offsets, A.i64p),
                           A.ptr(self.seq, A.u8p), A.ptr(self.qual, A.u8p))

    def seqs(self):
        return A.SeqBatch(self.seq[:self.offsets[-1]], self.offsets)

    def quals(self):
        return A.SeqBatch(self.qual[:self.offsets[-1]], self.offsets)


def merge_args(negative, up, down, upq, downq, up_start, down_start):
    up = up if isinstance(up, A.SeqBatch) else A.SeqBatch.from_strings(up)
    down = down if isinstance(down, A.SeqBatch) else A.SeqBatch.from_strings(down)
    upq = upq if isinstance(upq, A.SeqBatch) else A.SeqBatch.from_strings(upq)
    downq = downq if isinstance(downq, A.SeqBatch) else A.SeqBatch.from_strings(downq)
    assert (upq.offsets == up.offsets).all() and (downq.offsets == down.offsets).all(), "qualities must match the sequences' lengths"
    neg = np.ascontiguousarray(negative, dtype=np.uint8)
    us = np.ascontiguousarray(up_start, dtype=np.int32)
    ds = np.ascontiguousarray(down_start, dtype=np.int32)
    mb = A.MergeBatch(up.n, A.ptr(neg, A.u8p), up.c(), down.c(), A.ptr(upq.data, A.u8p), A.ptr(downq.data, A.u8p), A.ptr(us, A.i32p),
                      A.ptr(ds, A.i32p))
    return mb, (up, down, upq, downq, neg, us, ds)


class ProbeRangesHost:
    """hsq_probe_ranges: search window, contig and strand of every probe (…Identification.java:285-331)."""

    def __init__(self, contig, reverse_strand, start, stop):
        self.contig = np.ascontiguousarray(contig, dtype=np.int32)
        self.reverse_strand = np.ascontiguousarray(reverse_strand, dtype=np.uint8)
        self.start = np.ascontiguousarray(start, dtype=np.int32)
        self.stop = np.ascontiguousarray(stop, dtype=np.int32)
        self.n = len(self.contig)
        assert len(self.reverse_strand) == len(self.start) == len(self.stop) == self.n

    @classmethod
    def from_probe_table(cls, probes, contig, bases_inside_extension_window=None, bases_inside_ligation_window=None):
        """getProbeSearchBoundaries (…Identification.java:287-300): Probe.getStart()/getStop() are the min / max of the four
        primer coordinates (COMMON/utils/probeinfo/Probe.java:75-81); the optional header values shrink the window."""
        coords = np.stack([probes.lig_start, probes.lig_stop, probes.ext_start, probes.ext_stop]).astype(np.int64)
        start, stop = coords.min(axis=0), coords.max(axis=0)
        if bases_inside_extension_window is not None:
            start = start + int(bases_inside_extension_window)
        if bases_inside_ligation_window is not None:
            stop = stop - int(bases_inside_ligation_window)
        return cls(contig, probes.reverse_strand, start, stop)

    def c(self):
        return A.ProbeRanges(self.n, A.ptr(self.contig, A.i32p), A.ptr(self.reverse_strand, A.u8p), A.ptr(self.start, A.i32p),
                             A.ptr(self.stop, A.i32p))


class MappedRecordsHost:
    """hsq_mapped_records: the fields ReadToProbeAssigner.assignProbesToReads reads off a SAMRecord (…Identification.java:439-576).

    alt = None or (alt_offsets[n+1], alt_contig, alt_negative, alt_start, alt_stop)."""

    def __init__(self, contig, negative_strand, second_of_pair, start, end, read_index, alt=None):
        self.contig = np.ascontiguousarray(contig, dtype=np.int32)
        self.negative_strand = np.ascontiguousarray(negative_strand, dtype=np.uint8)
        self.second_of_pair = np.ascontiguousarray(second_of_pair, dtype=np.uint8)
        self.start = np.ascontiguousarray(start, dtype=np.int32)
        self.end = np.ascontiguousarray(end, dtype=np.int32)
        self.read_index = np.ascontiguousarray(read_index, dtype=np.int32)
        self.n = len(self.contig)
        if alt is None:
            self.alt_offsets = self.alt_contig = self.alt_negative = self.alt_start = self.alt_stop = None
        else:
            self.alt_offsets = np.ascontiguousarray(alt[0], dtype=np.int64)
            self.alt_contig = np.ascontiguousarray(alt[1], dtype=np.int32)
            self.alt_negative = np.ascontiguousarray(alt[2], dtype=np.uint8)
            self.alt_start = np.ascontiguousarray(alt[3], dtype=np.int32)
            self.alt_stop = np.ascontiguousarray(alt[4], dtype=np.int32)
            assert len(self.alt_offsets) == self.n + 1

    def c(self):
        return A.MappedRecords(self.n, A.ptr(self.contig, A.i32p), A.ptr(self.negative_strand, A.u8p), A.ptr(self.second_of_pair, A.u8p),
                               A.ptr(self.start, A.i32p), A.ptr(self.end, A.i32p), A.ptr(self.read_index, A.i32p),
                               A.ptr(self.alt_offsets, A.i64p), A.ptr(self.alt_contig, A.i32p), A.ptr(self.alt_negative, A.u8p),
                               A.ptr(self.alt_start, A.i32p), A.ptr(self.alt_stop, A.i32p))


class AssignResult:
    """hsq_assign_out with host arrays."""

    def __init__(self, n, max_probes=4):
        self.n, self.max_probes = n, max_probes
        self.n_probes = np.zeros(n, dtype=np.int32)
        self.probes = np.full((n, max_probes), -1, dtype=np.int32)
        self.mate = np.zeros(n, dtype=np.int32)
        self.n_common = np.zeros(n, dtype=np.int32)
        self.common = np.full((n, max_probes), -1, dtype=np.int32)
        self.overflow = np.zeros(n, dtype=np.uint8)

    def c(self):
        return A.AssignOut(self.max_probes, A.ptr(self.n_probes, A.i32p), A.ptr(self.probes, A.i32p), A.ptr(self.mate, A.i32p),
                           A.ptr(self.n_common, A.i32p), A.ptr(self.common, A.i32p), A.ptr(self.overflow, A.u8p))

    def probe_sets(self):
        return [tuple(self.probes[k, :self.n_probes[k]]) for k in range(self.n)]

    def common_sets(self):
        return [tuple(self.common[k, :self.n_common[k]]) for k in range(self.n)]


class FindResult:
    """hsq_find_out with host arrays."""

    def __init__(self, n, max_candidates=16):
        self.n, self.max_candidates = n, max_candidates
        self.lead_trimmed = np.zeros(n, dtype=np.int32)
        self.trail_trimmed = np.zeros(n, dtype=np.int32)
        self.has_n = np.zeros(n, dtype=np.uint8)
        self.n_candidates = np.zeros(n, dtype=np.int32)
        self.candidates = np.full((n, max_candidates), -1, dtype=np.int32)
        self.hits = np.zeros((n, max_candidates), dtype=np.int32)
        self.overflow = np.zeros(n, dtype=np.uint8)
        self.order_dependent = np.zeros(n, dtype=np.uint8)

    def c(self):
        return A.FindOut(self.max_candidates, A.ptr(self.lead_trimmed, A.i32p), A.ptr(self.trail_trimmed, A.i32p), A.ptr(self.has_n, A.u8p),
                         A.ptr(self.n_candidates, A.i32p), A.ptr(self.candidates, A.i32p), A.ptr(self.hits, A.i32p),
                         A.ptr(self.overflow, A.u8p), A.ptr(self.order_dependent, A.u8p))

    def candidate_lists(self):
        return [[(int(self.candidates[k, j]), int(self.hits[k, j])) for j in range(self.n_candidates[k])] for k in range(self.n)]


class FindMatchResult:
    """hsq_find_match_out with host arrays."""

    def __init__(self, n, max_candidates=16):
        self.n, self.max_candidates = n, max_candidates
        self.cand = FindResult(n, max_candidates)
        self.counts = np.zeros((n, max_candidates, 5), dtype=np.int32)
        self.accepted = np.zeros((n, max_candidates), dtype=np.uint8)
        self.n_matches = np.zeros(n, dtype=np.int32)
        self.cells = 0
        self._c = None

    def c(self):
        self._c = A.FindMatchOut(self.cand.c(), A.ptr(self.counts, A.i32p), A.ptr(self.accepted, A.u8p), A.ptr(self.n_matches, A.i32p), 0)
        return self._c

    def finish(self):
        self.cells = int(self._c.cells)
        return self


class FastqIndexResult:
    """hsq_fastq_index with host arrays: where header, bases and qualities of every record sit in the FASTQ text."""

    def __init__(self, cap, alloc=None):
        """alloc(nbytes) -> uint8 array, e.g. Context.pinned_array (default: pageable numpy memory)."""
        self.cap, self.n = int(cap), 0
        new = lambda dt: (np.zeros(cap, dtype=dt) if alloc is None else alloc(cap * np.dtype(dt).itemsize).view(dt))
        self.header_off, self.header_len = new(np.int64), new(np.int32)
        self.seq_off, self.seq_len = new(np.int64), new(np.int32)
        self.qual_off, self.qual_len = new(np.int64), new(np.int32)

    def c(self):
        return A.FastqIndex(self.cap, 0, A.ptr(self.header_off, A.i64p), A.ptr(self.header_len, A.i32p), A.ptr(self.seq_off, A.i64p),
                            A.ptr(self.seq_len, A.i32p), A.ptr(self.qual_off, A.i64p), A.ptr(self.qual_len, A.i32p))

    FIELDS = ("header_off", "header_len", "seq_off", "seq_len", "qual_off", "qual_len")

    def head(self):
        return {f: getattr(self, f)[:self.n] for f in self.FIELDS}


class SamIndexResult:
    """hsq_sam_index with host planes; `<field>_off`, `<field>_len` and the numeric columns as attributes after take()."""

    def __init__(self, cap, alloc=None):
        """alloc(nbytes) -> uint8 array, e.g. Context.pinned_array (default: pageable numpy memory)."""
        self.cap, self.n, self.n_header_lines = int(cap), 0, 0
        nf, nv = len(A.SAM_FIELDS), len(A.SAM_FIELDS) + len(A.SAM_VALUES)
        if alloc is None:
            self.off = np.zeros((nf, self.cap), dtype=np.int64)
            self.val = np.zeros((nv, self.cap), dtype=np.int32)
        else:
            self.off = alloc(nf * self.cap * 8).view(np.int64).reshape(nf, self.cap)
            self.val = alloc(nv * self.cap * 4).view(np.int32).reshape(nv, self.cap)

    def c(self):
        return A.SamIndex(self.cap, 0, 0, A.ptr(self.off, A.i64p), A.ptr(self.val, A.i32p))

    def take(self, oc):
        self.n, self.n_header_lines = int(oc.n), int(oc.n_header_lines)
        for f, name in enumerate(A.SAM_FIELDS):
            setattr(self, name + "_off", self.off[f, :self.n])
            setattr(self, name + "_len", self.val[f, :self.n])
        for v, name in enumerate(A.SAM_VALUES):
            setattr(self, name, self.val[len(A.SAM_FIELDS) + v, :self.n])
        return self

    def columns(self):
        return [n + "_off" for n in A.SAM_FIELDS] + [n + "_len" for n in A.SAM_FIELDS] + list(A.SAM_VALUES)


class SamRecordsResult:
    """hsq_sam_records_out with host arrays; records() gives the MappedRecordsHost hsq_assign_probes takes."""

    def __init__(self, cap, alt_cap):
        self.cap, self.alt_cap, self.n, self.n_alt = int(cap), int(alt_cap), 0, 0
        i32, u8 = (lambda m: np.zeros(m, dtype=np.int32)), (lambda m: np.zeros(m, dtype=np.uint8))
        self.sam_line, self.contig, self.start, self.end, self.read_index = i32(cap), i32(cap), i32(cap), i32(cap), i32(cap)
        self.negative_strand, self.second_of_pair = u8(cap), u8(cap)
        self.alt_offsets = np.zeros(cap + 1, dtype=np.int64)
        self.alt_contig, self.alt_start, self.alt_stop, self.alt_negative = i32(alt_cap), i32(alt_cap), i32(alt_cap), u8(alt_cap)

    def c(self):
        return A.SamRecordsOut(self.cap, self.alt_cap, 0, 0, A.ptr(self.sam_line, A.i32p), A.ptr(self.contig, A.i32p), A.ptr(self.negative_strand, A.u8p),
                               A.ptr(self.second_of_pair, A.u8p), A.ptr(self.start, A.i32p), A.ptr(self.end, A.i32p), A.ptr(self.read_index, A.i32p),
                               A.ptr(self.alt_offsets, A.i64p), A.ptr(self.alt_contig, A.i32p), A.ptr(self.alt_negative, A.u8p),
                               A.ptr(self.alt_start, A.i32p), A.ptr(self.alt_stop, A.i32p))

    def take(self, oc):
        self.n, self.n_alt = int(oc.n), int(oc.n_alt)
        return self

    COLUMNS = ("sam_line", "contig", "negative_strand", "second_of_pair", "start", "end", "read_index")
    ALT_COLUMNS = ("alt_contig", "alt_negative", "alt_start", "alt_stop")

    def records(self):
        n, m = self.n, self.n_alt
        return MappedRecordsHost(self.contig[:n], self.negative_strand[:n], self.second_of_pair[:n], self.start[:n], self.end[:n], self.read_index[:n],
                                 (self.alt_offsets[:n + 1], self.alt_contig[:m], self.alt_negative[:m], self.alt_start[:m], self.alt_stop[:m]))


class BamIndexResult:
    """hsq_bam_index with host planes for `cap` records (and room for the uncompressed stream when want_payload)."""

    def __init__(self, cap, payload_cap=0):
        self.cap = cap
        self.off = np.zeros((len(A.HSQ_BAM_FIELDS), max(cap, 1)), dtype=np.int64)
        self.val = np.zeros((len(A.HSQ_BAM_FIELDS) + len(A.HSQ_BAM_VALUES), max(cap, 1)), dtype=np.int32)
        self.payload = np.zeros(payload_cap, dtype=np.uint8) if payload_cap else None
        self.n = self.n_ref = self.text_bytes = self.first_record = self.payload_bytes = self.exact_walk = 0
        self._c = None

    def c(self):
        self._c = A.BamIndex(self.cap, 0, 0, 0, 0, 0, 0 if self.payload is None else len(self.payload), A.ptr(self.payload, A.u8p),
                             A.ptr(self.off, A.i64p), A.ptr(self.val, A.i32p), 0)
        return self._c

    def finish(self):
        c = self._c
        self.n, self.n_ref, self.text_bytes, self.first_record = int(c.n), int(c.n_ref), int(c.text_bytes), int(c.first_record)
        self.payload_bytes, self.exact_walk = int(c.payload_bytes), int(c.exact_walk)
        return self

    def columns(self):
        """dict: <field>_off / <field>_len for the six parts, and the eight fixed columns, each cut to n records"""
        out = {}
        for f, name in enumerate(A.HSQ_BAM_FIELDS):
            out[name + "_off"] = self.off[f, :self.n]
            out[name + "_len"] = self.val[f, :self.n]
        for v, name in enumerate(A.HSQ_BAM_VALUES):
            out[name] = self.val[len(A.HSQ_BAM_FIELDS) + v, :self.n]
        return out


class NameJoinHost:
    """hsq_name_join: what FastqAndBamFileMerger reads off the SAM records and the two FASTQ files
    (HSQ/process/FastqAndBamFileMerger.java:107-158,213-345)."""

    def __init__(self, sam_names, sam_flags, sam_seq_len, fq1_headers, fq2_headers, fq1_seq_len=None, fq2_seq_len=None, sam_qual_len=None,
                 by_index=False, trimming_skipped=False, check_trim=True, three_prime=False, r1_first=0, r1_stop=0, r2_first=0, r2_stop=0):
        sb = lambda x: x if isinstance(x, A.SeqBatch) else A.SeqBatch.from_strings(x)
        i32 = lambda x: None if x is None else np.ascontiguousarray(x, dtype=np.int32)
        self.sam_names, self.fq1_headers, self.fq2_headers = sb(sam_names), sb(fq1_headers), sb(fq2_headers)
        self.sam_flags, self.sam_seq_len, self.sam_qual_len = i32(sam_flags), i32(sam_seq_len), i32(sam_qual_len)
        self.fq1_seq_len, self.fq2_seq_len = i32(fq1_seq_len), i32(fq2_seq_len)
        self.params = (int(by_index), int(trimming_skipped), int(check_trim), int(three_prime), int(r1_first), int(r1_stop), int(r2_first),
                       int(r2_stop))
        self.n = self.sam_names.n
        assert len(self.sam_flags) == self.n and len(self.sam_seq_len) == self.n

    def c(self):
        return A.NameJoin(self.sam_names.c(), A.ptr(self.sam_flags, A.i32p), A.ptr(self.sam_seq_len, A.i32p), A.ptr(self.sam_qual_len, A.i32p),
                          self.fq1_headers.c(), self.fq2_headers.c(), A.ptr(self.fq1_seq_len, A.i32p), A.ptr(self.fq2_seq_len, A.i32p),
                          *self.params)


class NameJoinResult:
    """hsq_name_join_out with host arrays."""

    def __init__(self, n):
        self.n = n
        self.n_emitted = 0
        self.total_matching_pairs = 0
        self.sam_record = np.full(n, -1, dtype=np.int32)
        self.fastq_record = np.full(n, -1, dtype=np.int32)
        self.name_len = np.full(n, -1, dtype=np.int32)
        self.status = np.zeros(n, dtype=np.uint8)

    def c(self):
        return A.NameJoinOut(0, 0, A.ptr(self.sam_record, A.i32p), A.ptr(self.fastq_record, A.i32p), A.ptr(self.name_len, A.i32p),
                             A.ptr(self.status, A.u8p))

    def take(self, oc):
        self.n_emitted, self.total_matching_pairs = int(oc.n_emitted), int(oc.total_matching_pairs)
        return self

    def emitted(self):
        m = self.n_emitted
        return self.sam_record[:m], self.fastq_record[:m], self.name_len[:m]


def scorer(match=1, mismatch=-4, gap_extend=-1, gap_start=-6):
    """SimpleAlignmentScorer defaults (COMMON/alignment/SimpleAlignmentScorer.java:23-26)."""
    return A.Scorer(match, mismatch, gap_extend, gap_start)


byref = C.byref
