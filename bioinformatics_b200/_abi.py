"""ctypes mirror of include/hsq_b200.h (struct layouts + small packing helpers).

Only layouts live here; no compute.  Both the product binding (`_lib.py`) and the test-only oracle
binding (the ctypes module next to the CPU restatement) use these definitions so that the same Python objects can be handed to
either side.
"""
import ctypes as C

import numpy as np

HSQ_OK = 0
HSQ_ERR_INVALID_ARGUMENT = -1
HSQ_ERR_CUDA = -2
HSQ_ERR_NO_DEVICE = -3
HSQ_ERR_BAD_CHARACTER = -4
HSQ_ERR_UNSUPPORTED = -5
HSQ_ERR_OUT_OF_RANGE = -6
HSQ_ERR_OVERFLOW = -7
HSQ_ERR_NAME_MISMATCH = -8
HSQ_ERR_NOT_TRIMMED = -9
HSQ_ERR_NO_MATCHES = -10
HSQ_ERR_MALFORMED = -11
HSQ_PRIMER_NOT_FOUND = -2147483647
HSQ_ALIGN_PENALIZE_START = 1
HSQ_ALIGN_PENALIZE_END = 2
HSQ_COMM_ID_BYTES = 128

u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)


class Scorer(C.Structure):
    _fields_ = [("match", C.c_int32), ("mismatch", C.c_int32), ("gap_extend", C.c_int32), ("gap_start", C.c_int32)]


class Seqs(C.Structure):
    _fields_ = [("data", u8p), ("offsets", i64p), ("n", C.c_int64)]


class ProbeTable(C.Structure):
    _fields_ = [("n", C.c_int64), ("reverse_strand", u8p), ("ext_start", i32p), ("ext_stop", i32p), ("lig_start", i32p),
                ("lig_stop", i32p), ("target_start", i32p), ("target_stop", i32p), ("ext", Seqs), ("lig", Seqs),
                ("target", Seqs)]


class PairBatch(C.Structure):
    _fields_ = [("n", C.c_int64), ("probe_id", i32p), ("read_index", i32p), ("r1", Seqs), ("r2", Seqs), ("q1", u8p),
                ("q2", u8p)]


class ExtendOut(C.Structure):
    _fields_ = [("cigar_cap", C.c_int32), ("md_cap", C.c_int32), ("ok", u8p), ("primer_end", i32p), ("score", i32p),
                ("pos", i32p), ("read_start", i32p), ("read_stop", i32p), ("nm", i32p), ("tlen", i32p),
                ("cigar_len", i32p), ("cigar", u8p), ("md_len", i32p), ("md", u8p), ("overflow", u8p), ("pair_ok", u8p),
                ("cells", C.c_int64)]


class AlignOut(C.Structure):
    _fields_ = [("cols_cap", C.c_int32), ("md_cap", C.c_int32), ("score", i32p), ("n_cols", i32p), ("ref_aln", u8p),
                ("qry_aln", u8p), ("ops", u8p), ("first_q", i32p), ("last_q", i32p), ("first_r", i32p), ("last_r", i32p),
                ("matches", i32p), ("mismatches", i32p), ("non_terminal_gaps", i32p), ("edit_distance", i32p),
                ("md_len", i32p), ("md", u8p), ("rmd_len", i32p), ("rmd", u8p), ("overflow", u8p), ("cells", C.c_int64), ("find_counts", i32p)]


u32p = C.POINTER(C.c_uint32)
i16p = C.POINTER(C.c_int16)


class PackedReads(C.Structure):
    _fields_ = [("bits", C.c_int32), ("n", C.c_int64), ("data", u8p), ("len", i32p), ("n_exceptions", C.c_int64), ("exc_read", u32p),
                ("exc_pos", u32p), ("exc_char", u8p), ("data_bytes", C.c_int64), ("n_bases", C.c_int64)]


class PackedPairs(C.Structure):
    _fields_ = [("n", C.c_int64), ("probe_id", i32p), ("read_index", i32p), ("r1", PackedReads), ("r2", PackedReads), ("q1", u8p), ("q2", u8p),
                ("qsum1", i16p), ("qsum2", i16p)]


class ExtendPooledOut(C.Structure):
    _fields_ = [("cigar_cap", C.c_int32), ("md_cap", C.c_int32), ("ok", u8p), ("primer_end", i32p), ("score", i32p), ("pos", i32p),
                ("read_start", i32p), ("read_stop", i32p), ("nm", i32p), ("tlen", i32p), ("cigar_len", i32p), ("md_len", i32p),
                ("str_off", i64p), ("pool", u8p), ("pool_cap", C.c_int64), ("pool_bytes", C.c_int64), ("pair_ok", u8p), ("cells", C.c_int64)]


class Config(C.Structure):
    _fields_ = [("struct_bytes", C.c_int32), ("force_dp32", C.c_int32), ("force_dedup_radix", C.c_int32), ("extend_chunks", C.c_int32),
                ("uniform_chunks", C.c_int32), ("band", C.c_int32), ("trim_chunks", C.c_int32), ("l2_persist_rows", C.c_int32)]

    def __init__(self, **kw):
        super().__init__(**kw)
        self.struct_bytes = C.sizeof(Config)


class ProbeStats(C.Structure):
    _fields_ = [("n_probes", C.c_int64), ("unique_pairs", i32p), ("duplicate_pairs", i32p), ("max_pairs_per_uid", i32p)]


HSQ_BAM_FIELDS = ("record", "qname", "cigar", "seq", "qual", "tags")
HSQ_BAM_VALUES = ("refid", "pos", "flag", "mapq", "next_refid", "pnext", "tlen", "ref_len")  # planes 6..13 of val


class BamIndex(C.Structure):
    _fields_ = [("cap", C.c_int64), ("n", C.c_int64), ("n_ref", C.c_int64), ("text_bytes", C.c_int64), ("first_record", C.c_int64),
                ("payload_bytes", C.c_int64), ("payload_cap", C.c_int64), ("payload", u8p), ("off", i64p), ("val", i32p),
                ("exact_walk", C.c_int32)]


class UidReport(C.Structure):
    _fields_ = [("uid_len", C.c_int32), ("unique_by_position", i32p), ("weighted_by_position", i32p)]


class DedupExtendOut(C.Structure):
    _fields_ = [("is_best", u8p), ("group_size", i32p), ("stats", C.POINTER(ProbeStats)), ("unique_unable_to_extend", i32p),
                ("duplicate_unable_to_extend", i32p), ("n_extended", C.c_int64), ("extended_pair", i32p), ("extend", C.POINTER(ExtendOut)),
                ("extend_pooled", C.c_void_p)]


class RecordOut(C.Structure):
    _fields_ = [("flag", C.POINTER(C.c_uint16)), ("mate_pos", i32p), ("offsets", i64p), ("seq", u8p), ("qual", u8p)]


class MergeBatch(C.Structure):
    _fields_ = [("n", C.c_int64), ("negative_strand", u8p), ("up", Seqs), ("down", Seqs), ("up_qual", u8p), ("down_qual", u8p),
                ("up_start", i32p), ("down_start", i32p)]


class ProbeRanges(C.Structure):
    _fields_ = [("n", C.c_int64), ("contig", i32p), ("reverse_strand", u8p), ("start", i32p), ("stop", i32p)]


class MappedRecords(C.Structure):
    _fields_ = [("n", C.c_int64), ("contig", i32p), ("negative_strand", u8p), ("second_of_pair", u8p), ("start", i32p), ("end", i32p),
                ("read_index", i32p), ("alt_offsets", i64p), ("alt_contig", i32p), ("alt_negative", u8p), ("alt_start", i32p),
                ("alt_stop", i32p)]


class AssignOut(C.Structure):
    _fields_ = [("max_probes", C.c_int32), ("n_probes", i32p), ("probes", i32p), ("mate", i32p), ("n_common", i32p), ("common", i32p),
                ("overflow", u8p)]


class FindOut(C.Structure):
    _fields_ = [("max_candidates", C.c_int32), ("lead_trimmed", i32p), ("trail_trimmed", i32p), ("has_n", u8p), ("n_candidates", i32p),
                ("candidates", i32p), ("hits", i32p), ("overflow", u8p), ("order_dependent", u8p)]


class FindLimits(C.Structure):
    _fields_ = [("mismatches", C.c_int32), ("insertion_gaps", C.c_int32), ("insertion_bases", C.c_int32), ("deletion_gaps", C.c_int32),
                ("deletion_bases", C.c_int32)]


class FindMatchOut(C.Structure):
    _fields_ = [("cand", FindOut), ("counts", i32p), ("accepted", u8p), ("n_matches", i32p), ("cells", C.c_int64)]


class FastqIndex(C.Structure):
    _fields_ = [("cap", C.c_int64), ("n", C.c_int64), ("header_off", i64p), ("header_len", i32p), ("seq_off", i64p), ("seq_len", i32p),
                ("qual_off", i64p), ("qual_len", i32p)]


class SamIndex(C.Structure):
    _fields_ = [("cap", C.c_int64), ("n", C.c_int64), ("n_header_lines", C.c_int64), ("off", i64p), ("val", i32p)]


SAM_FIELDS = ("qname", "rname", "cigar", "seq", "qual", "tags", "xa")
SAM_VALUES = ("flag", "pos", "mapq", "pnext", "tlen", "ref_len")


class SamRecordsOut(C.Structure):
    _fields_ = [("cap", C.c_int64), ("alt_cap", C.c_int64), ("n", C.c_int64), ("n_alt", C.c_int64), ("sam_line", i32p), ("contig", i32p),
                ("negative_strand", u8p), ("second_of_pair", u8p), ("start", i32p), ("end", i32p), ("read_index", i32p), ("alt_offsets", i64p),
                ("alt_contig", i32p), ("alt_negative", u8p), ("alt_start", i32p), ("alt_stop", i32p)]


class CoordSortIn(C.Structure):
    _fields_ = [("n", C.c_int64), ("ref_index", i32p), ("pos", i32p), ("flags", i32p), ("mapq", i32p), ("mate_ref_index", i32p), ("mate_pos", i32p),
                ("tlen", i32p), ("names", Seqs)]


class NameJoin(C.Structure):
    _fields_ = [("sam_names", Seqs), ("sam_flags", i32p), ("sam_seq_len", i32p), ("sam_qual_len", i32p), ("fq1_headers", Seqs),
                ("fq2_headers", Seqs), ("fq1_seq_len", i32p), ("fq2_seq_len", i32p), ("by_index", C.c_int32), ("trimming_skipped", C.c_int32),
                ("check_trim", C.c_int32), ("three_prime", C.c_int32), ("r1_first", C.c_int32), ("r1_stop", C.c_int32),
                ("r2_first", C.c_int32), ("r2_stop", C.c_int32)]


class NameJoinOut(C.Structure):
    _fields_ = [("n_emitted", C.c_int64), ("total_matching_pairs", C.c_int64), ("sam_record", i32p), ("fastq_record", i32p),
                ("name_len", i32p), ("status", u8p)]


ASSIGN_UNPAIRED, ASSIGN_AMBIGUOUS_NAME, ASSIGN_IGNORED = -1, -2, -3


class Candidates(C.Structure):
    _fields_ = [("n", C.c_int64), ("r1", Seqs), ("r2", Seqs), ("cand_offsets", i64p), ("cand_probe", i32p)]


def ptr(a, typ):
    """Pointer to a numpy array's buffer (None -> NULL)."""
    if a is None:
        return C.cast(None, typ)
    return C.cast(a.ctypes.data, typ)


class SeqBatch:
    """Host-side ragged ASCII batch: `data` uint8[total], `offsets` int64[n+1]. Keeps arrays alive."""

    def __init__(self, data, offsets):
        self.data = np.ascontiguousarray(data, dtype=np.uint8)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        self.n = len(self.offsets) - 1

    @classmethod
    def from_strings(cls, strings):
        bs = [s.encode("ascii") if isinstance(s, str) else bytes(s) for s in strings]
        offsets = np.zeros(len(bs) + 1, dtype=np.int64)
        if bs:
            offsets[1:] = np.cumsum([len(b) for b in bs])
        data = np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(0, dtype=np.uint8)
        return cls(data, offsets)

    @classmethod
    def from_fixed(cls, arr2d):
        """uint8[n, L] -> batch of n strings of length L (no copy when contiguous)."""
        arr2d = np.ascontiguousarray(arr2d, dtype=np.uint8)
        n, L = arr2d.shape
        return cls(arr2d.reshape(-1), np.arange(n + 1, dtype=np.int64) * L)

    def take(self, idx):
        """Ragged gather: the strings idx[0], idx[1], ... as a new batch (vectorised)."""
        idx = np.asarray(idx, dtype=np.int64)
        if self.n > 0 and len(self.data) == self.n * int(self.offsets[1]) and (np.diff(self.offsets) == self.offsets[1]).all():
            L = int(self.offsets[1])  # equal lengths: a row gather
            return SeqBatch(self.data.reshape(self.n, L)[idx].reshape(-1), np.arange(len(idx) + 1, dtype=np.int64) * L)
        lens = (self.offsets[1:] - self.offsets[:-1])[idx]
        offsets = np.zeros(len(idx) + 1, dtype=np.int64)
        np.cumsum(lens, out=offsets[1:])
        src = np.repeat(self.offsets[:-1][idx] - offsets[:-1], lens) + np.arange(offsets[-1], dtype=np.int64)
        return SeqBatch(self.data[src], offsets)

    def string(self, k):
        return self.data[self.offsets[k]:self.offsets[k + 1]].tobytes().decode("ascii")

    def strings(self):
        return [self.string(k) for k in range(self.n)]

    def lengths(self):
        return np.diff(self.offsets)

    def c(self):
        if len(self.data) == 0:  # keep a valid pointer even for empty batches
            self._pad = np.zeros(1, dtype=np.uint8)
            return Seqs(ptr(self._pad, u8p), ptr(self.offsets, i64p), self.n)
        return Seqs(ptr(self.data, u8p), ptr(self.offsets, i64p), self.n)


def slot_strings(buf, lens, cap):
    """Decode an [n][cap] slot array + lengths into python strings (truncated at cap)."""
    out = []
    for k in range(len(lens)):
        ln = min(int(lens[k]), cap)
        out.append(buf[k * cap:k * cap + ln].tobytes().decode("ascii"))
    return out
