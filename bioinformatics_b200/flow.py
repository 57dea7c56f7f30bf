"""The data path of `HSQUtils dedup` between the merged BAM and the output records, as a chain of C-ABI calls (harness).

Mirrors PrimerReadExtensionAndPcrDuplicateIdentification (HSQ/process/…Identification.java):
  getReadToProbeAssignments            :302-402,439-576   -> assign_probes           (hsq_assign_probes)
  multi-probe tie-break                :704-765           -> resolve_multi_probe     (hsq_resolve_multi_probe)
  "no N in either read" filter         :815-821           -> flag_n_pairs            (hsq_flag_n_pairs)
  the per-probe task body              :1172-1216         -> dedup_extend_batch      (hsq_dedup_extend_batch)
  record building                      ExtendReadsToPrimer.java:131-181,247-284 -> extend_records (hsq_extend_records)

`be` is a bioinformatics_b200._lib.Context (probe table, scorer and probe ranges already set).  The tests run the same
function over an adapter of the CPU oracle and compare every output; nothing here computes on the host beyond index glue.
"""
import numpy as np

from .batches import PairBatchHost


def pair_table(be, records, fastq, ext_uid_len, lig_uid_len, max_probes=4):
    """Mapped records -> the pair table the probe tasks consume (probe id, read index, raw reads and qualities).

    records: MappedRecordsHost; fastq: PairBatchHost whose pair k holds the raw reads of FASTQ ordinal k (the merged BAM
    carries the raw FASTQ sequences, FastqAndBamFileMerger.java:386-396).  Returns (pairs, info)."""
    a = be.assign_probes(records, max_probes)
    if a.overflow.any():
        raise OverflowError("a read overlaps more than max_probes probe windows: call again with more slots")
    mate = a.mate
    k = np.nonzero((mate >= 0) & (records.second_of_pair == 0))[0]
    k = k[records.second_of_pair[mate[k]] == 1]  # a pair needs a first and a second of pair record (:809)
    read_index = records.read_index[k]
    n_common = a.n_common[k]
    probe = np.full(len(k), -1, dtype=np.int32)
    single = n_common == 1
    probe[single] = a.common[k[single], 0]
    multi = np.nonzero(n_common > 1)[0]
    if len(multi):
        ri = read_index[multi]
        cand_offsets = np.concatenate([[0], np.cumsum(n_common[multi])]).astype(np.int64)
        rows = a.common[k[multi]]
        cand_probe = rows[np.arange(rows.shape[1])[None, :] < n_common[multi][:, None]].astype(np.int32)
        best, _, _ = be.resolve_multi_probe(fastq.r1.take(ri), fastq.r2.take(ri), cand_offsets, cand_probe, ext_uid_len, lig_uid_len)
        probe[multi] = best  # -1: more than one best probe, the read is removed from all of them (:757-765)
    keep = probe >= 0
    ri = read_index[keep]
    pairs = PairBatchHost(probe[keep], fastq.r1.take(ri), fastq.r2.take(ri), fastq.q1.take(ri), fastq.q2.take(ri), ri)
    has_n = be.flag_n_pairs(pairs)
    info = {"pairs_seen": int(len(k)), "multi_probe_pairs": int(len(multi)), "unresolved": int((~keep).sum() - (n_common == 0).sum()),
            "without_probe": int((n_common == 0).sum()), "with_n": int(has_n.sum()), "assignment": a}
    return pairs.subset(np.nonzero(has_n == 0)[0]), info


def dedup_flow(be, records, fastq, ext_uid_len, lig_uid_len, mark_duplicates=False, max_probes=4, cigar_cap=64, md_cap=96):
    """Mapped records + raw reads -> dedup decisions, extension results and SAM record fields of the pairs that are written."""
    pairs, info = pair_table(be, records, fastq, ext_uid_len, lig_uid_len, max_probes)
    res = be.dedup_extend_batch(pairs, ext_uid_len, lig_uid_len, extend_duplicates=mark_duplicates, cigar_cap=cigar_cap, md_cap=md_cap)
    m = res.n_extended
    slot_pair = res.extended_pair[:m]
    written = pairs.subset(slot_pair)
    duplicate = (1 - res.is_best[slot_pair]).astype(np.uint8) if mark_duplicates else None  # setDuplicateReadFlag (:1203-1212)
    ext = res.extend.head(m)
    rec = be.extend_records(written, ext, duplicate=duplicate)
    return {"pairs": pairs, "info": info, "dedup": res, "extended_pairs": written, "extend": ext, "records": rec}


def find_flow(be, seqs, reads, quals, leading_quality=0, trailing_quality=0, allowed_mismatches=0, allowed_insertion_gaps=0,
              allowed_insertion_bases=0, allowed_deletion_gaps=0, allowed_deletion_bases=0, max_candidates=16, cols_cap=640, md_cap=640):
    """The per-entry decision of FastqFindAndTrimTool.findAndTrim (COMMON/fastqtool/FastqFindAndTrimTool.java:904-951):
    quality trim -> N test -> SimpleMapper candidates (find_candidates) -> one alignment of the trimmed read (reference) with
    every candidate sequence (query) under the default scorer -> the five count thresholds of the find settings.

    `be` must have had set_find_sequences(seqs, ...) called.  Returns the candidate stage's result, the (read, candidate) list,
    their alignments and `accepted` (one flag per listed alignment); matches of read k = the accepted rows with item_read == k,
    in candidate order (1 row: single match, more: multiple matches, none: no match :953-985)."""
    from ._abi import SeqBatch
    cand = be.find_candidates(reads, quals, leading_quality, trailing_quality, max_candidates=max_candidates)
    if cand.overflow.any():
        raise OverflowError("a read has more candidates than max_candidates: call again with more slots")
    item_read = np.repeat(np.arange(reads.n), cand.n_candidates)
    mask = np.arange(cand.max_candidates)[None, :] < cand.n_candidates[:, None]
    item_seq = cand.candidates[mask]
    # the trimmed reads: rows of `reads` cut to [lead, len - trail)
    a = reads.offsets[:-1] + cand.lead_trimmed
    b = reads.offsets[1:] - cand.trail_trimmed
    lens = (b - a)[item_read]
    off = np.zeros(len(item_read) + 1, dtype=np.int64)
    off[1:] = np.cumsum(lens)
    src = np.repeat(a[item_read] - off[:-1], lens) + np.arange(int(off[-1]))
    refs = SeqBatch(reads.data[src], off)
    aln = be.align_batch(refs, seqs.take(item_seq), flags=0, cols_cap=cols_cap, md_cap=md_cap)
    limits = np.array([allowed_mismatches, allowed_insertion_gaps, allowed_insertion_bases, allowed_deletion_gaps, allowed_deletion_bases])
    accepted = (aln.find_counts <= limits[None, :]).all(axis=1) if len(item_read) else np.zeros(0, dtype=bool)
    n_matches = np.bincount(item_read[accepted], minlength=reads.n)
    return {"candidates": cand, "item_read": item_read, "item_seq": item_seq, "alignments": aln, "accepted": accepted, "n_matches": n_matches}
