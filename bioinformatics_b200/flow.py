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


def ingest_flow(be, fq1_text, fq2_text, sam_text, trim=(0, 0, 0, 0), three_prime=False, trimming_skipped=False):
    """FASTQ texts + SAM text -> the merged records createBamFileWithDataFromRawFastqFiles writes
    (HSQ/process/FastqAndBamFileMerger.java:107-158, first configuration …Identification.java:150-160), as index glue over
    hsq_index_fastq, hsq_index_sam, hsq_gather_rows and hsq_join_names.  `be` needs index_fastq, index_sam, gather_rows and
    join_names (a Context, or the tests' adapter of the CPU oracle).  Returns a dict: per returned record the SAM line it came
    from, its new read name (= FASTQ ordinal, FastqSorter.java:298-302), the ML tag (mapped read length,
    FastqAndBamFileMerger.java:390-391), first-of-pair, the raw bases / qualities it now carries, and the SAM columns the
    read->probe assignment reads next."""
    from .batches import NameJoinHost
    i1, i2, sam = be.index_fastq(fq1_text), be.index_fastq(fq2_text), be.index_sam(sam_text)
    heads = [be.gather_rows(t, ix.header_off[:ix.n], ix.header_len[:ix.n]) for t, ix in ((fq1_text, i1), (fq2_text, i2))]
    names = be.gather_rows(sam_text, sam.qname_off, sam.qname_len)
    f1, s1, f2, s2 = trim
    j = NameJoinHost(names, sam.flag, sam.seq_len, heads[0], heads[1], i1.seq_len[:i1.n], i2.seq_len[:i2.n], sam_qual_len=sam.qual_len,
                     trimming_skipped=trimming_skipped, three_prime=three_prime, r1_first=f1, r1_stop=s1, r2_first=f2, r2_stop=s2)
    res = be.join_names(j)
    rec, fq, _ = res.emitted()
    first = (sam.flag[rec] & 0x40) != 0
    out = dict(sam_line=rec, read_index=fq, first_of_pair=first, mapped_read_length=sam.seq_len[rec], status=res.status,
               total_matching_pairs=res.total_matching_pairs, flag=sam.flag[rec], pos=sam.pos[rec], ref_len=sam.ref_len[rec],
               rname_off=sam.rname_off[rec], rname_len=sam.rname_len[rec], xa_off=sam.xa_off[rec], xa_len=sam.xa_len[rec])
    for key, text, ix, sel in (("one", fq1_text, i1, first), ("two", fq2_text, i2, ~first)):
        rows = fq[sel]
        out["seq_" + key] = be.gather_rows(text, ix.seq_off[rows], ix.seq_len[rows])
        out["qual_" + key] = be.gather_rows(text, ix.qual_off[rows], ix.qual_len[rows])
    return out


def fastq_pair_table(be, fq1_text, fq2_text):
    """The raw reads of both FASTQ files as the pair batch dedup_flow indexes by read index (pair k = FASTQ ordinal k)."""
    i1, i2 = be.index_fastq(fq1_text), be.index_fastq(fq2_text)
    n = min(i1.n, i2.n)
    rows = lambda t, off, ln: be.gather_rows(t, off[:n], ln[:n])
    return PairBatchHost(np.zeros(n, dtype=np.int32), rows(fq1_text, i1.seq_off, i1.seq_len), rows(fq2_text, i2.seq_off, i2.seq_len),
                         rows(fq1_text, i1.qual_off, i1.qual_len), rows(fq2_text, i2.qual_off, i2.qual_len), np.arange(n, dtype=np.int32))


def dedup_from_files(be, fq1_text, fq2_text, sam_text, contig_names, ext_uid_len, lig_uid_len, write_merged, **kw):
    """`HSQutils dedup` from its three input files to the output record fields: ingest_flow (merge the raw FASTQ data into the
    mapper's records, …Identification.java:150-160), the caller's writer (`write_merged(sam_text, merged) -> merged SAM text`, the
    htsjdk side: records renamed to their FASTQ ordinal), hsq_sam_mapped_records, then dedup_flow."""
    merged = ingest_flow(be, fq1_text, fq2_text, sam_text, trimming_skipped=True)
    merged_text = write_merged(sam_text, merged)
    records = be.sam_mapped_records(merged_text, contig_names).records()
    out = dedup_flow(be, records, fastq_pair_table(be, fq1_text, fq2_text), ext_uid_len, lig_uid_len, **kw)
    out["merged"], out["mapped_records"] = merged, records
    return out


def probe_trimming_information(probes, ext_uid_len, lig_uid_len, trim_primers=True, additional_ext_trim=0, additional_lig_trim=0):
    """FastqReadTrimmer.getProbeTrimmingInformation (HSQ/process/FastqReadTrimmer.java:78-127) from the probe table's primer and
    capture-target lengths (collectStatsFromProbeInformation :129-151): (r1_first, r1_stop, r2_first, r2_stop)."""
    ext, lig, target = probes.ext.lengths(), probes.lig.lengths(), probes.target.lengths()
    r1_first = additional_ext_trim + ext_uid_len + (int(ext.max()) if trim_primers else 0)
    r1_stop = int(target.min()) + (int(lig.min()) if trim_primers else 0)
    r2_first = additional_lig_trim + lig_uid_len + (int(lig.max()) if trim_primers else 0)
    r2_stop = int(target.min()) + (int(ext.min()) if trim_primers else 0)
    return r1_first, r1_stop, r2_first, r2_stop


def trim_flow(be, probes, fq1_text, fq2_text, ext_uid_len, lig_uid_len, three_prime=True, trim_primers=True):
    """`HSQutils trim` file to file (FastqReadTrimmer.trimReads :263-307 on both FASTQ files): returns (trimmed FASTQ one text, trimmed
    FASTQ two text, trimming information).  `be` needs trim_fastq (a Context, or the tests' adapter of the CPU oracle)."""
    info = probe_trimming_information(probes, ext_uid_len, lig_uid_len, trim_primers)
    t1, n1 = be.trim_fastq(fq1_text, info[0], info[1], three_prime)
    t2, n2 = be.trim_fastq(fq2_text, info[2], info[3], three_prime)
    return t1, t2, info
