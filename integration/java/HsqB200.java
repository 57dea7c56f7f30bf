/*
 * HsqB200 -- Panama FFM (JDK 22+) binding of include/hsq_b200.h for nimblegen_heatseq.
 *
 * NOT compiled or run in this repository's image (no JVM there); the same ABI is exercised from Python (ctypes) by
 * tests/.  It shows the reference-side binding a maintainer adds: one context per GPU, pinned staging buffers filled
 * from the Strings the reference already holds, one downcall per batch.  There is no JVM-side compute fallback: every
 * failure surfaces as an exception.
 */
package com.roche.heatseq.b200;

import java.lang.foreign.*;
import java.lang.invoke.MethodHandle;
import java.nio.charset.StandardCharsets;

import static java.lang.foreign.ValueLayout.*;

public final class HsqB200 implements AutoCloseable {

    public static final int HSQ_OK = 0, ERR_INVALID_ARGUMENT = -1, ERR_CUDA = -2, ERR_NO_DEVICE = -3, ERR_BAD_CHARACTER = -4,
            ERR_UNSUPPORTED = -5, ERR_OUT_OF_RANGE = -6, ERR_OVERFLOW = -7;
    public static final int PRIMER_NOT_FOUND = -2147483647;

    private static final Linker LINKER = Linker.nativeLinker();
    private static final SymbolLookup LIB =
            SymbolLookup.libraryLookup(System.getProperty("hsq.b200.lib", "libhsq_b200.so"), Arena.global());

    private static MethodHandle fn(String name, FunctionDescriptor d) {
        return LINKER.downcallHandle(LIB.find(name).orElseThrow(() -> new UnsatisfiedLinkError(name)), d);
    }

    // ---- struct layouts (field order = include/hsq_b200.h) ----------------------------------------------------------
    static final StructLayout SCORER = MemoryLayout.structLayout(JAVA_INT.withName("match"), JAVA_INT.withName("mismatch"),
            JAVA_INT.withName("gap_extend"), JAVA_INT.withName("gap_start"));
    static final StructLayout SEQS = MemoryLayout.structLayout(ADDRESS.withName("data"), ADDRESS.withName("offsets"), JAVA_LONG.withName("n"));
    static final StructLayout PROBE_TABLE = MemoryLayout.structLayout(JAVA_LONG.withName("n"), ADDRESS.withName("reverse_strand"),
            ADDRESS.withName("ext_start"), ADDRESS.withName("ext_stop"), ADDRESS.withName("lig_start"), ADDRESS.withName("lig_stop"),
            ADDRESS.withName("target_start"), ADDRESS.withName("target_stop"), SEQS.withName("ext"), SEQS.withName("lig"),
            SEQS.withName("target"));
    static final StructLayout PAIR_BATCH = MemoryLayout.structLayout(JAVA_LONG.withName("n"), ADDRESS.withName("probe_id"),
            ADDRESS.withName("read_index"), SEQS.withName("r1"), SEQS.withName("r2"), ADDRESS.withName("q1"), ADDRESS.withName("q2"));
    static final StructLayout EXTEND_OUT = MemoryLayout.structLayout(JAVA_INT.withName("cigar_cap"), JAVA_INT.withName("md_cap"),
            ADDRESS.withName("ok"), ADDRESS.withName("primer_end"), ADDRESS.withName("score"), ADDRESS.withName("pos"),
            ADDRESS.withName("read_start"), ADDRESS.withName("read_stop"), ADDRESS.withName("nm"), ADDRESS.withName("tlen"),
            ADDRESS.withName("cigar_len"), ADDRESS.withName("cigar"), ADDRESS.withName("md_len"), ADDRESS.withName("md"),
            ADDRESS.withName("overflow"), ADDRESS.withName("pair_ok"), JAVA_LONG.withName("cells"));
    static final StructLayout PROBE_STATS = MemoryLayout.structLayout(JAVA_LONG.withName("n_probes"), ADDRESS.withName("unique_pairs"),
            ADDRESS.withName("duplicate_pairs"), ADDRESS.withName("max_pairs_per_uid"));

    static final StructLayout PROBE_RANGES = MemoryLayout.structLayout(JAVA_LONG.withName("n"), ADDRESS.withName("contig"),
            ADDRESS.withName("reverse_strand"), ADDRESS.withName("start"), ADDRESS.withName("stop"));
    static final StructLayout MAPPED_RECORDS = MemoryLayout.structLayout(JAVA_LONG.withName("n"), ADDRESS.withName("contig"),
            ADDRESS.withName("negative_strand"), ADDRESS.withName("second_of_pair"), ADDRESS.withName("start"), ADDRESS.withName("end"),
            ADDRESS.withName("read_index"), ADDRESS.withName("alt_offsets"), ADDRESS.withName("alt_contig"), ADDRESS.withName("alt_negative"),
            ADDRESS.withName("alt_start"), ADDRESS.withName("alt_stop"));
    static final StructLayout ASSIGN_OUT = MemoryLayout.structLayout(JAVA_INT.withName("max_probes"), MemoryLayout.paddingLayout(4),
            ADDRESS.withName("n_probes"), ADDRESS.withName("probes"), ADDRESS.withName("mate"), ADDRESS.withName("n_common"),
            ADDRESS.withName("common"), ADDRESS.withName("overflow"));
    static final StructLayout DEDUP_EXTEND_OUT = MemoryLayout.structLayout(ADDRESS.withName("is_best"), ADDRESS.withName("group_size"),
            ADDRESS.withName("stats"), ADDRESS.withName("unique_unable_to_extend"), ADDRESS.withName("duplicate_unable_to_extend"),
            JAVA_LONG.withName("n_extended"), ADDRESS.withName("extended_pair"), ADDRESS.withName("extend"));
    static final StructLayout RECORD_OUT = MemoryLayout.structLayout(ADDRESS.withName("flag"), ADDRESS.withName("mate_pos"),
            ADDRESS.withName("offsets"), ADDRESS.withName("seq"), ADDRESS.withName("qual"));
    static final StructLayout MERGE_BATCH = MemoryLayout.structLayout(JAVA_LONG.withName("n"), ADDRESS.withName("negative_strand"),
            SEQS.withName("up"), SEQS.withName("down"), ADDRESS.withName("up_qual"), ADDRESS.withName("down_qual"),
            ADDRESS.withName("up_start"), ADDRESS.withName("down_start"));

    // ---- downcall handles --------------------------------------------------------------------------------------------
    private static final MethodHandle CTX_CREATE = fn("hsq_ctx_create", FunctionDescriptor.of(JAVA_INT, JAVA_INT, ADDRESS));
    private static final MethodHandle CTX_DESTROY = fn("hsq_ctx_destroy", FunctionDescriptor.ofVoid(ADDRESS));
    private static final MethodHandle LAST_ERROR = fn("hsq_last_error", FunctionDescriptor.of(ADDRESS, ADDRESS));
    private static final MethodHandle PINNED_ALLOC = fn("hsq_pinned_alloc", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS));
    private static final MethodHandle PINNED_FREE = fn("hsq_pinned_free", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle SET_SCORER = fn("hsq_set_scorer", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle SET_PROBES = fn("hsq_set_probes", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle TRIM = fn("hsq_trim_batch", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_INT,
            JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle DEDUP = fn("hsq_dedup_batch", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT,
            ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle EXTEND = fn("hsq_extend_batch", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));

    private static final MethodHandle SET_PROBE_RANGES = fn("hsq_set_probe_ranges", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle ASSIGN_PROBES = fn("hsq_assign_probes", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle DEDUP_EXTEND = fn("hsq_dedup_extend_batch", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT,
            JAVA_INT, JAVA_INT, ADDRESS));
    private static final MethodHandle EXTEND_SELECTED = fn("hsq_extend_selected", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS,
            ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle EXTEND_RECORDS = fn("hsq_extend_records", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS,
            ADDRESS, ADDRESS, JAVA_LONG));
    private static final MethodHandle FLAG_N = fn("hsq_flag_n_pairs", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle SET_FIND_SEQUENCES = fn("hsq_set_find_sequences", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT));
    private static final MethodHandle FIND_CANDIDATES = fn("hsq_find_candidates", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_INT,
            JAVA_INT, JAVA_INT, JAVA_DOUBLE, ADDRESS));
    private static final MethodHandle FIND_MATCHES = fn("hsq_find_matches", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_INT,
            JAVA_INT, JAVA_INT, JAVA_DOUBLE, ADDRESS, ADDRESS));
    private static final MethodHandle MERGE_PAIRS = fn("hsq_merge_pairs", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS,
            ADDRESS, JAVA_LONG));
    private static final MethodHandle INDEX_FASTQ = fn("hsq_index_fastq", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
    private static final MethodHandle TRIM_FASTQ = fn("hsq_trim_fastq", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, JAVA_INT, JAVA_INT,
            JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS));
    private static final MethodHandle INDEX_SAM = fn("hsq_index_sam", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
    private static final MethodHandle SAM_MAPPED_RECORDS = fn("hsq_sam_mapped_records", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS));
    private static final MethodHandle COORDINATE_ORDER = fn("hsq_coordinate_order", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle JOIN_NAMES = fn("hsq_join_names", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle GATHER_ROWS = fn("hsq_gather_rows", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS,
            ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, JAVA_LONG));
    private static final MethodHandle RESOLVE = fn("hsq_resolve_multi_probe", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT,
            JAVA_INT, ADDRESS, ADDRESS, ADDRESS));

    private final Arena arena = Arena.ofShared();
    private final MemorySegment ctx;

    public HsqB200(int device) {
        MemorySegment out = arena.allocate(ADDRESS);
        int st = invokeInt(CTX_CREATE, device, out);
        if (st != HSQ_OK) throw new HsqException(st, lastError(MemorySegment.NULL));  // e.g. ERR_NO_DEVICE: never fall back
        ctx = out.get(ADDRESS, 0);
    }

    /** SimpleAlignmentScorer's four numbers (COMMON/alignment/SimpleAlignmentScorer.java:28-31). */
    public void setScorer(int match, int mismatch, int gapExtend, int gapStart) {
        MemorySegment s = arena.allocate(SCORER);
        s.set(JAVA_INT, 0, match);
        s.set(JAVA_INT, 4, mismatch);
        s.set(JAVA_INT, 8, gapExtend);
        s.set(JAVA_INT, 12, gapStart);
        check(invokeInt(SET_SCORER, ctx, s));
    }

    /** Pinned host staging memory the caller fills straight from Strings (ISO_8859_1 bytes). */
    public MemorySegment pinned(long bytes) {
        MemorySegment out = arena.allocate(ADDRESS);
        check(invokeInt(PINNED_ALLOC, ctx, bytes, out));
        return out.get(ADDRESS, 0).reinterpret(bytes);
    }

    /** Ragged ASCII batch -> hsq_seqs inside `parent` at `offset`. */
    static void putSeqs(MemorySegment parent, long offset, MemorySegment data, MemorySegment offsets, long n) {
        parent.set(ADDRESS, offset, data);
        parent.set(ADDRESS, offset + 8, offsets);
        parent.set(JAVA_LONG, offset + 16, n);
    }

    /** ParsedProbeFile -> device resident probe table (fields of COMMON/utils/probeinfo/Probe.java:53-69). */
    public void setProbes(MemorySegment probeTable) { check(invokeInt(SET_PROBES, ctx, probeTable)); }

    /** FilterByUid.reduceReadsByProbeAndUid + findBestData for a batch spanning probes (HSQ/process/FilterByUid.java:71-297). */
    public void dedup(MemorySegment pairBatch, int extUidLen, int ligUidLen, MemorySegment isBest, MemorySegment groupSize,
                      MemorySegment probeStats) {
        check(invokeInt(DEDUP, ctx, pairBatch, extUidLen, ligUidLen, isBest, groupSize, probeStats));
    }

    /** ExtendReadsToPrimer.extendReadsToPrimers for a batch spanning probes (HSQ/process/ExtendReadsToPrimer.java:347-380). */
    public void extend(MemorySegment pairBatch, MemorySegment extendOut) { check(invokeInt(EXTEND, ctx, pairBatch, extendOut)); }

    /** The two range-map tables of getReadToProbeAssignments (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:313-331). */
    public void setProbeRanges(MemorySegment probeRanges) { check(invokeInt(SET_PROBE_RANGES, ctx, probeRanges)); }

    /** ReadToProbeAssigner.assignProbesToReads for a batch of SAM records (…Identification.java:439-576). */
    public void assignProbes(MemorySegment mappedRecords, MemorySegment assignOut) { check(invokeInt(ASSIGN_PROBES, ctx, mappedRecords, assignOut)); }

    /** The body of PrimerReadExtensionAndFilteringOfUniquePcrProbesTask.run for a batch spanning probes (…Identification.java:1172-1216). */
    public void dedupExtend(MemorySegment pairBatch, int extUidLen, int ligUidLen, boolean extendDuplicates, MemorySegment dedupExtendOut) {
        check(invokeInt(DEDUP_EXTEND, ctx, pairBatch, extUidLen, ligUidLen, extendDuplicates ? 1 : 0, dedupExtendOut));
    }

    /** extendReadsToPrimers on the pairs flagged by the caller (unique pairs of a reduction that ran on another GPU). */
    public void extendSelected(MemorySegment pairBatch, MemorySegment keep, MemorySegment nExtended, MemorySegment extendedPair,
                               MemorySegment extendOut) {
        check(invokeInt(EXTEND_SELECTED, ctx, pairBatch, keep, nExtended, extendedPair, extendOut));
    }

    /** FLAG / PNEXT / read string / qualities of the records createRecord + setSAMRecordsAsPair build (ExtendReadsToPrimer.java:131-181,247-284). */
    public void extendRecords(MemorySegment pairBatch, MemorySegment extendOut, MemorySegment duplicate, MemorySegment recordOut, long outCap) {
        check(invokeInt(EXTEND_RECORDS, ctx, pairBatch, extendOut, duplicate, recordOut, outCap));
    }

    /** hasN[k] = 1 when either read of pair k holds an N (…Identification.java:815-821): such pairs skip the probe tasks. */
    public void flagNPairs(MemorySegment pairBatch, MemorySegment hasN) { check(invokeInt(FLAG_N, ctx, pairBatch, hasN)); }

    /** The SimpleMapper index of FastqFindAndTrimTool (FastqFindAndTrimTool.java:231-237): k = 5, at most 15000 sequences per k-mer. */
    public void setFindSequences(MemorySegment seqs, int kmer, int maxRefsPerKmer) { check(invokeInt(SET_FIND_SEQUENCES, ctx, seqs, kmer, maxRefsPerKmer)); }

    /** Quality trim, N test and getBestCandidateReferences per read (FastqFindAndTrimTool.java:904-926); findOut = hsq_find_out. */
    public void findCandidates(MemorySegment reads, MemorySegment quals, int leadingQuality, int trailingQuality, int candidateLimit,
                               double minHitRatio, MemorySegment findOut) {
        check(invokeInt(FIND_CANDIDATES, ctx, reads, quals, leadingQuality, trailingQuality, candidateLimit, minHitRatio, findOut));
    }

    /** findCandidates + one alignment per candidate + the five limits (FastqFindAndTrimTool.java:926-951); limits = hsq_find_limits. */
    public void findMatches(MemorySegment reads, MemorySegment quals, int leadingQuality, int trailingQuality, int candidateLimit,
                            double minHitRatio, MemorySegment limits, MemorySegment findMatchOut) {
        check(invokeInt(FIND_MATCHES, ctx, reads, quals, leadingQuality, trailingQuality, candidateLimit, minHitRatio, limits, findMatchOut));
    }

    /** Record index of FASTQ text as FastqReader reads it (call sites FastqSorter.java:77,98); fastqIndex = hsq_fastq_index. */
    public void indexFastq(MemorySegment text, long nBytes, MemorySegment fastqIndex) { check(invokeInt(INDEX_FASTQ, ctx, text, nBytes, fastqIndex)); }

    /** FastqReadTrimmer.trimReads (FastqReadTrimmer.java:263-307), FASTQ text in, trimmed FASTQ text out; returns the bytes written. */
    public long trimFastq(MemorySegment text, long nBytes, int firstBaseToKeep, int lastBaseToKeep, boolean threePrime, MemorySegment outText, long outCap) {
        MemorySegment outBytes = arena.allocate(JAVA_LONG), nRecords = arena.allocate(JAVA_LONG);
        check(invokeInt(TRIM_FASTQ, ctx, text, nBytes, firstBaseToKeep, lastBaseToKeep, threePrime ? 1 : 0, outText, outCap, outBytes, nRecords));
        return outBytes.get(JAVA_LONG, 0);
    }

    /** Column index of SAM text (QNAME, FLAG, RNAME, POS, CIGAR reference length, SEQ, QUAL, XA); samIndex = hsq_sam_index. */
    public void indexSam(MemorySegment text, long nBytes, MemorySegment samIndex) { check(invokeInt(INDEX_SAM, ctx, text, nBytes, samIndex)); }

    /** SAM text -> the columns of hsq_mapped_records + the XA hits (…Identification.java:439-474, SAMRecordUtil.java:440-468);
     *  contigNames = hsq_seqs in the numbering of setProbeRanges, samRecordsOut = hsq_sam_records_out. */
    public void samMappedRecords(MemorySegment text, long nBytes, MemorySegment contigNames, MemorySegment samRecordsOut) {
        check(invokeInt(SAM_MAPPED_RECORDS, ctx, text, nBytes, contigNames, samRecordsOut));
    }

    /** Order of the final coordinate sort (…Identification.java:897, SAMRecordCoordinateComparator); coordSortIn = hsq_coord_sort_in, order = int32[n]. */
    public void coordinateOrder(MemorySegment coordSortIn, MemorySegment order) { check(invokeInt(COORDINATE_ORDER, ctx, coordSortIn, order)); }

    /** Both sorts and the merge loop of FastqAndBamFileMerger.createBamFileWithDataFromRawFastqFiles (FastqAndBamFileMerger.java:107-158,
     *  213-345); nameJoin = hsq_name_join, nameJoinOut = hsq_name_join_out.  The status codes map back to the exceptions of the loop:
     *  ERR_NAME_MISMATCH / ERR_NO_MATCHES -> IllegalStateException, ERR_NOT_TRIMMED -> UnableToMergeFastqAndBamFilesException. */
    public void joinNames(MemorySegment nameJoin, MemorySegment nameJoinOut) { check(invokeInt(JOIN_NAMES, ctx, nameJoin, nameJoinOut)); }

    /** Rows text[off[k] .. off[k] + len[k]) as a ragged batch: the raw reads / qualities storeFastqInfoInRecord copies (:386-396). */
    public int gatherRows(MemorySegment text, long nBytes, MemorySegment off, MemorySegment len, long n, MemorySegment outData,
                          MemorySegment outOffsets, long outCap) {
        int st = invokeInt(GATHER_ROWS, ctx, text, nBytes, off, len, n, outData, outOffsets, outCap);
        if (st != ERR_OVERFLOW) check(st);
        return st;
    }

    /** mergeSequences per item (…Identification.java:1362-1552); ERR_OVERFLOW = call again with the sizes in outOffsets. */
    public int mergePairs(MemorySegment mergeBatch, MemorySegment outSeq, MemorySegment outQual, MemorySegment outOffsets, long outCap) {
        int st = invokeInt(MERGE_PAIRS, ctx, mergeBatch, outSeq, outQual, outOffsets, outCap);
        if (st != ERR_OVERFLOW) check(st);
        return st;
    }

    /** The multi-probe tie-break (…Identification.java:704-765): bestProbe[k] = winning probe row or -1. */
    public void resolveMultiProbe(MemorySegment candidates, int extUidLen, int ligUidLen, MemorySegment bestProbe) {
        check(invokeInt(RESOLVE, ctx, candidates, extUidLen, ligUidLen, bestProbe, MemorySegment.NULL, MemorySegment.NULL));
    }

    /** FastqReadTrimmer.trim per read (HSQ/process/FastqReadTrimmer.java:309-328). */
    public void trim(MemorySegment reads, MemorySegment quals, int firstBaseToKeep, int lastBaseToKeep, boolean threePrime,
                     MemorySegment outSeq, MemorySegment outQual, MemorySegment outOffsets) {
        check(invokeInt(TRIM, ctx, reads, quals, firstBaseToKeep, lastBaseToKeep, threePrime ? 1 : 0, outSeq, outQual, outOffsets));
    }

    private static int invokeInt(MethodHandle h, Object... args) {
        try {
            return (int) h.invokeWithArguments(args);
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    private static String lastError(MemorySegment c) {
        try {
            return ((MemorySegment) LAST_ERROR.invoke(c)).reinterpret(4096).getString(0, StandardCharsets.ISO_8859_1);
        } catch (Throwable t) {
            return "unknown";
        }
    }

    private void check(int status) {
        if (status == HSQ_OK) return;
        String msg = lastError(ctx);
        switch (status) {  // the exceptions the Java implementation throws at the same places
            case ERR_BAD_CHARACTER: throw new IllegalArgumentException(msg);          // IupacNucleotideCode.getCodesFromString
            case ERR_OUT_OF_RANGE: throw new IndexOutOfBoundsException(msg);          // FastqReadTrimmer.java:310-312, ExtendReadsToPrimer.java:299
            case ERR_UNSUPPORTED: throw new UnsupportedOperationException(msg);
            default: throw new HsqException(status, msg);
        }
    }

    @Override
    public void close() {
        try {
            CTX_DESTROY.invoke(ctx);
        } catch (Throwable ignored) {
        }
        arena.close();
    }

    public static final class HsqException extends RuntimeException {
        public final int status;
        public HsqException(int status, String message) {
            super("hsq status " + status + ": " + message);
            this.status = status;
        }
    }
}
