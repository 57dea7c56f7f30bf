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
            ERR_UNSUPPORTED = -5, ERR_OUT_OF_RANGE = -6, ERR_OVERFLOW = -7, ERR_NAME_MISMATCH = -8, ERR_NOT_TRIMMED = -9, ERR_NO_MATCHES = -10,
            ERR_MALFORMED = -11;
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
            JAVA_LONG.withName("n_extended"), ADDRESS.withName("extended_pair"), ADDRESS.withName("extend"), ADDRESS.withName("extend_pooled"));
    static final StructLayout RECORD_OUT = MemoryLayout.structLayout(ADDRESS.withName("flag"), ADDRESS.withName("mate_pos"),
            ADDRESS.withName("offsets"), ADDRESS.withName("seq"), ADDRESS.withName("qual"));
    static final StructLayout MERGE_BATCH = MemoryLayout.structLayout(JAVA_LONG.withName("n"), ADDRESS.withName("negative_strand"),
            SEQS.withName("up"), SEQS.withName("down"), ADDRESS.withName("up_qual"), ADDRESS.withName("down_qual"),
            ADDRESS.withName("up_start"), ADDRESS.withName("down_start"));

    static final StructLayout ALIGN_OUT = MemoryLayout.structLayout(JAVA_INT.withName("cols_cap"), JAVA_INT.withName("md_cap"),
            ADDRESS.withName("score"), ADDRESS.withName("n_cols"), ADDRESS.withName("ref_aln"), ADDRESS.withName("qry_aln"), ADDRESS.withName("ops"),
            ADDRESS.withName("first_q"), ADDRESS.withName("last_q"), ADDRESS.withName("first_r"), ADDRESS.withName("last_r"),
            ADDRESS.withName("matches"), ADDRESS.withName("mismatches"), ADDRESS.withName("non_terminal_gaps"), ADDRESS.withName("edit_distance"),
            ADDRESS.withName("md_len"), ADDRESS.withName("md"), ADDRESS.withName("rmd_len"), ADDRESS.withName("rmd"), ADDRESS.withName("overflow"),
            JAVA_LONG.withName("cells"), ADDRESS.withName("find_counts"));
    static final StructLayout PACKED_READS = MemoryLayout.structLayout(JAVA_INT.withName("bits"), MemoryLayout.paddingLayout(4), JAVA_LONG.withName("n"),
            ADDRESS.withName("data"), ADDRESS.withName("len"), JAVA_LONG.withName("n_exceptions"), ADDRESS.withName("exc_read"),
            ADDRESS.withName("exc_pos"), ADDRESS.withName("exc_char"), JAVA_LONG.withName("data_bytes"), JAVA_LONG.withName("n_bases"));
    static final StructLayout PACKED_PAIRS = MemoryLayout.structLayout(JAVA_LONG.withName("n"), ADDRESS.withName("probe_id"), ADDRESS.withName("read_index"),
            PACKED_READS.withName("r1"), PACKED_READS.withName("r2"), ADDRESS.withName("q1"), ADDRESS.withName("q2"), ADDRESS.withName("qsum1"),
            ADDRESS.withName("qsum2"));
    static final StructLayout EXTEND_POOLED_OUT = MemoryLayout.structLayout(JAVA_INT.withName("cigar_cap"), JAVA_INT.withName("md_cap"),
            ADDRESS.withName("ok"), ADDRESS.withName("primer_end"), ADDRESS.withName("score"), ADDRESS.withName("pos"), ADDRESS.withName("read_start"),
            ADDRESS.withName("read_stop"), ADDRESS.withName("nm"), ADDRESS.withName("tlen"), ADDRESS.withName("cigar_len"), ADDRESS.withName("md_len"),
            ADDRESS.withName("str_off"), ADDRESS.withName("pool"), JAVA_LONG.withName("pool_cap"), JAVA_LONG.withName("pool_bytes"),
            ADDRESS.withName("pair_ok"), JAVA_LONG.withName("cells"));
    static final StructLayout CONFIG = MemoryLayout.structLayout(JAVA_INT.withName("struct_bytes"), JAVA_INT.withName("force_dp32"),
            JAVA_INT.withName("force_dedup_radix"), JAVA_INT.withName("extend_chunks"), JAVA_INT.withName("uniform_chunks"), JAVA_INT.withName("band"), JAVA_INT.withName("trim_chunks"));
    static final StructLayout DEDUP_KEY = MemoryLayout.structLayout(JAVA_LONG.withName("uid"), JAVA_LONG.withName("rank"), JAVA_INT.withName("probe"),
            JAVA_INT.withName("origin"));
    public static final int COMM_ID_BYTES = 128, ALIGN_PENALIZE_START = 1, ALIGN_PENALIZE_END = 2;

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

    // operator seam, packed batches, multi-GPU, configuration, measurement
    private static final MethodHandle ALIGN = fn("hsq_align_batch", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_INT, ADDRESS));
    private static final MethodHandle VARIABLE_LENGTH_UID = fn("hsq_variable_length_uid", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS,
            JAVA_INT, ADDRESS));
    private static final MethodHandle CTX_CREATE_EX = fn("hsq_ctx_create_ex", FunctionDescriptor.of(JAVA_INT, JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle CTX_CONFIGURE = fn("hsq_ctx_configure", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle VERSION = fn("hsq_version", FunctionDescriptor.of(ADDRESS));
    private static final MethodHandle PACKED_BYTES = fn("hsq_packed_bytes", FunctionDescriptor.of(JAVA_LONG, ADDRESS, JAVA_LONG, JAVA_INT));
    private static final MethodHandle PACK_READS = fn("hsq_pack_reads", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG,
            ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_INT));
    private static final MethodHandle QUALITY_SUMS = fn("hsq_quality_sums", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, JAVA_INT));
    private static final MethodHandle UNPACK_READS = fn("hsq_unpack_reads", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle EXTEND_PACKED = fn("hsq_extend_packed", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle DEDUP_PACKED = fn("hsq_dedup_packed", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS,
            ADDRESS, ADDRESS));
    private static final MethodHandle DEDUP_EXTEND_PACKED = fn("hsq_dedup_extend_packed", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT,
            JAVA_INT, JAVA_INT, ADDRESS));
    private static final MethodHandle EXTEND_STAGE = fn("hsq_extend_stage", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle EXTEND_RUN = fn("hsq_extend_run", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT));
    private static final MethodHandle EXTEND_FETCH = fn("hsq_extend_fetch", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle COMM_UNIQUE_ID = fn("hsq_comm_unique_id", FunctionDescriptor.of(JAVA_INT, ADDRESS));
    private static final MethodHandle COMM_INIT_RANK = fn("hsq_comm_init_rank", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT));
    private static final MethodHandle COMM_INIT_ALL = fn("hsq_comm_init_all", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT));
    private static final MethodHandle COMM_DESTROY = fn("hsq_comm_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
    private static final MethodHandle DEDUP_SHARDED = fn("hsq_dedup_sharded", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS,
            ADDRESS, ADDRESS));
    private static final MethodHandle DEDUP_EXTEND_SHARDED = fn("hsq_dedup_extend_sharded", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT,
            JAVA_INT, JAVA_INT, ADDRESS));
    private static final MethodHandle DEDUP_EXTEND_MULTI_GPU = fn("hsq_dedup_extend_multi_gpu", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS,
            JAVA_INT, JAVA_INT, JAVA_INT, ADDRESS));
    private static final MethodHandle DEDUP_KEYS = fn("hsq_dedup_keys", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, JAVA_LONG,
            JAVA_INT, ADDRESS));
    private static final MethodHandle DEDUP_REDUCE_KEYS = fn("hsq_dedup_reduce_keys", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, JAVA_INT,
            ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle ALIGN_SCORES_NOALLOC = fn("hsq_align_scores_noalloc", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS,
            JAVA_LONG, ADDRESS, ADDRESS, JAVA_LONG, JAVA_LONG, JAVA_INT, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle LAST_TIMING = fn("hsq_last_timing", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle LAST_SPANS = fn("hsq_last_spans", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle DEDUP_BATCH_REPORT = fn("hsq_dedup_batch_report",
            FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle PAIR_REASONS = fn("hsq_pair_reasons", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, JAVA_INT, ADDRESS));
    private static final MethodHandle UID_KEYS = fn("hsq_uid_keys", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, ADDRESS));
    private static final MethodHandle DEDUP_EXTEND_PACKED_KEYS = fn("hsq_dedup_extend_packed_keys",
            FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, ADDRESS));
    private static final MethodHandle INDEX_BAM = fn("hsq_index_bam", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));
    private static final MethodHandle BAND_STATS = fn("hsq_band_stats", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle TIMER_START = fn("hsq_timer_start", FunctionDescriptor.of(JAVA_INT, ADDRESS));
    private static final MethodHandle TIMER_STOP = fn("hsq_timer_stop", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle INT32_PEAK = fn("hsq_int32_peak", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_DOUBLE, ADDRESS));

    private final Arena arena = Arena.ofShared();
    private final MemorySegment ctx;

    /** Context with a configuration (hsq_ctx_create_ex); config = CONFIG layout or MemorySegment.NULL. */
    public HsqB200(int device, MemorySegment config) {
        MemorySegment out = arena.allocate(ADDRESS);
        int st = invokeInt(CTX_CREATE_EX, device, config, out);
        if (st != HSQ_OK) throw new HsqException(st, lastError(MemorySegment.NULL));
        ctx = out.get(ADDRESS, 0);
    }

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

    // ---- operator seam --------------------------------------------------------------------------------------------------
    /** One NeedlemanWunschGlobalAlignment per item (COMMON/alignment/NeedlemanWunschGlobalAlignment.java:44,57,85); alignOut = hsq_align_out.
     *  GpuNeedlemanWunschGlobalAlignment reads a row of it through the reference's getters. */
    public void alignBatch(MemorySegment refs, MemorySegment queries, MemorySegment scorer, int flags, MemorySegment alignOut) {
        check(invokeInt(ALIGN, ctx, refs, queries, scorer, flags, alignOut));
    }

    /** SAMRecordUtil.getVariableLengthUid per item (HSQ/utils/SAMRecordUtil.java:223-261); uidLen[k] = prefix length or -1 for null. */
    public void variableLengthUid(MemorySegment reads, MemorySegment primers, MemorySegment scorer, int flags, MemorySegment uidLen) {
        check(invokeInt(VARIABLE_LENGTH_UID, ctx, reads, primers, scorer, flags, uidLen));
    }

    // ---- configuration ---------------------------------------------------------------------------------------------------
    /** hsq_config: forceDp32 / forceDedupRadix / extendChunks / uniformChunks / band / trimChunks (0 = defaults). */
    public void configure(int forceDp32, int forceDedupRadix, int extendChunks, int uniformChunks, int band, int trimChunks) {
        MemorySegment c = arena.allocate(CONFIG);
        int[] v = {(int) CONFIG.byteSize(), forceDp32, forceDedupRadix, extendChunks, uniformChunks, band, trimChunks};
        for (int k = 0; k < v.length; k++) c.set(JAVA_INT, 4L * k, v[k]);
        check(invokeInt(CTX_CONFIGURE, ctx, c));
    }

    public static String version() {
        try {
            return ((MemorySegment) VERSION.invoke()).reinterpret(256).getString(0, StandardCharsets.ISO_8859_1);
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    // ---- packed read batches (2 / 4 bit per base) ---------------------------------------------------------------------------
    /** Bytes hsq_pack_reads writes for a ragged batch with these offsets. */
    public static long packedBytes(MemorySegment offsets, long n, int bits) {
        try {
            return (long) PACKED_BYTES.invoke(offsets, n, bits);
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    /** Host helper of the library: ASCII batch -> packed words + lengths + (2 bit) exception list; returns the status (ERR_OVERFLOW: more
     *  exceptions than excCap, nExceptions says how many). */
    public static int packReads(MemorySegment reads, int bits, MemorySegment dataOut, MemorySegment lenOut, long excCap, MemorySegment nExceptions,
                                MemorySegment excRead, MemorySegment excPos, MemorySegment excChar, int threads) {
        return invokeInt(PACK_READS, reads, bits, dataOut, lenOut, excCap, nExceptions, excRead, excPos, excChar, threads);
    }

    /** (short) sum(q - 33) per read, BamFileUtil.getQualityScore (HSQ/utils/BamFileUtil.java:74-85). */
    public static void qualitySums(MemorySegment quals, MemorySegment offsets, long n, MemorySegment out, int threads) {
        int st = invokeInt(QUALITY_SUMS, quals, offsets, n, out, threads);
        if (st != HSQ_OK) throw new HsqException(st, lastError(MemorySegment.NULL));
    }

    public void unpackReads(MemorySegment packedReads, MemorySegment outData, MemorySegment outOffsets) {
        check(invokeInt(UNPACK_READS, ctx, packedReads, outData, outOffsets));
    }

    /** extendReadsToPrimers for a packed batch, CIGAR / MD pooled; packedPairs = hsq_packed_pairs, pooledOut = hsq_extend_pooled_out. */
    public void extendPacked(MemorySegment packedPairs, MemorySegment pooledOut) { check(invokeInt(EXTEND_PACKED, ctx, packedPairs, pooledOut)); }

    public void dedupPacked(MemorySegment packedPairs, int extUidLen, int ligUidLen, MemorySegment isBest, MemorySegment groupSize,
                            MemorySegment probeStats) {
        check(invokeInt(DEDUP_PACKED, ctx, packedPairs, extUidLen, ligUidLen, isBest, groupSize, probeStats));
    }

    /** The probe task for a packed batch (results in dedupExtendOut.extend_pooled); sharded: collective over this context's communicator. */
    public void dedupExtendPacked(MemorySegment packedPairs, int extUidLen, int ligUidLen, boolean extendDuplicates, boolean sharded,
                                  MemorySegment dedupExtendOut) {
        check(invokeInt(DEDUP_EXTEND_PACKED, ctx, packedPairs, extUidLen, ligUidLen, extendDuplicates ? 1 : 0, sharded ? 1 : 0, dedupExtendOut));
    }

    // ---- resident batch: stage once, run, fetch ----------------------------------------------------------------------------
    public void extendStage(MemorySegment pairBatch) { check(invokeInt(EXTEND_STAGE, ctx, pairBatch)); }

    public void extendRun(int cigarCap, int mdCap) { check(invokeInt(EXTEND_RUN, ctx, cigarCap, mdCap)); }

    public void extendFetch(MemorySegment extendOut) { check(invokeInt(EXTEND_FETCH, ctx, extendOut)); }

    // ---- multi-GPU: the exchange runs inside the library over NCCL ----------------------------------------------------------
    /** One JVM that owns n GPUs (the reference's shape: probe tasks fanned out from one process, …Identification.java:836-846):
     *  pairs[k] / outs[k] (arrays of hsq_pair_batch / hsq_dedup_extend_out) belong to contexts[k]. */
    public static void dedupExtendMultiGpu(HsqB200[] contexts, MemorySegment pairBatches, int extUidLen, int ligUidLen, boolean extendDuplicates,
                                           MemorySegment dedupExtendOuts) {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment handles = a.allocate(ADDRESS, contexts.length);
            for (int k = 0; k < contexts.length; k++) handles.setAtIndex(ADDRESS, k, contexts[k].ctx);
            int st = invokeInt(DEDUP_EXTEND_MULTI_GPU, handles, contexts.length, pairBatches, extUidLen, ligUidLen, extendDuplicates ? 1 : 0, dedupExtendOuts);
            if (st != HSQ_OK) {
                StringBuilder msg = new StringBuilder();
                for (HsqB200 c : contexts) msg.append(lastError(c.ctx)).append(" | ");
                throw new HsqException(st, msg.toString());
            }
        }
    }

    /** Communicator over the contexts of this process (rank k = contexts[k]); dedupExtendMultiGpu does this on first use. */
    public static void commInitAll(HsqB200[] contexts) {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment handles = a.allocate(ADDRESS, contexts.length);
            for (int k = 0; k < contexts.length; k++) handles.setAtIndex(ADDRESS, k, contexts[k].ctx);
            int st = invokeInt(COMM_INIT_ALL, handles, contexts.length);
            if (st != HSQ_OK) throw new HsqException(st, lastError(contexts[0].ctx));
        }
    }

    /** One process per GPU: rank 0 makes the id (COMM_ID_BYTES bytes), the host program hands it to the others, every rank joins. */
    public static byte[] commUniqueId() {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment id = a.allocate(COMM_ID_BYTES);
            int st = invokeInt(COMM_UNIQUE_ID, id);
            if (st != HSQ_OK) throw new HsqException(st, lastError(MemorySegment.NULL));
            return id.toArray(JAVA_BYTE);
        }
    }

    public void commInitRank(byte[] id, int rank, int world) {
        MemorySegment seg = arena.allocate(COMM_ID_BYTES);
        seg.copyFrom(MemorySegment.ofArray(id));
        check(invokeInt(COMM_INIT_RANK, ctx, seg, rank, world));
    }

    public void commDestroy() { check(invokeInt(COMM_DESTROY, ctx)); }

    /** Collective: FilterByUid over the union of all ranks' pairs; answers for this rank's pairs, counters of the whole batch. */
    public void dedupSharded(MemorySegment pairBatch, int extUidLen, int ligUidLen, MemorySegment isBest, MemorySegment groupSize,
                             MemorySegment probeStats) {
        check(invokeInt(DEDUP_SHARDED, ctx, pairBatch, extUidLen, ligUidLen, isBest, groupSize, probeStats));
    }

    public void dedupExtendSharded(MemorySegment pairBatch, int extUidLen, int ligUidLen, boolean extendDuplicates, MemorySegment dedupExtendOut) {
        check(invokeInt(DEDUP_EXTEND_SHARDED, ctx, pairBatch, extUidLen, ligUidLen, extendDuplicates ? 1 : 0, dedupExtendOut));
    }

    /** The two halves of the sharded dedup for a host that runs its own exchange (device memory in / out). */
    public void dedupKeys(MemorySegment pairBatch, int extUidLen, int ligUidLen, long nProbes, int originBase, MemorySegment deviceKeysOut) {
        check(invokeInt(DEDUP_KEYS, ctx, pairBatch, extUidLen, ligUidLen, nProbes, originBase, deviceKeysOut));
    }

    public void dedupReduceKeys(MemorySegment deviceKeys, long n, int uidLen, MemorySegment deviceIsBest, MemorySegment deviceGroupSize,
                                MemorySegment probeStats) {
        check(invokeInt(DEDUP_REDUCE_KEYS, ctx, deviceKeys, n, uidLen, deviceIsBest, deviceGroupSize, probeStats));
    }

    /** The alignment DP on caller-owned DEVICE buffers without any allocation inside the library (no counterpart in the reference). */
    public void alignScoresNoAlloc(MemorySegment dRefData, MemorySegment dRefOffsets, long refBytes, MemorySegment dQryData, MemorySegment dQryOffsets,
                                   long qryBytes, long n, int maxRefLen, int maxQryLen, MemorySegment scorer, int flags, MemorySegment dScore,
                                   MemorySegment dNCols, MemorySegment dMatches, MemorySegment dMismatches, MemorySegment dStatus) {
        check(invokeInt(ALIGN_SCORES_NOALLOC, ctx, dRefData, dRefOffsets, refBytes, dQryData, dQryOffsets, qryBytes, n, maxRefLen, maxQryLen, scorer, flags,
                dScore, dNCols, dMatches, dMismatches, dStatus));
    }

    // ---- measurement -----------------------------------------------------------------------------------------------------
    /** Device time (ms) and kernel launches of the last batch call. */
    public float[] lastTiming() {
        MemorySegment ms = arena.allocate(JAVA_FLOAT), n = arena.allocate(JAVA_INT);
        check(invokeInt(LAST_TIMING, ctx, ms, n));
        return new float[] {ms.get(JAVA_FLOAT, 0), n.get(JAVA_INT, 0)};
    }

    /** Kernel-only spans (ms) of the last batch call; -1 where it has none. */
    public float[] lastSpans() {
        MemorySegment ms = arena.allocate(JAVA_FLOAT, 4);
        check(invokeInt(LAST_SPANS, ctx, ms));
        return ms.toArray(JAVA_FLOAT);
    }

    /** hsq_dedup_batch_report: hsq_dedup_batch + the UID composition tables (hsq_uid_report: int32 uid_len, two int32 tables
     *  [nProbes][uidLen][A, C, G, T]) that NucleotideCompositionUtil's strings are ratios of (FilterByUid.java:245-249). */
    public void dedupBatchReport(MemorySegment pairs, int extUidLen, int ligUidLen, MemorySegment isBest, MemorySegment groupSize,
                                 MemorySegment stats, MemorySegment uidReport) {
        check(invokeInt(DEDUP_BATCH_REPORT, ctx, pairs, extUidLen, ligUidLen, isBest, groupSize, stats, uidReport));
    }

    /** hsq_pair_reasons (host helper): one HSQ_PAIR_* code per pair of a probe task's outputs -- what ReadNameTracking's messages say. */
    public void pairReasons(MemorySegment dedupExtendOut, long n, boolean extendDuplicates, MemorySegment reason) {
        check(invokeInt(PAIR_REASONS, dedupExtendOut, n, extendDuplicates ? 1 : 0, reason));
    }

    /** hsq_uid_keys (host helper): one uint64 UID code per pair from the two packed read batches. */
    public void uidKeys(MemorySegment packedR1, MemorySegment packedR2, int extUidLen, int ligUidLen, MemorySegment out) {
        check(invokeInt(UID_KEYS, packedR1, packedR2, extUidLen, ligUidLen, out));
    }

    /** hsq_dedup_extend_packed_keys: hsq_dedup_extend_packed with the reduction fed by 20 bytes of keys per pair, the reads following in
     *  chunks under the kernels (same results). */
    public void dedupExtendPackedKeys(MemorySegment packedPairs, MemorySegment uidKeys, int extUidLen, int ligUidLen, boolean extendDuplicates,
                                      boolean sharded, MemorySegment out) {
        check(invokeInt(DEDUP_EXTEND_PACKED_KEYS, ctx, packedPairs, uidKeys, extUidLen, ligUidLen, extendDuplicates ? 1 : 0, sharded ? 1 : 0, out));
    }

    /** hsq_index_bam: record index of a level-0 BGZF BAM (the intermediate files of FastqAndBamFileMerger.java:131-133); `index` is an
     *  hsq_bam_index: seven int64 (cap, n, n_ref, text_bytes, first_record, payload_bytes, payload_cap), three pointers, one int32. */
    public void indexBam(MemorySegment file, long nBytes, MemorySegment index) { check(invokeInt(INDEX_BAM, ctx, file, nBytes, index)); }

    /** hsq_band_stats: {read ends aligned inside the band, of them redone with full matrices} of the last extend call. */
    public long[] bandStats() {
        MemorySegment a = arena.allocate(JAVA_LONG), b = arena.allocate(JAVA_LONG);
        check(invokeInt(BAND_STATS, ctx, a, b));
        return new long[] {a.get(JAVA_LONG, 0), b.get(JAVA_LONG, 0)};
    }

    public void timerStart() { check(invokeInt(TIMER_START, ctx)); }

    public float timerStop() {
        MemorySegment ms = arena.allocate(JAVA_FLOAT);
        check(invokeInt(TIMER_STOP, ctx, ms));
        return ms.get(JAVA_FLOAT, 0);
    }

    public double int32Peak(double seconds) {
        MemorySegment v = arena.allocate(JAVA_DOUBLE);
        check(invokeInt(INT32_PEAK, ctx, seconds, v));
        return v.get(JAVA_DOUBLE, 0);
    }

    MemorySegment handle() { return ctx; }

    Arena arena() { return arena; }

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
