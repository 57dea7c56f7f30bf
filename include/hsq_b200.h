/* hsq_b200.h -- C ABI of the B200-native HSQUtils hot path (trim / UID dedup / primer extension).
 *
 * This is the drop-in boundary.  The reference (NimbleGen/bioinformatics, pure Java) has no FFI;
 * the seams this ABI sits behind are ordinary Java entry points.  Each function cites the one it
 * replaces.  Roots under /root/reference:
 *   COMMON/ = bioinformatics_common/src/main/java/com/roche/sequencing/bioinformatics/common/
 *   HSQ/    = nimblegen_heatseq/src/main/java/com/roche/heatseq/
 *
 * Conventions
 *   - plain C, no C++/torch types; every function returns int32 status (0 = HSQ_OK, <0 = error);
 *     hsq_last_error(ctx) gives a message.  No exception crosses the boundary, nothing is dropped
 *     silently, and there is NO CPU fallback: without a usable sm_100 device hsq_ctx_create fails.
 *   - caller owns every input/output buffer (host memory; pinned via hsq_pinned_alloc is fastest),
 *     the library owns only the opaque context and its device scratch.
 *   - one context per GPU; a context is used by one thread at a time; contexts are independent.
 *   - sequences are ASCII (IUPAC letters, any case, '_' '.' '-' = gap) exactly as the Java side
 *     holds them in Strings; encoding to device codes happens on the GPU.
 *   - "read end" index m = 2*pair + mate (mate 0 = read one / extension side, 1 = read two).
 */
#ifndef HSQ_B200_H
#define HSQ_B200_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HSQ_OK 0
#define HSQ_ERR_INVALID_ARGUMENT (-1)
#define HSQ_ERR_CUDA (-2)
#define HSQ_ERR_NO_DEVICE (-3)
#define HSQ_ERR_BAD_CHARACTER (-4) /* a sequence holds a byte that is not an IUPAC code (Java: IllegalArgumentException) */
#define HSQ_ERR_UNSUPPORTED (-5)   /* input outside the accelerated domain (e.g. UID longer than 32 bases) */
#define HSQ_ERR_OUT_OF_RANGE (-6)  /* Java would throw IndexOutOfBounds / IllegalArgument (e.g. trim start >= read length) */
#define HSQ_ERR_OVERFLOW (-7)      /* an output slot was too small; the per-item overflow flags say which */
#define HSQ_ERR_NAME_MISMATCH (-8) /* FASTQ one / two read names differ at a visited position (IllegalStateException, FastqAndBamFileMerger.java:270-273) */
#define HSQ_ERR_NOT_TRIMMED (-9)   /* a SAM read is longer than its trimmed FASTQ read (UnableToMergeFastqAndBamFilesException, :366-371) */
#define HSQ_ERR_NO_MATCHES (-10)   /* no FASTQ pair matched any SAM record (IllegalStateException, :143-148) */
#define HSQ_ERR_MALFORMED (-11)    /* FASTQ text is not a sequence of 4-line records (htsjdk FastqReader throws SAMException) */

#define HSQ_PRIMER_NOT_FOUND (-2147483647) /* Java `null` from getPrimerEndIndexInRead */

#define HSQ_ALIGN_PENALIZE_START 1u /* IAlignmentScorer.shouldPenalizeStartingTerminalGaps */
#define HSQ_ALIGN_PENALIZE_END 2u   /* IAlignmentScorer.shouldPenalizeEndingTerminalGaps */

typedef struct hsq_ctx hsq_ctx;

/* SimpleAlignmentScorer's four numbers (COMMON/alignment/SimpleAlignmentScorer.java:28-31).
 * Integer valued in every reachable configuration (HSQ/cli/DeduplicationCli.java:288-327). */
typedef struct hsq_scorer {
    int32_t match, mismatch, gap_extend, gap_start;
} hsq_scorer;

/* Ragged batch of ASCII strings: string k is data[offsets[k] .. offsets[k+1]). */
typedef struct hsq_seqs {
    const uint8_t* data;
    const int64_t* offsets; /* n+1 entries */
    int64_t n;
} hsq_seqs;

/* Probe table = the fields of COMMON/utils/probeinfo/Probe.java:53-69 the hot path reads. */
typedef struct hsq_probe_table {
    int64_t n;
    const uint8_t* reverse_strand; /* 1 when probe_strand is '-' */
    const int32_t* ext_start;
    const int32_t* ext_stop;
    const int32_t* lig_start;
    const int32_t* lig_stop;
    const int32_t* target_start;
    const int32_t* target_stop;
    hsq_seqs ext, lig, target; /* primer and capture-target sequences as in the probe-info file */
} hsq_probe_table;

/* Read pairs as they reach PrimerReadExtensionAndFilteringOfUniquePcrProbesTask
 * (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:836): raw reads + qualities,
 * read name = FASTQ ordinal, assigned probe. */
typedef struct hsq_pair_batch {
    int64_t n;
    const int32_t* probe_id;   /* row of the probe table */
    const int32_t* read_index; /* Integer.parseInt(readName); may be NULL for hsq_extend_batch */
    hsq_seqs r1, r2;           /* raw read strings */
    const uint8_t* q1;         /* base-quality chars, laid out by r1.offsets; may be NULL for hsq_extend_batch */
    const uint8_t* q2;         /* laid out by r2.offsets */
} hsq_pair_batch;

/* Per read end results of ExtendReadsToPrimer.calculateDetailsForReadExtensionToPrimer
 * (HSQ/process/ExtendReadsToPrimer.java:200-245).  Arrays have 2*n entries (index 2*pair+mate). */
typedef struct hsq_extend_out {
    int32_t cigar_cap, md_cap; /* in: bytes per slot */
    uint8_t* ok;               /* 1 = ReadExtensionDetails != null */
    int32_t* primer_end;       /* getPrimerEndIndexInRead, HSQ_PRIMER_NOT_FOUND for null */
    int32_t* score;            /* capture-target alignment score (0 when that DP was not reached) */
    int32_t* pos;              /* alignmentStartInReference */
    int32_t* read_start;       /* first kept base of the raw read */
    int32_t* read_stop;        /* last kept base (inclusive) */
    int32_t* nm;               /* editDistance (NM tag) */
    int32_t* tlen;             /* inferred insert size (+-capture target length) */
    int32_t* cigar_len;
    uint8_t* cigar;            /* [2n][cigar_cap] standard CIGAR (M/I/D), not NUL terminated */
    int32_t* md_len;
    uint8_t* md;               /* [2n][md_cap] MD tag */
    uint8_t* overflow;         /* bit0: cigar slot too small, bit1: md slot too small */
    uint8_t* pair_ok;          /* [n] both mates extended (ExtendReadsToPrimer.java:364) */
    int64_t cells;             /* out: DP cells filled = sum len(q)*len(r) over the DPs the reference runs */
} hsq_extend_out;

/* Operator seam: one NeedlemanWunschGlobalAlignment per item
 * (COMMON/alignment/NeedlemanWunschGlobalAlignment.java:44,85,131,236,250; AlignmentPair.java). */
typedef struct hsq_align_out {
    int32_t cols_cap, md_cap;  /* in: bytes per slot */
    int32_t* score;            /* getAlignmentScore */
    int32_t* n_cols;           /* alignment length */
    uint8_t* ref_aln;          /* [n][cols_cap] reference alignment as characters (gap = '_') */
    uint8_t* qry_aln;          /* [n][cols_cap] query alignment */
    uint8_t* ops;              /* [n][cols_cap] expanded cigar: '=' 'X' 'I' 'D' (CigarStringUtil.java:61-83) */
    int32_t* first_q;          /* AlignmentPair.firstNonDeletionIndexInQuery */
    int32_t* last_q;
    int32_t* first_r;
    int32_t* last_r;
    int32_t* matches;          /* getNumberOfMatches */
    int32_t* mismatches;       /* getNumberOfMismatches */
    int32_t* non_terminal_gaps;/* getNumberOfNonTerminalGaps */
    int32_t* edit_distance;    /* AlignmentPair.getEditDistance (equality based) */
    int32_t* md_len;
    uint8_t* md;               /* [n][md_cap] getMismatchDetailsString */
    int32_t* rmd_len;
    uint8_t* rmd;              /* [n][md_cap] getReverseMismatchDetailsString */
    uint8_t* overflow;         /* bit0 cols, bit1 md, bit2 rmd */
    int64_t cells;             /* out */
    /* [n][5], may be NULL: the counters FastqFindAndTrimTool filters on (COMMON/fastqtool/FastqFindAndTrimTool.java:932-938),
     * taken over getAlignmentWithoutEndingAndBeginningQueryInserts (AlignmentPair.java:95-99): getNumberOfMismatches,
     * getNumberOfInsertionGapsRelativeToReference, getNumberOfInsertionsRelativeToReference,
     * getNumberOfDeletionGapsRelativeToReference, getNumberOfDeletionsRelativeToReference (AlignmentPair.java:190-256) */
    int32_t* find_counts;
} hsq_align_out;

/* Per probe counters of FilterByUid.reduceReadsByProbeAndUid (HSQ/process/FilterByUid.java:149-171,251). */
typedef struct hsq_probe_stats {
    int64_t n_probes;          /* in: rows in the arrays below */
    int32_t* unique_pairs;     /* number of UID groups */
    int32_t* duplicate_pairs;
    int32_t* max_pairs_per_uid;
} hsq_probe_stats;

/* Candidates of the multi-probe tie-break (…Identification.java:704-765): read k has candidate
 * probes cand_probe[cand_offsets[k] .. cand_offsets[k+1]). */
typedef struct hsq_candidates {
    int64_t n;                 /* reads */
    hsq_seqs r1, r2;
    const int64_t* cand_offsets;
    const int32_t* cand_probe;
} hsq_candidates;

/* Per read results of hsq_find_candidates; slot arrays have max_candidates entries per read. */
typedef struct hsq_find_out {
    int32_t max_candidates;  /* in: slots per read */
    int32_t* lead_trimmed;   /* trimReadBasedOnQuality: leading bases with quality < leading threshold */
    int32_t* trail_trimmed;  /* ... trailing bases with quality < trailing threshold, never past the leading trim */
    uint8_t* has_n;          /* 1: the trimmed read holds an N / n: the tool skips the entry (no candidates) */
    int32_t* n_candidates;   /* candidates of the read (<= max_candidates) */
    int32_t* candidates;     /* [n][max_candidates] rows of the sequences-to-find table, most hits first, ties by ascending row */
    int32_t* hits;           /* [n][max_candidates] tally of each candidate */
    uint8_t* overflow;       /* 1: more candidates than slots, call again with more slots */
    uint8_t* order_dependent;/* 1: the walk stopped on the min-ratio test inside a group of equal tallies that also holds
                              * passing entries, so the reference's list depends on its HashMap iteration order; the list
                              * here is the one for ascending-row order */
} hsq_find_out;

/* The five limits of the find settings (FastqToolFindSettings: allowed mismatch bases, insertion gaps / bases, deletion
 * gaps / bases; COMMON/fastqtool/FastqFindAndTrimTool.java:940-949). */
typedef struct hsq_find_limits {
    int32_t mismatches, insertion_gaps, insertion_bases, deletion_gaps, deletion_bases;
} hsq_find_limits;

/* Results of hsq_find_matches: the candidate stage's arrays plus, per candidate slot, the alignment's counters and verdict. */
typedef struct hsq_find_match_out {
    hsq_find_out cand;
    int32_t* counts;     /* [n][max_candidates][5]: mismatches, insertion gaps, insertion bases, deletion gaps, deletion bases */
    uint8_t* accepted;   /* [n][max_candidates] 1: all five counters within the limits (isAcceptable :947) */
    int32_t* n_matches;  /* matches.size() of the read: 0 no match, 1 single match, > 1 multiple matches (:953-985) */
    int64_t cells;       /* out: DP cells computed */
} hsq_find_match_out;

/* Key record of the sharded dedup: what one pair contributes to the (probe, UID) grouping.
 * uid = 2 bit per base of R1[0:u] ++ R2[0:v]; rank orders the pairs of a group best first:
 * (32767 - (short)qualitySum) << 32 | (readIndex ^ 0x80000000)  (FilterByUid.java:281-297); origin identifies the pair
 * on the rank that holds its reads. */
typedef struct hsq_dedup_key {
    uint64_t uid;
    uint64_t rank;
    int32_t probe;
    uint32_t origin;
} hsq_dedup_key; /* 24 bytes */

/* One mergeSequences call per item (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:1362-1552): the
 * upstream and downstream mate of a pair on the genome's forward strand with their 1-based alignment starts. */
typedef struct hsq_merge_batch {
    int64_t n;
    const uint8_t* negative_strand; /* 1: the merged read is reverse-complemented, its qualities reversed (:1544-1548) */
    hsq_seqs up, down;              /* upstream / downstream sequences */
    const uint8_t* up_qual;         /* laid out by up.offsets */
    const uint8_t* down_qual;       /* laid out by down.offsets */
    const int32_t* up_start;        /* upstreamAlignmentStart (>= 1) */
    const int32_t* down_start;      /* downstreamAlignmentStart (>= up_start when the two overlap) */
} hsq_merge_batch;

/* ---- coordinate read -> probe assignment ---------------------------------------------------- */
/* Search window of every probe, in probe-info-file order: what getProbeSearchBoundaries returns
 * (…Identification.java:285-297) and the contig / strand the probe sits on.  One RangeMap per (contig, strand)
 * (HSQ/process/RangeMap.java; built at …Identification.java:313-331). */
typedef struct hsq_probe_ranges {
    int64_t n;
    const int32_t* contig;         /* contig id of the probe (caller's numbering, >= 0) */
    const uint8_t* reverse_strand; /* 1 when probe_strand is '-' */
    const int32_t* start;          /* queryStart */
    const int32_t* stop;           /* queryStop */
} hsq_probe_ranges;

/* Mapped records as ReadToProbeAssigner.assignProbesToReads sees them (…Identification.java:439-545): both mates mapped
 * (the caller drops records with the unmapped / mate-unmapped flag, :446), read name = FASTQ ordinal. */
typedef struct hsq_mapped_records {
    int64_t n;
    const int32_t* contig;          /* contig id; -1 = a contig without any probe (probeRanges == null, :462): record ignored */
    const uint8_t* negative_strand; /* getReadNegativeStrandFlag */
    const uint8_t* second_of_pair;  /* getSecondOfPairFlag (the strand is flipped for read two, :453-455) */
    const int32_t* start;           /* getAlignmentStart */
    const int32_t* end;             /* getAlignmentEnd */
    const int32_t* read_index;      /* Integer.parseInt(readName) */
    /* alternative hits of record k (XA tag, SAMRecordUtil.getAlternativeHitsFromAttribute): alt_offsets[k] .. alt_offsets[k+1];
     * alt_offsets may be NULL when no record has any */
    const int64_t* alt_offsets;
    const int32_t* alt_contig;      /* -1 = contig without probes */
    const uint8_t* alt_negative;    /* alternativeHit.getStrand() == REVERSE */
    const int32_t* alt_start;
    const int32_t* alt_stop;
} hsq_mapped_records;

/* Per record results; slot arrays have max_probes entries per record. */
typedef struct hsq_assign_out {
    int32_t max_probes;    /* in: slots per record */
    int32_t* n_probes;     /* probes whose window contains the record or one of its alternative hits (<= max_probes) */
    int32_t* probes;       /* [n][max_probes] those probes (rows of hsq_probe_ranges), ascending */
    int32_t* mate;         /* record index of the mate; HSQ_ASSIGN_UNPAIRED / _AMBIGUOUS_NAME / _IGNORED */
    int32_t* n_common;     /* probes assigned to the PAIR = intersection of the two mates' sets (:478-483); 0 unless paired */
    int32_t* common;       /* [n][max_probes] ascending; identical for both mates */
    uint8_t* overflow;     /* 1: a set of this record or of its mate did not fit max_probes: its slots are incomplete, call
                            * again with more slots */
} hsq_assign_out;
#define HSQ_ASSIGN_UNPAIRED (-1)
#define HSQ_ASSIGN_AMBIGUOUS_NAME (-2) /* the read name occurs on four or more usable records */
#define HSQ_ASSIGN_IGNORED (-3)        /* the record's contig holds no probe */

/* ---- report inputs of a probe task beyond the three counters ---------------------------------------------------------- */
/* FilterByUid.java:149-171,245-252 hands NucleotideCompositionUtil the set of a probe's UIDs and the list of its pairs' UIDs; the
 * strings it formats are ratios of these counts.  Tables are [n_probes][uid_len][4] in the order A, C, G, T (the accelerated UID
 * domain, see hsq_dedup_batch).  The per-UID tallies of the same report (uid:count, :213-237) are the UID and group_size of the
 * pairs whose is_best is 1. */
typedef struct hsq_uid_report {
    int32_t uid_len;                /* out: ext_uid_len + lig_uid_len */
    int32_t* unique_by_position;    /* one UID per group */
    int32_t* weighted_by_position;  /* every pair's UID */
} hsq_uid_report;
/* hsq_dedup_batch plus the UID composition tables, counted on the device from the keys of the same call */
int32_t hsq_dedup_batch_report(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t ext_uid_len, int32_t lig_uid_len, uint8_t* is_best,
                               int32_t* group_size, hsq_probe_stats* stats, hsq_uid_report* report);

/* ---- one probe task, batched: dedup then extend (…Identification.java:1172-1234) ----------------------------------- */
typedef struct hsq_dedup_extend_out {
    uint8_t* is_best;                   /* [n] as hsq_dedup_batch */
    int32_t* group_size;                /* [n] may be NULL */
    hsq_probe_stats* stats;             /* per-probe counters of the reduction (required) */
    int32_t* unique_unable_to_extend;   /* [stats->n_probes] numberOfUniqueReadPairsUnableExtendPrimer (:1189); may be NULL */
    int32_t* duplicate_unable_to_extend;/* [stats->n_probes] numberOfDuplicateReadPairsUnableExtendPrimer (:1208); may be NULL */
    int64_t n_extended;                 /* out: pairs handed to the extension = slots used in `extend` */
    int32_t* extended_pair;             /* [n] out: extended_pair[j] = index of the pair whose results sit in slot j (ascending) */
    hsq_extend_out* extend;             /* arrays sized for n pairs (2n read ends); slots j < n_extended are written */
    struct hsq_extend_pooled_out* extend_pooled; /* hsq_dedup_extend_packed writes here instead of `extend` (same slot numbering) */
} hsq_dedup_extend_out;

/* Why a pair of a probe task was or was not written (the analogue of ReadNameTracking's messages, HSQ/process/ReadNameTracking.java:30-66,
 * FilterByUid.java:176-198, ...Identification.java:1185-1212), from the outputs of hsq_dedup_extend_batch / _packed.  Host helper, no device work.
 * extend_duplicates as passed to that call.  reason[k] for pair k: */
enum {
    HSQ_PAIR_WRITTEN_UNIQUE = 0,          /* best pair of its UID group, both reads extended */
    HSQ_PAIR_WRITTEN_DUPLICATE = 1,       /* --markDuplicates: written with the duplicate flag */
    HSQ_PAIR_DUPLICATE_DROPPED = 2,       /* not the best pair of its group, duplicates are not written */
    HSQ_PAIR_PRIMER_NOT_FOUND_ONE = 3,    /* extension primer not found in read one (getPrimerEndIndexInRead) */
    HSQ_PAIR_PRIMER_NOT_FOUND_TWO = 4,    /* ligation primer not found in read two */
    HSQ_PAIR_NOT_ALIGNED_ONE = 5,         /* read one did not align to the capture target (score <= 0 or too short, ExtendReadsToPrimer.java:212) */
    HSQ_PAIR_NOT_ALIGNED_TWO = 6
};
int32_t hsq_pair_reasons(const hsq_dedup_extend_out* out, int64_t n, int32_t extend_duplicates, uint8_t* reason);

/* ---- SAM record fields of extended pairs (SURVEY.md 8(a) A5 outputs, 8(f) f4) ---------------------------------------- */
/* What ExtendReadsToPrimer.extendReadPair / createRecord and SAMRecordUtil.setSAMRecordsAsPair put on the two records of a
 * pair beyond the fields of hsq_extend_out (ExtendReadsToPrimer.java:131-181,247-284; SAMRecordUtil.java:270-289).
 * Entry 2k is read one of pair k, 2k+1 read two.  Pairs whose pair_ok is 0 get empty strings, flag 0 and mate_pos 0
 * (the reference writes only pairs whose two reads extended, ExtendReadsToPrimer.java:364). */
typedef struct hsq_record_out {
    uint16_t* flag;     /* [2n] SAM FLAG: 0x1 paired | 0x2 proper pair | 0x10 read on - strand | 0x20 mate on - strand |
                         * 0x40 first / 0x80 second of pair | 0x400 duplicate */
    int32_t* mate_pos;  /* [2n] PNEXT: alignment start of the other record */
    int64_t* offsets;   /* [2n+1] byte range of each record in seq / qual */
    uint8_t* seq;       /* read strings: read[readStart..readStop], reverse-complemented on the - strand */
    uint8_t* qual;      /* base qualities of the same range, reversed on the - strand */
} hsq_record_out;

/* ---- N1: read batches packed to 2 or 4 bits per base (north_star subsystem 1) ------------------------------------------ */
/* The reference itself keeps reads packed (2 bit: COMMON/sequence/NucleotideCodeSequence.java:32-48; 5 bit IUPAC:
 * COMMON/sequence/IupacNucleotideCodeSequence.java:34-58); this is the form in which a host should hand large batches over
 * PCIe.  Read k occupies words(len[k]) = (len[k] * bits + 31) / 32 little-endian 32-bit words, reads back to back; base i of a
 * read sits at bits [bits * (i % (32 / bits)), +bits) of its word i / (32 / bits).
 *   2 bit: A = 0, C = 1, T = 2, G = 3 (bits 1-2 of the upper-case ASCII letter).  Every base that is not an upper- or
 *          lower-case A/C/G/T (N, other IUPAC codes, gap characters) is packed as 0 and listed in the exception arrays
 *          (ascending by read, then position) with its ASCII character; the device restores it.
 *   4 bit: the IUPAC nucleotide set (A = 1, C = 2, G = 4, T = 8, N = 15, gap = 0); 'U' is not representable
 *          (hsq_pack_reads fails with HSQ_ERR_UNSUPPORTED), no exception list.
 * Letter case is folded (the alignment is case blind, IupacNucleotideCode.getCodesFromString :86-104). */
#define HSQ_PACK_2BIT 2
#define HSQ_PACK_4BIT 4
typedef struct hsq_packed_reads {
    int32_t bits;             /* HSQ_PACK_2BIT or HSQ_PACK_4BIT */
    int64_t n;
    const uint8_t* data;      /* 4 * sum(words(len[k])) bytes */
    const int32_t* len;       /* [n] bases per read */
    int64_t n_exceptions;     /* 2 bit only */
    const uint32_t* exc_read; /* [n_exceptions] read index, ascending */
    const uint32_t* exc_pos;  /* [n_exceptions] base position inside the read */
    const uint8_t* exc_char;  /* [n_exceptions] the ASCII character at that position */
    /* optional totals (0 = the library adds up `len` on the host, a pass over n entries before the first copy can start): bytes of
     * `data` (hsq_packed_bytes) and sum of len.  When given they are checked on the device (a mismatch: HSQ_ERR_INVALID_ARGUMENT). */
    int64_t data_bytes;
    int64_t n_bases;
} hsq_packed_reads;

/* hsq_pair_batch with packed reads.  Qualities stay bytes, laid out back to back in read order (read k at the sum of the
 * lengths before it); NULL where a call does not read them (hsq_extend_packed). */
typedef struct hsq_packed_pairs {
    int64_t n;
    const int32_t* probe_id;
    const int32_t* read_index;
    hsq_packed_reads r1, r2;
    const uint8_t* q1;
    const uint8_t* q2;
    /* optional, hsq_dedup_packed only: (short) sum(q - 33) per read as hsq_quality_sums computes it (BamFileUtil.getQualityScore,
     * HSQ/utils/BamFileUtil.java:74-85).  When both are given q1 / q2 are not read and may be NULL: 2 bytes per read cross PCIe
     * instead of the quality string.  NULL (the default): the device sums q1 / q2. */
    const int16_t* qsum1;
    const int16_t* qsum2;
} hsq_packed_pairs;

/* Results of hsq_extend_packed: the per read end fields of hsq_extend_out (any array may be NULL: it is then not copied to
 * the host) with CIGAR and MD in one byte pool instead of fixed slots: the CIGAR of read end m is
 * pool[str_off[m] .. str_off[m] + cigar_len[m]), its MD follows directly.  The pool is filled in read end order.  If `pool` is
 * pinned host memory (hsq_pinned_alloc) the device writes it directly while later chunks still compute. */
typedef struct hsq_extend_pooled_out {
    int32_t cigar_cap, md_cap; /* in: bytes per device-side slot (0 = 64 / 96); a longer string fails the call with HSQ_ERR_OVERFLOW */
    uint8_t* ok;
    int32_t* primer_end;
    int32_t* score;
    int32_t* pos;
    int32_t* read_start;
    int32_t* read_stop;
    int32_t* nm;
    int32_t* tlen;
    int32_t* cigar_len;
    int32_t* md_len;
    int64_t* str_off;          /* [2n] */
    uint8_t* pool;
    int64_t pool_cap;          /* in: bytes available in pool */
    int64_t pool_bytes;        /* out: bytes used (needed, when the call fails with HSQ_ERR_OVERFLOW) */
    uint8_t* pair_ok;          /* [n] */
    int64_t cells;             /* out */
} hsq_extend_pooled_out;

/* Library behaviour that a caller (or a test) may change; zero-initialise for the defaults.  Replaces the environment
 * variables earlier builds read. */
typedef struct hsq_config {
    int32_t struct_bytes;       /* sizeof(hsq_config) of the caller's header */
    int32_t force_dp32;         /* 1: always run the 32-bit DP kernels (the packed u16x2 kernels are the default when the scores fit) */
    int32_t force_dedup_radix;  /* 1: dedup through the radix-sort cross-check path instead of the per-probe group-by */
    int32_t extend_chunks;      /* chunks of the host pipeline of hsq_extend_batch / hsq_extend_packed; 0 = automatic */
    int32_t uniform_chunks;     /* 1: equal chunk sizes instead of the ramp */
    int32_t band;               /* opt-in band half-width of the capture-target DP; 0 = full matrices.  Results stay those of the
                                 * full matrix: an alignment whose band is touched is re-run in full (see hsq_band_stats) */
    int32_t trim_chunks;        /* chunks of hsq_trim_batch's copy-in | kernel | copy-out pipeline; 0 = automatic (by size), 1 = none */
    int32_t l2_persist_rows;    /* 1: the boundary rows of the packed DP kernels are given an L2 persisting access window (experiment) */
} hsq_config;

/* ---- context ------------------------------------------------------------------------------ */
int32_t hsq_ctx_create(int32_t device, hsq_ctx** out);
/* as hsq_ctx_create, with a configuration (NULL = defaults) */
int32_t hsq_ctx_create_ex(int32_t device, const hsq_config* config, hsq_ctx** out);
/* changes the configuration of an existing context (takes effect with the next batch call) */
int32_t hsq_ctx_configure(hsq_ctx* ctx, const hsq_config* config);
void hsq_ctx_destroy(hsq_ctx* ctx);
const char* hsq_last_error(const hsq_ctx* ctx); /* ctx may be NULL: last create error */
const char* hsq_version(void);
int32_t hsq_pinned_alloc(hsq_ctx* ctx, size_t bytes, void** out);
int32_t hsq_pinned_free(hsq_ctx* ctx, void* p);
/* ApplicationSettings.getAlignmentScorer(): the user scorer for the capture-target DP. */
int32_t hsq_set_scorer(hsq_ctx* ctx, const hsq_scorer* scorer);
/* ParsedProbeFile -> device resident probe table (copied; caller may free its arrays). */
int32_t hsq_set_probes(hsq_ctx* ctx, const hsq_probe_table* probes);
/* Device time (ms, CUDA events on the context's stream) of the kernels of the last batch call,
 * and how many kernels it launched. */
int32_t hsq_last_timing(const hsq_ctx* ctx, float* kernel_ms, int32_t* launches);
/* Kernel-only device times (ms, CUDA events between the copies and the kernels) of the last batch call, -1 where the call has no
 * such span: hsq_trim_batch [0] length + scan + copy kernels; hsq_dedup_batch / hsq_dedup_packed / hsq_dedup_sharded /
 * hsq_dedup_extend_* [0] key generation, [1] per-probe bucketing + group-by (sharded: the exchange and the owner's reduction),
 * [2] compaction of the pairs to extend, [3] the extension kernels. */
int32_t hsq_last_spans(const hsq_ctx* ctx, float* ms4);
/* The opt-in band (hsq_config.band) in the last extend call: *banded = read ends whose capture-target DP ran inside the band,
 * *redone = those of them whose traceback left the diagonals |j - i| < band (or that were not traced back at all) and that were
 * therefore aligned again with full matrices.  Both 0 when the band is off.  Either pointer may be NULL. */
int32_t hsq_band_stats(const hsq_ctx* ctx, int64_t* banded, int64_t* redone);

/* ---- batch entry points (host buffers in, host buffers out) -------------------------------- */
/* new NeedlemanWunschGlobalAlignment(ref[k], qry[k], SimpleAlignmentScorer(scorer, flags)).getAlignmentPair() */
int32_t hsq_align_batch(hsq_ctx* ctx, const hsq_seqs* refs, const hsq_seqs* queries, const hsq_scorer* scorer,
                        uint32_t flags, hsq_align_out* out);
/* SAMRecordUtil.getVariableLengthUid (HSQ/utils/SAMRecordUtil.java:223-261; SURVEY row A10, dead in the CLI: allowVariableLengthUids is
 * false, HSQ/cli/DeduplicationCli.java:120): one NeedlemanWunschGlobalAlignment(read, primer, scorer) per item; uid_len[k] = length of the
 * read prefix in front of the first column holding a primer base, or -1 where the reference returns null (no such column, or insertions +
 * deletions + substitutions >= primer length / 4 -- with the reference's never-incremented insertsSinceLastInsertionOrMismatch, :238-250,
 * every non-match column counts).  The extension side passes the extension primer, the ligation side revcomp(ligation primer) (:215-217). */
int32_t hsq_variable_length_uid(hsq_ctx* ctx, const hsq_seqs* reads, const hsq_seqs* primers, const hsq_scorer* scorer, uint32_t flags,
                                int32_t* uid_len);
/* FastqReadTrimmer.trim(String,String,int,int,boolean) (HSQ/process/FastqReadTrimmer.java:309-328) per read.
 * out_offsets gets n+1 entries; out_seq/out_qual must hold the input's total bytes. */
int32_t hsq_trim_batch(hsq_ctx* ctx, const hsq_seqs* reads, const uint8_t* quals, int32_t first_base_to_keep,
                       int32_t last_base_to_keep, int32_t three_prime, uint8_t* out_seq, uint8_t* out_qual,
                       int64_t* out_offsets);
/* FilterByUid.reduceReadsByProbeAndUid + findBestData over every probe of the batch at once
 * (HSQ/process/FilterByUid.java:71-203,281-297).  is_best[k]=1: pair k represents its (probe,UID) group. */
/* Accelerated domain of the UID: upper-case A/C/G/T only, at most 32 bases in total.  The reference groups by the raw UID string
 * (FilterByUid.java:134-147), so "acgt" and "ACGT" are different groups there; the caller has already dropped pairs with an N / n
 * (…Identification.java:815-821).  A UID with any other byte -- a lower-case base included -- is REFUSED with HSQ_ERR_UNSUPPORTED rather than
 * grouped differently from the reference; a read shorter than its UID is HSQ_ERR_OUT_OF_RANGE (String.substring throws there). */
int32_t hsq_dedup_batch(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t ext_uid_len, int32_t lig_uid_len,
                        uint8_t* is_best, int32_t* group_size, hsq_probe_stats* stats);
/* ExtendReadsToPrimer.extendReadsToPrimers (HSQ/process/ExtendReadsToPrimer.java:347-380) for a batch that
 * may span probes.  Needs hsq_set_probes and hsq_set_scorer first. */
int32_t hsq_extend_batch(hsq_ctx* ctx, const hsq_pair_batch* pairs, hsq_extend_out* out);
/* The three phases of hsq_extend_batch, separately callable so a caller can keep a batch resident in HBM
 * (stage = host->device copies; run = encode + DP + traceback + CIGAR/MD kernels; fetch = device->host copies). */
int32_t hsq_extend_stage(hsq_ctx* ctx, const hsq_pair_batch* pairs);
int32_t hsq_extend_run(hsq_ctx* ctx, int32_t cigar_cap, int32_t md_cap);
int32_t hsq_extend_fetch(hsq_ctx* ctx, hsq_extend_out* out);
/* Multi-probe tie-break (…Identification.java:704-765): best_probe[k] = winning probe row or -1 when
 * the read must be removed from all probes.  matches/edit (per candidate, may be NULL) expose the two metrics. */
int32_t hsq_resolve_multi_probe(hsq_ctx* ctx, const hsq_candidates* cands, int32_t ext_uid_len, int32_t lig_uid_len,
                                int32_t* best_probe, int32_t* cand_matches, int32_t* cand_edit);


/* The body of PrimerReadExtensionAndFilteringOfUniquePcrProbesTask.run for every probe of the batch at once
 * (…Identification.java:1172-1216): reduceReadsByProbeAndUid, then extendReadsToPrimers on the unique pairs and -- only when
 * extend_duplicates != 0 (keepDuplicates or markDuplicates, :1203) -- on the duplicates too.  The reads are uploaded once;
 * the unique pairs are compacted on the device and the extension runs on them alone.  Results equal hsq_dedup_batch
 * followed by hsq_extend_batch on the pairs listed in extended_pair. */
int32_t hsq_dedup_extend_batch(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t ext_uid_len, int32_t lig_uid_len,
                               int32_t extend_duplicates, hsq_dedup_extend_out* out);

/* Record fields of the pairs of a batch that hsq_extend_batch (or hsq_dedup_extend_batch) extended: `ext` holds that
 * call's host results for the same pairs (ok, read_start, read_stop, pos, pair_ok are read).  duplicate[k] != 0 sets the
 * duplicate flag of pair k (isMarkedDuplicate, ExtendReadsToPrimer.java:266); NULL = none.  seq / qual have out_cap bytes;
 * when more are needed the call returns HSQ_ERR_OVERFLOW with out->offsets filled. */
int32_t hsq_extend_records(hsq_ctx* ctx, const hsq_pair_batch* pairs, const hsq_extend_out* ext, const uint8_t* duplicate,
                           hsq_record_out* out, int64_t out_cap);

/* ExtendReadsToPrimer.extendReadsToPrimers on the pairs with keep[k] != 0 only (e.g. the unique pairs of a dedup whose
 * reduction ran sharded over several GPUs): the batch is uploaded once, the flagged pairs are compacted on the device and
 * extended; slot j of `out` (sized for n pairs) belongs to pair extended_pair[j], j < *n_extended. */
int32_t hsq_extend_selected(hsq_ctx* ctx, const hsq_pair_batch* pairs, const uint8_t* keep, int64_t* n_extended,
                            int32_t* extended_pair, hsq_extend_out* out);

/* has_n[k] = 1 when read one or read two of pair k holds an 'N' or 'n': such pairs are not handed to the probe tasks
 * (readString.toLowerCase().contains("n"), …Identification.java:815-821). */
int32_t hsq_flag_n_pairs(hsq_ctx* ctx, const hsq_pair_batch* pairs, uint8_t* has_n);

/* SimpleMapper as FastqFindAndTrimTool builds it (COMMON/mapping/SimpleMapper.java:90-143,
 * COMMON/fastqtool/FastqFindAndTrimTool.java:231-237): every k-mer of every sequence to find at start positions
 * 0 .. size-k-1 (the last k-mer is not indexed, :93), reference and query spacing 1; a k-mer held by more than
 * max_refs_per_kmer sequences is dropped from the index (:135-138; <= 0 = no cap).  Sequences must hold A/C/G/T only
 * (NucleotideCode.getNucleotidesFromString throws otherwise: HSQ_ERR_BAD_CHARACTER) and be at least k long
 * (IllegalStateException :98: HSQ_ERR_OUT_OF_RANGE).  1 <= k <= 12. */
int32_t hsq_set_find_sequences(hsq_ctx* ctx, const hsq_seqs* seqs, int32_t kmer, int32_t max_refs_per_kmer);
/* The per-read front half of FastqFindAndTrimTool.findAndTrim (FastqFindAndTrimTool.java:904-926): quality trim
 * (trimReadBasedOnQuality :1188-1219, Phred+33), the N test on the trimmed read (:920) and
 * SimpleMapper.getBestCandidateReferences (SimpleMapper.java:154-217) with the given candidate limit and hit ratio
 * (10 and 0.5 in the tool).  The caller aligns each (trimmed read, candidate) with hsq_align_batch and applies the
 * find_counts thresholds (:930-951).  A trimmed read without N that holds anything but A/C/G/T fails the call with
 * HSQ_ERR_BAD_CHARACTER (the reference throws IllegalArgumentException on it). */
int32_t hsq_find_candidates(hsq_ctx* ctx, const hsq_seqs* reads, const uint8_t* quals, int32_t leading_quality,
                            int32_t trailing_quality, int32_t candidate_limit, double min_hit_ratio, hsq_find_out* out);

/* hsq_find_candidates followed, on the device, by one NeedlemanWunschGlobalAlignment(trimmed read, candidate) under the
 * default SimpleAlignmentScorer for every listed candidate and the five-limit test (FastqFindAndTrimTool.java:926-951):
 * the whole per-entry decision of findAndTrim in one call.  A read with more candidates than slots makes the call return
 * HSQ_ERR_OVERFLOW (its overflow flag is set; the listed candidates are still evaluated). */
int32_t hsq_find_matches(hsq_ctx* ctx, const hsq_seqs* reads, const uint8_t* quals, int32_t leading_quality, int32_t trailing_quality,
                         int32_t candidate_limit, double min_hit_ratio, const hsq_find_limits* limits, hsq_find_match_out* out);

/* mergeSequences per item (…Identification.java:1362-1552, behind the unregistered --mergePairs option): lead + overlap
 * (per position the base with the higher quality wins, ties to upstream; a gap is filled with N / '!') + trail.
 * out_offsets gets n+1 entries; out_seq/out_qual have out_cap bytes: when the merged total is larger the call returns
 * HSQ_ERR_OVERFLOW with out_offsets filled, so the caller can size the buffers and call again. */
int32_t hsq_merge_pairs(hsq_ctx* ctx, const hsq_merge_batch* pairs, uint8_t* out_seq, uint8_t* out_qual, int64_t* out_offsets,
                        int64_t out_cap);

/* Builds the per (contig, strand) range maps on the device: probes ordered by window start with a stable sort, as
 * RangeMap.getObjectsThatContainRangeInclusive does on first use (RangeMap.java:98-101). */
int32_t hsq_set_probe_ranges(hsq_ctx* ctx, const hsq_probe_ranges* ranges);
/* ReadToProbeAssigner.assignProbesToReads for a batch of records (…Identification.java:439-545) with
 * RangeMap.getObjectsThatContainRangeInclusive (RangeMap.java:92-153: binary search on the window start, then the leftward
 * walk that stops at the first window ending before the query -- the "no probe contains another" assumption is kept, as
 * are the search's landing rules among equal starts).  Records are paired by read name in input order, i.e. the
 * reference's outcome when one thread feeds the assigner (:470-540): the first two usable records of a name form the pair
 * (probe set = intersection of the mates' sets, :478-483), a third stays unpaired, and a name seen on four or more records
 * completes a second pair and is removed from the results altogether (:484-486, :392-400). */
int32_t hsq_assign_probes(hsq_ctx* ctx, const hsq_mapped_records* records, hsq_assign_out* out);

/* ---- f2: FASTQ ingest and the FASTQ <-> SAM name join in front of the dedup path ------------------------------------- */
/* Record index of FASTQ text as htsjdk's FastqReader reads it (htsjdk 2.9.0, git_superpom/pom.xml:72-75; the reference's call
 * sites: HSQ/utils/FastqSorter.java:77,98): four lines per record ('\n' or '\r\n'), '@' header, '+' quality header, equal
 * sequence / quality length, no blank line -- anything else fails the call with HSQ_ERR_MALFORMED.  Offsets are byte offsets
 * into `text`; header_off/len exclude the '@'.  Record k is the k-th record of the file: the index FastqSorter appends to
 * the header (FastqSorter.java:298-302) and the read index the dedup path breaks quality ties with.  More than `cap`
 * records: HSQ_ERR_OVERFLOW with `n` set. */
typedef struct hsq_fastq_index {
    int64_t cap; /* in: entries available in each array */
    int64_t n;   /* out: records in the text */
    int64_t* header_off;
    int32_t* header_len;
    int64_t* seq_off;
    int32_t* seq_len;
    int64_t* qual_off;
    int32_t* qual_len;
} hsq_fastq_index;
int32_t hsq_index_fastq(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, hsq_fastq_index* out);

/* FastqReadTrimmer.trimReads (HSQ/process/FastqReadTrimmer.java:263-307), the file loop of `HSQutils trim`, text to text: every
 * record of the FASTQ text keeps its header and quality header, bases and qualities are cut as hsq_trim_batch cuts them (:309-328),
 * and records are written the way htsjdk's BasicFastqWriter writes them ("@" header "\n" bases "\n+" quality header "\n"
 * qualities "\n").  first < 0 or last <= first: HSQ_ERR_INVALID_ARGUMENT (:264-270); a read not longer than `first`:
 * HSQ_ERR_OUT_OF_RANGE (:310-312); more than out_cap bytes: HSQ_ERR_OVERFLOW with *out_bytes set. */
int32_t hsq_trim_fastq(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, int32_t first, int32_t last, int32_t three_prime,
                       uint8_t* out_text, int64_t out_cap, int64_t* out_bytes, int64_t* n_records);

/* Field index of SAM text for the columns the dedup path reads off a SAMRecord (QNAME, FLAG, SEQ and QUAL lengths:
 * FastqAndBamFileMerger.java:222-330; RNAME, POS, strand, alignment end = POS + reference length of the CIGAR - 1 and the XA
 * tag: …Identification.java:439-576).  Grammar as htsjdk's SAMLineParser: '@' header lines first (counted, skipped), then
 * lines of at least 11 tab separated columns; FLAG, POS, MAPQ, PNEXT, TLEN are integers; CIGAR is '*' or <count><op> runs
 * (HSQ_ERR_MALFORMED otherwise).  Planes: off[f * cap + k], val[f * cap + k] = offset / length of string field f of
 * alignment line k (XA: the value of the first XA:Z: tag, length 0 when absent); val[v * cap + k] for the numeric columns. */
enum { HSQ_SAM_QNAME = 0, HSQ_SAM_RNAME = 1, HSQ_SAM_CIGAR = 2, HSQ_SAM_SEQ = 3, HSQ_SAM_QUAL = 4, HSQ_SAM_TAGS = 5, HSQ_SAM_XA = 6,
       HSQ_SAM_N_FIELDS = 7 };
enum { HSQ_SAM_FLAG = 7, HSQ_SAM_POS = 8, HSQ_SAM_MAPQ = 9, HSQ_SAM_PNEXT = 10, HSQ_SAM_TLEN = 11, HSQ_SAM_REF_LEN = 12, HSQ_SAM_N_VALUES = 13 };
typedef struct hsq_sam_index {
    int64_t cap;            /* in: alignment lines the planes can hold */
    int64_t n;              /* out: alignment lines */
    int64_t n_header_lines; /* out */
    int64_t* off;           /* [HSQ_SAM_N_FIELDS][cap] */
    int32_t* val;           /* [HSQ_SAM_N_VALUES][cap] */
} hsq_sam_index;
int32_t hsq_index_sam(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, hsq_sam_index* out);

/* Record index of a BAM file whose BGZF blocks were written at compression level 0 -- the way the reference writes its intermediate
 * BAMs (HSQ/process/FastqAndBamFileMerger.java:131-133; records added at :386-396, read back through htsjdk at
 * ...Identification.java:439-474).  At level 0 every BGZF block holds `stored` deflate blocks, so the record bytes are in the file as
 * they are: the library gathers them into one uncompressed stream on the device (kept there; copied to `payload` if that is not
 * NULL) and indexes its alignment records.  A deflate-compressed block: HSQ_ERR_UNSUPPORTED (no inflate here; the empty end-of-file
 * block is accepted).  Offsets are byte offsets into the uncompressed stream.  Planes: off[f * cap + k] / val[f * cap + k] =
 * offset / size of part f of record k -- RECORD: the whole record incl. its block_size word (bytes); QNAME (bytes, without the NUL);
 * CIGAR (operations, 4 bytes each); SEQ (bases, 4 bit each); QUAL (bytes); TAGS (bytes) -- and val[v * cap + k] for the fixed
 * columns, POS and PNEXT 1-based as in SAM text, REF_LEN = reference bases the CIGAR consumes (alignment end = POS + REF_LEN - 1).
 * More records than cap: HSQ_ERR_OVERFLOW with n set.  exact_walk: 1 when bytes inside some record looked like a record header, so
 * that the records had to be picked from the candidates by following the block_size chain (path marking by binary lifting on the
 * device; same result as a sequential reader's, a little slower than the direct test). */
enum { HSQ_BAM_RECORD = 0, HSQ_BAM_QNAME = 1, HSQ_BAM_CIGAR = 2, HSQ_BAM_SEQ = 3, HSQ_BAM_QUAL = 4, HSQ_BAM_TAGS = 5, HSQ_BAM_N_FIELDS = 6 };
enum { HSQ_BAM_REFID = 6, HSQ_BAM_POS = 7, HSQ_BAM_FLAG = 8, HSQ_BAM_MAPQ = 9, HSQ_BAM_NEXT_REFID = 10, HSQ_BAM_PNEXT = 11, HSQ_BAM_TLEN = 12,
       HSQ_BAM_REF_LEN = 13, HSQ_BAM_N_VALUES = 14 };
typedef struct hsq_bam_index {
    int64_t cap;            /* in: records the planes can hold */
    int64_t n;              /* out: alignment records */
    int64_t n_ref;          /* out: reference sequences of the header */
    int64_t text_bytes;     /* out: l_text; the header text sits at offset 8 of the stream */
    int64_t first_record;   /* out: offset of the first alignment record (= end of the reference table) */
    int64_t payload_bytes;  /* out: size of the uncompressed stream */
    int64_t payload_cap;    /* in: bytes available in payload */
    uint8_t* payload;       /* may be NULL */
    int64_t* off;           /* [HSQ_BAM_N_FIELDS][cap] */
    int32_t* val;           /* [HSQ_BAM_N_VALUES][cap] */
    int32_t exact_walk;     /* out */
} hsq_bam_index;
int32_t hsq_index_bam(hsq_ctx* ctx, const uint8_t* file, int64_t n_bytes, hsq_bam_index* out);

/* SAM text -> the columns hsq_assign_probes takes (hsq_mapped_records), as ReadToProbeAssigner.assignProbesToReads reads
 * them off the merged BAM's records (…Identification.java:439-474): records with the unmapped or mate-unmapped flag are dropped
 * (:446; a record that is not paired makes htsjdk throw there: HSQ_ERR_UNSUPPORTED), contig = index of RNAME in contig_names
 * (-1: not listed), end = POS + reference length of the CIGAR - 1, read_index = Integer.parseInt(QNAME) for records on a listed
 * contig (not a number: HSQ_ERR_BAD_CHARACTER, the IllegalStateException of :471-474).  Alternative hits are the XA entries as
 * AlternativeHit.fromAttributeText keeps them (HSQ/utils/SAMRecordUtil.java:440-468): four comma separated fields, the second
 * an integer (else the entry is skipped), stop = start + reference length of the entry's CIGAR - 1.  More records / hits than
 * cap / alt_cap: HSQ_ERR_OVERFLOW with n / n_alt set. */
typedef struct hsq_sam_records_out {
    int64_t cap, alt_cap; /* in */
    int64_t n, n_alt;     /* out */
    int32_t* sam_line;    /* [cap] alignment line (0-based, header lines not counted) of kept record k */
    int32_t* contig;
    uint8_t* negative_strand;
    uint8_t* second_of_pair;
    int32_t* start;
    int32_t* end;
    int32_t* read_index;
    int64_t* alt_offsets; /* [cap + 1] */
    int32_t* alt_contig;  /* [alt_cap] */
    uint8_t* alt_negative;
    int32_t* alt_start;
    int32_t* alt_stop;
} hsq_sam_records_out;
int32_t hsq_sam_mapped_records(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, const hsq_seqs* contig_names, hsq_sam_records_out* out);

/* FastqAndBamFileMerger.createBamFileWithDataFromRawFastqFiles (HSQ/process/FastqAndBamFileMerger.java:107-158) without the
 * file codecs: both sorts and the merge loop of MergedSamIterator (:213-345) for the two configurations the dedup command
 * uses.  by_index = 0 (…Identification.java:150-160): FASTQ records sorted by header + " _ " + ordinal
 * (FastqSorter.java:298-302,334-339), SAM records by (read name, first of pair first) (BamSorter.java:348-361), joined on
 * IlluminaFastQReadNameUtil.getUniqueIdForReadHeader (COMMON/utils/IlluminaFastQReadNameUtil.java:40-58).  by_index = 1
 * (…Identification.java:871-881): FASTQ in file order, SAM names are integers (BamSorter.java:322-346), joined on the
 * ordinal.  Records that compare equal keep input order.  A result that would depend on the merge loop meeting names out
 * of order (unique ids not ascending along a sorted side) fails with HSQ_ERR_UNSUPPORTED. */
typedef struct hsq_name_join {
    hsq_seqs sam_names;          /* QNAME per SAM record, input order */
    const int32_t* sam_flags;    /* FLAG: 0x40 first of pair, 0x100 not primary (dropped, :222-225) */
    const int32_t* sam_seq_len;  /* record.getReadString().length() */
    const int32_t* sam_qual_len; /* record.getBaseQualityString().length(); NULL = same as sam_seq_len */
    hsq_seqs fq1_headers, fq2_headers; /* read headers without '@', file order */
    const int32_t* fq1_seq_len;  /* read lengths (trim check); may be NULL when the check is off */
    const int32_t* fq2_seq_len;
    int32_t by_index;
    int32_t trimming_skipped, check_trim;                     /* :366-368 */
    int32_t three_prime, r1_first, r1_stop, r2_first, r2_stop; /* ProbeTrimmingInformation (FastqReadTrimmer.java:153-190) */
} hsq_name_join;
typedef struct hsq_name_join_out {
    int64_t n_emitted;            /* out: records the merged iterator returns, in its order */
    int64_t total_matching_pairs; /* out: getTotalMatches() (:343) */
    int32_t* sam_record;   /* [sam n] e-th returned record = input record sam_record[e] */
    int32_t* fastq_record; /* [sam n] ordinal, in FASTQ one (first of pair) or FASTQ two (otherwise), of the record whose
                            *         sequence, qualities and appended index the returned record takes (:303-325) */
    int32_t* name_len;     /* [sam n] length of the unique id of that position's FASTQ-one header (FASTQ_READ_NAME, :317) */
    uint8_t* status;       /* [sam n] by INPUT record: 0 not primary, 1 returned, 2 dropped */
} hsq_name_join_out;
int32_t hsq_join_names(hsq_ctx* ctx, const hsq_name_join* in, hsq_name_join_out* out);

/* out row e = text[off[e] .. off[e]+len[e]) for n rows (the raw read / quality strings storeFastqInfoInRecord puts into the
 * merged records, FastqAndBamFileMerger.java:386-396): out_offsets gets n+1 entries; more than out_cap bytes:
 * HSQ_ERR_OVERFLOW with out_offsets filled. */
int32_t hsq_gather_rows(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, const int64_t* off, const int32_t* len, int64_t n,
                        uint8_t* out_data, int64_t* out_offsets, int64_t out_cap);

/* Order of the final coordinate sort of the output records (BamSorter.sortBamFile(..., new SAMRecordCoordinateComparator()),
 * …Identification.java:897; htsjdk 2.9.0's comparator, source absent from /root/reference: reference index with -1 last, alignment
 * start, forward strand before reverse, then read name (String.compareTo), FLAG, MAPQ, mate reference index, mate alignment
 * start, inferred insert size as signed integers).  order[k] = input record at sorted position k; records that compare equal
 * keep input order (the reference's chunk merge refuses such records when they meet, BamSorter.java:153-165). */
typedef struct hsq_coord_sort_in {
    int64_t n;
    const int32_t* ref_index;
    const int32_t* pos;
    const int32_t* flags;
    const int32_t* mapq;
    const int32_t* mate_ref_index;
    const int32_t* mate_pos;
    const int32_t* tlen;
    hsq_seqs names;
} hsq_coord_sort_in;
int32_t hsq_coordinate_order(hsq_ctx* ctx, const hsq_coord_sort_in* in, int32_t* order);

/* ---- N1 entry points ---------------------------------------------------------------------------------------------------- */
/* Host helper (plain CPU code, no device): packs a ragged ASCII batch into the layout of hsq_packed_reads.  data_out must hold
 * hsq_packed_bytes(...) bytes, len_out n entries.  2 bit: at most exc_cap exceptions are written, *n_exceptions gets their
 * number (more than exc_cap: HSQ_ERR_OVERFLOW).  A byte that is not an IUPAC letter or gap character: HSQ_ERR_BAD_CHARACTER.
 * threads <= 0: all host cores.  hsq_last_error(NULL) holds the message of a failure. */
int64_t hsq_packed_bytes(const int64_t* offsets, int64_t n, int32_t bits);
int32_t hsq_pack_reads(const hsq_seqs* reads, int32_t bits, uint8_t* data_out, int32_t* len_out, int64_t exc_cap,
                       int64_t* n_exceptions, uint32_t* exc_read, uint32_t* exc_pos, uint8_t* exc_char, int32_t threads);
/* Host helper: out[k] = (short) sum over read k of (quality char - 33), Java short wrap-around included. */
int32_t hsq_quality_sums(const uint8_t* quals, const int64_t* offsets, int64_t n, int16_t* out, int32_t threads);
/* Inverse of hsq_pack_reads on the device (test / diagnostic entry point): upper-case ASCII of every read, ragged, with the
 * offsets it implies; out_data must hold sum(len) bytes, out_offsets n + 1 entries. */
int32_t hsq_unpack_reads(hsq_ctx* ctx, const hsq_packed_reads* reads, uint8_t* out_data, int64_t* out_offsets);
/* hsq_extend_batch for packed reads with pooled CIGAR / MD output: same results, about a third of the bytes over PCIe.
 * Chunks flow through the same H2D | prep | capture | D2H pipeline. */
int32_t hsq_extend_packed(hsq_ctx* ctx, const hsq_packed_pairs* pairs, hsq_extend_pooled_out* out);
/* hsq_dedup_batch for packed reads: the UID comes from the packed bases (an exception inside a UID: HSQ_ERR_UNSUPPORTED, as a
 * non-ACGT UID base is for hsq_dedup_batch), the quality sums from q1 / q2. */
int32_t hsq_dedup_packed(hsq_ctx* ctx, const hsq_packed_pairs* pairs, int32_t ext_uid_len, int32_t lig_uid_len, uint8_t* is_best,
                         int32_t* group_size, hsq_probe_stats* stats);

/* Host helper (no device): the UID code of every pair from its packed reads, as the key generation computes it -- R1[0:u] ++ R2[0:v],
 * 2 bit per base in the order A, C, G, T, first base most significant (u + v <= 32).  A base other than A/C/G/T inside a UID (an
 * exception of a 2-bit read, an IUPAC nibble of a 4-bit one): HSQ_ERR_UNSUPPORTED, as for hsq_dedup_batch; a read shorter than its
 * UID: HSQ_ERR_OUT_OF_RANGE. */
int32_t hsq_uid_keys(const hsq_packed_reads* r1, const hsq_packed_reads* r2, int32_t ext_uid_len, int32_t lig_uid_len, uint64_t* out);

/* hsq_dedup_extend_packed with the reduction fed by keys: uid_keys (hsq_uid_keys), pairs->qsum1 / qsum2 (hsq_quality_sums), probe_id and
 * read_index -- 20 bytes per pair -- go to the device first and are reduced; the reads then follow in chunks, each copied in while the
 * previous one is aligned, cut down to its unique pairs on the device and extended.  Same results, slot for slot; what changes is that
 * the copy of the reads (80 of the 100 bytes per pair) no longer waits in front of the kernels.  All four key inputs are required. */
int32_t hsq_dedup_extend_packed_keys(hsq_ctx* ctx, const hsq_packed_pairs* pairs, const uint64_t* uid_keys, int32_t ext_uid_len,
                                     int32_t lig_uid_len, int32_t extend_duplicates, int32_t sharded, hsq_dedup_extend_out* out);

/* hsq_dedup_extend_batch for a packed batch with pooled results (out->extend_pooled; out->extend is not used).  With host-made quality
 * sums (pairs->qsum1 / qsum2) about 100 bytes per pair cross PCIe instead of 624.  sharded != 0: the reduction is the collective of
 * hsq_dedup_extend_sharded over the context's communicator. */
int32_t hsq_dedup_extend_packed(hsq_ctx* ctx, const hsq_packed_pairs* pairs, int32_t ext_uid_len, int32_t lig_uid_len,
                                int32_t extend_duplicates, int32_t sharded, hsq_dedup_extend_out* out);

/* ---- sharded dedup: one process per GPU, exchange by the caller ------------------------------------------------------ */
/* First half of FilterByUid.reduceReadsByProbeAndUid for this rank's pairs (host buffers in): writes pairs->n key
 * records to DEVICE memory d_keys_out (origin = origin_base + pair index).  The caller routes the records to the rank
 * that owns each probe (NCCL all-to-all) ... */
int32_t hsq_dedup_keys(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t ext_uid_len, int32_t lig_uid_len, int64_t n_probes,
                       uint32_t origin_base, hsq_dedup_key* d_keys_out);
/* ... and the owner groups the n records it received (DEVICE memory) and picks the best pair of every (probe, UID)
 * group (findBestData, FilterByUid.java:281-297): d_is_best[k], d_group_size[k] (DEVICE memory) answer record k; the
 * per-probe counters (host arrays) cover the probes this rank owns. */
int32_t hsq_dedup_reduce_keys(hsq_ctx* ctx, const hsq_dedup_key* d_keys, int64_t n, int32_t uid_len, uint8_t* d_is_best,
                              int32_t* d_group_size, hsq_probe_stats* stats);

/* ---- multi-GPU behind the C ABI ------------------------------------------------------------------- */
/* The reference fans its probe tasks out from one JVM (…Identification.java:836-846, task body :1172-1234); here the contexts of a
 * job -- one per GPU, in one process or in one process per GPU -- share an NCCL communicator and the library does the exchange
 * itself: key records bucketed by owner (probe mod world) on the device, counts by ncclAllGather, payload by ncclGroupStart +
 * ncclSend / ncclRecv on device buffers, the owner's reduction, the answers back along the same routes.  Reads never move.  NCCL is
 * loaded on first use (dlopen of libnccl.so.2); without it these calls return HSQ_ERR_UNSUPPORTED. */
#define HSQ_COMM_ID_BYTES 128
/* one process per GPU: rank 0 makes an id, the host program hands it to the other ranks, every rank joins */
int32_t hsq_comm_unique_id(uint8_t* id_out /* [HSQ_COMM_ID_BYTES] */);
int32_t hsq_comm_init_rank(hsq_ctx* ctx, const uint8_t* id /* [HSQ_COMM_ID_BYTES] */, int32_t rank, int32_t world);
/* one process, n contexts on n different GPUs: ctxs[k] becomes rank k */
int32_t hsq_comm_init_all(hsq_ctx** ctxs, int32_t n);
int32_t hsq_comm_destroy(hsq_ctx* ctx);
/* hsq_dedup_batch over the union of all ranks' pairs.  Collective: every rank calls it with its own pairs (possibly none).
 * is_best / group_size answer this rank's pairs; stats hold the counters of the whole batch on every rank.  Exact ties of
 * (quality sum, read index) go to the pair that comes first in the rank-major concatenation of the ranks' batches, so the answer
 * does not depend on how the batch was split. */
int32_t hsq_dedup_sharded(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t ext_uid_len, int32_t lig_uid_len, uint8_t* is_best,
                          int32_t* group_size, hsq_probe_stats* stats);
/* hsq_dedup_extend_batch with the reduction over all ranks' pairs (collective); the extension runs on this rank's own pairs, the
 * per-probe counters (stats, unable-to-extend) are those of the whole batch. */
int32_t hsq_dedup_extend_sharded(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t ext_uid_len, int32_t lig_uid_len,
                                 int32_t extend_duplicates, hsq_dedup_extend_out* out);
/* One call for one process that owns n GPUs: pairs[k] / outs[k] belong to ctxs[k]; a host thread per context runs
 * hsq_dedup_extend_sharded; the communicator is created on first use (hsq_comm_init_all).  On failure the message is in
 * hsq_last_error of the context that failed. */
int32_t hsq_dedup_extend_multi_gpu(hsq_ctx** ctxs, int32_t n, const hsq_pair_batch* pairs, int32_t ext_uid_len, int32_t lig_uid_len,
                                   int32_t extend_duplicates, hsq_dedup_extend_out* outs);

/* ---- an entry point that allocates nothing -------------------------------------------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The DP of hsq_align_batch -- one
 * NeedlemanWunschGlobalAlignment(ref[k], qry[k], SimpleAlignmentScorer(scorer, flags)) per item -- with scalar results only:
 * getAlignmentScore, alignment length, getNumberOfMatches, getNumberOfMismatches.  Unlike every other entry point it calls no
 * cudaMalloc / cudaFree / allocating routine: its scratch (encoded inputs, pointer words, boundary rows) is one 64 MiB file-scope
 * __device__ array that every GPU loading the library holds until the process exits.
 *   - ALL data pointers are DEVICE pointers owned by the caller: ASCII bases (ragged, offsets[0] = 0, offsets[n] = *_bytes), four
 *     int32 outputs of n entries, and one int32 status word (0 = ok, else the HSQ_ERR_* code of the first problem found on the device:
 *     HSQ_ERR_BAD_CHARACTER, or HSQ_ERR_INVALID_ARGUMENT for offsets that disagree with *_bytes / a sequence longer than declared).
 *   - limits (beyond them the call returns HSQ_ERR_UNSUPPORTED and launches nothing): ref_bytes, qry_bytes <=
 *     HSQ_NOALLOC_MAX_BYTES_PER_SIDE; (max_ref_len + 1) * (ceil(max_qry_len / 16) + 1) <= HSQ_NOALLOC_MAX_WORDS (160 x 160 bases fit).
 *   - the scratch array is shared by all calls on a GPU and keeps the last call's contents (nothing in it needs clearing; the status
 *     word is cleared by the call).  Calls that overlap on the same GPU -- two contexts, two threads -- race on it: serialise them.
 *   - the first call on a device looks the array's address up (cudaGetSymbolAddress, which may load the module): make one warm-up call
 *     before any region that forbids it.  The call returns after synchronising the context's stream. */
#define HSQ_NOALLOC_MAX_BYTES_PER_SIDE (4 << 20)
#define HSQ_NOALLOC_MAX_WORDS 1792
int32_t hsq_align_scores_noalloc(hsq_ctx* ctx, const uint8_t* d_ref_data, const int64_t* d_ref_offsets, int64_t ref_bytes,
                                 const uint8_t* d_qry_data, const int64_t* d_qry_offsets, int64_t qry_bytes, int64_t n, int32_t max_ref_len,
                                 int32_t max_qry_len, const hsq_scorer* scorer, uint32_t flags, int32_t* d_score, int32_t* d_n_cols,
                                 int32_t* d_matches, int32_t* d_mismatches, int32_t* d_status);

/* ---- measurement helpers ------------------------------------------------------------------- */
/* CUDA-event stopwatch on the context's stream (the stream every kernel of this library is launched on). */
int32_t hsq_timer_start(hsq_ctx* ctx);
int32_t hsq_timer_stop(hsq_ctx* ctx, float* ms);
/* Sustained INT32 issue rate (lane-ops/s) of the device from an IMAD/VIMNMX/LOP3 micro-kernel run for `seconds`:
 * the roofline denominator of the DP kernels. */
int32_t hsq_int32_peak(hsq_ctx* ctx, double seconds, double* int_ops_per_s);

#ifdef __cplusplus
}
#endif
#endif /* HSQ_B200_H */
