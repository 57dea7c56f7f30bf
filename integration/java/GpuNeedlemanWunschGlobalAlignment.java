/*
 * GpuNeedlemanWunschGlobalAlignment -- the operator seam of bioinformatics_common on the GPU.
 *
 * Source compatible with com.roche.sequencing.bioinformatics.common.alignment.NeedlemanWunschGlobalAlignment
 * (COMMON/alignment/NeedlemanWunschGlobalAlignment.java:44,57 constructors; getters :85,119,131,236,243,250,257,381,388,392,412,430,438,442):
 * same constructors, same getters, same results, but the matrix is filled and traced back by hsq_align_batch and every getter reads
 * one row of its hsq_align_out.  It lives in the reference's own package because AlignmentPair's constructor and
 * CigarString.FromNonSummarizedCigarString are package private there (AlignmentPair.java:39, CigarString.java:39).
 *
 * Two ways to use it:
 *   - one at a time, like the reference: new GpuNeedlemanWunschGlobalAlignment(ref, query[, scorer]); the first getter runs a batch of one
 *     (a few hundred microseconds of launch and copy: correct, but only the batch form is fast);
 *   - batched: GpuNeedlemanWunschGlobalAlignment.alignAll(refs, queries, scorer) -- one downcall for the whole list; this is what the
 *     reference's call sites are given through HsqBatcher (ExtendReadsToPrimer.java:211,293; ...Identification.java:722,731,1306;
 *     SAMRecordUtil.java:225; FastqFindAndTrimTool.java:930).
 * Only SimpleAlignmentScorer instances with integer scores are accelerated; anything else throws UnsupportedOperationException
 * (no JVM-side compute fallback, by design).
 *
 * NOT compiled in this repository's image (no JVM there).
 */
package com.roche.sequencing.bioinformatics.common.alignment;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.nio.charset.StandardCharsets;
import java.util.ArrayList;
import java.util.List;

import com.roche.heatseq.b200.HsqB200;
import com.roche.heatseq.b200.HsqBatcher;
import com.roche.sequencing.bioinformatics.common.sequence.ICode;
import com.roche.sequencing.bioinformatics.common.sequence.ISequence;
import com.roche.sequencing.bioinformatics.common.sequence.IupacNucleotideCode;
import com.roche.sequencing.bioinformatics.common.sequence.IupacNucleotideCodeSequence;
import com.roche.sequencing.bioinformatics.common.utils.StringUtil;

import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_INT;

public class GpuNeedlemanWunschGlobalAlignment {

    private final ISequence referenceSequence;
    private final ISequence querySequence;
    private final IAlignmentScorer alignmentScorer;

    // one row of hsq_align_out (filled by alignAll, or lazily by a batch of one)
    private boolean computed;
    private int score, matches, mismatches;
    private AlignmentPair alignment;

    public GpuNeedlemanWunschGlobalAlignment(ISequence referenceSequence, ISequence querySequence, IAlignmentScorer alignmentScorer) {
        this.referenceSequence = referenceSequence;
        this.querySequence = querySequence;
        this.alignmentScorer = alignmentScorer;
        HsqBatcher.requireSimple(alignmentScorer);  // fail at construction, where the caller can still see which scorer it passed
    }

    public GpuNeedlemanWunschGlobalAlignment(ISequence referenceSequence, ISequence querySequence) {
        this(referenceSequence, querySequence, new SimpleAlignmentScorer());
    }

    /** All alignments of a list in one downcall. */
    public static List<GpuNeedlemanWunschGlobalAlignment> alignAll(List<? extends ISequence> references, List<? extends ISequence> queries,
                                                                   IAlignmentScorer scorer) {
        if (references.size() != queries.size()) throw new IllegalArgumentException("references and queries differ in size");
        List<GpuNeedlemanWunschGlobalAlignment> out = new ArrayList<>(references.size());
        for (int k = 0; k < references.size(); k++) out.add(new GpuNeedlemanWunschGlobalAlignment(references.get(k), queries.get(k), scorer));
        fill(out, scorer);
        return out;
    }

    private static void fill(List<GpuNeedlemanWunschGlobalAlignment> items, IAlignmentScorer scorer) {
        int n = items.size();
        if (n == 0) return;
        int[] sc = HsqBatcher.requireSimple(scorer);  // match, mismatch, gapExtend, gapStart, flags
        int maxCols = 1;
        for (GpuNeedlemanWunschGlobalAlignment a : items) maxCols = Math.max(maxCols, a.referenceSequence.size() + a.querySequence.size());
        HsqBatcher batcher = HsqBatcher.get();
        synchronized (batcher.alignLock()) {  // a context is used by one thread at a time
            HsqB200 gpu = batcher.alignContext();
            try (Arena arena = Arena.ofConfined()) {
                List<String> r = new ArrayList<>(n), q = new ArrayList<>(n);
                for (GpuNeedlemanWunschGlobalAlignment a : items) {
                    r.add(a.referenceSequence.toString());
                    q.add(a.querySequence.toString());
                }
                MemorySegment refs = HsqBatcher.seqs(arena, r);
                MemorySegment qrys = HsqBatcher.seqs(arena, q);
                MemorySegment scorerSeg = HsqBatcher.scorer(arena, sc);
                // the getters below need score, the two gapped strings, matches and mismatches; the other columns of hsq_align_out may be
                // NULL (the library does not copy what is not asked for), MD / CIGAR strings are derived from the AlignmentPair exactly
                // as the reference derives them, so their code path stays the reference's own
                HsqBatcher.AlignOut o = new HsqBatcher.AlignOut(arena, n, maxCols, 2 * maxCols + 8);  // MD slots are sized even though they are not fetched: an MD that does not fit its device slot fails the call
                gpu.alignBatch(refs, qrys, scorerSeg, sc[4], o.struct);
                for (int k = 0; k < n; k++) {
                    GpuNeedlemanWunschGlobalAlignment a = items.get(k);
                    int cols = o.nCols.getAtIndex(JAVA_INT, k);
                    a.score = o.score.getAtIndex(JAVA_INT, k);
                    a.matches = o.matches.getAtIndex(JAVA_INT, k);
                    a.mismatches = o.mismatches.getAtIndex(JAVA_INT, k);
                    a.alignment = new AlignmentPair(slot(o.refAln, k, maxCols, cols), slot(o.qryAln, k, maxCols, cols));
                    a.computed = true;
                }
            }
        }
    }

    private static ISequence slot(MemorySegment slots, int k, int cap, int len) {
        byte[] b = slots.asSlice((long) k * cap, len).toArray(JAVA_BYTE);
        return new IupacNucleotideCodeSequence(new String(b, StandardCharsets.ISO_8859_1));  // '_' parses back to GAP
    }

    private void ensure() {
        if (!computed) {
            List<GpuNeedlemanWunschGlobalAlignment> one = new ArrayList<>(1);
            one.add(this);
            fill(one, alignmentScorer);
        }
    }

    // ---- the reference's getters -------------------------------------------------------------------------------------------------
    public AlignmentPair getAlignmentPair() {  // :85
        ensure();
        return alignment;
    }

    public double getLengthNormalizedAlignmentScore() {  // :119 (0 / 0 = NaN when both sequences are empty, as there)
        ISequence referenceAlignment = getAlignmentPair().getReferenceAlignment();
        return ((double) getAlignmentScore()) / ((double) referenceAlignment.size());
    }

    public double getAlignmentScore() {  // :131
        ensure();
        return score;
    }

    public IAlignmentScorer getAlignmentScorer() { return alignmentScorer; }

    public ISequence getReferenceSequence() { return referenceSequence; }

    public ISequence getQuerySequence() { return querySequence; }

    public CigarString getCigarString() { return getAlignmentPair().getCigarString(); }  // :236

    public CigarString getReverseCigarString() { return getAlignmentPair().getReverseCigarString(); }  // :243

    public String getMismatchDetailsString() { return getAlignmentPair().getMismatchDetailsString(); }  // :250

    public String getReverseMismatchDetailsString() { return getAlignmentPair().getReverseMismatchDetailsString(); }  // :257

    public int getEditDistance() { return getCigarString().getEditDistance(); }  // :381

    public String getMismatchDetails() { return CigarStringUtil.getMismatchDetailsString(getAlignmentPair()); }  // :388

    public int getNumberOfNucleotidesOutsideOfReference() {  // :392
        ISequence referenceAlignment = getAlignmentPair().getReferenceAlignment();
        int i = 0;
        while (i < referenceAlignment.size() && referenceAlignment.getCodeAt(i).equals(IupacNucleotideCode.GAP)) i++;
        int outside = i;
        i = referenceAlignment.size() - 1;
        while (i >= 0 && referenceAlignment.getCodeAt(i).equals(IupacNucleotideCode.GAP)) i--;
        return outside + (referenceAlignment.size() - i - 1);
    }

    public int getIndexOfFirstMatchInReference() {  // :412 (first non-gap column of the QUERY alignment; hsq_align_out.first_q, -1 if none)
        ISequence queryAlignment = getAlignmentPair().getQueryAlignment();
        for (int index = 0; index < queryAlignment.size(); index++) {
            ICode code = queryAlignment.getCodeAt(index);
            if (!code.matches(IupacNucleotideCode.GAP)) return index;
        }
        return -1;
    }

    public double getRatioOfQueryMatchesToQueryLength() {  // :430
        String expanded = getCigarString().getCigarString(false, true);
        double m = StringUtil.countMatches(expanded, "" + CigarStringUtil.CIGAR_SEQUENCE_MATCH);
        return m / (double) querySequence.size();
    }

    public int getNumberOfMismatches() {  // :438 (the reference reads the field directly: NPE before getAlignmentPair(); here it computes)
        ensure();
        return mismatches;
    }

    public int getNumberOfMatches() {  // :442
        ensure();
        return matches;
    }
}
