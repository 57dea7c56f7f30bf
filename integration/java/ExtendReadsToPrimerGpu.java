/*
 * ExtendReadsToPrimerGpu -- the body a maintainer gives ExtendReadsToPrimer.extendReadsToPrimers (HSQ/process/ExtendReadsToPrimer.java:347-380):
 * signature unchanged; getPrimerEndIndexInRead (:286-345), calculateDetailsForReadExtensionToPrimer (:200-245) and the strand logic of
 * extendReadPair (:96-198) run in hsq_extend_batch through HsqBatcher; the SAMRecord construction (createRecord :247-284,
 * SAMRecordUtil.setSAMRecordsAsPair) stays the reference's own code, fed with the per read end results.
 * NOT compiled in this repository's image (no JVM there).
 */
package com.roche.heatseq.process;

import java.util.ArrayList;
import java.util.List;

import com.roche.heatseq.b200.HsqBatcher;
import com.roche.heatseq.objects.ExtendReadResults;
import com.roche.heatseq.objects.IReadPair;
import com.roche.heatseq.objects.ReadPair;
import com.roche.heatseq.utils.SAMRecordUtil;
import com.roche.sequencing.bioinformatics.common.alignment.CigarString;
import com.roche.sequencing.bioinformatics.common.alignment.IAlignmentScorer;
import com.roche.sequencing.bioinformatics.common.sequence.ISequence;
import com.roche.sequencing.bioinformatics.common.sequence.IupacNucleotideCodeSequence;
import com.roche.sequencing.bioinformatics.common.utils.StringUtil;
import com.roche.sequencing.bioinformatics.common.utils.probeinfo.Probe;
import com.roche.sequencing.bioinformatics.common.utils.probeinfo.ProbeHeaderInformation;
import com.roche.sequencing.bioinformatics.common.utils.probeinfo.Strand;

import htsjdk.samtools.SAMRecord;

final class ExtendReadsToPrimerGpu {
    private ExtendReadsToPrimerGpu() {}

    static ExtendReadResults extendReadsToPrimers(ProbeHeaderInformation probeHeaderInformation, Probe probe, List<IReadPair> readPairs,
            IAlignmentScorer alignmentScorer) {
        HsqBatcher batcher = HsqBatcher.get();
        HsqBatcher.requireSimple(alignmentScorer);  // the user scorer was given to the contexts once (HsqBatcher.setUserScorer)
        int row = batcher.rowOf(probe);
        List<HsqBatcher.Pair> pairs = new ArrayList<>(readPairs.size());
        for (IReadPair p : readPairs)
            pairs.add(new HsqBatcher.Pair(row, Integer.parseInt(p.getReadName()), p.getSequenceOne(), p.getSequenceTwo(), p.getSequenceOneQualityString(),
                    p.getSequenceTwoQualityString()));

        HsqBatcher.ExtendAnswer a = batcher.extend(pairs);

        List<IReadPair> extended = new ArrayList<>(), unable = new ArrayList<>();
        boolean oneReverse = probe.getProbeStrand() == Strand.REVERSE;
        for (int k = 0; k < readPairs.size(); k++) {
            IReadPair in = readPairs.get(k);
            if (a.pairOk()[k] == 0) {  // :364: a pair is kept only when both mates extended
                unable.add(in);
                continue;
            }
            SAMRecord one = record(in, probe, a, 2 * k, in.getSequenceOne(), in.getSequenceOneQualityString(), oneReverse, in.getOneMappingQuality(), true);
            SAMRecord two = record(in, probe, a, 2 * k + 1, in.getSequenceTwo(), in.getSequenceTwoQualityString(), !oneReverse, in.getTwoMappingQuality(), false);
            SAMRecordUtil.setSAMRecordsAsPair(one, two);  // :183-185
            extended.add(new ReadPair(one, two, in.getExtensionUid(), in.getLigationUid(), probe.getCaptureTargetSequence(), probe.getProbeId(), true, true));
        }
        return new ExtendReadResults(extended, unable);
    }

    /** createRecord (:247-284) from one read end's results: the kept slice of the raw read, reverse complemented on the - strand (:131-142,
     *  :166-178; hsq_extend_records does the same on the device for hosts that want the strings from the GPU as well). */
    private static SAMRecord record(IReadPair in, Probe probe, HsqBatcher.ExtendAnswer a, int e, String read, String quals, boolean negative, int mapq,
            boolean first) {
        ISequence seq = new IupacNucleotideCodeSequence(read).subSequence(a.readStart()[e], a.readStop()[e]);
        String q = quals.substring(a.readStart()[e], a.readStop()[e] + 1);
        if (negative) {
            seq = seq.getReverseCompliment();
            q = StringUtil.reverse(q);
        }
        return ExtendReadsToPrimer.createRecord(in.getSamHeader(), in.getReadName(), in.getReadGroup(), negative, CigarString.fromStandard(a.cigar()[e]),
                a.md()[e], a.nm()[e], a.pos()[e], seq.toString(), q, probe.getSequenceName(), mapq, a.tlen()[e], in.isMarkedDuplicate(),
                in.getExtensionUid(), in.getLigationUid(), probe.getProbeId(), in.isBestPairDuplicateGroup(),
                SAMRecordUtil.getAlternativeHitsFromAttribute(first ? in.getRecord() : in.getMateRecord()));
    }
}
