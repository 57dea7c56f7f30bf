/*
 * FilterByUidGpu -- the body a maintainer gives FilterByUid.reduceReadsByProbeAndUid (HSQ/process/FilterByUid.java:71-255): signature
 * unchanged, grouping and findBestData (:134-203, :281-297) done by hsq_dedup_batch through HsqBatcher.  Everything around the
 * reduction stays the reference's own code: building the IReadPair list (:77-128), the report lines (:209-243), the UID composition
 * strings (:245-252).  Variable-length UIDs (allowVariableLengthUids, dead in the CLI) take hsq_variable_length_uid.
 * NOT compiled in this repository's image (no JVM there).
 */
package com.roche.heatseq.process;

import java.util.ArrayList;
import java.util.HashMap;
import java.util.List;
import java.util.Map;
import java.util.Set;

import com.roche.heatseq.b200.HsqBatcher;
import com.roche.heatseq.objects.IReadPair;
import com.roche.heatseq.objects.ReadPair;
import com.roche.heatseq.objects.SAMRecordPair;
import com.roche.heatseq.objects.UidReductionResultsForAProbe;
import com.roche.heatseq.qualityreport.ProbeProcessingStats;
import com.roche.heatseq.qualityreport.ReportManager;
import com.roche.heatseq.utils.SAMRecordUtil;
import com.roche.sequencing.bioinformatics.common.alignment.IAlignmentScorer;
import com.roche.sequencing.bioinformatics.common.sequence.ISequence;
import com.roche.sequencing.bioinformatics.common.sequence.IupacNucleotideCodeSequence;
import com.roche.sequencing.bioinformatics.common.utils.probeinfo.Probe;

import htsjdk.samtools.SAMRecord;

final class FilterByUidGpu {
    private FilterByUidGpu() {}

    static UidReductionResultsForAProbe reduceReadsByProbeAndUid(Probe probe, Map<Integer, SAMRecordPair> readIndexToRecordsMap, ReportManager reportManager,
            boolean allowVariableLengthUids, int expectedExtensionUidLength, int expectedLigationUidLength, IAlignmentScorer alignmentScorer,
            Set<ISequence> distinctUids, List<ISequence> uids, boolean markDuplicates, boolean useStrictReadToProbeMatching) {
        if (allowVariableLengthUids)  // per-pair UID lengths: group on the host by the strings hsq_variable_length_uid cuts (kept out of this listing)
            throw new UnsupportedOperationException("variable length UIDs: use HsqB200.variableLengthUid + the reference's grouping");
        HsqBatcher batcher = HsqBatcher.get();
        int row = batcher.rowOf(probe);

        // :77-128 as in the reference: the pairs with both mates, their UIDs (fixed length: parseUidFromReadOne / Two)
        List<IReadPair> datas = new ArrayList<>();
        List<HsqBatcher.Pair> pairs = new ArrayList<>();
        for (SAMRecordPair recordPair : readIndexToRecordsMap.values()) {
            SAMRecord record = recordPair.getFirstOfPairRecord(), mate = recordPair.getSecondOfPairRecord();
            if (record == null || mate == null) continue;
            String extensionUid = SAMRecordUtil.parseUidFromReadOne(record.getReadString(), expectedExtensionUidLength);
            String ligationUid = SAMRecordUtil.parseUidFromReadTwo(mate.getReadString(), expectedLigationUidLength);
            if (extensionUid == null || ligationUid == null) continue;
            ISequence full = new IupacNucleotideCodeSequence(extensionUid);
            full.append(new IupacNucleotideCodeSequence(ligationUid));
            distinctUids.add(full);
            synchronized (uids) {
                uids.add(full);
            }
            IReadPair pair = new ReadPair(record, mate, extensionUid, ligationUid, probe.getCaptureTargetSequence(), probe.getProbeId(), false, false);
            datas.add(pair);
            pairs.add(new HsqBatcher.Pair(row, Integer.parseInt(pair.getReadName()), record.getReadString(), mate.getReadString(),
                    record.getBaseQualityString(), mate.getBaseQualityString()));
        }

        // :134-203 + :281-297 on the GPU: one flag and one group size per pair, the three counters per probe
        HsqBatcher.DedupAnswer answer = batcher.dedup(pairs, expectedExtensionUidLength, expectedLigationUidLength);

        List<IReadPair> uniqueReadPairs = new ArrayList<>(), duplicateReadPairs = new ArrayList<>();
        Map<String, Integer> sizeByUid = new HashMap<>();  // for the report lines (:209-243) and the weighted UID lists (:149-171)
        for (int k = 0; k < datas.size(); k++) {
            IReadPair pair = datas.get(k);
            sizeByUid.put(pair.getExtensionUid() + pair.getLigationUid(), answer.groupSize()[k]);
            if (answer.isBest()[k] != 0) {
                pair.setAsBestPairInUidGroup();
                uniqueReadPairs.add(pair);
            } else {
                if (markDuplicates) pair.markAsDuplicate();
                duplicateReadPairs.add(pair);
            }
        }
        int[] counters = answer.probeCounters().getOrDefault(row, new int[3]);  // unique, duplicate, max pairs per UID
        // the report and composition strings are the reference's own code fed from sizeByUid (FilterByUid.java:209-252)
        ProbeProcessingStats stats = FilterByUid.statsFor(probe, reportManager, sizeByUid, counters[1], counters[0], counters[2]);
        return new UidReductionResultsForAProbe(stats, uniqueReadPairs, duplicateReadPairs);
    }
}
