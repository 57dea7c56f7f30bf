/*
 * HsqBatcher -- turns the reference's one-probe-at-a-time calls into GPU batches.
 *
 * The reference runs one PrimerReadExtensionAndFilteringOfUniquePcrProbesTask per probe on a pool of <= numProcessors threads
 * (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:836-846, task body :1172-1234).  The replacements of
 * FilterByUid.reduceReadsByProbeAndUid and ExtendReadsToPrimer.extendReadsToPrimers (FilterByUidGpu / ExtendReadsToPrimerGpu in this
 * directory) keep their signatures; inside, each pool thread hands its probe's pairs to this batcher and waits on a ticket.  One flusher
 * thread per process owns the HsqB200 contexts (a context is used by one thread at a time), concatenates the waiting requests into one
 * hsq_pair_batch in pinned memory and makes one downcall per batch: hsq_dedup_batch / hsq_extend_batch on one GPU,
 * hsq_dedup_extend_multi_gpu-style sharding (hsq_comm_init_all + hsq_dedup_sharded) when hsq.b200.devices > 1.
 *
 * A batch is flushed when it holds >= hsq.b200.batchPairs pairs (default 1 M), when every registered pool thread is waiting, or after
 * hsq.b200.flushMillis without a new request.  Exceptions of a batch are rethrown on every waiting caller's thread, so the reference's
 * catch-and-log around a task (:1229-1231) behaves as before.  There is no JVM-side compute fallback.
 *
 * NOT compiled in this repository's image (no JVM there).
 */
package com.roche.heatseq.b200;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.nio.charset.StandardCharsets;
import java.util.ArrayList;
import java.util.HashMap;
import java.util.List;
import java.util.Map;
import java.util.concurrent.CompletableFuture;
import java.util.concurrent.LinkedBlockingQueue;
import java.util.concurrent.TimeUnit;
import java.util.concurrent.atomic.AtomicInteger;

import com.roche.sequencing.bioinformatics.common.alignment.IAlignmentScorer;
import com.roche.sequencing.bioinformatics.common.alignment.SimpleAlignmentScorer;
import com.roche.sequencing.bioinformatics.common.sequence.ISequence;
import com.roche.sequencing.bioinformatics.common.sequence.IupacNucleotideCode;
import com.roche.sequencing.bioinformatics.common.utils.probeinfo.Probe;
import com.roche.sequencing.bioinformatics.common.utils.probeinfo.Strand;

import static java.lang.foreign.ValueLayout.*;

public final class HsqBatcher implements AutoCloseable {

    private static volatile HsqBatcher instance;

    public static HsqBatcher get() {
        HsqBatcher b = instance;
        if (b == null) {
            synchronized (HsqBatcher.class) {
                if (instance == null) instance = new HsqBatcher(Integer.getInteger("hsq.b200.devices", 1));
                b = instance;
            }
        }
        return b;
    }

    /** One pair as the probe task holds it: raw reads, qualities, FASTQ ordinal (Integer.parseInt(readName)), probe row. */
    public record Pair(int probeRow, int readIndex, String r1, String r2, String q1, String q2) {}

    /** Per pair answers of the reduction + the counters of every probe touched by the request. */
    public record DedupAnswer(byte[] isBest, int[] groupSize, Map<Integer, int[]> probeCounters /* row -> unique, duplicate, max per UID */) {}

    /** Per read end answers of the extension (index 2 * pair + mate). */
    public record ExtendAnswer(byte[] ok, int[] pos, int[] readStart, int[] readStop, int[] nm, int[] tlen, String[] cigar, String[] md, byte[] pairOk) {}

    private enum Kind { DEDUP, EXTEND }

    private static final class Request {
        final Kind kind;
        final List<Pair> pairs;
        final int extUid, ligUid;
        final CompletableFuture<Object> done = new CompletableFuture<>();
        Request(Kind kind, List<Pair> pairs, int extUid, int ligUid) {
            this.kind = kind; this.pairs = pairs; this.extUid = extUid; this.ligUid = ligUid;
        }
    }

    private final HsqB200[] gpus;
    private final LinkedBlockingQueue<Request> queue = new LinkedBlockingQueue<>();
    private final AtomicInteger poolThreads = new AtomicInteger(), waiting = new AtomicInteger();
    private final Thread flusher;
    private final int batchPairs = Integer.getInteger("hsq.b200.batchPairs", 1 << 20);
    private final long flushMillis = Long.getLong("hsq.b200.flushMillis", 5);
    private final Map<String, Integer> probeRow = new HashMap<>();
    private int nProbes;
    private volatile boolean closed;
    private final Object alignLock = new Object();

    private HsqBatcher(int devices) {
        gpus = new HsqB200[devices];
        for (int d = 0; d < devices; d++) gpus[d] = new HsqB200(d);  // throws when there is no B200: never fall back
        if (devices > 1) HsqB200.commInitAll(gpus);
        flusher = new Thread(this::flushLoop, "hsq-b200-flusher");
        flusher.setDaemon(true);
        flusher.start();
    }

    // ---- scorer ----------------------------------------------------------------------------------------------------------------
    /** match, mismatch, gapExtend, gapStart, terminal-gap flags of a SimpleAlignmentScorer (its fields are private: read through the
     *  interface, COMMON/alignment/SimpleAlignmentScorer.java:60-96); any other scorer, or a non-integer score: UnsupportedOperationException. */
    public static int[] requireSimple(IAlignmentScorer s) {
        if (!(s instanceof SimpleAlignmentScorer)) throw new UnsupportedOperationException("only SimpleAlignmentScorer is accelerated: " + s.getClass());
        double[] v = {s.getMatchScore(IupacNucleotideCode.A, IupacNucleotideCode.A), s.getMatchScore(IupacNucleotideCode.A, IupacNucleotideCode.C),
                s.getGapScore(1), s.getGapStartScore()};
        int[] out = new int[5];
        for (int k = 0; k < 4; k++) {
            if (v[k] != Math.rint(v[k]) || Math.abs(v[k]) > 1 << 20) throw new UnsupportedOperationException("non-integer alignment score " + v[k]);
            out[k] = (int) v[k];
        }
        out[4] = (s.shouldPenalizeStartingTerminalGaps() ? HsqB200.ALIGN_PENALIZE_START : 0) | (s.shouldPenalizeEndingTerminalGaps() ? HsqB200.ALIGN_PENALIZE_END : 0);
        return out;
    }

    public static MemorySegment scorer(Arena a, int[] sc) {
        MemorySegment s = a.allocate(HsqB200.SCORER);
        for (int k = 0; k < 4; k++) s.set(JAVA_INT, 4L * k, sc[k]);
        return s;
    }

    // ---- buffers ---------------------------------------------------------------------------------------------------------------
    /** hsq_seqs over the ISO_8859_1 bytes of `strings` (ragged: data + n + 1 offsets), allocated in `a`. */
    public static MemorySegment seqs(Arena a, List<String> strings) {
        long total = 0;
        for (String s : strings) total += s.length();
        MemorySegment data = a.allocate(Math.max(total, 1)), off = a.allocate(JAVA_LONG, strings.size() + 1L);
        long at = 0;
        for (int k = 0; k < strings.size(); k++) {
            off.setAtIndex(JAVA_LONG, k, at);
            byte[] b = strings.get(k).getBytes(StandardCharsets.ISO_8859_1);
            MemorySegment.copy(b, 0, data, JAVA_BYTE, at, b.length);
            at += b.length;
        }
        off.setAtIndex(JAVA_LONG, strings.size(), at);
        MemorySegment s = a.allocate(HsqB200.SEQS);
        HsqB200.putSeqs(s, 0, data, off, strings.size());
        return s;
    }

    public static List<String> strings(List<? extends ISequence> seqs) {
        List<String> out = new ArrayList<>(seqs.size());
        for (ISequence s : seqs) out.add(s.toString());
        return out;
    }

    /** hsq_align_out with the columns GpuNeedlemanWunschGlobalAlignment reads; the others stay NULL and are not copied. */
    public static final class AlignOut {
        public final MemorySegment struct, score, nCols, refAln, qryAln, matches, mismatches, overflow;
        public AlignOut(Arena a, int n, int colsCap, int mdCap) {
            struct = a.allocate(HsqB200.ALIGN_OUT);  // zero filled: every pointer NULL
            score = a.allocate(JAVA_INT, n); nCols = a.allocate(JAVA_INT, n); matches = a.allocate(JAVA_INT, n); mismatches = a.allocate(JAVA_INT, n);
            refAln = a.allocate((long) n * colsCap); qryAln = a.allocate((long) n * colsCap); overflow = a.allocate(n);
            struct.set(JAVA_INT, 0, colsCap);
            struct.set(JAVA_INT, 4, mdCap);
            set("score", score); set("n_cols", nCols); set("ref_aln", refAln); set("qry_aln", qryAln); set("matches", matches);
            set("mismatches", mismatches); set("overflow", overflow);
        }
        private void set(String field, MemorySegment v) {
            struct.set(ADDRESS, HsqB200.ALIGN_OUT.byteOffset(java.lang.foreign.MemoryLayout.PathElement.groupElement(field)), v);
        }
    }

    public Object alignLock() { return alignLock; }

    /** The context the operator seam uses (GPU 0; guarded by alignLock, and by the flusher holding the same lock while it runs a batch). */
    public HsqB200 alignContext() { return gpus[0]; }

    // ---- probes ----------------------------------------------------------------------------------------------------------------
    /** ParsedProbeFile -> device resident probe table on every GPU; row = position in `probes` (fields of Probe.java:53-69). */
    public synchronized void registerProbes(List<Probe> probes) {
        try (Arena a = Arena.ofConfined()) {
            int n = probes.size();
            MemorySegment t = a.allocate(HsqB200.PROBE_TABLE), rev = a.allocate(Math.max(n, 1));
            MemorySegment[] cols = new MemorySegment[6];
            for (int c = 0; c < 6; c++) cols[c] = a.allocate(JAVA_INT, Math.max(n, 1));
            List<String> ext = new ArrayList<>(n), lig = new ArrayList<>(n), target = new ArrayList<>(n);
            probeRow.clear();
            for (int k = 0; k < n; k++) {
                Probe p = probes.get(k);
                probeRow.put(p.getProbeId(), k);
                rev.set(JAVA_BYTE, k, (byte) (p.getProbeStrand() == Strand.REVERSE ? 1 : 0));
                int[] v = {p.getExtensionPrimerStart(), p.getExtensionPrimerStop(), p.getLigationPrimerStart(), p.getLigationPrimerStop(),
                        p.getCaptureTargetStart(), p.getCaptureTargetStop()};
                for (int c = 0; c < 6; c++) cols[c].setAtIndex(JAVA_INT, k, v[c]);
                ext.add(p.getExtensionPrimerSequence().toString());
                lig.add(p.getLigationPrimerSequence().toString());
                target.add(p.getCaptureTargetSequence().toString());
            }
            t.set(JAVA_LONG, 0, n);
            t.set(ADDRESS, 8, rev);
            for (int c = 0; c < 6; c++) t.set(ADDRESS, 16 + 8L * c, cols[c]);
            MemorySegment.copy(seqs(a, ext), 0, t, 64, 24);
            MemorySegment.copy(seqs(a, lig), 0, t, 88, 24);
            MemorySegment.copy(seqs(a, target), 0, t, 112, 24);
            synchronized (alignLock) {
                for (HsqB200 g : gpus) g.setProbes(t);
            }
            nProbes = n;
        }
    }

    public int rowOf(Probe p) {
        Integer r = probeRow.get(p.getProbeId());
        if (r == null) throw new IllegalStateException("probe " + p.getProbeId() + " was not registered with HsqBatcher.registerProbes");
        return r;
    }

    public void setUserScorer(IAlignmentScorer s) {
        int[] sc = requireSimple(s);
        synchronized (alignLock) {
            for (HsqB200 g : gpus) g.setScorer(sc[0], sc[1], sc[2], sc[3]);
        }
    }

    // ---- tickets ---------------------------------------------------------------------------------------------------------------
    /** Pool threads announce themselves so that "every pool thread is waiting" can flush a short batch at once. */
    public void enterPoolThread() { poolThreads.incrementAndGet(); }

    public void leavePoolThread() { poolThreads.decrementAndGet(); }

    public DedupAnswer dedup(List<Pair> pairs, int extUidLen, int ligUidLen) { return (DedupAnswer) await(new Request(Kind.DEDUP, pairs, extUidLen, ligUidLen)); }

    public ExtendAnswer extend(List<Pair> pairs) { return (ExtendAnswer) await(new Request(Kind.EXTEND, pairs, 0, 0)); }

    private Object await(Request r) {
        if (closed) throw new IllegalStateException("HsqBatcher is closed");
        queue.add(r);
        waiting.incrementAndGet();
        try {
            return r.done.join();  // CompletionException carries the batch's exception to this caller's thread
        } finally {
            waiting.decrementAndGet();
        }
    }

    // ---- the flusher -----------------------------------------------------------------------------------------------------------
    private void flushLoop() {
        List<Request> open = new ArrayList<>();
        long openPairs = 0;
        while (!closed) {
            try {
                Request r = queue.poll(flushMillis, TimeUnit.MILLISECONDS);
                if (r != null) {
                    // a batch holds one kind of request with one UID layout; anything else closes it first
                    if (!open.isEmpty() && (open.get(0).kind != r.kind || open.get(0).extUid != r.extUid || open.get(0).ligUid != r.ligUid)) {
                        run(open);
                        open = new ArrayList<>();
                        openPairs = 0;
                    }
                    open.add(r);
                    openPairs += r.pairs.size();
                }
                boolean everyoneWaits = poolThreads.get() > 0 && waiting.get() >= poolThreads.get() && queue.isEmpty();
                if (!open.isEmpty() && (openPairs >= batchPairs || everyoneWaits || r == null)) {
                    run(open);
                    open = new ArrayList<>();
                    openPairs = 0;
                }
            } catch (InterruptedException e) {
                Thread.currentThread().interrupt();
                break;
            }
        }
        for (Request r : open) r.done.completeExceptionally(new IllegalStateException("HsqBatcher closed"));
    }

    private void run(List<Request> batch) {
        try (Arena a = Arena.ofConfined()) {
            List<Pair> all = new ArrayList<>();
            for (Request r : batch) all.addAll(r.pairs);
            int n = all.size();
            // hsq_pair_batch.  (Large installations fill pinned buffers from hsq_pinned_alloc once and reuse them, and hand the reads
            // over packed -- hsq_pack_reads + hsq_dedup_extend_packed, a third of the PCIe bytes; the Arena keeps this listing short.)
            MemorySegment pb = a.allocate(HsqB200.PAIR_BATCH), probe = a.allocate(JAVA_INT, Math.max(n, 1)), index = a.allocate(JAVA_INT, Math.max(n, 1));
            List<String> r1 = new ArrayList<>(n), r2 = new ArrayList<>(n), q1 = new ArrayList<>(n), q2 = new ArrayList<>(n);
            for (int k = 0; k < n; k++) {
                Pair p = all.get(k);
                probe.setAtIndex(JAVA_INT, k, p.probeRow());
                index.setAtIndex(JAVA_INT, k, p.readIndex());
                r1.add(p.r1()); r2.add(p.r2()); q1.add(p.q1()); q2.add(p.q2());
            }
            MemorySegment s1 = seqs(a, r1), s2 = seqs(a, r2), sq1 = seqs(a, q1), sq2 = seqs(a, q2);
            pb.set(JAVA_LONG, 0, n);
            pb.set(ADDRESS, 8, probe);
            pb.set(ADDRESS, 16, index);
            MemorySegment.copy(s1, 0, pb, 24, 24);
            MemorySegment.copy(s2, 0, pb, 48, 24);
            pb.set(ADDRESS, 72, sq1.get(ADDRESS, 0));  // qualities are laid out by the reads' offsets (equal lengths per read)
            pb.set(ADDRESS, 80, sq2.get(ADDRESS, 0));
            synchronized (alignLock) {
                if (batch.get(0).kind == Kind.DEDUP) runDedup(a, batch, pb, n);
                else runExtend(a, batch, pb, n);
            }
        } catch (Throwable t) {
            for (Request r : batch) r.done.completeExceptionally(t);
        }
    }

    private void runDedup(Arena a, List<Request> batch, MemorySegment pb, int n) {
        MemorySegment best = a.allocate(Math.max(n, 1)), size = a.allocate(JAVA_INT, Math.max(n, 1));
        MemorySegment stats = a.allocate(HsqB200.PROBE_STATS), u = a.allocate(JAVA_INT, nProbes), d = a.allocate(JAVA_INT, nProbes), m = a.allocate(JAVA_INT, nProbes);
        stats.set(JAVA_LONG, 0, nProbes);
        stats.set(ADDRESS, 8, u); stats.set(ADDRESS, 16, d); stats.set(ADDRESS, 24, m);
        int ext = batch.get(0).extUid, lig = batch.get(0).ligUid;
        // one GPU: hsq_dedup_batch.  Several: the batch is cut into one slice per context and every context's own host thread calls
        // hsq_dedup_sharded (collective; the library exchanges the key records over NCCL) -- see HsqB200.dedupExtendMultiGpu for the
        // one-call form of the whole probe task
        if (gpus.length == 1) gpus[0].dedup(pb, ext, lig, best, size, stats);
        else ShardedCalls.dedup(gpus, pb, n, ext, lig, best, size, stats);
        int at = 0;
        for (Request r : batch) {
            int k = r.pairs.size();
            byte[] b = best.asSlice(at, k).toArray(JAVA_BYTE);
            int[] g = size.asSlice(4L * at, 4L * k).toArray(JAVA_INT);
            Map<Integer, int[]> counters = new HashMap<>();
            for (Pair p : r.pairs)
                counters.computeIfAbsent(p.probeRow(), row -> new int[] {u.getAtIndex(JAVA_INT, row), d.getAtIndex(JAVA_INT, row), m.getAtIndex(JAVA_INT, row)});
            r.done.complete(new DedupAnswer(b, g, counters));
            at += k;
        }
    }

    private void runExtend(Arena a, List<Request> batch, MemorySegment pb, int n) {
        final int cigarCap = 64, mdCap = 96;
        long n2 = 2L * Math.max(n, 1);
        MemorySegment o = a.allocate(HsqB200.EXTEND_OUT);
        String[] i32 = {"primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen", "cigar_len", "md_len"};
        Map<String, MemorySegment> col = new HashMap<>();
        for (String f : i32) col.put(f, a.allocate(JAVA_INT, n2));
        MemorySegment ok = a.allocate(n2), overflow = a.allocate(n2), pairOk = a.allocate(Math.max(n, 1)), cigar = a.allocate(n2 * cigarCap), md = a.allocate(n2 * mdCap);
        o.set(JAVA_INT, 0, cigarCap);
        o.set(JAVA_INT, 4, mdCap);
        java.util.function.BiConsumer<String, MemorySegment> put =
                (f, v) -> o.set(ADDRESS, HsqB200.EXTEND_OUT.byteOffset(java.lang.foreign.MemoryLayout.PathElement.groupElement(f)), v);
        for (String f : i32) put.accept(f, col.get(f));
        put.accept("ok", ok); put.accept("overflow", overflow); put.accept("pair_ok", pairOk); put.accept("cigar", cigar); put.accept("md", md);
        // extension shards by read pair with no exchange: with several GPUs the batch is cut into slices (ShardedCalls.extend)
        if (gpus.length == 1) gpus[0].extend(pb, o);
        else ShardedCalls.extend(gpus, pb, o, n, cigarCap, mdCap);
        int at = 0;
        for (Request r : batch) {
            int k = r.pairs.size();
            String[] cg = new String[2 * k], mds = new String[2 * k];
            for (int e = 0; e < 2 * k; e++) {
                long row = 2L * at + e;
                cg[e] = new String(cigar.asSlice(row * cigarCap, col.get("cigar_len").getAtIndex(JAVA_INT, row)).toArray(JAVA_BYTE), StandardCharsets.ISO_8859_1);
                mds[e] = new String(md.asSlice(row * mdCap, col.get("md_len").getAtIndex(JAVA_INT, row)).toArray(JAVA_BYTE), StandardCharsets.ISO_8859_1);
            }
            final int first = at;
            java.util.function.Function<String, int[]> ints = f -> col.get(f).asSlice(8L * first, 8L * k).toArray(JAVA_INT);
            r.done.complete(new ExtendAnswer(ok.asSlice(2L * at, 2L * k).toArray(JAVA_BYTE), ints.apply("pos"), ints.apply("read_start"), ints.apply("read_stop"),
                    ints.apply("nm"), ints.apply("tlen"), cg, mds, pairOk.asSlice(at, k).toArray(JAVA_BYTE)));
            at += k;
        }
    }

    @Override
    public void close() {
        closed = true;
        flusher.interrupt();
        for (HsqB200 g : gpus) g.close();
    }
}
