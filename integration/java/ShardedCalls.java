/*
 * ShardedCalls -- one host thread per GPU context for the collective entry points (hsq_dedup_sharded) and for the exchange-free
 * extension over slices of one batch.  The library does the NCCL exchange itself (csrc/hsq_capi_comm.inl); Java only starts one
 * thread per context because a collective needs every rank inside the call at the same time.
 * NOT compiled in this repository's image (no JVM there).
 */
package com.roche.heatseq.b200;

import java.lang.foreign.Arena;
import java.lang.foreign.MemorySegment;
import java.util.ArrayList;
import java.util.List;

import static java.lang.foreign.ValueLayout.*;

final class ShardedCalls {
    private ShardedCalls() {}

    interface Call { void run(int rank, long first, long count) throws Throwable; }

    /** Runs call(rank, first pair, pair count) on one thread per rank over equal slices of n pairs; rethrows the first failure. */
    static void overSlices(int ranks, long n, Call call) {
        List<Thread> threads = new ArrayList<>();
        Throwable[] failure = new Throwable[ranks];
        for (int r = 0; r < ranks; r++) {
            final int rank = r;
            final long first = n * r / ranks, count = n * (r + 1) / ranks - first;
            Thread t = new Thread(() -> {
                try {
                    call.run(rank, first, count);
                } catch (Throwable e) {
                    failure[rank] = e;
                }
            }, "hsq-b200-rank-" + r);
            threads.add(t);
            t.start();
        }
        for (Thread t : threads) {
            try {
                t.join();
            } catch (InterruptedException e) {
                Thread.currentThread().interrupt();
            }
        }
        for (Throwable e : failure) if (e != null) throw new IllegalStateException(e);
    }

    /** hsq_pair_batch window [first, first + count) of `pb` (offsets may start anywhere: the library rebases them). */
    static MemorySegment window(Arena a, MemorySegment pb, long first, long count) {
        MemorySegment w = a.allocate(HsqB200.PAIR_BATCH);
        w.copyFrom(pb);
        w.set(JAVA_LONG, 0, count);
        w.set(ADDRESS, 8, MemorySegment.ofAddress(pb.get(ADDRESS, 8).address() + 4 * first));
        w.set(ADDRESS, 16, MemorySegment.ofAddress(pb.get(ADDRESS, 16).address() + 4 * first));
        for (long seqs : new long[] {24, 48}) {  // r1, r2: data stays, the offsets pointer moves
            w.set(ADDRESS, seqs + 8, MemorySegment.ofAddress(pb.get(ADDRESS, seqs + 8).address() + 8 * first));
            w.set(JAVA_LONG, seqs + 16, count);
        }
        return w;
    }

    private static void shift(MemorySegment dst, MemorySegment src, String field, long bytes) {
        long at = HsqB200.EXTEND_OUT.byteOffset(java.lang.foreign.MemoryLayout.PathElement.groupElement(field));
        MemorySegment p = src.get(ADDRESS, at);
        if (p.address() != 0) dst.set(ADDRESS, at, MemorySegment.ofAddress(p.address() + bytes));
    }

    /** FilterByUid over the whole batch with the reduction sharded over the contexts' communicator. */
    static void dedup(HsqB200[] gpus, MemorySegment pb, long n, int ext, int lig, MemorySegment best, MemorySegment size, MemorySegment stats) {
        overSlices(gpus.length, n, (rank, first, count) -> {
            try (Arena a = Arena.ofConfined()) {
                // every rank gets the counters of the whole batch; rank 0 writes the caller's arrays, the others scratch copies
                MemorySegment st = stats;
                if (rank != 0) {
                    long np = stats.get(JAVA_LONG, 0);
                    st = a.allocate(HsqB200.PROBE_STATS);
                    st.set(JAVA_LONG, 0, np);
                    for (int c = 0; c < 3; c++) st.set(ADDRESS, 8 + 8L * c, a.allocate(JAVA_INT, np));
                }
                gpus[rank].dedupSharded(window(a, pb, first, count), ext, lig, best.asSlice(first), size.asSlice(4 * first), st);
            }
        });
    }

    /** extendReadsToPrimers over slices of one batch: no exchange, every context writes its own part of the caller's arrays. */
    static void extend(HsqB200[] gpus, MemorySegment pb, MemorySegment out, long n, int cigarCap, int mdCap) {
        overSlices(gpus.length, n, (rank, first, count) -> {
            try (Arena a = Arena.ofConfined()) {
                MemorySegment o = a.allocate(HsqB200.EXTEND_OUT);
                o.copyFrom(out);
                String[] perEnd4 = {"primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen", "cigar_len", "md_len"};  // int32 per read end
                for (String f : perEnd4) shift(o, out, f, 8 * first);
                shift(o, out, "ok", 2 * first);
                shift(o, out, "overflow", 2 * first);
                shift(o, out, "cigar", 2 * first * cigarCap);
                shift(o, out, "md", 2 * first * mdCap);
                shift(o, out, "pair_ok", first);
                gpus[rank].extend(window(a, pb, first, count), o);
            }
        });
    }
}
