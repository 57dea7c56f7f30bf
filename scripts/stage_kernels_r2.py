"""One pass over the HBM-bound stage calls at the bench's shapes (development aid, and the command behind profiles/launches_stages_r2_*.csv):
   ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv python scripts/stage_kernels_r2.py [pairs]
Without ncu it prints the CUDA-event spans of each call (hsq_last_spans)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bioinformatics_b200 import synth
from bioinformatics_b200._lib import Context, pack_reads, quality_sums
from bioinformatics_b200.batches import PackedPairsHost

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
shapes = (("c3: 10 k probes, 14 bp UID", 10000, 14), ("c4: 200 k probes, 10 bp UID", 200000, 10), ("small panel: 1 k probes (every probe split)", 1000, 10))
with Context(0) as ctx:
    for name, n_probes, u in shapes:
        probes = synth.make_probes(n_probes)
        base, _ = synth.drop_n_pairs(synth.make_pairs(probes, min(n, 500_000), read_len=150, ext_uid_len=u, seed=11))
        pairs = synth.tile_pairs(base, max(1, n // base.n), ctx.pinned_array)
        for rep in range(2):
            best, gs, st = ctx.dedup_batch(pairs, u, 0, n_probes)
        print("%-45s dedup_batch %8d pairs: call %.3f ms, keygen %.3f ms, group %.3f ms" % ((name, pairs.n, ctx.last_timing()[0]) + tuple(ctx.last_spans()[:2])), flush=True)
        if n_probes == 10000:
            pk = PackedPairsHost(pairs.probe_id, pack_reads(pairs.r1, 2, ctx.pinned_array), pack_reads(pairs.r2, 2, ctx.pinned_array), None, None,
                                 pairs.read_index, quality_sums(pairs.q1), quality_sums(pairs.q2))
            b2, _, _ = ctx.dedup_packed(pk, u, 0, n_probes)
            assert (b2 == best).all()
            print("%-45s dedup_packed (host sums): call %.3f ms, keygen %.3f ms, group %.3f ms" % ((name, ctx.last_timing()[0]) + tuple(ctx.last_spans()[:2])), flush=True)
            tout = (ctx.pinned_array(pairs.r1.data.nbytes), ctx.pinned_array(pairs.r1.data.nbytes), ctx.pinned_array(8 * (pairs.n + 1)).view(np.int64))
            for chunks in (1, 0):
                ctx.configure(trim_chunks=chunks)
                for sb, qb in ((pairs.r1, pairs.q1), (pairs.r2, pairs.q2)):
                    ctx.trim_batch(sb, qb, u + 22, 150 - 23, True, out=tout)
                    print("%-45s trim_batch (trim_chunks %d, pinned in and out): call %.3f ms, spans %s" % (name, chunks, ctx.last_timing()[0], ctx.last_spans()[:2]), flush=True)
            ctx.configure()
