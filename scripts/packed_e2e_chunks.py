"""hsq_extend_packed host to host for several chunk counts of its pipeline (development aid): python scripts/packed_e2e_chunks.py [pairs]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bioinformatics_b200 import synth
from bioinformatics_b200._lib import Context, pack_reads
from bioinformatics_b200.batches import PackedPairsHost, PooledExtendResult

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
probes = synth.make_probes(1000)
base = synth.make_pairs(probes, min(n, 1_250_000), with_quals=False)
with Context(0) as ctx:
    ctx.set_probes(probes)
    pin = ctx.pinned_array
    pairs = synth.tile_pairs(base, max(1, n // base.n), pin) if n > base.n else base
    n = pairs.n
    pp = PackedPairsHost(pairs.probe_id, pack_reads(pairs.r1, 2, pin), pack_reads(pairs.r2, 2, pin))
    slim = PooledExtendResult(n, 2 * n * 24, fields=("ok", "pair_ok", "pos", "read_start", "read_stop", "nm", "cigar_len", "md_len", "str_off"), alloc=pin)
    ctx.extend_stage(pairs)
    ctx.extend_run(32, 64)
    ctx.timer_start()
    for _ in range(3):
        ctx.extend_run(32, 64)
    print("resident: %.2f ms" % (ctx.timer_stop() / 3))
    for chunks, uniform in ((0, 0), (6, 0), (8, 0), (10, 0), (16, 0), (20, 0), (8, 1), (12, 1), (16, 1)):
        ctx.configure(extend_chunks=chunks, uniform_chunks=uniform)
        best = 1e9
        for it in range(4):
            t = time.perf_counter(); ctx.extend_packed(pp, out=slim); best = min(best, time.perf_counter() - t)
        print("chunks %2d uniform %d: %.2f ms host to host (%d launches)" % (chunks, uniform, best * 1e3, ctx.last_timing()[1]), flush=True)
