"""hsq_extend_packed against hsq_extend_batch on the same pairs (development aid; bench.py is the contract):
   python scripts/packed_kernels.py [pairs] [chunks]     -- wall / CUDA-event time per call
   ncu --metrics gpu__time_duration.sum --clock-control none --csv python scripts/packed_kernels.py 500000   -- launch list"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bioinformatics_b200 import synth
from bioinformatics_b200._lib import Context, pack_reads
from bioinformatics_b200.batches import ExtendResult, PackedPairsHost, PooledExtendResult

n = int(sys.argv[1]) if len(sys.argv) > 1 else 500000
chunks = int(sys.argv[2]) if len(sys.argv) > 2 else 0
probes = synth.make_probes(1000)
pairs = synth.make_pairs(probes, n, with_quals=False)
with Context(0, config={"extend_chunks": chunks}) as ctx:
    ctx.set_probes(probes)
    pin = ctx.pinned_array
    for sb in (pairs.r1, pairs.r2):
        d = pin(sb.data.nbytes); d[:] = sb.data; sb.data = d
        o = pin(sb.offsets.nbytes).view(np.int64); o[:] = sb.offsets; sb.offsets = o
    t = time.perf_counter()
    pp = PackedPairsHost(pairs.probe_id, pack_reads(pairs.r1, 2, pin), pack_reads(pairs.r2, 2, pin))
    print("pack %.1f ms" % ((time.perf_counter() - t) * 1e3))
    slim = PooledExtendResult(n, 2 * n * 24, fields=("ok", "pair_ok", "pos", "read_start", "read_stop", "nm", "cigar_len", "md_len", "str_off"), alloc=pin)
    out = ExtendResult(n, 24, 40)
    for f in ExtendResult.FIELDS_I32 + ("ok", "overflow", "pair_ok", "cigar", "md"):
        a = getattr(out, f); q = pin(a.nbytes).view(a.dtype); setattr(out, f, q)
    for it in range(3):
        t = time.perf_counter(); ctx.extend_packed(pp, out=slim); w = time.perf_counter() - t
        print("packed: call %.2f ms (events %.2f ms, %d launches), %d + %d bytes" % ((w * 1e3,) + ctx.last_timing() + (pp.r1.nbytes + pp.r2.nbytes + 4 * n, slim.nbytes())))
        t = time.perf_counter(); ctx.extend_batch(pairs, out=out); w = time.perf_counter() - t
        print("ascii : call %.2f ms (events %.2f ms, %d launches)" % ((w * 1e3,) + ctx.last_timing()))
