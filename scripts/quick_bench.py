"""Quick device timing of hsq_extend_batch (development aid; bench.py is the contract)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bioinformatics_b200 import synth
from bioinformatics_b200._lib import Context
from bioinformatics_b200.batches import ExtendResult

n = int(sys.argv[1]) if len(sys.argv) > 1 else 500000
sortp = len(sys.argv) > 2 and sys.argv[2] == "sorted"
probes = synth.make_probes(1000)
t = time.time()
pairs = synth.make_pairs(probes, n, with_quals=False, sort_by_probe=sortp)
print("gen %.1fs" % (time.time() - t))
with Context(0) as ctx:
    ctx.set_probes(probes)
    out = ExtendResult(n)
    for it in range(4):
        t = time.time()
        ctx.extend_batch(pairs, out=out)
        wall = time.time() - t
        ms, launches = ctx.last_timing()
        print("iter %d: call %.1f ms (events %.1f ms, %d launches) cells %.3e -> %.1f GCUPS e2e, pair_ok %.3f" %
              (it, wall * 1e3, ms, launches, out.cells, out.cells / wall / 1e9, out.pair_ok.mean()))
