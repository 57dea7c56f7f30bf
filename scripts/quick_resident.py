"""Resident timing of hsq_extend_run (development aid).  Usage: quick_resident.py [pairs] [read_len] [band] [l2_persist_rows]
HSQ_B200_LIB selects an alternative build of the library (experiments)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bioinformatics_b200 import synth, _lib
if os.environ.get("HSQ_B200_LIB"):
    _lib.SO_PATH = os.environ["HSQ_B200_LIB"]
from bioinformatics_b200._lib import Context, HsqError
from bioinformatics_b200.batches import ExtendResult

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 150
probes = synth.make_probes(1000, target_len=(L - 50, L + 10))
pairs = synth.make_pairs(probes, n, read_len=L, with_quals=False)
band = int(sys.argv[3]) if len(sys.argv) > 3 else 0
persist = int(sys.argv[4]) if len(sys.argv) > 4 else 0
with Context(0, config={"band": band, "l2_persist_rows": persist}) as ctx:
    ctx.set_probes(probes)
    out = ExtendResult(n, 32, 64)
    ctx.extend_stage(pairs)
    for _ in range(3):
        ctx.extend_run(32, 64)
    try:
        ctx.extend_fetch(out)
    except HsqError as e:  # experiment builds may produce wrong strings that overflow their slots: the timing is still valid
        print("fetch:", e)
    ctx.timer_start()
    for _ in range(5):
        ctx.extend_run(32, 64)
    ms = ctx.timer_stop() / 5
    print("%s: %d pairs L=%d band %d: %.2f ms/step, %.0f GCUPS, pair_ok %.3f, band stats %s" % (os.path.basename(_lib.SO_PATH), n, L, band, ms, out.cells / ms / 1e6, out.pair_ok.mean(), ctx.band_stats()))
