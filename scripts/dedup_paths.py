"""hsq_dedup_batch device time (copies included) with the segment path and with the radix path (hsq_config.force_dedup_radix) on the shapes of
BASELINE.json configs[2] / configs[3] (development aid):  python scripts/dedup_paths.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests.test_full_size import _dedup_problem
from bioinformatics_b200._lib import Context

for n, n_probes in ((2_000_000, 2_000), (2_000_000, 200_000), (2_000_000, 20)):
    pairs = _dedup_problem(n, n_probes, 14, 150, seed=1)[0]
    res = {}
    for mode in ("segment", "radix"):
        with Context(0, config={"force_dedup_radix": 1 if mode == "radix" else 0}) as ctx:
            for _ in range(2):
                best, gs, st = ctx.dedup_batch(pairs, 14, 0, n_probes)
            ctx.lib.hsq_timer_start(ctx.h)
            res[mode] = (ctx.last_timing(), int(best.sum()), int(gs.max()))
    print(n, n_probes, res, flush=True)
