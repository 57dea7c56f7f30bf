"""BASELINE.json configs[4]: read-length sweep 100..300 bp, GCUPS against the measured INT32 issue rate (one GPU per
process; under torchrun every rank runs the sweep on its GPU and rank 0 prints the max-over-ranks time).
The band axis of that config has a single point, "full": the reference fills full matrices and its free end gaps make
any band lossy, so this implementation does not band.
Usage: python scripts/sweep_lengths.py [pairs_per_length]   -> one JSON line per length."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from bioinformatics_b200 import synth  # noqa: E402
from bioinformatics_b200._lib import Context  # noqa: E402
from bioinformatics_b200.batches import ExtendResult  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
rank = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
dist = None
if world > 1:
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))

with Context(rank) as ctx:
    peak = ctx.int32_peak(1.0)
    for L in (100, 150, 200, 250, 300):
        probes = synth.make_probes(1000, target_len=(L - 50, L + 10))
        pairs = synth.make_pairs(probes, n, read_len=L, ext_uid_len=10, with_quals=False, seed=synth.SEED_READS + rank)
        ctx.set_probes(probes)
        out = ExtendResult(n, 32, 64)
        ctx.extend_stage(pairs)
        for _ in range(3):
            ctx.extend_run(32, 64)
        ctx.extend_fetch(out)
        ctx.timer_start()
        for _ in range(5):
            ctx.extend_run(32, 64)
        ms = ctx.timer_stop() / 5
        cells = float(out.cells)
        if dist:
            t = torch.tensor([ms, cells], dtype=torch.float64, device="cuda")
            mx = t.clone()
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            ms, cells = float(mx[0]), float(t[1])
        if rank == 0:
            print(json.dumps({"read_len": L, "band": "full", "n_gpus": world, "pairs_per_gpu": n, "GCUPS": cells / ms / 1e6,
                              "ms_per_step": ms, "pair_ok": float(out.pair_ok.mean()),
                              "frac_of_measured_int32_peak": cells / world * 12 / (ms * 1e-3) / peak}), flush=True)
if dist:
    dist.destroy_process_group()
