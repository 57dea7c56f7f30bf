"""BASELINE.json configs[4]: read length 100..300 bp x band {8, 16, 32, 64, full}: GCUPS against the measured INT32 issue rate, the
share of read ends the band had to hand to the full matrices, and the number of results that differ from the full-matrix run
(one GPU per process; under torchrun every rank runs the sweep on its GPU and rank 0 prints the max-over-ranks time).
GCUPS counts the cells the reference fills (full matrices) whatever the band computes.
Usage: python scripts/sweep_lengths.py [pairs_per_length] [indel_rate]  -> one JSON line per (length, band)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from bioinformatics_b200 import synth  # noqa: E402
from bioinformatics_b200._lib import Context  # noqa: E402
from bioinformatics_b200.batches import ExtendResult  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
indel = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0005
rank = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
dist = None
if world > 1:
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))

FIELDS = ("primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen", "cigar_len", "md_len", "ok", "pair_ok")

with Context(rank) as ctx:
    peak = ctx.int32_peak(1.0)
    for L in (100, 150, 200, 250, 300):
        probes = synth.make_probes(1000, target_len=(L - 50, L + 10))
        pairs = synth.make_pairs(probes, n, read_len=L, ext_uid_len=10, with_quals=False, seed=synth.SEED_READS + rank, ins_rate=indel,
                                 del_rate=indel)
        ctx.set_probes(probes)
        full = None
        for band in (0, 8, 16, 32, 64):
            ctx.configure(band=band)
            out = ExtendResult(n, 32, 64)
            ctx.extend_stage(pairs)
            for _ in range(3):
                ctx.extend_run(32, 64)
            ctx.extend_fetch(out)
            banded, redone = ctx.band_stats()
            ctx.timer_start()
            for _ in range(5):
                ctx.extend_run(32, 64)
            ms = ctx.timer_stop() / 5
            cells = float(out.cells)
            differing = 0
            if band == 0:
                full = out
                full_strings = [np.array(out.cigars(), dtype=object), np.array(out.mds(), dtype=object)]
            else:  # read ends whose result is not the full-matrix one
                bad = np.zeros(2 * n, dtype=bool)
                for f in FIELDS:
                    a, b = getattr(out, f), getattr(full, f)
                    ne = a != b
                    if len(ne) == n:
                        ne = np.repeat(ne, 2)
                    bad |= ne
                bad |= np.array(out.cigars(), dtype=object) != full_strings[0]
                bad |= np.array(out.mds(), dtype=object) != full_strings[1]
                differing = int(bad.sum()) + (0 if out.cells == full.cells else 1)
            if dist:
                t = torch.tensor([ms, cells, banded, redone, differing], dtype=torch.float64, device="cuda")
                mx = t.clone()
                dist.all_reduce(mx, op=dist.ReduceOp.MAX)
                dist.all_reduce(t, op=dist.ReduceOp.SUM)
                ms, cells, banded, redone, differing = float(mx[0]), float(t[1]), int(t[2]), int(t[3]), int(t[4])
            if rank == 0:
                print(json.dumps({"read_len": L, "band": band if band else "full", "n_gpus": world, "pairs_per_gpu": n, "indel_rate": indel,
                                  "GCUPS": cells / ms / 1e6, "ms_per_step": ms, "pair_ok": float(out.pair_ok.mean()),
                                  "banded_read_ends": banded, "redone_in_full": redone,
                                  "fallback_rate": (redone / banded) if banded else None, "read_ends_differing_from_full": differing,
                                  "frac_of_measured_int32_peak": cells / world * 12 / (ms * 1e-3) / peak}), flush=True)
        ctx.configure()
if dist:
    dist.destroy_process_group()
