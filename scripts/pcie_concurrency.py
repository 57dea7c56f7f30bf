"""N ranks, one per GPU, each copying pinned host memory to its GPU and back at the same time (what every host-to-host number of this
repository is made of): per-rank and aggregate GB/s, H2D alone, D2H alone, both directions together.
   python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 scripts/pcie_concurrency.py [GB per direction]
Prints one JSON line on rank 0 (profiles/pcie_concurrency_r2_*.json)."""
import json, os, sys, time
import torch
import torch.distributed as dist

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
gb = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(gb * (1 << 30))
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
h_in.fill_(7)
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.full((n,), 3, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def run(h2d, d2h, reps=5):
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    moved = reps * n * ((1 if h2d else 0) + (1 if d2h else 0))
    return {"per_rank_gb_s": moved / dt / 1e9, "aggregate_gb_s": moved * world / dt / 1e9}


run(True, True, 1)
res = {"n_gpus": world, "gb_per_direction_per_rank": gb, "h2d": run(True, False), "d2h": run(False, True), "both": run(True, True)}
try:
    res["numa"] = os.popen("nvidia-smi topo -m 2>/dev/null | head -12").read()
    res["host_cpus"] = os.cpu_count()
except Exception:
    pass
if rank == 0:
    print(json.dumps(res), flush=True)
if world > 1:
    dist.destroy_process_group()
