"""The sharded dedup on real GPUs over NCCL (needs >= 2 GPUs; skipped otherwise): every rank's answers equal the
single-process FilterByUid restatement on the union of the ranks' pairs."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, ret):
    import torch
    import torch.distributed as dist
    from bioinformatics_b200 import sharded, synth
    from bioinformatics_b200._lib import Context
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        probes = synth.make_probes(61, seed=3)
        pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 40000, read_len=100, ext_uid_len=6, lig_uid_len=2, seed=4))
        mine = np.arange(rank, pairs.n, world)
        with Context(rank) as ctx:
            best, gsize, stats = sharded.dedup_sharded(sharded.DeviceBackend(ctx), pairs.subset(mine), 6, 2, probes.n)
            merged = sharded.merge_owner_stats(stats, probes.n)
        ret[rank] = (mine, best, gsize, merged)
    finally:
        dist.destroy_process_group()


def test_sharded_dedup_over_nccl():
    import torch
    import torch.multiprocessing as mp
    from bioinformatics_b200 import synth
    from oracle import oracle as O
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    probes = synth.make_probes(61, seed=3)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 40000, read_len=100, ext_uid_len=6, lig_uid_len=2, seed=4))
    eb, egs, est = O.dedup_batch(pairs, 6, 2, probes.n)
    got_b = np.zeros(pairs.n, dtype=np.uint8)
    got_g = np.zeros(pairs.n, dtype=np.int32)
    for r in range(world):
        mine, best, gsize, merged = ret[r]
        got_b[mine] = best
        got_g[mine] = gsize
        assert (merged[0] == est.unique_pairs).all() and (merged[2] == est.max_pairs_per_uid).all()
    assert (got_b == eb).all() and (got_g == egs).all()
