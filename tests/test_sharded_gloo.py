"""The multi-GPU dedup exchange (bioinformatics_b200/sharded.py) at world_size 2 over gloo on the CPU.

The two device calls are replaced by the oracle's CPU stand-ins, so what is tested here is the host logic: routing of key
records to the probe's owner, the variable-split all-to-all, the way back, and that the result equals the single-process
FilterByUid restatement on the whole batch whatever the partition."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bioinformatics_b200 import sharded, synth
from oracle import oracle as O


class OracleBackend:
    def keys(self, pairs, u, v, n_probes, origin_base):
        k = O.dedup_keys(pairs, u, v, origin_base)
        return torch.from_numpy(k.view(np.uint8).reshape(-1, sharded.KEY_BYTES).copy())

    def reduce(self, keys, uid_len, n_probes):
        rec = keys.contiguous().numpy().reshape(-1).view(O.KEY_DTYPE)
        best, gsize, stats = O.dedup_reduce_keys(rec, n_probes)
        return torch.from_numpy(best), torch.from_numpy(gsize), stats


def _workload(ties=False):
    probes = synth.make_probes(37, seed=3)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 3000, read_len=80, ext_uid_len=5, lig_uid_len=2, seed=4))
    pairs.q1.data[::5] = ord("5")  # quality-sum ties -> read-index tie-break
    if ties:  # repeated read indexes and equal qualities: exact ties, broken by the pair's place in the whole batch (origin)
        pairs.read_index[:] = np.random.default_rng(1).integers(0, 40, size=pairs.n)
        pairs.q1.data[:] = ord("I")
        pairs.q2.data[:] = ord("I")
    return probes, pairs


def _worker(rank, world, port, partition, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        probes, pairs = _workload(ties=partition == "ties")
        if partition == "interleaved":
            mine = np.arange(rank, pairs.n, world)
        elif partition in ("blocks", "ties"):  # rank-major concatenation = the original order, so origins equal pair indexes
            mine = np.array_split(np.arange(pairs.n), world)[rank]
        else:  # everything on rank 0: the other rank sends nothing and still owns probes
            mine = np.arange(pairs.n) if rank == 0 else np.arange(0)
        sub = pairs.subset(mine)
        best, gsize, stats = sharded.dedup_sharded(OracleBackend(), sub, 5, 2, probes.n)
        merged = sharded.merge_owner_stats(stats, probes.n)
        ret[rank] = (mine, best, gsize, merged)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("partition", ["interleaved", "blocks", "one_rank", "ties"])
def test_sharded_dedup_matches_single_process(partition):
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, partition, ret), nprocs=world, join=True)
    probes, pairs = _workload(ties=partition == "ties")
    eb, egs, est = O.dedup_batch(pairs, 5, 2, probes.n)
    got_b = np.zeros(pairs.n, dtype=np.uint8)
    got_g = np.zeros(pairs.n, dtype=np.int32)
    for r in range(world):
        mine, best, gsize, merged = ret[r]
        got_b[mine] = best
        got_g[mine] = gsize
        assert (merged[0] == est.unique_pairs).all() and (merged[1] == est.duplicate_pairs).all()
        assert (merged[2] == est.max_pairs_per_uid).all()
    assert (got_b == eb).all() and (got_g == egs).all()


def test_key_stand_in_matches_oracle_single_rank():
    """The numpy key/reduce stand-ins agree with the C++ restatement of FilterByUid (pins the test infrastructure)."""
    probes, pairs = _workload()
    best, gsize, stats = O.dedup_reduce_keys(O.dedup_keys(pairs, 5, 2), probes.n)
    eb, egs, est = O.dedup_batch(pairs, 5, 2, probes.n)
    assert (best == eb).all() and (gsize == egs).all() and (stats.unique_pairs == est.unique_pairs).all()
