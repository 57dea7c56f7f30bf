"""the multi-GPU dedup behind the C ABI (hsq_comm_*, hsq_dedup_sharded, hsq_dedup_extend_sharded,
hsq_dedup_extend_multi_gpu).  The exchange runs inside the library over NCCL; these tests run it with as many GPUs as the box has
(world = 1 on a single-GPU box: the whole collective path with self send / receive) and compare with the single-process
FilterByUid / ExtendReadsToPrimer restatement (oracle) on the union of the ranks' pairs."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

U, V = 6, 2


def _workload(ties=False):
    from bioinformatics_b200 import synth
    probes = synth.make_probes(61, seed=3)
    pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, 24000, read_len=100, ext_uid_len=U, lig_uid_len=V, seed=4))
    pairs.q1.data[::5] = ord("5")
    if ties:
        pairs.read_index[:] = np.random.default_rng(2).integers(0, 300, size=pairs.n)
        pairs.q1.data[:] = ord("I")
        pairs.q2.data[:] = ord("I")
    return probes, pairs


def _expected(probes, pairs, mark):
    """Oracle: dedup of the whole batch, then extension of the unique pairs (all pairs when duplicates are marked)."""
    from oracle import oracle as O
    eb, egs, est = O.dedup_batch(pairs, U, V, probes.n)
    idx = np.arange(pairs.n) if mark else np.nonzero(eb)[0]
    ext = O.extend_batch(probes, pairs.subset(idx), cigar_cap=64, md_cap=96, nthreads=8)
    return eb, egs, est, idx, ext


def _check_rank(res, mine, exp, probes, pairs, mark):
    """One rank's hsq_dedup_extend_out against the oracle on the whole batch."""
    eb, egs, est, idx, ext = exp
    assert np.array_equal(res.is_best, eb[mine]) and np.array_equal(res.group_size, egs[mine])
    for f in ("unique_pairs", "duplicate_pairs", "max_pairs_per_uid"):
        assert np.array_equal(getattr(res.stats, f), getattr(est, f)), f
    m = res.n_extended
    slots = mine[res.extended_pair[:m]]  # global pair of every slot
    want = mine if mark else mine[eb[mine] == 1]
    assert np.array_equal(slots, want)
    pos_of = {int(g): k for k, g in enumerate(idx.tolist())}
    rows = np.array([pos_of[int(g)] for g in slots.tolist()], dtype=np.int64)
    ends = np.stack([2 * rows, 2 * rows + 1], axis=1).reshape(-1)
    for f in ("ok", "primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen", "cigar_len", "md_len"):
        assert np.array_equal(getattr(res.extend, f)[:2 * m], getattr(ext, f)[ends]), f
    assert np.array_equal(res.extend.pair_ok[:m], ext.pair_ok[rows])
    got_c, exp_c = res.extend.head(m).cigars(), ext.cigars()
    assert got_c == [exp_c[e] for e in ends.tolist()]
    # unable-to-extend counters are those of the whole batch
    unable_u = np.zeros(probes.n, dtype=np.int64)
    unable_d = np.zeros(probes.n, dtype=np.int64)
    bad = ext.pair_ok == 0
    np.add.at(unable_u, pairs.probe_id[idx[bad & (eb[idx] == 1)]], 1)
    np.add.at(unable_d, pairs.probe_id[idx[bad & (eb[idx] == 0)]], 1)
    assert np.array_equal(res.unique_unable_to_extend, unable_u) and np.array_equal(res.duplicate_unable_to_extend, unable_d)


def _partition(n, world, kind):
    if kind == "interleaved":
        return [np.arange(r, n, world) for r in range(world)]
    if kind == "one_rank":  # everything on rank 0: the other ranks own probes and send nothing
        return [np.arange(n) if r == 0 else np.arange(0) for r in range(world)]
    return np.array_split(np.arange(n), world)


@pytest.mark.parametrize("kind,mark,ties", [("interleaved", False, False), ("blocks", True, False), ("one_rank", False, False), ("blocks", False, True)])
def test_dedup_extend_multi_gpu_one_process(kind, mark, ties):
    """hsq_dedup_extend_multi_gpu: one process, one context per GPU, the exchange inside the library."""
    import torch
    from bioinformatics_b200._lib import Context, dedup_extend_multi_gpu
    world = min(torch.cuda.device_count(), 4)
    probes, pairs = _workload(ties)
    exp = _expected(probes, pairs, mark)
    parts = _partition(pairs.n, world, kind)
    ctxs = [Context(r) for r in range(world)]
    try:
        for c in ctxs:
            c.set_probes(probes)
        outs = dedup_extend_multi_gpu(ctxs, [pairs.subset(p) for p in parts], U, V, extend_duplicates=mark, cigar_cap=64, md_cap=96)
        for res, mine in zip(outs, parts):
            _check_rank(res, mine, exp, probes, pairs, mark)
        # a second call reuses the communicator; hsq_dedup_sharded alone gives the same flags
        if world == 1:
            b, g, st = ctxs[0].dedup_sharded(pairs, U, V)
            assert np.array_equal(b, exp[0]) and np.array_equal(g, exp[1]) and np.array_equal(st.unique_pairs, exp[2].unique_pairs)
    finally:
        for c in ctxs:
            c.close()


def _worker(rank, world, port, kind, mark, ret):
    import torch.distributed as dist
    from bioinformatics_b200._lib import Context
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)  # only carries the communicator id
    try:
        probes, pairs = _workload()
        mine = _partition(pairs.n, world, kind)[rank]
        with Context(rank) as ctx:
            ctx.set_probes(probes)
            ctx.comm_init_torch()
            res = ctx.dedup_extend_sharded(pairs.subset(mine), U, V, extend_duplicates=mark, cigar_cap=64, md_cap=96)
            b, g, st = ctx.dedup_sharded(pairs.subset(mine), U, V)
            assert np.array_equal(b, res.is_best) and np.array_equal(g, res.group_size)
            # the keys-first form of the same collective call on packed reads: same flags, same slots, same positions
            from bioinformatics_b200._lib import pack_reads, quality_sums, uid_keys
            from bioinformatics_b200.batches import PackedPairsHost
            sub = pairs.subset(mine)
            pk = PackedPairsHost(sub.probe_id, pack_reads(sub.r1, 2), pack_reads(sub.r2, 2), None, None, sub.read_index, quality_sums(sub.q1),
                                 quality_sums(sub.q2))
            kf = ctx.dedup_extend_packed_keys(pk, uid_keys(pk.r1, pk.r2, U, V), U, V, extend_duplicates=mark, sharded=True)
            m = res.n_extended
            assert kf.n_extended == m and np.array_equal(kf.is_best, res.is_best) and np.array_equal(kf.group_size, res.group_size)
            assert np.array_equal(kf.extended_pair[:m], res.extended_pair[:m]) and np.array_equal(kf.pooled.pos[:2 * m], res.extend.pos[:2 * m])
            assert np.array_equal(kf.stats.unique_pairs, res.stats.unique_pairs)
            assert np.array_equal(kf.unique_unable_to_extend, res.unique_unable_to_extend)
        for o in (res, res.extend, res.stats):  # ctypes mirrors do not pickle
            for a in ("_c", "_stats_c", "_ext_c"):
                if hasattr(o, a):
                    setattr(o, a, None)
        ret[rank] = res
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind,mark", [("interleaved", False), ("one_rank", True)])
def test_dedup_extend_sharded_one_process_per_gpu(kind, mark):
    """hsq_comm_init_rank + hsq_dedup_extend_sharded with one process per GPU (the launch shape of bench.py under torchrun)."""
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, kind, mark, ret), nprocs=world, join=True)
    probes, pairs = _workload()
    exp = _expected(probes, pairs, mark)
    for r, mine in enumerate(_partition(pairs.n, world, kind)):
        _check_rank(ret[r], mine, exp, probes, pairs, mark)


def test_sharded_call_without_communicator_is_an_error():
    from bioinformatics_b200 import _abi as A
    from bioinformatics_b200._lib import Context, HsqError
    probes, pairs = _workload()
    with Context(0) as ctx:
        ctx.set_probes(probes)
        with pytest.raises(HsqError) as e:
            ctx.dedup_sharded(pairs.head(10), U, V)
        assert e.value.status == A.HSQ_ERR_INVALID_ARGUMENT
