"""Multi-GPU UID dedup: one process per GPU, reads never move, one exchange of 24-byte key records.

  rank r holds some read pairs (any partition of the batch).
  1. keys      : every rank turns its pairs into key records (probe, uid, rank, origin)      -- hsq_dedup_keys
  2. exchange  : a record goes to the rank that owns its probe, owner = probe mod world_size   -- all_to_all_single
  3. reduce    : the owner sorts its records and picks the best pair of every (probe, UID)     -- hsq_dedup_reduce_keys
  4. return    : (is_best, group_size) travel back to the rank that holds the reads            -- all_to_all_single
Per-probe counters stay with the owner, as in the reference where one task owns one probe
(PrimerReadExtensionAndPcrDuplicateIdentification.java:1172).  The result does not depend on the partition because
findBestData's order (quality sum descending, read index ascending; FilterByUid.java:281-297) is total.

The two compute steps are passed in as a backend: `DeviceBackend` (the CUDA library; tensors stay on the GPU and the
exchange runs over NCCL) is the product path.  tests/ supply a CPU stand-in built on the oracle so that the exchange
logic runs at world_size 2 over gloo without a GPU.
"""
import numpy as np
import torch
import torch.distributed as dist

KEY_BYTES = 24


def owner_of(probe, world):
    return probe % world


class DeviceBackend:
    """Key generation and group reduction on this rank's GPU through the C ABI (no CPU fallback)."""

    def __init__(self, ctx):
        self.ctx = ctx
        self.device = torch.device("cuda", ctx.device)

    def keys(self, pairs, u, v, n_probes, origin_base):
        out = torch.empty((pairs.n, KEY_BYTES), dtype=torch.uint8, device=self.device)
        torch.cuda.synchronize(self.device)
        self.ctx.dedup_keys(pairs, u, v, n_probes, origin_base, out.data_ptr())
        return out

    def reduce(self, keys, uid_len, n_probes):
        n = keys.shape[0]
        keys = keys.contiguous()
        best = torch.empty(n, dtype=torch.uint8, device=self.device)
        gsize = torch.empty(n, dtype=torch.int32, device=self.device)
        torch.cuda.synchronize(self.device)
        stats = self.ctx.dedup_reduce_keys(keys.data_ptr(), n, uid_len, best.data_ptr(), gsize.data_ptr(), n_probes)
        return best, gsize, stats


def _probe_column(keys):
    """int32 probe field (bytes 16..19) of an (n, 24) uint8 key tensor."""
    if keys.shape[0] == 0:
        return torch.empty(0, dtype=torch.int32, device=keys.device)
    return keys[:, 16:20].contiguous().view(torch.int32).reshape(-1)


def _exchange(t, send_counts, recv_counts, group):
    out = torch.empty((int(sum(recv_counts)),) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    dist.all_to_all_single(out, t.contiguous(), output_split_sizes=list(recv_counts), input_split_sizes=list(send_counts),
                           group=group)
    return out


def dedup_sharded(backend, pairs, ext_uid_len, lig_uid_len, n_probes, group=None):
    """Dedup of the union of every rank's `pairs`.  Returns (is_best, group_size) for this rank's pairs (numpy, in the
    order of `pairs`) and the owner's per-probe counters (meaningful for probes with owner_of(probe) == rank)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    # origin = the pair's place in the rank-major concatenation of all ranks' pairs: the last tie-break of the reduction, so the
    # answer does not depend on the partition even when read indexes repeat
    dev = backend.device if hasattr(backend, "device") else "cpu"
    n_all = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(n_all, torch.tensor([pairs.n], dtype=torch.int64, device=dev), group=group)
    origin_base = int(sum(int(c.item()) for c in n_all[:rank]))
    keys = backend.keys(pairs, ext_uid_len, lig_uid_len, n_probes, origin_base)
    n = keys.shape[0]
    dest = owner_of(_probe_column(keys).to(torch.int64), world)
    order = torch.argsort(dest, stable=True)
    send_counts = torch.bincount(dest, minlength=world)
    # counts first (world x world), then the payload
    all_counts = [torch.empty_like(send_counts) for _ in range(world)]
    dist.all_gather(all_counts, send_counts, group=group)
    recv_counts = [int(c[rank]) for c in all_counts]
    send_list = [int(c) for c in send_counts]
    recv_keys = _exchange(keys[order], send_list, recv_counts, group)
    best_o, gsize_o, stats = backend.reduce(recv_keys, ext_uid_len + lig_uid_len, n_probes)
    # answers go back along the same routes, then into the order of `pairs`
    best_back = _exchange(best_o, recv_counts, send_list, group)
    gsize_back = _exchange(gsize_o, recv_counts, send_list, group)
    best = torch.empty(n, dtype=torch.uint8, device=keys.device)
    gsize = torch.empty(n, dtype=torch.int32, device=keys.device)
    best[order] = best_back
    gsize[order] = gsize_back
    return best.cpu().numpy(), gsize.cpu().numpy(), stats


def merge_owner_stats(stats, n_probes, group=None):
    """All ranks' per-probe counters combined (each probe is counted on exactly one rank, so a sum is exact)."""
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    t = torch.from_numpy(np.stack([stats.unique_pairs, stats.duplicate_pairs, stats.max_pairs_per_uid]).astype(np.int64)).to(dev)
    dist.all_reduce(t, group=group)
    return t.cpu().numpy()
