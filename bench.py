#!/usr/bin/env python
"""bench.py -- the measurement contract.

Workload (BASELINE.json configs[1]): batched primer x read alignment microbench, 10 M read ends (5 M pairs,
2 x 150 bp, 10 bp UID) against their matched probes (1 k probes), per GPU.  One step = one pass of the hot path
(ExtendReadsToPrimer: primer DP + capture-target DP + traceback + CIGAR/MD) over that batch.

  value  : GCUPS with the batch already resident in HBM (hsq_extend_run: encode + DP kernels), CUDA events.
  e2e    : GCUPS through hsq_extend_batch with pinned HOST buffers (H2D + kernels + D2H inside the timed region).
  --impl reference : the CPU restatement of the reference algorithm (oracle/, "port"; the Java reference cannot run
                     here: no JVM in the image) on all host threads, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "primer-alignment GCUPS (ExtendReadsToPrimer DP cells/s, full matrices)"
N_PROBES = 1000
READ_LEN = 150
UID = 10
CIGAR_CAP, MD_CAP = 24, 40
OPS_PER_CELL = 12  # algorithmic INT32 ops per DP cell
TRAFFIC_PER_PAIR_BYTES = (4.690506e9 + 6.299373e9) / 500000  # measured DRAM traffic of capture2_kernel per read pair (ncu --set full, r2_h)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pairs", type=int, default=5_000_000, help="read pairs per GPU (10 M read ends)")
    ap.add_argument("--cpu-sample-pairs", type=int, default=0, help="0 = size the CPU sample for ~15 s")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--stage-base-pairs", type=int, default=1_250_000, help="generated base batch of the stage shapes (tiled up to c3 / c4 size)")
    ap.add_argument("--c3-pairs", type=int, default=10_000_000, help="configs[2]: pairs of the dedup shape (N = 1)")
    ap.add_argument("--c4-pairs", type=int, default=12_500_000, help="configs[3]: pairs per GPU of the whole-exome shape")
    ap.add_argument("--stage-cpu-pairs", type=int, default=500_000, help="pairs of the CPU-port sample of the stages")
    ap.add_argument("--no-c3", action="store_true")
    ap.add_argument("--no-c4", action="store_true")
    ap.add_argument("--no-stages", action="store_true", help="skip the trim/dedup side measurements")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []   # (arrival time, line)
        self.t_mark = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def wait_first(self, timeout=5.0):
        """nvidia-smi needs a few hundred ms to come up (more on an 8-GPU box): start it before the warm-up and wait for a sample."""
        t0 = time.perf_counter()
        while self.proc and not self.lines and time.perf_counter() - t0 < timeout:
            time.sleep(0.02)

    def mark(self):
        """The timed region starts now: samples that arrive from here on are the ones reported."""
        self.t_mark = time.perf_counter()

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        if self.t_mark is not None and not any(t >= self.t_mark for t, _ in self.lines):
            t0 = time.perf_counter()  # region shorter than one sampling period: take the sample that is about to arrive as well
            n0 = len(self.lines)
            while len(self.lines) == n0 and time.perf_counter() - t0 < 0.3:
                time.sleep(0.01)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [ln for t, ln in self.lines if self.t_mark is None or t >= self.t_mark]
        note = None
        if self.t_mark is not None and len(inside) <= 1 and len(self.lines) >= 2 and not any(t >= self.t_mark for t, _ in self.lines[:-1]):
            inside = [ln for _, ln in self.lines[-2:]]  # region shorter than the 50 ms sampling period: the samples on either side of it
            note = "timed region shorter than the sampling period: the samples just before and just after it"
        for ln in inside:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        out = {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
               "reasons": sorted(reasons), "samples": len(sm)}
        if note:
            out["note"] = note
        return out


def make_workload(n_pairs, rank, alloc=None, base_pairs=1_250_000):
    """configs[1]'s batch.  Generation costs ~10 s per million pairs, so batches above base_pairs are tiles of one generated base with
    the first two UID bases rotated per tile (synth.tile_pairs): same probes, same noise model, different molecules."""
    from bioinformatics_b200 import synth
    probes = synth.make_probes(N_PROBES)
    tiles = max(1, min(16, -(-n_pairs // base_pairs)))
    base = synth.make_pairs(probes, -(-n_pairs // tiles), read_len=READ_LEN, ext_uid_len=UID, seed=synth.SEED_READS + rank, with_quals=False)
    pairs = synth.tile_pairs(base, tiles, alloc) if (tiles > 1 or alloc) else base
    return probes, pairs


def cpu_baseline(probes, pairs, sample_pairs, threads, keep=None):
    """Oracle (port of the reference algorithm) timed on the host cores on a bounded sample of the workload."""
    from oracle import oracle as O
    O.build()
    if sample_pairs <= 0:
        probe_n = min(pairs.n, 4000)
        t = time.perf_counter()
        r = O.extend_batch(probes, pairs.head(probe_n), nthreads=threads)
        rate = r.cells / (time.perf_counter() - t)
        per_pair = r.cells / probe_n
        sample_pairs = int(max(probe_n, min(pairs.n, 15.0 * rate / per_pair)))
    sub = pairs.head(sample_pairs)
    t = time.perf_counter()
    r = O.extend_batch(probes, sub, nthreads=threads)
    dt = time.perf_counter() - t
    if keep is not None:
        keep.append(r)
    return {"value": r.cells / dt / 1e9, "unit": "GCUPS", "cores": threads, "kind": "port",
            "sample": "%d read pairs (%d read ends, %.3e cells) of the same synthetic batch, %.1f s" %
                      (sample_pairs, 2 * sample_pairs, r.cells, dt),
            "pairs_per_s": sample_pairs / dt}, dt, r.cells


HBM_BYTES = {  # algorithmic bytes per pair of the HBM-bound stage kernels (L = read length)
    "trim": lambda L: 8 * L,                 # 4L read (2 mates x bases + qualities) + <= 4L written
    "keygen": lambda L: 2 * L + 16 + 24,     # 2L qualities + UID bases read, one 24 B key record written
    "group": lambda L: 48 + 25,              # sort/group: 2 x 24 B per record moved once; reduce: 24 B read + 1 B written
}


def _hbm_peak():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (burst, torch copy)"
    except Exception:
        return 6554.9, "B200_PROFILING.md fallback (no MEASURED_PEAKS.json on this box)"


def _frac(bytes_per_pair, n, ms, peak_gbs):
    if not ms or ms <= 0:
        return None
    return bytes_per_pair * n / (ms * 1e-3) / 1e9 / peak_gbs


def stage_numbers(ctx, args, rank, world, dist, barrier, max_over_ranks):
    """BASELINE.json's metric leads with "read-pairs/sec trim+dedup": this leg measures that path at the shapes the metric is quoted
    on, through the C ABI with pinned host buffers, outside the headline's timed region.
      c3 (N = 1): `HSQUtils dedup` shape, 10 M pairs, 14 bp UID, 10 k probes: hsq_dedup_batch / hsq_dedup_packed / hsq_dedup_extend_batch.
      c4 (every N): whole-exome shape, 12.5 M pairs per GPU on a 200 k-probe panel: trim -> dedup (sharded inside the library over
        NCCL at N > 1: hsq_dedup_sharded / hsq_dedup_extend_sharded) -> extension of the unique pairs (all pairs with --markDuplicates).
    Each stage: read pairs/s host to host, kernel-only CUDA-event spans with their fraction of the HBM roofline (hsq_last_spans), the
    CPU port timed on a bounded sample at N = 1, and at N > 1 the sharded answer checked against the oracle on a gathered sample."""
    import torch
    from bioinformatics_b200 import synth
    from bioinformatics_b200._lib import pack_reads, quality_sums
    from bioinformatics_b200.batches import DedupExtendResult, ExtendResult, PackedPairsHost
    hbm, hbm_src = _hbm_peak()
    out = {"hbm_peak_gbs": hbm, "hbm_peak_source": hbm_src, "host_buffers": "pinned"}
    base_n = max(1000, args.stage_base_pairs)

    def pinned_result(n, n_probes):
        res = DedupExtendResult(n, n_probes, CIGAR_CAP, MD_CAP)
        for f in ExtendResult.FIELDS_I32 + ("ok", "overflow", "pair_ok", "cigar", "md"):
            a = getattr(res.extend, f)
            setattr(res.extend, f, ctx.pinned_array(a.nbytes).view(a.dtype))
        for f in ("is_best", "group_size", "extended_pair"):
            a = getattr(res, f)
            setattr(res, f, ctx.pinned_array(a.nbytes).view(a.dtype))
        return res

    def timed(fn, sync_ranks=True):
        """One warm-up call, one timed call; host-to-host seconds (max over ranks when the call is collective)."""
        fn()
        barrier()
        t0 = time.perf_counter()
        r = fn()
        dt = time.perf_counter() - t0
        if sync_ranks:
            barrier()
            dt = max_over_ranks(dt)
        return dt, r

    def dedup_entry(n, dt, spans, extra=None):
        e = {"read_pairs_per_s": n * world / dt, "ms_host_to_host": dt * 1e3, "keygen_kernel_ms": spans[0], "group_kernels_ms": spans[1],
             "keygen_hbm_frac": _frac(HBM_BYTES["keygen"](READ_LEN), n, spans[0], hbm), "group_hbm_frac": _frac(HBM_BYTES["group"](READ_LEN), n, spans[1], hbm),
             "algorithmic_bytes_per_pair": {"keygen": HBM_BYTES["keygen"](READ_LEN), "group": HBM_BYTES["group"](READ_LEN)}}
        e.update(extra or {})
        return e

    # ---------------- c3: HSQUtils dedup, 10 M pairs, 14 bp UID, 10 k probes (one GPU) ----------------
    if world == 1 and not args.no_c3:
        U3, NP3 = 14, 10000
        probes3 = synth.make_probes(NP3)
        base3, _ = synth.drop_n_pairs(synth.make_pairs(probes3, base_n, read_len=READ_LEN, ext_uid_len=U3, seed=synth.SEED_READS + 300))
        tiles = max(1, min(16, round(args.c3_pairs / base3.n)))
        pairs3 = synth.tile_pairs(base3, tiles, ctx.pinned_array)
        n3 = pairs3.n
        c3 = {"workload": "configs[2]: %d read pairs (%d tiles of a %d-pair base, N pairs dropped), 2x%dbp, %dbp UID, %d probes" %
                          (n3, tiles, base3.n, READ_LEN, U3, NP3), "pairs": int(n3)}
        dt, (best, gs, st) = timed(lambda: ctx.dedup_batch(pairs3, U3, 0, NP3))
        c3["dedup"] = dedup_entry(n3, dt, ctx.last_spans(), {"api": "hsq_dedup_batch (ASCII reads + qualities: %d B/pair over PCIe)" % (4 * READ_LEN + 24),
                                                              "unique_pairs": int(best.sum())})
        pk = PackedPairsHost(pairs3.probe_id, pack_reads(pairs3.r1, 2, ctx.pinned_array), pack_reads(pairs3.r2, 2, ctx.pinned_array),
                             pairs3.q1.data, pairs3.q2.data, pairs3.read_index)
        dt, (b2, g2, _) = timed(lambda: ctx.dedup_packed(pk, U3, 0, NP3))
        same_packed = bool(np.array_equal(b2, best) and np.array_equal(g2, gs))
        c3["dedup_packed"] = dedup_entry(n3, dt, ctx.last_spans(), {"api": "hsq_dedup_packed (reads 2 bit/base + quality bytes: ~%d B/pair over PCIe)" %
                                                                            (2 * READ_LEN + 2 * 40 + 16), "equals_hsq_dedup_batch": same_packed})
        t0 = time.perf_counter()
        qs1, qs2 = quality_sums(pairs3.q1), quality_sums(pairs3.q2)
        qs_s = time.perf_counter() - t0
        pks = PackedPairsHost(pairs3.probe_id, pk.r1, pk.r2, None, None, pairs3.read_index, qs1, qs2)
        dt, (b3, g3, _) = timed(lambda: ctx.dedup_packed(pks, U3, 0, NP3))
        c3["dedup_packed_host_quality_sums"] = dedup_entry(n3, dt, ctx.last_spans(), {
            "api": "hsq_dedup_packed with hsq_quality_sums made on the host (opt-in: 2 B per read instead of the quality string, ~%d B/pair over PCIe)" % (2 * 40 + 20),
            "equals_hsq_dedup_batch": bool(np.array_equal(b3, best) and np.array_equal(g3, gs)), "host_quality_sums_s": qs_s})
        ctx.set_probes(probes3)
        res3 = pinned_result(n3, NP3)
        for mark in (False, True):
            dt, r = timed(lambda: ctx.dedup_extend_batch(pairs3, U3, 0, extend_duplicates=mark, cigar_cap=CIGAR_CAP, md_cap=MD_CAP, out=res3))
            sp = ctx.last_spans()
            c3["dedup_extend_mark_duplicates" if mark else "dedup_extend"] = {
                "read_pairs_per_s": n3 / dt, "ms_host_to_host": dt * 1e3, "pairs_extended": int(r.n_extended),
                "pairs_extended_ok": int(r.extend.pair_ok[:r.n_extended].sum()), "keygen_kernel_ms": sp[0], "group_kernels_ms": sp[1],
                "compaction_kernels_ms": sp[2], "extend_kernels_ms": sp[3], "gcups_extend_kernels": r.extend.cells / (sp[3] * 1e-3) / 1e9 if sp[3] else None,
                "api": "hsq_dedup_extend_batch"}
        same_fused = bool(np.array_equal(res3.is_best, best) and int(best.sum()) == int(st.unique_pairs.sum()))
        # CPU port on a bounded sample + GPU == oracle on that sample
        from oracle import oracle as O
        O.build()
        sample = base3.head(min(base3.n, args.stage_cpu_pairs))
        t0 = time.perf_counter()
        eb, egs, est = O.dedup_batch(sample, U3, 0, NP3)
        cpu_dedup = time.perf_counter() - t0
        uniq = sample.subset(np.nonzero(eb)[0])
        threads = os.cpu_count() or 1
        t0 = time.perf_counter()
        eext = O.extend_batch(probes3, uniq.head(min(uniq.n, 60000)), nthreads=threads)
        cpu_ext = (time.perf_counter() - t0) * uniq.n / max(1, min(uniq.n, 60000))
        gb, gg, gst = ctx.dedup_batch(sample, U3, 0, NP3)
        c3["cpu_port"] = {"dedup_read_pairs_per_s": sample.n / cpu_dedup, "dedup_extend_read_pairs_per_s": sample.n / (cpu_dedup + cpu_ext),
                          "cores": {"dedup": 1, "extend": threads}, "kind": "port",
                          "sample": "%d pairs of the base batch: FilterByUid restatement on 1 thread (%.1f s); extension of its unique pairs timed on %d of "
                                    "them with %d threads and scaled" % (sample.n, cpu_dedup, min(uniq.n, 60000), threads)}
        c3["parity"] = {"gpu_equals_cpu_oracle_on_sample": bool(np.array_equal(gb, eb) and np.array_equal(gg, egs) and
                                                                 np.array_equal(gst.unique_pairs, est.unique_pairs) and np.array_equal(gst.duplicate_pairs, est.duplicate_pairs)),
                        "packed_equals_ascii_at_full_size": same_packed, "fused_call_equals_dedup_at_full_size": same_fused,
                        "unique_equals_sum_of_probe_counters": bool(int(best.sum()) == int(st.unique_pairs.sum())),
                        "tiles_have_equal_unique_counts": bool(len(set(best.reshape(tiles, -1).sum(axis=1).tolist())) == 1)}
        assert all(c3["parity"].values()), "c3 parity failed: %s" % c3["parity"]
        out["c3"] = c3
        out["_c3_state"] = (probes3, base3, res3)
        del pairs3, pk, pks

    # ---------------- c4: 12.5 M pairs per GPU on the 200 k-probe panel: trim -> dedup -> extend ----------------
    if not args.no_c4:
        U4, NP4 = UID, 200000
        probes4 = synth.make_probes(NP4)
        base4, _ = synth.drop_n_pairs(synth.make_pairs(probes4, base_n, read_len=READ_LEN, ext_uid_len=U4, seed=synth.SEED_READS + 400 + rank))
        tiles = max(1, min(16, round(args.c4_pairs / base4.n)))
        pairs4 = synth.tile_pairs(base4, tiles, ctx.pinned_array)
        pairs4.read_index = (pairs4.read_index.astype(np.int64) % (1 << 27) + rank * (1 << 27)).astype(np.int32)  # FASTQ ordinals are global
        ri = ctx.pinned_array(pairs4.read_index.nbytes).view(np.int32)
        ri[:] = pairs4.read_index
        pairs4.read_index = ri
        n4 = pairs4.n
        c4 = {"workload": "configs[3] per GPU: %d read pairs (%d tiles of a %d-pair base), 2x%dbp, %dbp UID, %d probes; %d GPUs = %d pairs" %
                          (n4, tiles, base4.n, READ_LEN, U4, NP4, world, n4 * world), "pairs_per_gpu": int(n4)}
        ctx.set_probes(probes4)
        if world > 1:
            ctx.comm_init_torch()
        # trim: both mates of the first tiles (its rate does not depend on the batch beyond ~1 M reads), offsets as HSQUtils trim derives them
        nt = min(n4, 2 * base4.n)
        head = pairs4.head(nt)
        first, last = U4 + 22, READ_LEN - 23
        trim_out = (ctx.pinned_array(head.r1.data.nbytes), ctx.pinned_array(head.r1.data.nbytes), ctx.pinned_array(8 * (nt + 1)).view(np.int64))
        spans = []

        def trim_both():
            spans.clear()
            for sb, qb in ((head.r1, head.q1), (head.r2, head.q2)):
                ctx.trim_batch(sb, qb, first, last, True, out=trim_out)
                sp = ctx.last_spans()
                spans.append(sp[0] + (sp[1] if sp[1] is not None else 0.0))
        ctx.configure(trim_chunks=1)   # one chunk: copy in, kernels, copy out in sequence -> kernel-only spans (lengths + scan, copy kernel)
        dt1, _ = timed(trim_both)
        k_ms = sum(spans)
        ctx.configure()                # automatic: the chunk pipeline overlaps the two PCIe directions
        dt, _ = timed(trim_both)
        c4["trim"] = {"read_pairs_per_s": nt * world / dt, "ms_host_to_host": dt * 1e3, "ms_host_to_host_one_chunk": dt1 * 1e3, "pairs": int(nt),
                      "kernel_ms": k_ms, "hbm_frac": _frac(HBM_BYTES["trim"](READ_LEN), nt, k_ms, hbm),
                      "algorithmic_bytes_per_pair": HBM_BYTES["trim"](READ_LEN),
                      "api": "hsq_trim_batch on both mates (copy-in | kernel | copy-out pipelined in chunks; kernel_ms from a one-chunk run)"}
        trim_ms_per_pair = dt * 1e3 / nt
        # dedup alone
        if world == 1:
            dt, (best, gs, st) = timed(lambda: ctx.dedup_batch(pairs4, U4, 0, NP4))
            mode = "hsq_dedup_batch"
        else:
            dt, (best, gs, st) = timed(lambda: ctx.dedup_sharded(pairs4, U4, 0, NP4))
            mode = "hsq_dedup_sharded: key records bucketed by owner on the device, ncclAllGather of counts, ncclSend/ncclRecv all-to-all, owner reduces, answers back"
        c4["dedup"] = dedup_entry(n4, dt, ctx.last_spans(), {"api": mode, "unique_pairs_this_rank": int(best.sum()),
                                                              "unique_pairs_all_ranks": int(st.unique_pairs.sum())})
        res4 = pinned_result(n4, NP4)
        for mark in (False, True):
            if world == 1:
                fn = lambda: ctx.dedup_extend_batch(pairs4, U4, 0, extend_duplicates=mark, cigar_cap=CIGAR_CAP, md_cap=MD_CAP, out=res4)
                mode = "hsq_dedup_extend_batch"
            else:
                fn = lambda: ctx.dedup_extend_sharded(pairs4, U4, 0, extend_duplicates=mark, cigar_cap=CIGAR_CAP, md_cap=MD_CAP, out=res4)
                mode = "hsq_dedup_extend_sharded (exchange inside the library over NCCL), extension on each rank's own pairs"
            dt, r = timed(fn)
            sp = ctx.last_spans()
            key = "dedup_extend_mark_duplicates" if mark else "dedup_extend"
            c4[key] = {"read_pairs_per_s": n4 * world / dt, "ms_host_to_host": dt * 1e3, "pairs_extended_this_rank": int(r.n_extended),
                       "pairs_extended_ok_this_rank": int(r.extend.pair_ok[:r.n_extended].sum()), "keygen_kernel_ms": sp[0], "group_or_exchange_ms": sp[1],
                       "compaction_kernels_ms": sp[2], "extend_kernels_ms": sp[3], "api": mode}
            ms = trim_ms_per_pair * n4 + dt * 1e3
            c4["trim_plus_" + key] = {"read_pairs_per_s": n4 * world / (ms * 1e-3), "ms_host_to_host": ms,
                                      "note": "the two HSQUtils commands back to back: trim of both mates (rate of the trim entry above applied to the whole "
                                              "batch) + the probe task; the mapper that runs between them in a real run is not part of this path"}
        # the same probe task in the N1 form: reads 2 bit/base + host-made quality sums in, integer columns + CIGAR/MD pool out
        from bioinformatics_b200.batches import PooledExtendResult
        t0 = time.perf_counter()
        pk4 = PackedPairsHost(pairs4.probe_id, pack_reads(pairs4.r1, 2, ctx.pinned_array), pack_reads(pairs4.r2, 2, ctx.pinned_array), None, None,
                              pairs4.read_index, quality_sums(pairs4.q1), quality_sums(pairs4.q2))
        pack_s = time.perf_counter() - t0
        POOLED_FIELDS = ("ok", "pair_ok", "pos", "read_start", "read_stop", "nm", "cigar_len", "md_len", "str_off")
        resp = DedupExtendResult(n4, NP4, pooled=PooledExtendResult(n4, 2 * n4 * 24, fields=POOLED_FIELDS, alloc=ctx.pinned_array))
        for f in ("is_best", "group_size", "extended_pair"):
            a = getattr(resp, f)
            setattr(resp, f, ctx.pinned_array(a.nbytes).view(a.dtype))
        packed_same = True
        for mark in (False, True):
            dt, r = timed(lambda: ctx.dedup_extend_packed(pk4, U4, 0, extend_duplicates=mark, sharded=world > 1, out=resp))
            sp = ctx.last_spans()
            key = "dedup_extend_packed_mark_duplicates" if mark else "dedup_extend_packed"
            h2d_b = pk4.r1.nbytes + pk4.r2.nbytes + 4 * n4 * 2 + 2 * n4 * 2
            d2h_b = resp.pooled.nbytes() * r.n_extended // max(n4, 1) + 5 * n4 + 4 * r.n_extended
            c4[key] = {"read_pairs_per_s": n4 * world / dt, "ms_host_to_host": dt * 1e3, "pairs_extended_this_rank": int(r.n_extended),
                       "keygen_kernel_ms": sp[0], "group_or_exchange_ms": sp[1], "compaction_kernels_ms": sp[2], "extend_kernels_ms": sp[3],
                       "h2d_bytes_per_pair": h2d_b / n4, "d2h_bytes_per_pair": d2h_b / n4,
                       "api": "hsq_dedup_extend_packed (sharded=%d): reads 2 bit/base + hsq_quality_sums in, columns %s + CIGAR/MD pool out" % (world > 1, POOLED_FIELDS)}
            ms = trim_ms_per_pair * n4 + dt * 1e3
            c4["trim_plus_" + key] = {"read_pairs_per_s": n4 * world / (ms * 1e-3), "ms_host_to_host": ms}
            if mark:  # res4 still holds the ASCII call's mark-duplicates run: same flags, same positions
                m = r.n_extended
                packed_same = bool(packed_same and m == res4.n_extended and np.array_equal(resp.is_best, res4.is_best) and
                                   np.array_equal(resp.pooled.pos[:2 * m], res4.extend.pos[:2 * m]) and
                                   np.array_equal(resp.pooled.cigar_len[:2 * m], res4.extend.cigar_len[:2 * m]) and
                                   np.array_equal(resp.pooled.pair_ok[:m], res4.extend.pair_ok[:m]))
        if world == 1:
            # Two batches in flight on two contexts of the same GPU (one host thread each): a call is copy in -> kernels -> copy out in
            # sequence, so a second context's copies run under the first one's kernels.  This is how a host with more than one batch
            # queued (HsqBatcher: one flusher per context) is meant to drive a GPU.  Each context gets the first half of the batch.
            import threading
            from bioinformatics_b200._lib import Context as _Ctx
            nh = n4 // 2
            half = pairs4.head(nh)
            pkh = PackedPairsHost(half.probe_id, pack_reads(half.r1, 2, ctx.pinned_array), pack_reads(half.r2, 2, ctx.pinned_array), None, None,
                                  half.read_index, quality_sums(half.q1), quality_sums(half.q2))
            ctx_b = _Ctx(rank)
            ctx_b.set_probes(probes4)
            outs = []
            for c_ in (ctx, ctx_b):
                o = DedupExtendResult(nh, NP4, pooled=PooledExtendResult(nh, 2 * nh * 24, fields=POOLED_FIELDS, alloc=c_.pinned_array))
                for f in ("is_best", "group_size", "extended_pair"):
                    a = getattr(o, f)
                    setattr(o, f, c_.pinned_array(a.nbytes).view(a.dtype))
                outs.append(o)

            CALLS = 4  # per context, back to back: after the first call the two contexts run out of step (copies under kernels)

            def run_calls(c_, o):
                for _ in range(CALLS):
                    c_.dedup_extend_packed(pkh, U4, 0, extend_duplicates=False, out=o)

            def both():
                th = [threading.Thread(target=run_calls, args=(c_, o)) for c_, o in zip((ctx, ctx_b), outs)]
                for t in th:
                    t.start()
                for t in th:
                    t.join()
            dt1, r1_ = timed(lambda: run_calls(ctx, outs[0]))
            dt1 /= CALLS
            dt2, _ = timed(both)
            dt2 /= CALLS
            m_ = outs[0].n_extended
            same2 = bool(outs[1].n_extended == m_ and np.array_equal(outs[0].is_best, outs[1].is_best) and
                         np.array_equal(outs[0].pooled.pos[:2 * m_], outs[1].pooled.pos[:2 * m_]) and
                         np.array_equal(outs[0].pooled.str_off[:2 * m_], outs[1].pooled.str_off[:2 * m_]))
            c4["dedup_extend_packed_two_contexts"] = {
                "read_pairs_per_s": 2 * nh / dt2, "ms_host_to_host": dt2 * 1e3, "pairs_per_context": int(nh),
                "one_context_same_batch_read_pairs_per_s": nh / dt1, "both_contexts_equal": same2,
                "calls_per_context": CALLS,
                "api": "hsq_dedup_extend_packed on two contexts of one GPU from two host threads, half-size batches back to back: the copies of "
                       "one context's call run under the kernels of the other's"}
            del pkh, outs
            ctx_b.close()
        # keys first: the reduction from 20 bytes per pair, then the reads in chunks under the kernels (same results)
        from bioinformatics_b200._lib import uid_keys as _uid_keys
        t0 = time.perf_counter()
        keys4 = _uid_keys(pk4.r1, pk4.r2, U4, 0)
        keys_s = time.perf_counter() - t0
        pk_keys = ctx.pinned_array(keys4.nbytes).view(np.uint64)
        pk_keys[:] = keys4
        for nm_ in ("qsum1", "qsum2"):
            a_ = getattr(pk4, nm_)
            p_ = ctx.pinned_array(a_.nbytes).view(a_.dtype)
            p_[:] = a_
            setattr(pk4, nm_, p_)
        resk = DedupExtendResult(n4, NP4, pooled=PooledExtendResult(n4, 2 * n4 * 24, fields=POOLED_FIELDS, alloc=ctx.pinned_array))
        for f in ("is_best", "group_size", "extended_pair"):
            a = getattr(resk, f)
            setattr(resk, f, ctx.pinned_array(a.nbytes).view(a.dtype))
        keys_same = True
        for mark in (False, True):
            dt, r = timed(lambda: ctx.dedup_extend_packed_keys(pk4, pk_keys, U4, 0, extend_duplicates=mark, sharded=world > 1, out=resk))
            key = "dedup_extend_packed_keys_mark_duplicates" if mark else "dedup_extend_packed_keys"
            c4[key] = {"read_pairs_per_s": n4 * world / dt, "ms_host_to_host": dt * 1e3, "pairs_extended_this_rank": int(r.n_extended),
                       "api": "hsq_dedup_extend_packed_keys (sharded=%d): uid keys + quality sums + probe + read index first (20 B/pair), reduction, "
                              "then the packed reads in chunks copied in under the extension of the previous chunk" % (world > 1)}
            ms = trim_ms_per_pair * n4 + dt * 1e3
            c4["trim_plus_" + key] = {"read_pairs_per_s": n4 * world / (ms * 1e-3), "ms_host_to_host": ms}
            ctx.dedup_extend_packed(pk4, U4, 0, extend_duplicates=mark, sharded=world > 1, out=resp)
            m = resp.n_extended
            keys_same = bool(keys_same and r.n_extended == m and np.array_equal(resk.is_best, resp.is_best) and
                             np.array_equal(resk.extended_pair[:m], resp.extended_pair[:m]) and
                             all(np.array_equal(getattr(resk.pooled, f)[:2 * m], getattr(resp.pooled, f)[:2 * m])
                                 for f in ("pos", "read_start", "read_stop", "nm", "cigar_len", "md_len", "ok")) and
                             np.array_equal(resk.pooled.str_off[:2 * m], resp.pooled.str_off[:2 * m]) and
                             resk.pooled.pool_bytes == resp.pooled.pool_bytes and
                             np.array_equal(resk.pooled.pool[:resk.pooled.pool_bytes], resp.pooled.pool[:resp.pooled.pool_bytes]))
        c4["uid_keys_host_s"] = keys_s
        del resk
        c4["packing"] = {"host_s": pack_s, "note": "hsq_pack_reads + hsq_quality_sums over the batch (host helpers, all cores), outside the timed calls: "
                                                     "packed reads + quality sums are the input format of this path"}
        del pk4, resp
        c4["parity"] = {"fused_flags_equal_dedup": bool(np.array_equal(res4.is_best, best)), "packed_call_equals_ascii_call": packed_same, "keys_first_call_equals_packed_call": keys_same,
                        "unique_equals_sum_of_probe_counters": bool(_sum_ranks(dist, int(best.sum())) == int(st.unique_pairs.sum()))}
        # the sharded answer against the oracle on a sample gathered to rank 0
        if world > 1:
            S = max(1, 1_000_000 // world)
            smp = pairs4.head(S)
            sb_, sg_, _ = ctx.dedup_sharded(smp, U4, 0, NP4)

            def gather(a):
                t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
                parts = [torch.empty_like(t) for _ in range(world)]
                dist.all_gather(parts, t)
                return np.concatenate([x.cpu().numpy() for x in parts])
            cols = [gather(x) for x in (smp.probe_id, smp.read_index, smp.r1.data[:S * READ_LEN], smp.r2.data[:S * READ_LEN],
                                        smp.q1.data[:S * READ_LEN], smp.q2.data[:S * READ_LEN], sb_, sg_)]
            if rank == 0:
                from oracle import oracle as O
                from bioinformatics_b200 import _abi as A
                from bioinformatics_b200.batches import PairBatchHost
                O.build()
                fx = lambda a: A.SeqBatch.from_fixed(a.reshape(-1, READ_LEN))
                union = PairBatchHost(cols[0], fx(cols[2]), fx(cols[3]), fx(cols[4]), fx(cols[5]), cols[1])
                eb, egs, _ = O.dedup_batch(union, U4, 0, NP4)
                c4["parity"]["sharded_dedup_equals_oracle"] = bool(np.array_equal(cols[6], eb) and np.array_equal(cols[7], egs))
                c4["parity"]["sharded_sample"] = "%d pairs per rank x %d ranks gathered to rank 0, FilterByUid restatement on the union" % (S, world)
        else:
            from oracle import oracle as O
            O.build()
            threads = os.cpu_count() or 1
            sample = base4.head(min(base4.n, args.stage_cpu_pairs))
            t0 = time.perf_counter()
            O.trim_batch(sample.r1, sample.q1, first, last, True)
            O.trim_batch(sample.r2, sample.q2, first, last, True)
            cpu_trim = time.perf_counter() - t0
            t0 = time.perf_counter()
            eb, egs, est = O.dedup_batch(sample, U4, 0, NP4)
            cpu_dedup = time.perf_counter() - t0
            uniq = sample.subset(np.nonzero(eb)[0])
            m = min(uniq.n, 60000)
            t0 = time.perf_counter()
            O.extend_batch(probes4, uniq.head(m), nthreads=threads)
            cpu_ext = (time.perf_counter() - t0) * uniq.n / max(1, m)
            gb, gg, gst = ctx.dedup_batch(sample, U4, 0, NP4)
            c4["cpu_port"] = {"trim_read_pairs_per_s": sample.n / cpu_trim, "dedup_read_pairs_per_s": sample.n / cpu_dedup,
                              "dedup_extend_read_pairs_per_s": sample.n / (cpu_dedup + cpu_ext),
                              "trim_plus_dedup_extend_read_pairs_per_s": sample.n / (cpu_trim + cpu_dedup + cpu_ext),
                              "cores": {"trim": 1, "dedup": 1, "extend": threads}, "kind": "port",
                              "sample": "%d pairs of the base batch; trim and FilterByUid restatements on 1 thread (%.2f s, %.2f s); extension of the unique "
                                        "pairs timed on %d of them with %d threads and scaled" % (sample.n, cpu_trim, cpu_dedup, m, threads)}
            c4["parity"]["gpu_equals_cpu_oracle_on_sample"] = bool(np.array_equal(gb, eb) and np.array_equal(gg, egs) and
                                                                   np.array_equal(gst.unique_pairs, est.unique_pairs))
        if rank == 0:
            assert all(v for v in c4["parity"].values() if isinstance(v, bool)), "c4 parity failed: %s" % c4["parity"]
        out["c4"] = c4
        del pairs4, res4
    if world == 1 and "_c3_state" in out:
        probes3, base3, res3 = out.pop("_c3_state")
        ctx.set_probes(probes3)
        sub = base3.head(min(base3.n, 1_000_000))
        for sb in (sub.r1, sub.r2, sub.q1, sub.q2):
            d = ctx.pinned_array(sb.data.nbytes)
            d[:] = sb.data
            sb.data = d
        res = ctx.dedup_extend_batch(sub, 14, 0, extend_duplicates=True, cigar_cap=CIGAR_CAP, md_cap=MD_CAP)
        out.update(next_row_numbers(ctx, probes3, sub, res))
    out.pop("_c3_state", None)
    return out


def _sum_ranks(dist, x):
    if dist is None:
        return x
    import torch
    t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t)
    return int(t.item())


def next_row_numbers(ctx, probes, pairs, res):
    """SURVEY 8(f) rows at N = 1, host buffers in, host buffers out, each with the CPU oracle timed on a bounded sample beside it:
    f1 read->probe assignment, f4 record fields, f3 candidate stage (+ the alignment filter of the listed candidates), f2 text ingest + name join."""
    from bioinformatics_b200 import synth, flow
    from bioinformatics_b200 import _abi as A
    from oracle import oracle as O
    O.build()
    out = {}

    def timed(fn, reps=2):
        fn()
        best = 1e30
        for _ in range(reps):
            t0 = time.perf_counter()
            r = fn()
            best = min(best, time.perf_counter() - t0)
        return best, r

    # f1: two records per pair, 5 % with an alternative hit on both mates
    ranges, records = synth.make_mapped_records(probes, pairs, seed=9)
    ctx.set_probe_ranges(ranges)
    dt, a = timed(lambda: ctx.assign_probes(records, 4))
    sub = min(200000, records.n - 1)
    from bioinformatics_b200.batches import MappedRecordsHost
    ao = records.alt_offsets
    head = MappedRecordsHost(records.contig[:sub], records.negative_strand[:sub], records.second_of_pair[:sub], records.start[:sub],
                             records.end[:sub], records.read_index[:sub],
                             (ao[:sub + 1], records.alt_contig[:ao[sub]], records.alt_negative[:ao[sub]], records.alt_start[:ao[sub]],
                              records.alt_stop[:ao[sub]]))
    t0 = time.perf_counter()
    O.assign_probes(ranges, head, 4)
    dt_cpu = time.perf_counter() - t0
    out["assign_probes"] = {"records_per_s": records.n / dt, "ms_host_to_host": dt * 1e3, "records": int(records.n),
                            "paired_records": int((a.mate >= 0).sum()), "cpu_port_records_per_s": sub / dt_cpu,
                            "cpu_sample": "%d records, 1 thread" % sub,
                            "algorithmic_bytes_per_record": 22 + 8 + 4 * 4 * 2 + 13}
    # f4: record fields of the pairs the fused call extended (res still holds the mark-duplicates run: every pair)
    m = res.n_extended
    written = pairs if m == pairs.n else pairs.subset(res.extended_pair[:m])
    ext = res.extend.head(m)
    dup = (1 - res.is_best[res.extended_pair[:m]]).astype(np.uint8)
    from bioinformatics_b200.batches import RecordResult
    rec = RecordResult(m, int(len(written.r1.data) + len(written.r2.data)))
    for f in ("flag", "mate_pos", "offsets", "seq", "qual"):
        q = ctx.pinned_array(getattr(rec, f).nbytes).view(getattr(rec, f).dtype)
        setattr(rec, f, q)
    dt, rec = timed(lambda: ctx.extend_records(written, ext, duplicate=dup, out=rec))
    ms_dev = ctx.last_timing()[0]
    sub = min(m, 100000)
    t0 = time.perf_counter()
    O.extend_records(probes, written.head(sub), duplicate=dup[:sub])
    dt_cpu = time.perf_counter() - t0
    out["extend_records"] = {"read_pairs_per_s": m / dt, "ms_host_to_host": dt * 1e3, "ms_device_incl_copies": ms_dev, "pairs": int(m),
                             "record_bytes": int(rec.offsets[-1]), "cpu_port_read_pairs_per_s": sub / dt_cpu,
                             "cpu_sample": "%d pairs, 1 thread; the port re-runs the extension to get the record fields" % sub}
    # f3: the primers of the panel as sequences to find, read one of every pair as the FASTQ entries
    n_find = min(probes.n, 1000)  # 2000 sequences to find, as in round 1
    seqs = A.SeqBatch.from_strings(probes.ext.strings()[:n_find] + probes.lig.strings()[:n_find])
    ctx.set_find_sequences(seqs, 5, 15000)
    dt, c = timed(lambda: ctx.find_candidates(pairs.r1, pairs.q1, 20, 20, max_candidates=16))
    sub = min(4000, pairs.n)
    t0 = time.perf_counter()
    O.find_candidates(seqs, A.SeqBatch(pairs.r1.data, pairs.r1.offsets[:sub + 1]), A.SeqBatch(pairs.q1.data, pairs.q1.offsets[:sub + 1]),
                      leading_quality=20, trailing_quality=20, max_candidates=16)
    dt_cpu = time.perf_counter() - t0
    out["find_candidates"] = {"reads_per_s": pairs.n / dt, "ms_host_to_host": dt * 1e3, "reads": int(pairs.n), "sequences_to_find": int(seqs.n),
                              "candidates": int(c.n_candidates.sum()), "overflowed_reads": int(c.overflow.sum()),
                              "cpu_port_reads_per_s": sub / dt_cpu, "cpu_sample": "%d reads, 1 thread" % sub}
    sub = min(250000, pairs.n)
    rs, qs = A.SeqBatch(pairs.r1.data, pairs.r1.offsets[:sub + 1]), A.SeqBatch(pairs.q1.data, pairs.q1.offsets[:sub + 1])
    limits = (2, 1, 1, 1, 1)
    dt, f = timed(lambda: ctx.find_matches(rs, qs, 20, 20, limits, max_candidates=64), reps=1)
    ms_dev = ctx.last_timing()[0]
    sub_cpu = min(1500, sub)
    rc_, qc_ = A.SeqBatch(pairs.r1.data, pairs.r1.offsets[:sub_cpu + 1]), A.SeqBatch(pairs.q1.data, pairs.q1.offsets[:sub_cpu + 1])

    class _Port:  # the same decision over the CPU port: candidates, then one alignment per candidate
        def find_candidates(self, reads, quals, lq, tq, max_candidates=16):
            return O.find_candidates(seqs, reads, quals, leading_quality=lq, trailing_quality=tq, max_candidates=max_candidates)

        def align_batch(self, refs, qrys, flags=0, cols_cap=640, md_cap=640):
            return O.align_batch(refs, qrys, flags=flags, cols_cap=cols_cap, md_cap=md_cap, nthreads=1)

    t0 = time.perf_counter()
    fo = flow.find_flow(_Port(), seqs, rc_, qc_, 20, 20, *limits, max_candidates=64)
    dt_cpu = time.perf_counter() - t0
    same = bool((fo["n_matches"] == f.n_matches[:sub_cpu]).all())
    out["find_matches"] = {"reads_per_s": sub / dt, "ms_host_to_host": dt * 1e3, "ms_device_incl_copies": ms_dev, "reads": sub,
                           "alignments": int(f.cand.n_candidates.sum()), "gcups": f.cells / dt / 1e9,
                           "single_match_reads": int((f.n_matches == 1).sum()), "multiple_match_reads": int((f.n_matches > 1).sum()),
                           "cpu_port_reads_per_s": sub_cpu / dt_cpu, "cpu_sample": "%d reads, 1 thread" % sub_cpu,
                           "equals_cpu_port_on_sample": same,
                           "note": "hsq_find_matches: candidates + one alignment per candidate + the five limits in one call; slot arrays "
                                   "(64 per read) returned to pageable host memory"}
    # f2: FASTQ / SAM text -> record indexes -> the FASTQ <-> SAM name join (first dedup configuration), pageable host buffers
    n_ing = 250000 if pairs.n >= 250000 else 20000
    case = synth.make_ingest_case(n_ing, read_len=150, seed=31)
    sam_text = synth.sam_text_of(case, seed=32)
    from bioinformatics_b200.batches import NameJoinHost
    f1_, s1_, f2_, s2_ = case["trim"]
    join = NameJoinHost(case["sam_names"], case["sam_flags"], case["sam_seq_len"], case["fq1_headers"], case["fq2_headers"], case["fq_seq_len"],
                        case["fq_seq_len"], r1_first=f1_, r1_stop=s1_, r2_first=f2_, r2_stop=s2_)
    ing = {"read_pairs": n_ing, "sam_lines": int(join.n), "host_buffers": "text and index arrays in pinned memory"}

    def pinned_text(t):
        a = ctx.pinned_array(len(t))
        a[:] = np.frombuffer(t, dtype=np.uint8)
        return a

    for key, gpu_fn, cpu_fn, text in (("index_fastq", ctx.index_fastq, O.index_fastq, case["fq1_text"]), ("index_sam", ctx.index_sam, O.index_sam, sam_text)):
        ptext = pinned_text(text)
        from bioinformatics_b200.batches import FastqIndexResult, SamIndexResult
        res_out = (FastqIndexResult(n_ing + 8, ctx.pinned_array) if key == "index_fastq" else SamIndexResult(join.n + 8, ctx.pinned_array))
        dt, g = timed(lambda: gpu_fn(ptext, out=res_out))
        ms_dev = ctx.last_timing()[0]
        t0 = time.perf_counter()
        c = cpu_fn(text)
        dt_cpu = time.perf_counter() - t0
        cols = g.FIELDS if key == "index_fastq" else g.columns()
        same = g.n == c.n and all(np.array_equal(getattr(g, f)[:g.n], getattr(c, f)[:c.n]) for f in cols)
        ing[key] = {"records_per_s": g.n / dt, "text_gb_per_s_host_to_host": len(text) / dt / 1e9, "ms_host_to_host": dt * 1e3,
                    "ms_device_incl_copies": ms_dev, "text_bytes": len(text), "records": int(g.n), "cpu_port_records_per_s": c.n / dt_cpu,
                    "cpu_sample": "the same text, 1 thread", "equals_cpu_port": bool(same),
                    "algorithmic_bytes_per_record": len(text) / max(g.n, 1) + (36 if key == "index_fastq" else 108)}
    # config 1 file to file: FastqReadTrimmer.trimReads on FASTQ text (both buffers pinned)
    ptext = pinned_text(case["fq1_text"])
    pout = ctx.pinned_array(len(ptext) + 2)
    dt, (g, nrec) = timed(lambda: ctx.trim_fastq(ptext, 32, 127, True, out=pout))
    ms_dev = ctx.last_timing()[0]
    t0 = time.perf_counter()
    c, _ = O.trim_fastq(ptext, 32, 127, True)
    dt_cpu = time.perf_counter() - t0
    ing["trim_fastq"] = {"records_per_s": nrec / dt, "text_gb_per_s_host_to_host": len(ptext) / dt / 1e9, "ms_host_to_host": dt * 1e3,
                         "ms_device_incl_copies": ms_dev, "text_bytes_in": len(ptext), "text_bytes_out": int(len(g)), "records": int(nrec),
                         "cpu_port_records_per_s": nrec / dt_cpu, "cpu_sample": "the same text, 1 thread", "equals_cpu_port": bool(np.array_equal(g, c)),
                         "algorithmic_bytes_per_record": (len(ptext) + len(g)) / max(nrec, 1)}
    # BASELINE.json configs[4], band axis at 2 x 150 bp (the full length x band grid: scripts/sweep_lengths.py, profiles/band_sweep_r2_*.jsonl):
    # resident GCUPS of the capture stage with the opt-in band, the share of read ends redone with full matrices, results against band 0
    from bioinformatics_b200.batches import ExtendResult
    n_band = min(pairs.n, 1_000_000)
    bp = pairs.head(n_band)
    ctx.set_probes(probes)
    band_rows, full = [], None
    for band in (0, 8, 16, 32, 64):
        ctx.configure(band=band)
        ro = ExtendResult(n_band, 32, 64)
        ctx.extend_stage(bp)
        ctx.extend_run(32, 64)
        ctx.extend_fetch(ro)
        banded, redone = ctx.band_stats()
        ctx.timer_start()
        for _ in range(3):
            ctx.extend_run(32, 64)
        ms = ctx.timer_stop() / 3
        if band == 0:
            full = ro
        same = all(np.array_equal(getattr(ro, f), getattr(full, f)) for f in ("primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen",
                                                                              "cigar_len", "md_len", "ok", "pair_ok")) and \
            ro.cigars() == full.cigars() and ro.mds() == full.mds() and ro.cells == full.cells
        band_rows.append({"band": band if band else "full", "GCUPS_resident": ro.cells / ms / 1e6, "ms": ms, "read_ends_banded": banded,
                          "read_ends_redone_in_full": redone, "fallback_rate": (redone / banded) if banded else None,
                          "equals_full_matrices": bool(same)})
    ctx.configure()
    out["band_150bp"] = {"pairs": n_band, "rows": band_rows,
                         "api": "hsq_config.band: staircase fill + score bound + in-band traceback, every other read end redone by the full kernel"}
    # f2, binary side: record index of a level-0 BGZF BAM (the reference's intermediate files, FastqAndBamFileMerger.java:131-133)
    from tests import bamio
    rngb = np.random.default_rng(13)
    n_bam = 200000 if pairs.n >= 250000 else 20000
    seq150 = "ACGT" * 37 + "AC"
    recs = [dict(name="%d" % (k // 2), flag=99 if k % 2 == 0 else 147, ref_id=int(k % 22), pos=int(rngb.integers(0, 10 ** 8)), mapq=60,
                 cigar=[("M", 150)], seq=seq150, qual=bytes([30 + k % 10] * 150), next_ref_id=int(k % 22), next_pos=int(rngb.integers(0, 10 ** 8)),
                 tlen=300, tags=b"NMC\x01") for k in range(n_bam)]
    stream, _ = bamio.bam_stream(b"@HD\tVN:1.4\n", [("chr%d" % c, 10 ** 8) for c in range(1, 23)], recs)
    bam = pinned_text(bamio.bgzf(stream))
    dt, g = timed(lambda: ctx.index_bam(bam, cap=n_bam + 8))
    ms_dev, ms_kernels = ctx.last_timing()[0], ctx.last_spans()[0]
    t0 = time.perf_counter()
    c = O.bam_index(bam.tobytes())
    dt_cpu = time.perf_counter() - t0
    gc = g.columns()
    ing["index_bam"] = {"records_per_s": g.n / dt, "file_gb_per_s_host_to_host": len(bam) / dt / 1e9, "ms_host_to_host": dt * 1e3,
                        "ms_device_incl_copies": ms_dev, "kernel_ms": ms_kernels, "file_bytes": int(len(bam)), "records": int(g.n),
                        "exact_walk": int(g.exact_walk), "cpu_port_records_per_s": c["n"] / dt_cpu,
                        "cpu_sample": "the same file, 1 thread, the oracle's Python parser (zlib + struct)",
                        "equals_cpu_port": bool(g.n == c["n"] and all(np.array_equal(v, c[k]) for k, v in gc.items())),
                        "algorithmic_bytes_per_record": 2 * len(bam) / max(g.n, 1) + 6 * 8 + 14 * 4}
    contigs = ["chr%d" % c for c in range(1, 23)]
    merged_text = pinned_text(synth.sam_text_of(case, seed=32, index_names=True))  # the merged BAM's records: names are FASTQ ordinals
    dt, g = timed(lambda: ctx.sam_mapped_records(merged_text, contigs))
    ms_dev = ctx.last_timing()[0]
    t0 = time.perf_counter()
    c = O.sam_mapped_records(merged_text, contigs)
    dt_cpu = time.perf_counter() - t0
    same = g.n == c.n and g.n_alt == c.n_alt and all(np.array_equal(getattr(g, f)[:g.n], getattr(c, f)[:c.n]) for f in g.COLUMNS) and \
        all(np.array_equal(getattr(g, f)[:g.n_alt], getattr(c, f)[:c.n_alt]) for f in g.ALT_COLUMNS)
    ing["sam_mapped_records"] = {"records_per_s": g.n / dt, "text_gb_per_s_host_to_host": len(merged_text) / dt / 1e9, "ms_host_to_host": dt * 1e3,
                                 "ms_device_incl_copies": ms_dev, "records": int(g.n), "alternative_hits": int(g.n_alt),
                                 "cpu_port_records_per_s": c.n / dt_cpu, "cpu_sample": "the same text, 1 thread", "equals_cpu_port": bool(same)}
    for f in ("sam_names", "fq1_headers", "fq2_headers"):  # name bytes from pinned memory
        b = getattr(join, f)
        setattr(join, f, A.SeqBatch(pinned_text(b.data.tobytes()), b.offsets))
    dt, g = timed(lambda: ctx.join_names(join))
    ms_dev = ctx.last_timing()[0]
    t0 = time.perf_counter()
    c = O.join_names(join)
    dt_cpu = time.perf_counter() - t0
    same = g.n_emitted == c.n_emitted and all(np.array_equal(a, b) for a, b in zip(g.emitted(), c.emitted())) and np.array_equal(g.status, c.status)
    ing["join_names"] = {"sam_records_per_s": join.n / dt, "ms_host_to_host": dt * 1e3, "ms_device_incl_copies": ms_dev,
                         "returned_records": int(g.n_emitted), "matching_pairs": int(g.total_matching_pairs),
                         "cpu_port_sam_records_per_s": join.n / dt_cpu, "cpu_sample": "the same records, 1 thread (two std::stable_sort + the merge loop)",
                         "equals_cpu_port": bool(same)}
    # f4 output side: order of the final coordinate sort for the records of the extended pairs (names = read indexes as strings)
    rng = np.random.default_rng(5)
    n_out = 1000000 if pairs.n >= 250000 else 50000
    cols = (rng.integers(0, 22, size=n_out), rng.integers(1, 10 ** 8, size=n_out), rng.choice([99, 147, 83, 163], size=n_out), np.full(n_out, 60),
            rng.integers(0, 22, size=n_out), rng.integers(1, 10 ** 8, size=n_out), rng.integers(-400, 400, size=n_out))
    names = A.SeqBatch.from_strings(["%d" % k for k in rng.integers(0, 10 ** 7, size=n_out).tolist()])
    dt, g = timed(lambda: ctx.coordinate_order(*cols, names))
    ms_dev = ctx.last_timing()[0]
    sub = min(200000, n_out)
    t0 = time.perf_counter()
    c = O.coordinate_order(*[x[:sub] for x in cols], A.SeqBatch(names.data, names.offsets[:sub + 1]))
    dt_cpu = time.perf_counter() - t0
    out["coordinate_order"] = {"records_per_s": n_out / dt, "ms_host_to_host": dt * 1e3, "ms_device_incl_copies": ms_dev, "records": n_out,
                               "cpu_port_records_per_s": sub / dt_cpu, "cpu_sample": "%d records, 1 thread (std::stable_sort)" % sub,
                               "sorted": bool((np.diff(cols[0][g].astype(np.int64) * 10 ** 9 + cols[1][g]) >= 0).all())}
    out["ingest"] = ing
    return out


def run_reference(args, out_stream):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    probes, pairs = make_workload(min(args.pairs, 400_000), 0)
    base, _, _ = cpu_baseline(probes, pairs, args.cpu_sample_pairs or 0, threads)
    sample = args.cpu_sample_pairs or int(base["sample"].split()[0])
    sub = pairs.head(min(sample, pairs.n))
    from oracle import oracle as O
    for _ in range(min(args.warmup, 1)):
        O.extend_batch(probes, sub.head(2000), nthreads=threads)
    times, cells = [], 0
    steps = max(1, min(args.steps, 3))
    for _ in range(steps):
        t = time.perf_counter()
        r = O.extend_batch(probes, sub, nthreads=threads)
        times.append(time.perf_counter() - t)
        cells = r.cells
    dt = sum(times) / len(times)
    v = cells / dt / 1e9
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "GCUPS", "n_gpus": args.gpus, "steps": steps,
            "warmup": min(args.warmup, 1), "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": "configs[1]: primer x read alignment microbench, 2x%dbp, %d probes; CPU step = bounded sample of %d pairs"
                                   % (READ_LEN, N_PROBES, sub.n), "l2": "n/a (CPU)"},
            "cpu_baseline": {"value": v, "unit": "GCUPS", "cores": threads, "kind": "port",
                             "sample": "%d read pairs per step; C++ restatement of the reference algorithm (no JVM in the image)" % sub.n},
            "e2e": {"value": v, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "read_pairs_per_s": sub.n / dt}
    print(json.dumps(line), file=out_stream, flush=True)


def claim_stdout():
    """The contract is ONE JSON line on stdout, but libraries write there too (NCCL prints its version banner to stdout).
    Everything written to fd 1 from here on goes to stderr; the returned file object is the real stdout."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


def main():
    args = parse()
    out_stream = claim_stdout()
    if args.impl == "reference":
        return run_reference(args, out_stream)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    import torch
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist_mod
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod

    from bioinformatics_b200._lib import Context
    from bioinformatics_b200.batches import ExtendResult
    from bioinformatics_b200 import _abi as A

    ctx = Context(local)
    probes, pairs = make_workload(args.pairs, rank, ctx.pinned_array)  # pinned host buffers for the e2e legs
    n = pairs.n
    ctx.set_probes(probes)

    def pin(arr):
        p = ctx.pinned_array(arr.nbytes).view(arr.dtype)
        p[:] = arr.reshape(-1)
        return p
    out = ExtendResult(n, CIGAR_CAP, MD_CAP)
    for f in ExtendResult.FIELDS_I32 + ("ok", "overflow", "pair_ok", "cigar", "md"):
        setattr(out, f, pin(getattr(out, f)))
    h2d = pairs.r1.data.nbytes + pairs.r2.data.nbytes + pairs.r1.offsets.nbytes + pairs.r2.offsets.nbytes + pairs.probe_id.nbytes
    d2h = sum(getattr(out, f).nbytes for f in ExtendResult.FIELDS_I32 + ("ok", "overflow", "pair_ok", "cigar", "md"))

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- resident leg: batch staged once, K timed passes of the kernels --------------------------------------
    ctx.extend_stage(pairs)
    launches_per_step = 0
    sampler = ClockSampler(local)
    sampler.start()
    sampler.wait_first()
    for _ in range(max(3, args.warmup)):
        ctx.extend_run(CIGAR_CAP, MD_CAP)
        launches_per_step = ctx.last_timing()[1]
    ctx.extend_fetch(out)
    cells = out.cells
    assert out.pair_ok.mean() > 0.9, "synthetic batch should extend"
    CHECK_FIELDS = ExtendResult.FIELDS_I32 + ("ok", "overflow", "pair_ok")

    def digest(o):
        import hashlib
        h = hashlib.sha256()
        for f in CHECK_FIELDS:
            h.update(np.ascontiguousarray(getattr(o, f)).tobytes())
        for txt, ln, cap in ((o.cigar, o.cigar_len, CIGAR_CAP), (o.md, o.md_len, MD_CAP)):  # slot bytes up to their length
            m = np.arange(cap)[None, :] < np.minimum(ln, cap)[:, None]
            h.update(np.where(m, np.asarray(txt).reshape(-1, cap), 0).astype(np.uint8).tobytes())
        return h.hexdigest()
    digest_resident = digest(out)
    barrier()
    sampler.mark()
    ctx.timer_start()
    for _ in range(args.steps):
        ctx.extend_run(CIGAR_CAP, MD_CAP)
    ms_total = ctx.timer_stop()
    barrier()
    clocks = sampler.stop()
    ms_step = max_over_ranks(ms_total / args.steps)
    total_cells = sum_over_ranks(float(cells))
    value = total_cells / (ms_step * 1e-3) / 1e9

    # ---- e2e leg: host buffers in, host buffers out ------------------------------------------------------------
    # (a) the N1 path (north_star subsystem 1): reads packed to 2 bit/base (+ N exceptions) in pinned buffers, results as integer
    #     columns + one CIGAR/MD pool written straight into pinned memory: hsq_extend_packed.  Packing is the host's input
    #     format (hsq_pack_reads, timed separately below), not part of the step.
    # (b) the ASCII path of round 1 (hsq_extend_batch: 1 B/base in, fixed 24 + 40 B string slots out), reported as e2e_ascii.
    from bioinformatics_b200._lib import pack_reads
    from bioinformatics_b200.batches import PackedPairsHost, PooledExtendResult
    t0 = time.perf_counter()
    packed = PackedPairsHost(pairs.probe_id, pack_reads(pairs.r1, 2, ctx.pinned_array), pack_reads(pairs.r2, 2, ctx.pinned_array))
    pack_s = time.perf_counter() - t0
    pool_cap = 2 * n * 24
    E2E_FIELDS = ("ok", "pair_ok", "pos", "read_start", "read_stop", "nm", "cigar_len", "md_len", "str_off")
    full = PooledExtendResult(n, pool_cap, alloc=ctx.pinned_array)
    ctx.extend_packed(packed, out=full)  # warm-up + parity: every column, every string and the pool layout against the resident run
    lens = np.asarray(out.cigar_len) + np.asarray(out.md_len)
    both = np.concatenate([np.asarray(out.cigar).reshape(-1, CIGAR_CAP), np.asarray(out.md).reshape(-1, MD_CAP)], axis=1)
    mask = np.concatenate([np.arange(CIGAR_CAP)[None, :] < np.asarray(out.cigar_len)[:, None],
                           np.arange(MD_CAP)[None, :] < np.asarray(out.md_len)[:, None]], axis=1)
    packed_equal = all(np.array_equal(getattr(full, f), getattr(out, f)) for f in ExtendResult.FIELDS_I32 + ("ok", "pair_ok")) and \
        full.cells == cells and full.pool_bytes == int(lens.sum()) and np.array_equal(full.pool[:full.pool_bytes], both[mask]) and \
        np.array_equal(full.str_off, np.concatenate([[0], np.cumsum(lens.astype(np.int64))[:-1]]))
    del both, mask
    slim = PooledExtendResult(n, pool_cap, fields=E2E_FIELDS, alloc=ctx.pinned_array)
    ctx.extend_packed(packed, out=slim)
    barrier()
    t0 = time.perf_counter()
    e2e_ms_dev = 0.0
    for _ in range(args.steps):
        ctx.extend_packed(packed, out=slim)
        e2e_ms_dev += ctx.last_timing()[0]
    barrier()
    e2e_wall = (time.perf_counter() - t0) / args.steps
    e2e_s = max_over_ranks(max(e2e_wall, e2e_ms_dev / args.steps * 1e-3))
    e2e_value = total_cells / e2e_s / 1e9
    h2d_packed = packed.r1.nbytes + packed.r2.nbytes + packed.probe_id.nbytes
    d2h_packed = slim.nbytes()
    packed_equal = packed_equal and all(np.array_equal(getattr(slim, f), getattr(full, f)) for f in E2E_FIELDS) and \
        np.array_equal(slim.pool[:slim.pool_bytes], full.pool[:full.pool_bytes])

    for _ in range(1):
        ctx.extend_batch(pairs, out=out)
    barrier()
    t0 = time.perf_counter()
    asc_ms_dev = 0.0
    for _ in range(args.steps):
        ctx.extend_batch(pairs, out=out)
        asc_ms_dev += ctx.last_timing()[0]
    barrier()
    asc_wall = (time.perf_counter() - t0) / args.steps
    asc_s = max_over_ranks(max(asc_wall, asc_ms_dev / args.steps * 1e-3))
    # size-independent property at the full batch size: the chunked host pipelines return what the resident run returned
    parity = {"chunked_e2e_equals_resident": bool(digest(out) == digest_resident and out.cells == cells),
              "packed_e2e_equals_resident": bool(packed_equal)}
    assert parity["chunked_e2e_equals_resident"], "hsq_extend_batch (chunked) differs from stage/run/fetch"
    assert parity["packed_e2e_equals_resident"], "hsq_extend_packed differs from stage/run/fetch"

    stages = None
    if not args.no_stages:
        stages = stage_numbers(ctx, args, rank, world, dist, barrier, max_over_ranks)
    line = None
    if rank == 0:
        # the INT32 micro-kernel runs inside its own clock-sampler window: its rate is only meaningful beside the clock it ran at
        ps = ClockSampler(local)
        ps.start()
        ps.wait_first()
        ps.mark()
        peak = ctx.int32_peak(2.0)
        peak_clocks = ps.stop()
        achieved = cells * OPS_PER_CELL / (ms_total / args.steps * 1e-3)
        sms = torch.cuda.get_device_properties(local).multi_processor_count
        sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
        issue_limit = sms * 128 * sm_hz  # 4 warp instructions per clock per SM x 32 lanes, at the SM clock sampled in the timed region
        nominal = 148 * 128 * 1.965e9
        line = {"metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
                "data": "synthetic",
                "config": {"workload": "configs[1]: batched primer x read alignment microbench, %d read ends (%d pairs) per GPU, 2x%dbp, "
                                       "%dbp UID, %d probes, default scorer 1/-4/-1/-6" % (2 * n, n, READ_LEN, UID, N_PROBES),
                           "cells_per_step_per_gpu": int(cells), "l2": "inputs (%.1f GB) larger than L2; no flush needed" % (h2d / 1e9),
                           "cigar_cap": CIGAR_CAP, "md_cap": MD_CAP},
                "e2e": {"value": e2e_value, "unit": "GCUPS", "h2d_bytes_per_step": int(h2d_packed), "d2h_bytes_per_step": int(d2h_packed),
                        "read_pairs_per_s": n * world / e2e_s, "ms_per_step": e2e_s * 1e3,
                        "api": "hsq_extend_packed: reads 2 bit/base + N exception list in pinned host memory, results = integer columns %s + one "
                               "CIGAR/MD byte pool written into pinned host memory" % (E2E_FIELDS,),
                        "pack_host_gb_per_s": (pairs.r1.data.nbytes + pairs.r2.data.nbytes) / pack_s / 1e9,
                        "pack_note": "hsq_pack_reads (host helper, all cores) over the %.2f GB of ASCII reads, once, outside the step: packed reads "
                                     "are the input format" % ((pairs.r1.data.nbytes + pairs.r2.data.nbytes) / 1e9)},
                "e2e_ascii": {"value": total_cells / asc_s / 1e9, "unit": "GCUPS", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                              "read_pairs_per_s": n * world / asc_s, "ms_per_step": asc_s * 1e3,
                              "api": "hsq_extend_batch: ASCII reads, fixed %d + %d byte CIGAR / MD slots (the round-1 path)" % (CIGAR_CAP, MD_CAP)},
                "gpu_launches": int(launches_per_step * args.steps),
                "read_pairs_per_s": n * world / (ms_step * 1e-3),
                "roofline": {"bound": "int32_alu", "achieved": achieved / 1e12, "peak": issue_limit / 1e12, "unit": "Tiop/s",
                             "frac": achieved / issue_limit,
                             "peak_source": "INT32 issue limit at the SM clock sampled during the timed region: %d SMs x 128 lanes x %.0f MHz "
                                            "(MEASURED_PEAKS.json holds no INT32 figure)" % (sms, sm_hz / 1e6),
                             "micro_kernel": {"peak": peak / 1e12, "unit": "Tiop/s", "clocks": peak_clocks,
                                              "warp_instr_per_clk_per_sm": peak / 32.0 / sms / ((peak_clocks.get("sm_mhz") or 1965.0) * 1e6),
                                              "frac_of_it": achieved / peak,
                                              "note": "hsq_int32_peak: 2 s of 4 IMAD + 4 VIADDMNMX independent chains per thread, 64 warps per SM, "
                                                      "timed inside its own clock-sampler window"},
                             "issued": {"instructions_per_cell_pair": 11, "alu_pipe": 6, "fma_pipe": 5,
                                        "note": "the packed u16x2 cell issues 6 ALU-pipe + 5 FMA-pipe instructions per TWO cells (SASS count, "
                                                "profiles/README_r2.md); the 12 algorithmic ops per cell of SURVEY 8(d) are the scalar statement "
                                                "of the recurrence, so `frac` is an algorithmic-throughput fraction, not issue-slot utilisation "
                                                "(ncu: profiles/ncu_full_capture2_kernel_r2_*.txt)"},
                             "traffic": TRAFFIC_PER_PAIR_BYTES * n,
                             "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum of capture2_kernel from the ncu --set full capture "
                                             "at 500 k pairs (profiles/ncu_full_capture2_kernel_r2_h.txt: 4.69 + 6.30 GB) scaled to this "
                                             "batch: scratch (2-bit pointers, boundary rows), not algorithmic bytes; the kernel is ALU bound",
                             "note": "12 algorithmic INT32 ops per DP cell x cells per step / CUDA-event time of hsq_extend_run "
                                     "(capture2_kernel + primer2_kernel are ~93%% of it); nominal 148x128x1.965 GHz = %.1f Tiop/s" % (nominal / 1e12)},
                "clocks": clocks, "stages": stages, "parity": parity}
    if rank == 0 and not args.no_cpu_baseline and world == 1:
        threads = os.cpu_count() or 1
        kept = []
        base, _, _ = cpu_baseline(probes, pairs, args.cpu_sample_pairs, threads, keep=kept)
        line["cpu_baseline"] = base
        # the oracle's results on the sample against the GPU's results on the same read pairs (bit-exact)
        exp, k2 = kept[0], 2 * kept[0].n
        same = all(np.array_equal(getattr(exp, f), getattr(out, f)[:(k2 if f != "pair_ok" else k2 // 2)]) for f in CHECK_FIELDS)
        for name, cap_g in (("cigar", CIGAR_CAP), ("md", MD_CAP)):
            ln = getattr(exp, name + "_len")
            cap_e = getattr(exp, name + "_cap")
            c = min(cap_e, cap_g)
            m = np.arange(c)[None, :] < np.minimum(ln, c)[:, None]
            a_e = np.asarray(getattr(exp, name)).reshape(-1, cap_e)[:, :c]
            a_g = np.asarray(getattr(out, name)).reshape(-1, cap_g)[:k2, :c]
            same = same and bool((a_e[m] == a_g[m]).all())
        parity["gpu_equals_cpu_oracle_on_sample"] = {"read_pairs": int(exp.n), "equal": bool(same)}
        assert same, "GPU results differ from the CPU oracle on the bench sample"
    elif rank == 0:
        line["cpu_baseline"] = None
    ctx.close()
    if dist:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), file=out_stream, flush=True)


if __name__ == "__main__":
    main()
