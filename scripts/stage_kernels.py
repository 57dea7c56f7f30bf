"""One pass over every stage outside the headline DP kernels (trim, dedup, N filter, assign, records, find), for ncu:
   ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv python scripts/stage_kernels.py [pairs]
Development aid; prints the CUDA-event time of each call (copies included) when run without ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bioinformatics_b200 import synth, _abi as A
from bioinformatics_b200._lib import Context

n = int(sys.argv[1]) if len(sys.argv) > 1 else 500000
U = 10
probes = synth.make_probes(1000)
pairs, _ = synth.drop_n_pairs(synth.make_pairs(probes, n, read_len=150, ext_uid_len=U))
ranges, records = synth.make_mapped_records(probes, pairs, seed=9)
seqs = A.SeqBatch.from_strings(probes.ext.strings() + probes.lig.strings())
with Context(0) as ctx:
    ctx.set_probes(probes)
    ctx.set_probe_ranges(ranges)
    ctx.set_find_sequences(seqs, 5, 15000)
    def show(name):
        print("%-22s %8.3f ms (%d launches)" % ((name,) + ctx.last_timing()))
    ctx.trim_batch(pairs.r1, pairs.q1, U + 22, 150 - 23, True); show("trim_batch (R1)")
    ctx.flag_n_pairs(pairs); show("flag_n_pairs")
    ctx.dedup_batch(pairs, U, 0, probes.n); show("dedup_batch")
    ctx.assign_probes(records, 4); show("assign_probes")
    res = ctx.dedup_extend_batch(pairs, U, 0, extend_duplicates=False, cigar_cap=24, md_cap=40); show("dedup_extend_batch")
    m = res.n_extended
    ctx.extend_records(pairs.subset(res.extended_pair[:m]), res.extend.head(m)); show("extend_records")
    ctx.find_candidates(pairs.r1, pairs.q1, 20, 20, max_candidates=16); show("find_candidates")
    k = min(n, 100000)
    ctx.find_matches(A.SeqBatch(pairs.r1.data, pairs.r1.offsets[:k + 1]), A.SeqBatch(pairs.q1.data, pairs.q1.offsets[:k + 1]), 20, 20, (2, 1, 1, 1, 1),
                     max_candidates=64); show("find_matches (100 k)")
