"""One pass over the f2 ingest calls (FASTQ index, SAM index, name join, row gather), for ncu:
   ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv python scripts/ingest_kernels.py [pairs]
Development aid; prints the CUDA-event time of each call (copies included) when run without ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bioinformatics_b200 import synth, flow
from bioinformatics_b200._lib import Context
from bioinformatics_b200.batches import NameJoinHost

n = int(sys.argv[1]) if len(sys.argv) > 1 else 250000
case = synth.make_ingest_case(n, read_len=150, seed=31)
sam_text = synth.sam_text_of(case, seed=32)
f1, s1, f2, s2 = case["trim"]
join = NameJoinHost(case["sam_names"], case["sam_flags"], case["sam_seq_len"], case["fq1_headers"], case["fq2_headers"], case["fq_seq_len"],
                    case["fq_seq_len"], r1_first=f1, r1_stop=s1, r2_first=f2, r2_stop=s2)
with Context(0) as ctx:
    def show(name):
        print("%-22s %8.3f ms (%d launches)" % ((name,) + ctx.last_timing()))
    for rep in range(2):  # first pass allocates
        ix = ctx.index_fastq(case["fq1_text"]); show("index_fastq (%d MB)" % (len(case["fq1_text"]) >> 20))
        ctx.trim_fastq(case["fq1_text"], 32, 127, True); show("trim_fastq")
        sx = ctx.index_sam(sam_text); show("index_sam (%d MB)" % (len(sam_text) >> 20))
        r = ctx.join_names(join); show("join_names (%d SAM records)" % join.n)
        ctx.gather_rows(case["fq1_text"], ix.seq_off[:ix.n], ix.seq_len[:ix.n]); show("gather_rows")
    print("returned", r.n_emitted, "pairs", r.total_matching_pairs)
