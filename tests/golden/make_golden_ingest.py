"""Writes tests/golden/ingest_vectors.json: small seeded f2 / f4 inputs (FASTQ text, SAM text, name join, coordinate sort) with the
oracle's outputs frozen -- the reference's tests hold nothing for these steps, so a silent change of the oracle or of the
CUDA path shows up as a diff here.  Run from the repo root:  python tests/golden/make_golden_ingest.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from bioinformatics_b200 import synth  # noqa: E402
from bioinformatics_b200.batches import NameJoinHost  # noqa: E402
from oracle import oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
CONTIGS = ["chr%d" % c for c in range(1, 23)]


def vectors():
    case = synth.make_ingest_case(60, read_len=36, seed=77, trim=(4, 30, 2, 30), old_style=True)
    sam = synth.sam_text_of(case, seed=78, xa_rate=0.4)
    merged = synth.sam_text_of(case, seed=78, xa_rate=0.4, index_names=True)
    v = {"fq1_text": case["fq1_text"].decode(), "fq2_text": case["fq2_text"].decode(), "sam_text": sam.decode(), "merged_sam_text": merged.decode(),
         "contigs": CONTIGS, "trim": list(case["trim"])}
    ix = O.index_fastq(case["fq1_text"])
    v["fastq_index"] = {f: ix.head()[f].tolist() for f in ix.FIELDS}
    t, n = O.trim_fastq(case["fq1_text"], 4, 30, True)
    v["trim_fastq"] = {"first": 4, "last": 30, "three_prime": True, "text": t.tobytes().decode(), "records": n}
    sx = O.index_sam(sam)
    v["sam_index"] = {c: getattr(sx, c).tolist() for c in sx.columns()}
    v["sam_index"]["n_header_lines"] = sx.n_header_lines
    r = O.sam_mapped_records(merged, CONTIGS)
    v["sam_mapped_records"] = {c: getattr(r, c)[:r.n].tolist() for c in r.COLUMNS}
    v["sam_mapped_records"]["alt_offsets"] = r.alt_offsets[:r.n + 1].tolist()
    v["sam_mapped_records"].update({c: getattr(r, c)[:r.n_alt].tolist() for c in r.ALT_COLUMNS})
    f1, s1, f2, s2 = case["trim"]
    j = NameJoinHost(case["sam_names"], case["sam_flags"], case["sam_seq_len"], case["fq1_headers"], case["fq2_headers"], case["fq_seq_len"],
                     case["fq_seq_len"], r1_first=f1, r1_stop=s1, r2_first=f2, r2_stop=s2)
    res = O.join_names(j)
    v["join"] = {"sam_names": case["sam_names"], "sam_flags": case["sam_flags"].tolist(), "sam_seq_len": case["sam_seq_len"].tolist(),
                 "fq1_headers": case["fq1_headers"], "fq2_headers": case["fq2_headers"], "fq_seq_len": case["fq_seq_len"].tolist(),
                 "sam_record": res.emitted()[0].tolist(), "fastq_record": res.emitted()[1].tolist(), "name_len": res.emitted()[2].tolist(),
                 "status": res.status.tolist(), "total_matching_pairs": res.total_matching_pairs}
    rng = np.random.default_rng(79)
    m = 80
    cols = [rng.integers(-1, 3, m), rng.integers(1, 30, m), rng.choice([99, 147, 83, 163], m), rng.choice([0, 60], m), rng.integers(-1, 3, m),
            rng.integers(1, 30, m), rng.integers(-5, 6, m)]
    names = ["%d" % k for k in rng.integers(0, 25, m)]
    v["coordinate_order"] = {"columns": [c.tolist() for c in cols], "names": names, "order": O.coordinate_order(*cols, names).tolist()}
    return v


if __name__ == "__main__":
    with open(os.path.join(HERE, "ingest_vectors.json"), "w") as f:
        json.dump(vectors(), f, indent=0, sort_keys=True)
        f.write("\n")
    print("wrote", os.path.join(HERE, "ingest_vectors.json"))
