"""Writes tests/golden/*.json.

reference_kats.json  -- the reference's OWN known-answer vectors for this path, transcribed from its test sources
                        (file:line in each entry); these pin the oracle.
oracle_vectors.json  -- small seeded inputs with the oracle's outputs frozen, for the functions the reference's tests do
                        not pin: a change of the oracle or of the CUDA path shows up as a diff here.
Run from the repo root:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from bioinformatics_b200 import synth  # noqa: E402
from oracle import oracle as O  # noqa: E402
from tests import helpers as H  # noqa: E402
from tests.test_oracle_kat import MD_KATS, T2_QRY, T2_REF, T3_QRY, T3_REF, T4_REF  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def reference_kats():
    T = "bioinformatics_common/src/test/java/com/roche/sequencing/bioinformatics/common/alignment/"
    k = {"md": [], "md_trimmed_with_full_cigar": [], "other": []}
    lines = ["34-41", "44-51", "54-61", "64-72"]
    for (ref, qry, md), ln in zip(MD_KATS, lines):
        k["md"].append({"source": T + "CigarStringUtilTest.java:" + ln, "reference": ref, "query": qry, "md": md})
    k["md_trimmed_with_full_cigar"] = [
        {"source": T + "CigarStringUtilTest.java:75-84", "reference": T3_REF, "query": T3_QRY, "reverse": False,
         "md": "0^TCCCCCCCAGCTTGTGGAGCCT0C127"},
        {"source": T + "CigarStringUtilTest.java:87-96", "reference": T4_REF, "query": T3_QRY, "reverse": False,
         "md": "0^TCCCCCCCAGCTTGTGGAGCCT0C64^CAA63"},
        {"source": T + "CigarStringUtilTest.java:99-109", "reference": T2_REF[::-1], "query": T2_QRY[::-1], "reverse": True,
         "md": "0^GTCCATCGCCACTGGGATGG11C118"},
    ]
    k["other"] = [
        {"source": T + "CigarStringUtilTest.java:112-120", "what": "edit distance", "reference": "CATTCCG", "query": "CACTCCG",
         "expected": 1},
        {"source": T + "AlignmentTest.java:34-40", "what": "getIndexOfFirstMatchInReference",
         "reference": "ACTGAAATTTGGGCCCAGTCTAAANNNNTTTTT", "query": "ATTTGGGC", "expected": 6},
        {"source": "nimblegen_heatseq/src/test/java/com/roche/heatseq/process/FastqReadTrimmerTest.java:35-40", "what": "trim",
         "read": "ACGTACGTACGTACGTACGT", "qual": "ABCDEFGHIJKLMNOPQRST", "first": 5, "last": 14, "expected_length": 10},
    ]
    return k


def oracle_vectors():
    rng = np.random.default_rng(20240503)
    v = {}
    refs, qrys = H.align_cases(rng, 64, max_len=70)
    a = O.align_batch(refs, qrys)
    v["align"] = {"refs": refs, "qrys": qrys, "scorer": [1, -4, -1, -6], "flags": 0,
                  "score": a.score.tolist(), "ops": a.op_strings(), "md": a.mds(), "rmd": a.rmds(),
                  "edit_distance": a.edit_distance.tolist(), "matches": a.matches.tolist()}
    a = O.align_batch(refs, qrys, flags=1)
    v["align_penalize_start"] = {"score": a.score.tolist(), "ops": a.op_strings()}
    probes = synth.make_probes(12, seed=41)
    pairs = synth.make_pairs(probes, 240, read_len=150, ext_uid_len=10, seed=42, sub_rate=0.02, ins_rate=0.004, del_rate=0.004)
    e = O.extend_batch(probes, pairs, cigar_cap=64, md_cap=96)
    v["extend"] = {"n_probes": 12, "probe_seed": 41, "n_pairs": 240, "read_len": 150, "uid": 10, "read_seed": 42,
                   "rates": [0.02, 0.004, 0.004], "r1": pairs.r1.strings(), "r2": pairs.r2.strings(),
                   "probe_id": pairs.probe_id.tolist(), "ok": e.ok.tolist(), "primer_end": e.primer_end.tolist(),
                   "score": e.score.tolist(), "pos": e.pos.tolist(), "read_start": e.read_start.tolist(),
                   "read_stop": e.read_stop.tolist(), "nm": e.nm.tolist(), "tlen": e.tlen.tolist(), "cigar": e.cigars(),
                   "md": e.mds(), "cells": int(e.cells)}
    clean, _ = synth.drop_n_pairs(pairs)
    b, g, st = O.dedup_batch(clean, 10, 0, probes.n)
    v["dedup"] = {"note": "pairs of 'extend' without the N-containing ones", "u": 10, "v": 0, "n": int(clean.n),
                  "is_best": b.tolist(), "group_size": g.tolist(), "unique_pairs": st.unique_pairs.tolist(),
                  "max_pairs_per_uid": st.max_pairs_per_uid.tolist()}
    quads = [("ACGTTGCATGCAAGTC", "TTACGTTGCATGCAAGTCAGGCTTAACGT"), ("GGCATTACCAGTAGCATT", "ACGGCATTACCAGTAGCATTCCAGGTTAAC"),
             ("ACGTTGCATGCAAGTC", "GGGGGGGGGGGGGGGGGGGGGGGGGGGGG")]
    v["primer_end"] = [{"primer": p, "read": r, "end": O.primer_end(p, r)[0], "weighted": O.primer_end(p, r)[1]} for p, r in quads]
    return v


if __name__ == "__main__":
    O.build()
    json.dump(reference_kats(), open(os.path.join(HERE, "reference_kats.json"), "w"), indent=1)
    json.dump(oracle_vectors(), open(os.path.join(HERE, "oracle_vectors.json"), "w"), indent=0)
    print("written", os.listdir(HERE))
