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
    from tests.test_merge import KATS as MERGE_KATS
    names = ["standardOneTest :30-39", "gapOneTest :41-50", "disagreementOneTest :52-61", "noOverlapOneTest :63-72", "onlyOverlapOneTest :74-83",
             "longerUpstreamOneTest :85-93"]
    k["merge"] = [{"source": "nimblegen_heatseq/src/test/java/com/roche/heatseq/process/PrimerReadExtensionAndPcrDuplicateIdentificationTest.java " + nm,
                   "negative": c[0], "upstream": c[1], "downstream": c[2], "upstream_quality": c[3], "downstream_quality": c[4],
                   "upstream_start": c[5], "downstream_start": c[6], "merged": c[7], "merged_quality": c[8]} for c, nm in zip(MERGE_KATS, names)]
    k["range_map_inputs"] = [  # printed, not asserted, by the reference (RangeMap.java:234-254); answers hand-traced from :92-153
        {"source": "nimblegen_heatseq/src/main/java/com/roche/heatseq/process/RangeMap.java:234-242 (go3)",
         "windows": [[2145, 2284], [6462, 6601], [8617, 8756], [8712, 8851]], "query": [8741, 8787], "hand_traced": [3]},
        {"source": "nimblegen_heatseq/src/main/java/com/roche/heatseq/process/RangeMap.java:244-252 (go4)",
         "windows": [[1, 10], [2, 11], [3, 12], [4, 13]], "query": [5, 7], "hand_traced": [0, 1, 2, 3]}]
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
    # SAM record fields of the extended pairs of 'extend' (first 60 pairs, with qualities; every third pair a duplicate)
    sub = pairs.head(60)
    dup = (np.arange(60) % 3 == 0).astype(np.uint8)
    rec = O.extend_records(probes, sub, duplicate=dup)
    v["records"] = {"note": "first 60 pairs of 'extend'", "q1": sub.q1.strings(), "q2": sub.q2.strings(), "duplicate": dup.tolist(),
                    "flag": rec.flag.tolist(), "mate_pos": rec.mate_pos.tolist(), "seq": rec.seqs().strings(), "qual": rec.quals().strings()}
    # coordinate read -> probe assignment on a seeded problem
    from tests.test_assign import _random_problem
    ranges, records = _random_problem(np.random.default_rng(20240504), 40, 300, 4)
    a = O.assign_probes(ranges, records, max_probes=4)
    v["assign"] = {"ranges": {"contig": ranges.contig.tolist(), "reverse_strand": ranges.reverse_strand.tolist(), "start": ranges.start.tolist(),
                              "stop": ranges.stop.tolist()},
                   "records": {"contig": records.contig.tolist(), "negative_strand": records.negative_strand.tolist(),
                               "second_of_pair": records.second_of_pair.tolist(), "start": records.start.tolist(), "end": records.end.tolist(),
                               "read_index": records.read_index.tolist(), "alt_offsets": records.alt_offsets.tolist(),
                               "alt_contig": records.alt_contig.tolist(), "alt_negative": records.alt_negative.tolist(),
                               "alt_start": records.alt_start.tolist(), "alt_stop": records.alt_stop.tolist()},
                   "max_probes": 4, "probes": [list(map(int, t)) for t in a.probe_sets()], "mate": a.mate.tolist(),
                   "common": [list(map(int, t)) for t in a.common_sets()], "overflow": a.overflow.tolist()}
    return v


if __name__ == "__main__":
    O.build()
    json.dump(reference_kats(), open(os.path.join(HERE, "reference_kats.json"), "w"), indent=1)
    json.dump(oracle_vectors(), open(os.path.join(HERE, "oracle_vectors.json"), "w"), indent=0)
    print("written", os.listdir(HERE))
