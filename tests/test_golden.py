"""Committed golden vectors (tests/golden/, generator: make_golden.py).

reference_kats.json : the reference's own known-answer vectors -> pin the oracle (CPU) and the CUDA path (GPU).
oracle_vectors.json : frozen oracle outputs for the functions no reference test pins -> regression net for both sides.
"""
import json
import os

import numpy as np
import pytest

from bioinformatics_b200 import _abi as A
from bioinformatics_b200 import synth
from bioinformatics_b200.batches import PairBatchHost
from oracle import oracle as O

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KATS = json.load(open(os.path.join(HERE, "reference_kats.json")))
VEC = json.load(open(os.path.join(HERE, "oracle_vectors.json")))


def _extend_inputs():
    e = VEC["extend"]
    probes = synth.make_probes(e["n_probes"], seed=e["probe_seed"])
    pairs = synth.make_pairs(probes, e["n_pairs"], read_len=e["read_len"], ext_uid_len=e["uid"], seed=e["read_seed"],
                             sub_rate=e["rates"][0], ins_rate=e["rates"][1], del_rate=e["rates"][2])
    # the fixture carries the reads themselves: the generator must still produce them, else the fixture is stale
    assert pairs.r1.strings() == e["r1"] and pairs.r2.strings() == e["r2"] and pairs.probe_id.tolist() == e["probe_id"]
    return probes, pairs


def _check_align(a, flags_key="align"):
    g = VEC[flags_key]
    assert a.score.tolist() == g["score"] and a.op_strings() == g["ops"]
    if "md" in g:
        assert a.mds() == g["md"] and a.rmds() == g["rmd"]
        assert a.edit_distance.tolist() == g["edit_distance"] and a.matches.tolist() == g["matches"]


def _check_extend(e):
    g = VEC["extend"]
    for f in ("ok", "primer_end", "score", "pos", "read_start", "read_stop", "nm", "tlen"):
        assert getattr(e, f).tolist() == g[f], f
    assert e.cigars() == g["cigar"] and e.mds() == g["md"] and int(e.cells) == g["cells"]


def _check_dedup(b, gs, st):
    g = VEC["dedup"]
    assert b.tolist() == g["is_best"] and gs.tolist() == g["group_size"]
    assert st.unique_pairs.tolist() == g["unique_pairs"] and st.max_pairs_per_uid.tolist() == g["max_pairs_per_uid"]


# ---- CPU: the oracle against the committed vectors -------------------------------------------------------------------
def test_oracle_reference_kats():
    for k in KATS["md"]:
        assert O.align_batch([k["reference"]], [k["query"]]).mds()[0] == k["md"], k["source"]
    for k in KATS["md_trimmed_with_full_cigar"]:
        assert O.md_trimmed_with_full_cigar(k["reference"], k["query"], k["reverse"]) == k["md"], k["source"]
    ed, first, trim = KATS["other"]
    assert O.cigar_edit_distance(ed["reference"], ed["query"]) == ed["expected"]
    assert O.index_of_first_match_in_reference(first["reference"], first["query"]) == first["expected"]
    s, q = O.trim_batch([trim["read"]], [trim["qual"]], trim["first"], trim["last"], True)
    assert len(s.strings()[0]) == trim["expected_length"] == len(q.strings()[0])


def test_oracle_frozen_vectors():
    g = VEC["align"]
    _check_align(O.align_batch(g["refs"], g["qrys"]))
    _check_align(O.align_batch(g["refs"], g["qrys"], flags=1), "align_penalize_start")
    probes, pairs = _extend_inputs()
    _check_extend(O.extend_batch(probes, pairs, cigar_cap=64, md_cap=96))
    clean, _ = synth.drop_n_pairs(pairs)
    _check_dedup(*O.dedup_batch(clean, 10, 0, probes.n))
    for k in VEC["primer_end"]:
        end, w = O.primer_end(k["primer"], k["read"])
        assert end == k["end"] and abs(w - k["weighted"]) < 1e-12


def _records_inputs():
    probes, pairs = _extend_inputs()
    g = VEC["records"]
    sub = pairs.head(60)
    sub.q1 = A.SeqBatch.from_strings(g["q1"])
    sub.q2 = A.SeqBatch.from_strings(g["q2"])
    return probes, sub, np.array(g["duplicate"], dtype=np.uint8)


def _check_records(rec):
    g = VEC["records"]
    assert rec.flag.tolist() == g["flag"] and rec.mate_pos.tolist() == g["mate_pos"]
    assert rec.seqs().strings() == g["seq"] and rec.quals().strings() == g["qual"]


def _assign_inputs():
    from bioinformatics_b200.batches import MappedRecordsHost, ProbeRangesHost
    g = VEC["assign"]
    r, m = g["ranges"], g["records"]
    ranges = ProbeRangesHost(r["contig"], r["reverse_strand"], r["start"], r["stop"])
    records = MappedRecordsHost(m["contig"], m["negative_strand"], m["second_of_pair"], m["start"], m["end"], m["read_index"],
                                (m["alt_offsets"], m["alt_contig"], m["alt_negative"], m["alt_start"], m["alt_stop"]))
    return ranges, records


def _check_assign(a):
    g = VEC["assign"]
    assert a.overflow.tolist() == g["overflow"] and a.mate.tolist() == g["mate"]
    ok = [k for k in range(a.n) if not g["overflow"][k]]
    assert [list(map(int, a.probe_sets()[k])) for k in ok] == [g["probes"][k] for k in ok]
    assert [list(map(int, a.common_sets()[k])) for k in ok] == [g["common"][k] for k in ok]


def test_oracle_frozen_records_assign_and_merge_kats():
    probes, sub, dup = _records_inputs()
    _check_records(O.extend_records(probes, sub, duplicate=dup))
    ranges, records = _assign_inputs()
    _check_assign(O.assign_probes(ranges, records, max_probes=VEC["assign"]["max_probes"]))
    for k in KATS["merge"]:
        s, q = O.merge_pairs([k["negative"]], [k["upstream"]], [k["downstream"]], [k["upstream_quality"]], [k["downstream_quality"]],
                             [k["upstream_start"]], [k["downstream_start"]])
        assert s.strings() == [k["merged"]] and q.strings() == [k["merged_quality"]], k["source"]
    from bioinformatics_b200.batches import MappedRecordsHost, ProbeRangesHost
    for k in KATS["range_map_inputs"]:
        n = len(k["windows"])
        ranges = ProbeRangesHost(np.zeros(n), np.zeros(n), [w[0] for w in k["windows"]], [w[1] for w in k["windows"]])
        rec = MappedRecordsHost([0], [0], [0], [k["query"][0]], [k["query"][1]], [0])
        assert list(map(int, O.assign_probes(ranges, rec).probe_sets()[0])) == k["hand_traced"], k["source"]


@pytest.mark.gpu
def test_cuda_frozen_records_assign_and_merge_kats():
    from bioinformatics_b200._lib import Context
    with Context(0) as ctx:
        probes, sub, dup = _records_inputs()
        ctx.set_probes(probes)
        ext = ctx.extend_batch(sub, cigar_cap=64, md_cap=96)
        _check_records(ctx.extend_records(sub, ext, duplicate=dup))
        ranges, records = _assign_inputs()
        ctx.set_probe_ranges(ranges)
        _check_assign(ctx.assign_probes(records, max_probes=VEC["assign"]["max_probes"]))
        for k in KATS["merge"]:
            s, q = ctx.merge_pairs([k["negative"]], [k["upstream"]], [k["downstream"]], [k["upstream_quality"]], [k["downstream_quality"]],
                                   [k["upstream_start"]], [k["downstream_start"]])
            assert s.strings() == [k["merged"]] and q.strings() == [k["merged_quality"]], k["source"]


# ---- GPU: the CUDA path against the same vectors ---------------------------------------------------------------------
@pytest.mark.gpu
def test_cuda_reference_kats_and_frozen_vectors():
    from bioinformatics_b200._lib import Context
    with Context(0) as ctx:
        for k in KATS["md"]:
            assert ctx.align_batch([k["reference"]], [k["query"]]).mds()[0] == k["md"], k["source"]
        ed, first, trim = KATS["other"]
        a = ctx.align_batch([ed["reference"], first["reference"]], [ed["query"], first["query"]])
        assert int(a.edit_distance[0]) == ed["expected"] and int(a.first_q[1]) == first["expected"]
        s, q = ctx.trim_batch([trim["read"]], [trim["qual"]], trim["first"], trim["last"], True)
        assert len(s.strings()[0]) == trim["expected_length"]
        g = VEC["align"]
        _check_align(ctx.align_batch(g["refs"], g["qrys"]))
        _check_align(ctx.align_batch(g["refs"], g["qrys"], flags=1), "align_penalize_start")
        probes, pairs = _extend_inputs()
        ctx.set_probes(probes)
        _check_extend(ctx.extend_batch(pairs, cigar_cap=64, md_cap=96))
        clean, _ = synth.drop_n_pairs(pairs)
        _check_dedup(*ctx.dedup_batch(clean, 10, 0, probes.n))


# ---- f2 / f4 frozen vectors (tests/golden/ingest_vectors.json, generator: tests/golden/make_golden_ingest.py) ------------------------------
def _ingest_checks(be):
    import json
    import os
    from bioinformatics_b200.batches import NameJoinHost
    v = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ingest_vectors.json")))
    fq1, sam, merged = v["fq1_text"].encode(), v["sam_text"].encode(), v["merged_sam_text"].encode()
    ix = be.index_fastq(fq1)
    assert all(ix.head()[f].tolist() == v["fastq_index"][f] for f in ix.FIELDS)
    t, n = be.trim_fastq(fq1, v["trim_fastq"]["first"], v["trim_fastq"]["last"], v["trim_fastq"]["three_prime"])
    assert n == v["trim_fastq"]["records"] and t.tobytes().decode() == v["trim_fastq"]["text"]
    sx = be.index_sam(sam)
    assert sx.n_header_lines == v["sam_index"]["n_header_lines"] and all(getattr(sx, c).tolist() == v["sam_index"][c] for c in sx.columns())
    r = be.sam_mapped_records(merged, v["contigs"])
    want = v["sam_mapped_records"]
    assert all(getattr(r, c)[:r.n].tolist() == want[c] for c in r.COLUMNS) and r.alt_offsets[:r.n + 1].tolist() == want["alt_offsets"]
    assert all(getattr(r, c)[:r.n_alt].tolist() == want[c] for c in r.ALT_COLUMNS)
    j = v["join"]
    f1, s1, f2, s2 = v["trim"]
    res = be.join_names(NameJoinHost(j["sam_names"], j["sam_flags"], j["sam_seq_len"], j["fq1_headers"], j["fq2_headers"], j["fq_seq_len"], j["fq_seq_len"],
                                     r1_first=f1, r1_stop=s1, r2_first=f2, r2_stop=s2))
    assert [x.tolist() for x in res.emitted()] == [j["sam_record"], j["fastq_record"], j["name_len"]]
    assert res.status.tolist() == j["status"] and res.total_matching_pairs == j["total_matching_pairs"]
    c = v["coordinate_order"]
    assert be.coordinate_order(*c["columns"], c["names"]).tolist() == c["order"]


def test_oracle_frozen_ingest_vectors():
    import types
    _ingest_checks(types.SimpleNamespace(index_fastq=O.index_fastq, trim_fastq=O.trim_fastq, index_sam=O.index_sam, sam_mapped_records=O.sam_mapped_records,
                                         join_names=O.join_names, coordinate_order=O.coordinate_order))


@pytest.mark.gpu
def test_cuda_frozen_ingest_vectors():
    from bioinformatics_b200._lib import Context
    with Context(0) as ctx:
        _ingest_checks(ctx)
