"""hsq_extend_records (SURVEY.md 8(a) A5 outputs / 8(f) f4): read string, base qualities, FLAG and PNEXT of the two SAM
records of each extended pair -- ExtendReadsToPrimer.extendReadPair / createRecord
(nimblegen_heatseq/src/main/java/com/roche/heatseq/process/ExtendReadsToPrimer.java:131-181,247-284) and
SAMRecordUtil.setSAMRecordsAsPair (…/utils/SAMRecordUtil.java:270-289).  Unpinned by the reference (no test constructs these
records); the oracle's strings are checked here against the definition on hand-made pairs, the CUDA path against the oracle."""
import numpy as np
import pytest

from bioinformatics_b200 import synth
from bioinformatics_b200.batches import PairBatchHost, ProbeTableHost, scorer
from oracle import oracle as O

COMP = bytes.maketrans(b"ACGTUNRYKMSWBDHV", b"TGCAANYRMKSWVHDB")


def _revcomp(s):
    return s.encode().translate(COMP)[::-1].decode()


def _two_probe_table():
    # one forward and one reverse probe around the same target; primers 18 bp, target 60 bp
    rng = np.random.default_rng(4)
    def rnd(n):
        return "".join(rng.choice(list("ACGT"), n))
    ext, lig, target = rnd(18), rnd(18), rnd(60)
    return ext, lig, target


def test_oracle_records_forward_probe_by_definition():
    ext, lig, target = _two_probe_table()
    uid = "ACGTAC"
    # forward probe: read one = UID ++ ext primer ++ target ++ lig primer; read two = revcomp(lig) ++ revcomp(target) ++ revcomp(ext)
    r1 = uid + ext + target + lig
    r2 = _revcomp(lig) + _revcomp(target) + _revcomp(ext)
    q1 = "".join(chr(33 + (i % 40)) for i in range(len(r1)))
    q2 = "".join(chr(34 + (i % 39)) for i in range(len(r2)))
    probes = ProbeTableHost([0], ext_start=[1000], ext_stop=[1017], lig_start=[1078], lig_stop=[1095], target_start=[1018], target_stop=[1077],
                            ext=[ext], lig=[lig], target=[target])
    pairs = PairBatchHost([0], [r1], [r2], [q1], [q2], read_index=[0])
    ext_out = O.extend_batch(probes, pairs)
    assert ext_out.pair_ok.tolist() == [1]
    rec = O.extend_records(probes, pairs, duplicate=[1])
    a1, b1 = ext_out.read_start[0], ext_out.read_stop[0]
    a2, b2 = ext_out.read_start[1], ext_out.read_stop[1]
    assert (a1, b1) == (len(uid) + len(ext), len(r1) - 1) and (a2, b2) == (len(lig), len(r2) - 1)
    # read one on the + strand: the slice as is; read two on the - strand: reverse complement, qualities reversed
    assert rec.seqs().strings() == [r1[a1:b1 + 1], _revcomp(r2[a2:b2 + 1])]
    assert rec.quals().strings() == [q1[a1:b1 + 1], q2[a2:b2 + 1][::-1]]
    # paired | proper | mate reverse | first | duplicate ; paired | proper | read reverse | second | duplicate
    assert rec.flag.tolist() == [0x1 | 0x2 | 0x20 | 0x40 | 0x400, 0x1 | 0x2 | 0x10 | 0x80 | 0x400]
    assert rec.mate_pos.tolist() == [int(ext_out.pos[1]), int(ext_out.pos[0])]
    assert rec.seqs().strings()[1] == ext + target  # the - strand record reads along the reference


def test_oracle_records_u_base_and_failed_pair():
    ext, lig, target = _two_probe_table()
    r1 = "ACGTAC" + ext + target + lig
    r2 = _revcomp(lig) + _revcomp(target) + _revcomp(ext)
    # reverse probe (its sequences are given on its own strand, so the reads look the same): read one is now on the - strand,
    # read two on the + strand, and a U in read two comes back as T (complemented on entry and again after the slice:
    # U -> A -> T, ExtendReadsToPrimer.java:104,170) while a U in read one on the - strand becomes A
    probes = ProbeTableHost([1], ext_start=[1078], ext_stop=[1095], lig_start=[1000], lig_stop=[1017], target_start=[1018], target_stop=[1077],
                            ext=[ext], lig=[lig], target=[target])
    k2 = r2.index("T", len(lig) + 5)
    r2u = r2[:k2] + "U" + r2[k2 + 1:]
    k1 = r1.index("T", 6 + len(ext) + 5)
    r1u = r1[:k1] + "U" + r1[k1 + 1:]
    pairs = PairBatchHost([0, 0], [r1u, "ACGT" * 20], [r2u, r2], ["I" * len(r1), "I" * 80], ["I" * len(r2), "I" * len(r2)], read_index=[0, 1])
    ext_out = O.extend_batch(probes, pairs)
    rec = O.extend_records(probes, pairs)
    assert ext_out.pair_ok.tolist() == [1, 0]
    assert rec.flag.tolist()[2:] == [0, 0] and rec.offsets[4] == rec.offsets[2]  # the failed pair has no records
    a1, a2 = ext_out.read_start[0], ext_out.read_start[1]
    assert rec.seqs().strings()[:2] == [_revcomp(r1u[a1:]), r2u[a2:].replace("U", "T")]
    assert "U" not in rec.seqs().strings()[0] and _revcomp(r1u[a1:]) == _revcomp(r1[a1:])
    assert rec.flag.tolist()[:2] == [0x1 | 0x2 | 0x10 | 0x40, 0x1 | 0x2 | 0x20 | 0x80]


@pytest.mark.gpu
@pytest.mark.parametrize("read_len,dup", [(150, True), (100, False)])
def test_cuda_records_match_oracle(read_len, dup):
    from bioinformatics_b200._lib import Context
    probes = synth.make_probes(200, seed=7 + read_len)
    pairs = synth.make_pairs(probes, 5000, read_len=read_len, ext_uid_len=8, seed=13 + read_len, sub_rate=0.02, ins_rate=0.003, del_rate=0.003)
    rng = np.random.default_rng(6)
    # lower case, IUPAC codes and U inside reads (the record strings are canonical upper case)
    for sb in (pairs.r1, pairs.r2):
        at = rng.choice(len(sb.data), len(sb.data) // 50, replace=False)
        sb.data[at] = rng.choice(np.frombuffer(b"acgtURYN", dtype=np.uint8), len(at))
    duplicate = (rng.random(pairs.n) < 0.3) if dup else None
    want = O.extend_records(probes, pairs, duplicate=duplicate)
    with Context(0) as ctx:
        ctx.set_probes(probes)
        ctx.set_scorer(scorer())
        ext = ctx.extend_batch(pairs, cigar_cap=96, md_cap=128)
        got = ctx.extend_records(pairs, ext, duplicate=duplicate)
        assert ext.pair_ok.sum() > 1000
        assert (got.offsets == want.offsets).all()
        assert (got.flag == want.flag).all() and (got.mate_pos == want.mate_pos).all()
        total = got.offsets[-1]
        assert (got.seq[:total] == want.seq[:total]).all() and (got.qual[:total] == want.qual[:total]).all()
        # too small a buffer: HSQ_ERR_OVERFLOW with the offsets filled
        from bioinformatics_b200._lib import HsqError
        with pytest.raises(HsqError) as e:
            ctx.extend_records(pairs, ext, out_cap=10)
        assert e.value.status == -7
