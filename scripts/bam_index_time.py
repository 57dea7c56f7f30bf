import sys, time, numpy as np
sys.path.insert(0, '.')
from bioinformatics_b200._lib import Context
from tests import bamio
from oracle import oracle as O
rngb = np.random.default_rng(13)
n_bam = 200000
seq150 = "ACGT" * 37 + "AC"
recs = [dict(name="%d" % (k // 2), flag=99 if k % 2 == 0 else 147, ref_id=int(k % 22), pos=int(rngb.integers(0, 10 ** 8)), mapq=60,
             cigar=[("M", 150)], seq=seq150, qual=bytes([30 + k % 10] * 150), next_ref_id=int(k % 22), next_pos=int(rngb.integers(0, 10 ** 8)),
             tlen=300, tags=b"NMC\x01") for k in range(n_bam)]
stream, _ = bamio.bam_stream(b"@HD\tVN:1.4\n", [("chr%d" % c, 10 ** 8) for c in range(1, 23)], recs)
data = bamio.bgzf(stream)
with Context(0) as ctx:
    bam = ctx.pinned_array(len(data)); bam[:] = np.frombuffer(data, dtype=np.uint8)
    for _ in range(3):
        t0 = time.perf_counter(); g = ctx.index_bam(bam, cap=n_bam + 8); dt = time.perf_counter() - t0
    print("records", g.n, "exact_walk", g.exact_walk, "host-to-host ms %.2f kernels ms %.3f launches %d" % (dt * 1e3, ctx.last_spans()[0], ctx.last_timing()[1]))
    c = O.bam_index(data)
    print("equal", g.n == c["n"] and all(np.array_equal(v, c[k]) for k, v in g.columns().items()))
