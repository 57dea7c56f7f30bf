"""Seeded synthetic HEAT-Seq-shaped probes and read pairs; numpy only, vectorised.

Probe layout follows the probe-info columns of COMMON/utils/probeinfo/ProbeFileUtil.java:45-46.  Read
orientation follows what ExtendReadsToPrimer expects (HSQ/process/ExtendReadsToPrimer.java:104-165):
  R1 = UID_ext ++ extPrimer ++ target ++ ligPrimer ++ filler
  R2 = UID_lig ++ revcomp(ligPrimer) ++ revcomp(target) ++ revcomp(extPrimer) ++ filler
(the same construction serves '+' and '-' probes, because the reference complements both the probe
sequences and the read for the '-' strand).
"""
import numpy as np

from . import _abi as A
from .batches import PairBatchHost, ProbeTableHost

SEED_PROBES = 20240501
SEED_READS = 20240502
_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.zeros(256, dtype=np.uint8)
for a, b in zip(b"ACGTN", b"TGCAN"):
    _COMP[a] = b


def revcomp_bytes(arr):
    return _COMP[arr[::-1]]


def make_probes(n_probes, seed=SEED_PROBES, target_len=(100, 160), primer_len=(18, 25)):
    rng = np.random.default_rng(seed)
    le = rng.integers(primer_len[0], primer_len[1] + 1, n_probes)
    ll = rng.integers(primer_len[0], primer_len[1] + 1, n_probes)
    lt = rng.integers(target_len[0], target_len[1] + 1, n_probes)
    rev = (np.arange(n_probes) % 2).astype(np.uint8)

    def rand_batch(lens):
        off = np.zeros(n_probes + 1, dtype=np.int64)
        off[1:] = np.cumsum(lens)
        return A.SeqBatch(_ACGT[rng.integers(0, 4, int(off[-1]))], off)

    ext, lig, target = rand_batch(le), rand_batch(ll), rand_batch(lt)
    per_contig = max(1, (n_probes + 21) // 22)
    base = (np.arange(n_probes) % per_contig) * 1000 + 10000
    chrom = ["chr%d" % (1 + k // per_contig) for k in range(n_probes)]
    left_len = np.where(rev == 1, ll, le)
    right_len = np.where(rev == 1, le, ll)
    left_start = base
    left_stop = base + left_len - 1
    t_start = left_stop + 1
    t_stop = t_start + lt - 1
    right_start = t_stop + 1
    right_stop = right_start + right_len - 1
    ext_start = np.where(rev == 1, right_start, left_start)
    ext_stop = np.where(rev == 1, right_stop, left_stop)
    lig_start = np.where(rev == 1, left_start, right_start)
    lig_stop = np.where(rev == 1, left_stop, right_stop)
    ids = ["probe_%d" % k for k in range(n_probes)]
    return ProbeTableHost(rev, ext_start, ext_stop, lig_start, lig_stop, t_start, t_stop, ext, lig, target, ids, chrom)


def _templates(probes, read_len, rng):
    """Per-probe molecule templates for both mates, padded with random filler to read_len + slack."""
    n = probes.n
    width = read_len + 8
    t1 = _ACGT[rng.integers(0, 4, (n, width))]
    t2 = _ACGT[rng.integers(0, 4, (n, width))]
    for k in range(n):
        e = probes.ext.data[probes.ext.offsets[k]:probes.ext.offsets[k + 1]]
        t = probes.target.data[probes.target.offsets[k]:probes.target.offsets[k + 1]]
        g = probes.lig.data[probes.lig.offsets[k]:probes.lig.offsets[k + 1]]
        m1 = np.concatenate([e, t, g])[:width]
        m2 = revcomp_bytes(np.concatenate([e, t, g]))[:width]
        t1[k, :len(m1)] = m1
        t2[k, :len(m2)] = m2
    return t1, t2


def _noisy_copy(tmpl_rows, body_len, rng, sub_rate, ins_rate, del_rate):
    """Rows of templates -> body_len-long noisy copies (at most one insertion and one deletion per read)."""
    n, width = tmpl_rows.shape
    j = np.arange(body_len, dtype=np.int32)[None, :]
    p_del = np.where(rng.random(n) < del_rate * body_len, rng.integers(0, body_len, n), 1 << 20).astype(np.int32)[:, None]
    p_ins = np.where(rng.random(n) < ins_rate * body_len, rng.integers(0, body_len, n), 1 << 20).astype(np.int32)[:, None]
    src = j + (j >= p_del).astype(np.int32) - (j > p_ins).astype(np.int32)
    np.clip(src, 0, width - 1, out=src)
    out = np.take_along_axis(tmpl_rows, src, axis=1)
    ins_rows = np.nonzero(p_ins[:, 0] < body_len)[0]
    out[ins_rows, p_ins[ins_rows, 0]] = _ACGT[rng.integers(0, 4, len(ins_rows))]
    sub = rng.random((n, body_len)) < sub_rate
    k = int(sub.sum())
    if k:
        cur = out[sub]
        idx = np.searchsorted(_ACGT, cur)  # A,C,G,T are ascending ASCII
        out[sub] = _ACGT[(idx + rng.integers(1, 4, k)) % 4]
    return out


_QTABLE = None


def _qual_table():
    global _QTABLE
    if _QTABLE is None:
        rng = np.random.default_rng(7)
        q = np.round(np.clip(rng.normal(36, 4, 4096), 2, 41)).astype(np.uint8) + 33
        _QTABLE = q
    return _QTABLE


def make_pairs(probes, n_pairs, read_len=150, ext_uid_len=10, lig_uid_len=0, seed=SEED_READS, mean_family=3.0,
               sub_rate=0.005, ins_rate=0.0005, del_rate=0.0005, n_rate=0.001, with_quals=True, chunk=1 << 20,
               sort_by_probe=False):
    """Returns PairBatchHost with fixed-length reads.  Families: size ~ 1 + Geometric(mean mean_family)."""
    rng = np.random.default_rng(seed)
    t1, t2 = _templates(probes, read_len, rng)
    # molecules until we have n_pairs reads
    n_mol_guess = int(n_pairs / (1.0 + mean_family) * 1.2) + 16
    fam = 1 + rng.geometric(1.0 / (mean_family + 1.0), n_mol_guess) - 1
    fam = np.maximum(fam, 1)
    while fam.sum() < n_pairs:
        fam = np.concatenate([fam, np.maximum(1, rng.geometric(1.0 / (mean_family + 1.0), n_mol_guess))])
    n_mol = int(np.searchsorted(np.cumsum(fam), n_pairs) + 1)
    fam = fam[:n_mol]
    mol_probe = rng.integers(0, probes.n, n_mol).astype(np.int32)
    mol_uid1 = _ACGT[rng.integers(0, 4, (n_mol, max(1, ext_uid_len)))][:, :ext_uid_len]
    mol_uid2 = _ACGT[rng.integers(0, 4, (n_mol, max(1, lig_uid_len)))][:, :lig_uid_len]
    mol_of_read = np.repeat(np.arange(n_mol), fam)[:n_pairs]
    if sort_by_probe:
        order = np.argsort(mol_probe[mol_of_read], kind="stable")
        mol_of_read = mol_of_read[order]
    else:
        mol_of_read = mol_of_read[rng.permutation(n_pairs)]
    probe_id = mol_probe[mol_of_read]
    r1 = np.empty((n_pairs, read_len), dtype=np.uint8)
    r2 = np.empty((n_pairs, read_len), dtype=np.uint8)
    for b in range(0, n_pairs, chunk):
        e = min(n_pairs, b + chunk)
        mo = mol_of_read[b:e]
        r1[b:e, :ext_uid_len] = mol_uid1[mo]
        r2[b:e, :lig_uid_len] = mol_uid2[mo]
        r1[b:e, ext_uid_len:] = _noisy_copy(t1[probe_id[b:e]], read_len - ext_uid_len, rng, sub_rate, ins_rate, del_rate)
        r2[b:e, lig_uid_len:] = _noisy_copy(t2[probe_id[b:e]], read_len - lig_uid_len, rng, sub_rate, ins_rate, del_rate)
    n_n = int(n_pairs * n_rate)
    if n_n:
        rows = rng.integers(0, n_pairs, n_n)
        colsn = rng.integers(0, read_len, n_n)
        mate = rng.integers(0, 2, n_n)
        r1[rows[mate == 0], colsn[mate == 0]] = ord("N")
        r2[rows[mate == 1], colsn[mate == 1]] = ord("N")
    q1 = q2 = None
    if with_quals:
        tbl = _qual_table()
        q1 = A.SeqBatch.from_fixed(tbl[rng.integers(0, len(tbl), (n_pairs, read_len))])
        q2 = A.SeqBatch.from_fixed(tbl[rng.integers(0, len(tbl), (n_pairs, read_len))])
    return PairBatchHost(probe_id, A.SeqBatch.from_fixed(r1), A.SeqBatch.from_fixed(r2), q1, q2,
                         np.arange(n_pairs, dtype=np.int32))


def tile_pairs(base, tiles, alloc=None):
    """A batch of tiles * base.n pairs from one generated base batch of fixed-length reads, for the large bench shapes (generation
    costs ~10 s per million pairs): tile t is the base with the first two bases of read one (the start of the extension UID) rotated
    by t % 4 and t // 4 letters, so every tile carries its own molecules -- same probes, same family sizes, disjoint UID groups
    (tiles <= 16) -- and the FASTQ ordinals continue (read_index + t * (largest base ordinal + 1)).  alloc(nbytes) -> uint8 array (e.g. pinned memory)."""
    assert 1 <= tiles <= 16
    alloc = alloc or (lambda nbytes: np.empty(nbytes, dtype=np.uint8))
    n = base.n
    L1, L2 = len(base.r1.data) // max(n, 1), len(base.r2.data) // max(n, 1)

    def rep(sb, L):
        if sb is None:
            return None
        out = alloc(tiles * n * L)[:tiles * n * L].reshape(tiles, n, L)
        out[:] = sb.data.reshape(1, n, L)
        return out

    r1, r2, q1, q2 = rep(base.r1, L1), rep(base.r2, L2), rep(base.q1, L1), rep(base.q2, L2)
    idx = np.searchsorted(_ACGT, base.r1.data.reshape(n, L1)[:, :2])
    acgt = _ACGT[np.minimum(idx, 3)] == base.r1.data.reshape(n, L1)[:, :2]
    for t in range(1, tiles):
        for col, rot in ((0, t % 4), (1, t // 4)):
            r1[t, :, col] = np.where(acgt[:, col], _ACGT[(idx[:, col] + rot) % 4], r1[0, :, col])
    probe = alloc(4 * tiles * n).view(np.int32)[:tiles * n].reshape(tiles, n)
    probe[:] = base.probe_id[None, :]
    ri = alloc(4 * tiles * n).view(np.int32)[:tiles * n].reshape(tiles, n)
    ri[:] = base.read_index[None, :].astype(np.int64) + (np.arange(tiles, dtype=np.int64) * (int(base.read_index.max()) + 1 if n else 0))[:, None]

    def sb_of(a, L):
        if a is None:
            return None
        off = alloc(8 * (tiles * n + 1)).view(np.int64)[:tiles * n + 1]
        off[:] = np.arange(tiles * n + 1, dtype=np.int64) * L
        return A.SeqBatch(a.reshape(-1), off)

    return PairBatchHost(probe.reshape(-1), sb_of(r1, L1), sb_of(r2, L2), sb_of(q1, L1), sb_of(q2, L2), ri.reshape(-1))


def drop_n_pairs(pairs):
    """The caller-side filter of …Identification.java:815-821: pairs with an 'N'/'n' in either mate are not deduplicated."""
    L1 = pairs.r1.lengths()
    L2 = pairs.r2.lengths()
    if not ((L1 == L1[0]).all() and (L2 == L2[0]).all()):
        keep = [k for k in range(pairs.n) if "n" not in (pairs.r1.string(k) + pairs.r2.string(k)).lower()]
        return pairs.subset(keep), np.asarray(keep)
    a = pairs.r1.data.reshape(pairs.n, -1)
    b = pairs.r2.data.reshape(pairs.n, -1)
    bad = ((a | 0x20) == ord("n")).any(axis=1) | ((b | 0x20) == ord("n")).any(axis=1)
    keep = np.nonzero(~bad)[0]

    def take(sb):
        return None if sb is None else A.SeqBatch.from_fixed(sb.data.reshape(pairs.n, -1)[keep])

    return PairBatchHost(pairs.probe_id[keep], take(pairs.r1), take(pairs.r2), take(pairs.q1), take(pairs.q2),
                         pairs.read_index[keep]), keep


def make_mapped_records(probes, pairs, seed=5, alt_rate=0.05, lone_rate=0.02, off_panel_rate=0.02, repeat_names=3):
    """Mapped records (what a mapper would report for `pairs`) for the read -> probe assignment, in shuffled order.

    Each pair gives a read-one record on the probe's strand and a read-two record on the other one
    (…Identification.java:452-455), both inside the probe's window.  alt_rate of the pairs carry an alternative hit (XA) inside
    another probe's window on both mates, so the pair has two candidate probes; lone_rate lose their read two; off_panel_rate map
    to a contig without probes; repeat_names read names occur on four records.  Returns (ProbeRangesHost, MappedRecordsHost)."""
    from .batches import MappedRecordsHost, ProbeRangesHost
    rng = np.random.default_rng(seed)
    names = sorted(set(probes.chromosomes), key=lambda c: int(c[3:]))
    contig_of = {c: k for k, c in enumerate(names)}
    pcontig = np.array([contig_of[c] for c in probes.chromosomes], dtype=np.int32)
    ranges = ProbeRangesHost.from_probe_table(probes, pcontig)
    n = pairs.n
    p = pairs.probe_id.astype(np.int64)
    rev = probes.reverse_strand[p].astype(np.uint8)

    def inside(pid, m):
        a = ranges.start[pid] + rng.integers(0, 6, m)
        b = ranges.stop[pid] - rng.integers(0, 6, m)
        return a.astype(np.int32), b.astype(np.int32)

    a1, b1 = inside(p, n)
    a2, b2 = inside(p, n)
    contig = np.concatenate([pcontig[p], pcontig[p]])
    neg = np.concatenate([rev, 1 - rev]).astype(np.uint8)
    second = np.concatenate([np.zeros(n, np.uint8), np.ones(n, np.uint8)])
    start, end = np.concatenate([a1, a2]), np.concatenate([b1, b2])
    ridx = np.concatenate([pairs.read_index, pairs.read_index]).astype(np.int32)
    has_alt = rng.random(n) < alt_rate
    other = (p + 1 + rng.integers(0, max(1, probes.n - 1), n)) % probes.n
    alt_count = np.concatenate([has_alt, has_alt]).astype(np.int64)
    oa, ob = inside(other, n)
    orev = probes.reverse_strand[other].astype(np.uint8)
    alt_contig_full = np.concatenate([pcontig[other], pcontig[other]])
    alt_neg_full = np.concatenate([orev, 1 - orev]).astype(np.uint8)
    alt_a_full, alt_b_full = np.concatenate([oa, oa]), np.concatenate([ob, ob])
    off = np.concatenate([rng.random(n) < off_panel_rate] * 2)
    contig = np.where(off, -1, contig).astype(np.int32)
    keep = np.ones(2 * n, dtype=bool)
    keep[n:] = rng.random(n) >= lone_rate
    if repeat_names and n > 2 * repeat_names:  # a read name on four records: removed from every probe (:392-400)
        twice = rng.choice(n, repeat_names, replace=False)
        extra = np.concatenate([twice, twice + n])
        order = np.concatenate([np.nonzero(keep)[0], extra])
    else:
        order = np.nonzero(keep)[0]
    order = order[rng.permutation(len(order))]
    cnt = alt_count[order]
    alt_off = np.zeros(len(order) + 1, dtype=np.int64)
    alt_off[1:] = np.cumsum(cnt)
    src = order[cnt > 0]
    alt = (alt_off, alt_contig_full[src], alt_neg_full[src], alt_a_full[src], alt_b_full[src])
    return ranges, MappedRecordsHost(contig[order], neg[order], second[order], start[order], end[order], ridx[order], alt)


def write_probe_info_file(probes, path, ext_uid_len=10, lig_uid_len=0):
    """Probe-info TSV with the 13 named columns the reference parses (ProbeFileUtil.java:45-46) and its #key=value header."""
    cols = ["probe_id", "chromosome", "probe_strand", "ext_start", "ext_stop", "ext_sequence", "lig_start", "lig_stop",
            "lig_sequence", "target_start", "target_stop", "target_sequence", "annotation"]
    with open(path, "w") as f:
        f.write("#extension_uid=%d ligation_uid=%d genome=synthetic\n" % (ext_uid_len, lig_uid_len))
        f.write("\t".join(cols) + "\n")
        for k in range(probes.n):
            f.write("\t".join([probes.probe_ids[k], probes.chromosomes[k], "-" if probes.reverse_strand[k] else "+",
                               str(probes.ext_start[k]), str(probes.ext_stop[k]), probes.ext.string(k),
                               str(probes.lig_start[k]), str(probes.lig_stop[k]), probes.lig.string(k),
                               str(probes.target_start[k]), str(probes.target_stop[k]), probes.target.string(k), ""]) + "\n")


def make_ingest_case(n_pairs, read_len=150, seed=7, mapped_frac=0.9, lone_frac=0.03, nonprimary_frac=0.02, repeat_frac=0.01, orphan_frac=0.01,
                     old_style=False, trim=(10, 139, 0, 139), three_prime=False, shuffle=True, crlf=False, final_newline=True):
    """Synthetic input of `HSQutils dedup` as files hold it: two FASTQ texts in one (random) file order and the
    SAM-side columns FastqAndBamFileMerger reads (QNAME, FLAG, read length), for the records a mapper would emit from the
    trimmed reads: both mates of most pairs, some lone mates, not-primary lines, repeated lines and names the FASTQ files
    do not hold.  Names follow Casava 1.8 ("<id> 1:N:0:1") or the older "<id>#0/1" style
    (COMMON/utils/IlluminaFastQReadNameUtil.java:30-33).  Returns a dict."""
    rng = np.random.default_rng(seed)
    xy = rng.choice(100_000_000, size=n_pairs, replace=False)
    tile = 1101 + (xy % 16)
    ids = ["M01077:35:000000000-A3J96:1:%d:%d:%d" % (t, v // 10000, v % 10000) for t, v in zip(tile.tolist(), xy.tolist())]
    if old_style:
        ids = [s + "#0" for s in ids]
        h1 = [s + "/1" for s in ids]
        h2 = [s + "/2" for s in ids]
    else:
        h1 = [s + " 1:N:0:1" for s in ids]
        h2 = [s + " 2:N:0:1" for s in ids]
    lens = np.full(n_pairs, read_len, dtype=np.int32)
    bases = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=(2, n_pairs, read_len))]
    quals = (33 + rng.integers(2, 42, size=(2, n_pairs, read_len))).astype(np.uint8)
    eol = b"\r\n" if crlf else b"\n"

    def text(headers, b, q):
        parts = []
        for k, h in enumerate(headers):
            parts += [b"@", h.encode(), eol, b[k].tobytes(), eol, b"+", eol, q[k].tobytes(), eol]
        t = b"".join(parts)
        return t if final_newline or not t else t[:-len(eol)]

    fq1_text, fq2_text = text(h1, bases[0], quals[0]), text(h2, bases[1], quals[1])
    # SAM side: names as a mapper writes them (the unique id), trimmed lengths
    f1, s1, f2, s2 = trim
    exp1 = (min(s1, read_len - 1) if three_prime else read_len - 1) + 1 - f1
    exp2 = (min(s2, read_len - 1) if three_prime else read_len - 1) + 1 - f2
    names, flags, slen = [], [], []
    u = rng.random(n_pairs)
    for k in range(n_pairs):
        if u[k] >= mapped_frac:
            continue
        lone = rng.random() < lone_frac
        which = (1, 2) if not lone else ((1,) if rng.random() < 0.5 else (2,))
        for m in which:
            fl = (0x41 if m == 1 else 0x81) | (0x10 if rng.random() < 0.5 else 0)
            names.append(ids[k]); flags.append(fl); slen.append(exp1 if m == 1 else exp2)
            if rng.random() < nonprimary_frac:
                names.append(ids[k]); flags.append(fl | 0x100); slen.append(exp1 if m == 1 else exp2)
            if rng.random() < repeat_frac:
                names.append(ids[k]); flags.append(fl | 0x800); slen.append(exp1 if m == 1 else exp2)
    for k in range(int(orphan_frac * n_pairs)):
        names.append(("A00000:%d" if k % 2 else "Z99999:%d") % k); flags.append(0x41 if k % 3 else 0x81); slen.append(exp1)
    order = rng.permutation(len(names)) if shuffle else np.arange(len(names))
    return dict(n_pairs=n_pairs, ids=ids, fq1_headers=h1, fq2_headers=h2, fq1_text=fq1_text, fq2_text=fq2_text, fq_seq_len=lens,
                bases=bases, quals=quals, sam_names=[names[i] for i in order], sam_flags=np.asarray(flags, dtype=np.int32)[order],
                sam_seq_len=np.asarray(slen, dtype=np.int32)[order], trim=trim, three_prime=three_prime)


def sam_text_of(case, seed=3, n_contigs=22, xa_rate=0.05, crlf=False, final_newline=True, index_names=False):
    """SAM text for the records of make_ingest_case: header lines, then one alignment line per record with the trimmed read, a CIGAR
    with an occasional insertion / deletion / soft clip, mate columns and NM / MD / XA tags (as `bwa mem` writes them).
    index_names: QNAME = the FASTQ ordinal of the read, as in the merged BAM the dedup path works on (FastqSorter.java:298-302)."""
    rng = np.random.default_rng(seed)
    eol = "\r\n" if crlf else "\n"
    f1, _, f2, _ = case["trim"]
    row_of = {s: k for k, s in enumerate(case["ids"])}
    lines = ["@HD\tVN:1.5\tSO:unsorted"] + ["@SQ\tSN:chr%d\tLN:%d" % (c + 1, 50_000_000) for c in range(n_contigs)] + ["@PG\tID:bwa\tPN:bwa"]
    n = len(case["sam_names"])
    kind = rng.integers(0, 8, size=n)
    pos = rng.integers(1, 40_000_000, size=n)
    contig = rng.integers(1, n_contigs + 1, size=n)
    for k in range(n):
        name, fl, ln = case["sam_names"][k], int(case["sam_flags"][k]), int(case["sam_seq_len"][k])
        row = row_of.get(name)
        mate = 0 if fl & 0x40 else 1
        if row is None:
            seq, qual = "A" * ln, "I" * ln
        else:
            first = f1 if mate == 0 else f2
            seq = case["bases"][mate][row][first:first + ln].tobytes().decode()
            qual = case["quals"][mate][row][first:first + ln].tobytes().decode()
        if kind[k] == 0 and ln > 20:
            cigar = "%dM1I%dM" % (10, ln - 11)
        elif kind[k] == 1 and ln > 20:
            cigar = "%dM2D%dM" % (12, ln - 12)
        elif kind[k] == 2 and ln > 20:
            cigar = "5S%dM" % (ln - 5)
        else:
            cigar = "%dM" % ln
        tags = ["NM:i:%d" % (kind[k] % 3), "MD:Z:%d" % ln, "AS:i:%d" % ln]
        if rng.random() < xa_rate:
            hits = ["chr%d,%s%d,%dM,1;" % (int(rng.integers(1, n_contigs + 3)), "+-"[(k + h) % 2], int(pos[k]) + 7 * h, ln - h)
                    for h in range(int(rng.integers(1, 4)))]
            tags.insert(int(rng.integers(0, 4)), "XA:Z:" + "".join(hits))
        if index_names:
            name = str(row if row is not None else case["n_pairs"] + k)
        lines.append("\t".join([name, str(fl), "chr%d" % contig[k], str(int(pos[k])), str(int(rng.integers(0, 61))), cigar, "=",
                                str(int(pos[k]) + 100), str(int(rng.integers(-400, 400))), seq, qual] + tags))
    t = eol.join(lines) + (eol if final_newline else "")
    return t.encode()


def dedup_input_files(probes, fastq, records, seed=11):
    """The three inputs of `HSQutils dedup` as text for a synthetic run: FASTQ one / two (ordinal k = pair k of `fastq`) and the
    mapper's SAM text for `records` (make_mapped_records), read names = Illumina-style ids instead of ordinals, unsorted.
    Returns (fq1_text, fq2_text, sam_text, contig_names) with contig_names in the numbering of the probe ranges."""
    rng = np.random.default_rng(seed)
    n = fastq.n
    xy = rng.choice(90_000_000, size=n, replace=False)
    ids = ["M01077:35:000000000-A3J96:1:%d:%d:%d" % (1101 + v % 16, v // 10000, v % 10000) for v in xy.tolist()]

    def fq_text(reads, quals, tag):
        parts = []
        for k in range(n):
            parts += [b"@", ids[k].encode(), tag, b"\n", reads.data[reads.offsets[k]:reads.offsets[k + 1]].tobytes(), b"\n+\n",
                      quals.data[quals.offsets[k]:quals.offsets[k + 1]].tobytes(), b"\n"]
        return b"".join(parts)

    fq1, fq2 = fq_text(fastq.r1, fastq.q1, b" 1:N:0:1"), fq_text(fastq.r2, fastq.q2, b" 2:N:0:1")
    contigs = sorted(set(probes.chromosomes), key=lambda c: int(c[3:]))
    cname = lambda c: contigs[c] if c >= 0 else "chrUn_random"
    lines = ["@HD\tVN:1.5\tSO:unsorted"] + ["@SQ\tSN:%s\tLN:250000000" % c for c in contigs + ["chrUn_random"]]
    ao = records.alt_offsets
    for r in range(records.n):
        k = int(records.read_index[r])
        second = bool(records.second_of_pair[r])
        flag = 0x1 | 0x2 | (0x10 if records.negative_strand[r] else 0x20) | (0x80 if second else 0x40)
        reads, quals = (fastq.r2, fastq.q2) if second else (fastq.r1, fastq.q1)
        seq = reads.data[reads.offsets[k]:reads.offsets[k + 1]].tobytes().decode()
        qual = quals.data[quals.offsets[k]:quals.offsets[k + 1]].tobytes().decode()
        span = int(records.end[r]) - int(records.start[r]) + 1
        tags = ["NM:i:0"]
        if ao is not None and ao[r + 1] > ao[r]:
            tags.append("XA:Z:" + "".join("%s,%s%d,%dM,0;" % (cname(int(records.alt_contig[a])), "-" if records.alt_negative[a] else "+",
                                                            int(records.alt_start[a]), int(records.alt_stop[a]) - int(records.alt_start[a]) + 1)
                                          for a in range(int(ao[r]), int(ao[r + 1]))))
        lines.append("\t".join([ids[k], str(flag), cname(int(records.contig[r])), str(int(records.start[r])), "60", "%dM" % span, "=",
                                str(int(records.start[r])), "0", seq, qual] + tags))
    return fq1, fq2, ("\n".join(lines) + "\n").encode(), contigs
