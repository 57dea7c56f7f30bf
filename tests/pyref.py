"""TEST INFRASTRUCTURE: a second, independent statement of the reference's hot path in plain Python, written from the Java sources
(not from oracle/): the rows for which the reference's own tests pin nothing (A3-A7, A9).  tests/test_pyref.py checks
that the C++ oracle agrees with it on seeded cases; the CUDA path is then checked against the oracle.

Roots under /root/reference:
  COMMON/ = bioinformatics_common/src/main/java/com/roche/sequencing/bioinformatics/common/
  HSQ/    = nimblegen_heatseq/src/main/java/com/roche/heatseq/
Sequences are Python strings of IUPAC letters (upper case; '_' = GAP), as IupacNucleotideCodeSequence holds them after
getCodesFromString (COMMON/sequence/IupacNucleotideCode.java:86-104: upper-cased, '_' '.' '-' -> GAP).
"""
GAP = "_"
_SETS = {"A": "A", "C": "C", "G": "G", "T": "T", "U": "T", "R": "AG", "Y": "CT", "S": "GC", "W": "AT", "K": "GT", "M": "AC", "B": "CGT", "D": "AGT",
         "H": "ACT", "V": "ACG", "N": "ACTG", GAP: ""}
# complimentAbbreviation of every enum constant (IupacNucleotideCode.java:31-39)
_COMP = {"A": "T", "C": "G", "G": "C", "T": "A", "U": "A", "R": "Y", "Y": "R", "S": "S", "W": "W", "K": "M", "M": "K", "B": "V", "D": "H", "H": "D",
         "V": "B", "N": "N", GAP: GAP}


def codes(s):
    """new IupacNucleotideCodeSequence(String)"""
    return "".join(GAP if ch in "_.-" else ch.upper() for ch in s)


def matches(a, b):
    """IupacNucleotideCode.matches (:185-200): equals, else any shared nucleotide (GAP has none)."""
    if a == b:
        return True
    return any(x in _SETS[b] for x in _SETS[a])


def complement(s):
    return "".join(_COMP[c] for c in s)


def reverse(s):
    return s[::-1]


def sub_sequence(s, start, end):
    """IupacNucleotideCodeSequence.subSequence(start, end): inclusive, end clamped to size - 1 (:103-117)."""
    end = min(end, len(s) - 1)
    return s[start:end + 1] if end >= start else ""


class Scorer:
    """SimpleAlignmentScorer (COMMON/alignment/SimpleAlignmentScorer.java:23-96)."""

    def __init__(self, match=1, mismatch=-4, gap_extend=-1, gap_start=-6, penalize_start=False, penalize_end=False):
        self.match, self.mismatch, self.gap_extend, self.gap_start = match, mismatch, gap_extend, gap_start
        self.penalize_start, self.penalize_end = penalize_start, penalize_end

    def gap_score(self, index_in_gap):  # getGapScore: 0 for the first position of a gap, the extension penalty after it
        return 0 if index_in_gap == 0 else self.gap_extend

    def match_score(self, a, b):
        return self.match if matches(a, b) else self.mismatch


def needleman_wunsch(reference, query, sc=None):
    """NeedlemanWunschGlobalAlignment.getAlignmentPair (:85-114) with initializeTraceabilityMatrix / completeCell (:61-78,261-376).
    Returns (score, reference alignment, query alignment)."""
    sc = sc or Scorer()
    nq, nr = len(query), len(reference)
    S = [[0] * (nr + 1) for _ in range(nq + 1)]
    src = [[None] * (nr + 1) for _ in range(nq + 1)]   # source cell as (row, column)
    ih = [[None] * (nr + 1) for _ in range(nq + 1)]    # indexInHorizontalGap
    iv = [[None] * (nr + 1) for _ in range(nq + 1)]    # indexInVerticalGap
    for r in range(nq + 1):
        for c in range(nr + 1):
            if r == 0 and c != 0:
                S[r][c] = sc.gap_start + (c - 1) * (sc.gap_score(0) - 1) if sc.penalize_start else 0
                src[r][c] = (r, c - 1)
            elif c == 0 and r != 0:
                S[r][c] = sc.gap_start + (r - 1) * (sc.gap_score(0) - 1) if sc.penalize_start else 0
                src[r][c] = (r - 1, c)
    for r in range(1, nq + 1):
        bottom = r == nq
        qc = query[r - 1]
        Sr, Sa = S[r], S[r - 1]
        for c in range(1, nr + 1):
            end_of_row = c == nr
            sa = src[r - 1][c]
            vertical_cont = sa is not None and sa[1] == c
            v = Sa[c]
            if not end_of_row or sc.penalize_end:
                if vertical_cont:
                    assert iv[r - 1][c] is not None
                    v += sc.gap_score(iv[r - 1][c] + 1)
                else:
                    v += sc.gap_start + sc.gap_score(0)
            sl = src[r][c - 1]
            horizontal_cont = sl is not None and sl[0] == r
            h = Sr[c - 1]
            if not bottom or sc.penalize_end:
                if horizontal_cont:
                    assert ih[r][c - 1] is not None
                    h += sc.gap_score(ih[r][c - 1] + 1)
                else:
                    h += sc.gap_start + sc.gap_score(0)
            d = Sa[c - 1] + sc.match_score(qc, reference[c - 1])
            m = max(h, v, d)
            Sr[c] = m
            if m == h:
                src[r][c] = (r, c - 1)
                ih[r][c] = ih[r][c - 1] + 1 if ih[r][c - 1] is not None else 0
            elif m == v:
                src[r][c] = (r - 1, c)
                iv[r][c] = iv[r - 1][c] + 1 if iv[r - 1][c] is not None else 0
            else:
                src[r][c] = (r - 1, c - 1)
    a1, a2 = [], []
    r, c = nq, nr
    while src[r][c] is not None:
        pr, pc = src[r][c]
        a2.append(query[r - 1] if r - pr == 1 else GAP)
        a1.append(reference[c - 1] if c - pc == 1 else GAP)
        r, c = pr, pc
    return S[nq][nr], "".join(reversed(a1)), "".join(reversed(a2))


class AlignmentPair:
    """COMMON/alignment/AlignmentPair.java"""

    def __init__(self, ref_aln, qry_aln):
        self.ref, self.qry = ref_aln, qry_aln
        n = len(qry_aln)
        k = 0
        while k < n and qry_aln[k] == GAP:
            k += 1
        self.first_q = k
        k = n - 1
        while k >= 0 and qry_aln[k] == GAP:
            k -= 1
        self.last_q = k
        k = 0
        while k < len(ref_aln) and ref_aln[k] == GAP:
            k += 1
        self.first_r = k
        k = n - 1  # the reference starts this scan from the QUERY alignment's size (:58)
        while k >= 0 and ref_aln[k] == GAP:
            k -= 1
        self.last_r = k

    def sub(self, a, b):
        return AlignmentPair(sub_sequence(self.ref, a, b), sub_sequence(self.qry, a, b))

    def without_ending_query_inserts(self):
        return self.sub(0, self.last_q)

    def without_ending_and_beginning_query_inserts(self):
        return self.sub(self.first_q, self.last_q)

    def without_ending_and_beginning_reference_inserts(self):
        return self.sub(self.first_r, self.last_r)

    def expanded_cigar(self):
        """CigarStringUtil.getCigarString (:61-83)"""
        out = []
        for r, q in zip(self.ref, self.qry):
            if matches(r, q):
                out.append("=")
            elif r == GAP:
                out.append("I")
            elif q == GAP:
                out.append("D")
            else:
                out.append("X")
        return "".join(out)

    def standard_cigar(self, reversed_=False):
        e = self.expanded_cigar()
        if reversed_:
            e = e[::-1]
        return summarize(e.replace("X", "M").replace("=", "M"))

    def md(self):
        return mismatch_details(self.ref, self.qry, self.expanded_cigar())

    def reverse_md(self):
        return mismatch_details(self.ref[::-1], self.qry[::-1], self.expanded_cigar()[::-1])

    def edit_distance(self):
        return sum(1 for r, q in zip(self.ref, self.qry) if r != q)

    def n_mismatches(self):
        return sum(1 for r, q in zip(self.ref, self.qry) if r != GAP and q != GAP and not matches(r, q))

    def n_matches(self):
        return sum(1 for r, q in zip(self.ref, self.qry) if r != GAP and q != GAP and matches(r, q))

    def n_insertions(self):
        return sum(1 for r in self.ref if r == GAP)

    def n_deletions(self):
        return sum(1 for q in self.qry if q == GAP)

    def n_non_terminal_gaps(self):
        t = self.without_ending_and_beginning_reference_inserts().without_ending_and_beginning_query_inserts()
        return t.n_insertions() + t.n_deletions()


def summarize(cigar):
    """CigarStringUtil.summarizeCigarString (:206-229)"""
    out, k = [], 0
    while k < len(cigar):
        j = k
        while j < len(cigar) and cigar[j] == cigar[k]:
            j += 1
        out.append("%d%s" % (j - k, cigar[k]))
        k = j
    return "".join(out)


def reference_length(expanded):
    return sum(1 for ch in expanded if ch != "I")


def mismatch_details(ref_aln, qry_aln, expanded):
    """CigarStringUtil.getMismatchDetailsString(reference alignment, query alignment, cigar) (:98-160, entries :163-204)."""
    entries = []  # [type, count, sequence]
    index = 0
    for token in _tokens(summarize(expanded)):
        number, ch = token
        last = entries[-1] if entries else None
        if ch in "M=":
            if last is not None and last[0] in "M=":
                last[1] += number
            else:
                entries.append([ch, number, None])
        elif ch == "I":
            pass
        elif ch in "DX":
            entries.append([ch, number, sub_sequence(ref_aln, index, index + number - 1)])
        else:
            raise AssertionError(ch)
        index += number
    out = []
    prev = None
    for e in entries:
        is_match = e[0] in "M="
        if prev is None:
            if not is_match:
                out.append("0")
        elif prev[0] not in "M=" and not is_match:
            out.append("0")
        if is_match:
            out.append(str(e[1]))
        elif e[0] == "D":
            out.append("^" + e[2])
        else:
            out.append("0".join(e[2]))
        prev = e
    if prev is not None and prev[0] not in "M=":
        out.append("0")
    return "".join(out)


def _tokens(summarized):
    num = ""
    for ch in summarized:
        if ch.isdigit():
            num += ch
        else:
            yield int(num), ch
            num = ""


# ---- HSQ/process/ExtendReadsToPrimer.java ----------------------------------------------------------------------------------------
PRIMER_ALIGNMENT_BUFFER = 25
LENGTH_THRESHOLD_FOR_EXTENDED_READ = 10
WEIGHTED_ALIGNMENT_SCORE_THRESHOLD = 4.0


def primer_end_index_in_read(primer, read):
    """getPrimerEndIndexInRead (:286-345); None where the reference returns null."""
    read = sub_sequence(read, 0, min(len(read) - 1, len(primer) + PRIMER_ALIGNMENT_BUFFER))
    sc = Scorer(1, -4, -1, -6, True, False)
    _, read_aln, primer_aln = needleman_wunsch(read, primer, sc)
    k = 0
    while primer_aln[k] == GAP:  # (:296-300) IndexOutOfBounds when the primer alignment is all gaps
        k += 1
    rr, rp = read_aln[::-1], primer_aln[::-1]
    passed = False
    end = 0
    weighted = 0.0
    first = -1
    for i in range(len(rr)):
        if not passed and not matches(rp[i], GAP):
            passed = True
            first = i
        if passed and not matches(rr[i], GAP):
            end += 1
            if matches(rr[i], rp[i]):
                weighted += 0.95 ** float(i - first)  # Math.pow(0.95, distance)
    if weighted < WEIGHTED_ALIGNMENT_SCORE_THRESHOLD:
        return None
    return end - 1


def read_extension_details(primer, adjacent, capture, read, is_reversed, sc):
    """calculateDetailsForReadExtensionToPrimer (:200-245) -> dict or None."""
    e = primer_end_index_in_read(primer, read)
    out = {"primer_end": e, "score": 0}
    if e is None or not (0 <= e < len(read)):
        return None, out
    start = e + 1
    rest = sub_sequence(read, start, len(read))
    score, ra, qa = needleman_wunsch(capture, rest, sc)
    out["score"] = score
    n_cols = len(ra)
    normalized = float(score) / float(n_cols) if n_cols else float("nan")
    if not normalized > 0:
        return None, out
    aln = AlignmentPair(ra, qa).without_ending_query_inserts()
    read_length = len(rest) - 1
    if not read_length > LENGTH_THRESHOLD_FOR_EXTENDED_READ:
        return None, out
    if is_reversed:
        cigar = aln.standard_cigar(True)
        md = aln.reverse_md()
        pos = adjacent - reference_length(aln.expanded_cigar())
    else:
        cigar = aln.standard_cigar(False)
        md = aln.md()
        pos = adjacent + 1
    return {"pos": pos, "read_start": start, "read_stop": start + read_length, "cigar": cigar, "md": md, "nm": aln.edit_distance()}, out


def extend_read_pair(probe, r1, r2, sc=None):
    """extendReadPair (:96-198) up to the fields of the two records.  probe: dict with reverse (bool), ext, lig, target (strings),
    ext_start, ext_stop, lig_start, lig_stop.  Returns a list of two per-mate dicts (ok, primer_end, score, pos, read_start,
    read_stop, nm, tlen, cigar, md) and pair_ok."""
    sc = sc or Scorer()
    seq1, seq2 = codes(r1), complement(codes(r2))  # sequenceTwo = sequenceTwo.getCompliment() (:104)
    one_rev = probe["reverse"]
    two_rev = not one_rev
    adjacent = probe["ext_start"] if one_rev else probe["ext_stop"]
    target, ext, lig = codes(probe["target"]), codes(probe["ext"]), codes(probe["lig"])
    if one_rev:
        d1, x1 = read_extension_details(complement(ext), adjacent, complement(target), complement(seq1), one_rev, sc)
    else:
        d1, x1 = read_extension_details(ext, adjacent, target, seq1, one_rev, sc)
    if d1 is not None:  # the adjacent position of read two only moves when read one extended (:133-144)
        adjacent = probe["lig_stop"] if one_rev else probe["lig_start"]
    if one_rev:
        d2, x2 = read_extension_details(complement(reverse(lig)), adjacent, complement(reverse(target)), complement(seq2), two_rev, sc)
    else:
        d2, x2 = read_extension_details(reverse(lig), adjacent, reverse(target), seq2, two_rev, sc)
    tl = len(target)
    out = []
    for d, x, rev in ((d1, x1, one_rev), (d2, x2, two_rev)):
        m = {"ok": d is not None, "primer_end": x["primer_end"], "score": x["score"], "pos": 0, "read_start": 0, "read_stop": 0, "nm": 0, "tlen": 0,
             "cigar": "", "md": ""}
        if d is not None:
            m.update(d)
            m["tlen"] = -tl if rev else tl
        out.append(m)
    return out, d1 is not None and d2 is not None


# ---- HSQ/process/FilterByUid.java ---------------------------------------------------------------------------------------------------
def java_short(x):
    return ((x + 32768) % 65536) - 32768


def quality_score(q):
    """BamFileUtil.getQualityScore (HSQ/utils/BamFileUtil.java:74-85): short arithmetic."""
    s = 0
    for ch in q:
        s = java_short(s + java_short(ord(ch) - 33))
    return s


def reduce_by_uid(pairs, ext_uid_len, lig_uid_len):
    """FilterByUid.reduceReadsByProbeAndUid for the pairs of ONE probe (:134-203) with findBestData (:281-297).
    pairs: list of (read_index, r1, r2, q1, q2) in iteration order.  Returns (best flags, group sizes, unique, duplicates, max group)."""
    groups = {}
    for k, (ri, r1, r2, q1, q2) in enumerate(pairs):
        uid = r1[:ext_uid_len] + r2[:lig_uid_len]
        groups.setdefault(uid, []).append(k)
    best = [0] * len(pairs)
    size = [0] * len(pairs)
    for uid, members in groups.items():
        best_k, best_score, best_id = None, None, None
        for k in members:
            ri, r1, r2, q1, q2 = pairs[k]
            score = java_short(quality_score(q1) + quality_score(q2))
            if best_k is None or score > best_score or (score == best_score and ri < best_id):
                best_k, best_score, best_id = k, score, ri
        best[best_k] = 1
        for k in members:
            size[k] = len(members)
    n_groups = len(groups)
    return best, size, n_groups, len(pairs) - n_groups, max((len(m) for m in groups.values()), default=0)


# ---- multi-probe tie-break (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:704-765) -----------------------------
def tie_break(r1, r2, probes, ext_uid_len, lig_uid_len):
    """probes: list of (ext, target, lig) strings of the candidate probes.  Returns (index of the single best probe or -1, matches, edits)."""
    most, lowest = 0, None
    best = set()
    ms, es = [], []
    for k, (ext, target, lig) in enumerate(probes):
        probe_seq = codes(ext) + codes(target) + codes(lig)
        first = codes(r1)[ext_uid_len:]
        _, ra, qa = needleman_wunsch(first, probe_seq)
        a1 = AlignmentPair(ra, qa)
        edit = a1.first_q + a1.n_mismatches() + a1.n_non_terminal_gaps()
        second = codes(r2)[lig_uid_len:]
        _, ra, qa = needleman_wunsch(second, complement(reverse(probe_seq)))
        a2 = AlignmentPair(ra, qa)
        edit += a2.first_q + a2.n_mismatches() + a2.n_non_terminal_gaps()
        m = a1.n_matches() + a2.n_matches()
        ms.append(m)
        es.append(edit)
        if m > most:
            best = {k}
            lowest, most = edit, m
        elif m == most:
            if lowest is not None and edit == lowest:
                best.add(k)
            elif lowest is None or edit < lowest:
                best = {k}
                lowest, most = edit, m
    return (next(iter(best)) if len(best) == 1 else -1), ms, es


# ---- A10: HSQ/utils/SAMRecordUtil.java:223-261 ------------------------------------------------------------------------------------
def variable_length_uid(read, primer, sc=None):
    """getVariableLengthUid: the read prefix in front of the first primer-aligned column, None where the reference returns null.
    (insertsSinceLastInsertionOrMismatch is never incremented there, :238-250: every deletion column counts against the cutoff.)"""
    sc = sc or Scorer()
    _, ra, qa = needleman_wunsch(codes(read), codes(primer), sc)
    end = -1
    for i, ch in enumerate(qa):  # getIndexOfFirstMatchInReference reads the QUERY alignment (:412-428)
        if not matches(ch, GAP):
            end = i
            break
    uid = read[:end] if end >= 0 else None
    ins = dele = sub = 0
    since = 0
    for ch in AlignmentPair(ra, qa).expanded_cigar():
        if ch == "X":
            sub += 1
            since = 0
        elif ch == "I":
            ins += 1
            since = 0
        elif ch == "D":
            dele += 1
    dele -= since
    if ins + dele + sub >= len(codes(primer)) // 4:
        uid = None
    return uid
