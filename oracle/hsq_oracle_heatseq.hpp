// hsq_oracle_heatseq.hpp -- TEST INFRASTRUCTURE ONLY.  See hsq_oracle_core.hpp for the header note.
// Restatement of ExtendReadsToPrimer, FilterByUid, FastqReadTrimmer.trim and the multi-probe tie-break.
#pragma once
#include "hsq_oracle_align.hpp"
#include <map>

namespace hsqo {

static const int PRIMER_ALIGNMENT_BUFFER = 25;                 // HSQ/process/ExtendReadsToPrimer.java:54
static const int LENGTH_THRESHOLD_FOR_EXTENDED_READ = 10;      // :56
static const double WEIGHTED_ALIGNMENT_SCORE_THRESHOLD = 4.0;  // :58
static const int PRIMER_NOT_FOUND = -0x7fffffff;               // stands for the Java `null`

// getPrimerEndIndexInRead (HSQ/process/ExtendReadsToPrimer.java:286-345).  Returns PRIMER_NOT_FOUND for null.
// *weighted_out receives the FP64 gate sum (for near-threshold diagnostics).
inline int primer_end_index_in_read(const seq_t& primer, const seq_t& read_full, NwWorkspace* ws, int64_t* cells,
                                    double* weighted_out = nullptr) {
    seq_t read = seq_sub(read_full, 0, std::min((int)read_full.size() - 1, (int)primer.size() + PRIMER_ALIGNMENT_BUFFER));  // :288
    Scorer sc;  // defaults 1,-4,-1,-6 with penalizeStart=true, penalizeEnd=false (:290-291)
    sc.penalize_start = true;
    sc.penalize_end = false;
    NwResult r = nw_align(read, primer, sc, ws);  // reference = read prefix, query = primer (:293)
    if (cells) *cells += r.cells;
    const seq_t& read_aln = r.pair.ref_aln;
    const seq_t& primer_aln = r.pair.qry_aln;
    {  // :297-301 -- throws IndexOutOfBounds when the primer alignment is all gaps (empty primer)
        size_t k = 0;
        while (true) {
            if (k >= primer_aln.size()) throw std::out_of_range("primer alignment is all gaps");
            if (primer_aln[k] != GAP) break;
            k++;
        }
    }
    bool passed = false;
    int n = (int)read_aln.size();
    int primer_end = 0;
    double weighted = 0;
    int first_non_gap = -1;
    for (int i = 0; i < n; i++) {  // i indexes the REVERSED alignments (:305-333)
        code_t rc = read_aln[n - 1 - i], pc = primer_aln[n - 1 - i];
        if (!passed && pc != GAP) { passed = true; first_non_gap = i; }
        if (passed && rc != GAP) {
            primer_end++;
            if (code_matches(rc, pc)) weighted += std::pow(0.95, (double)(i - first_non_gap));
        }
    }
    if (weighted_out) *weighted_out = weighted;
    if (weighted < WEIGHTED_ALIGNMENT_SCORE_THRESHOLD) return PRIMER_NOT_FOUND;  // :335-337
    return primer_end - 1;                                                          // :339-342
}

struct ReadEndDetails {
    bool ok = false;
    int primer_end = PRIMER_NOT_FOUND;
    int score = 0;  // DP#2 score (0 when DP#2 was not run)
    int pos = 0, read_start = 0, read_stop = 0, nm = 0;
    std::string cigar;  // standard (M/I/D) summarized
    std::string md;
};

// calculateDetailsForReadExtensionToPrimer (HSQ/process/ExtendReadsToPrimer.java:200-245)
inline ReadEndDetails extend_read_end(const seq_t& primer, int adjacent_pos, const seq_t& capture, const seq_t& read,
                                      bool is_reversed, const Scorer& user, NwWorkspace* ws, int64_t* cells) {
    ReadEndDetails d;
    int e = primer_end_index_in_read(primer, read, ws, cells);
    d.primer_end = e;
    bool primer_ok = (e != PRIMER_NOT_FOUND) && e >= 0 && e < (int)read.size();  // :206
    if (!primer_ok) return d;
    int start_in_read = e + 1;
    seq_t rw = seq_sub(read, start_in_read, (int)read.size());  // :210 (end clamps to size-1)
    NwResult r = nw_align(capture, rw, user, ws);                 // :211
    if (cells) *cells += r.cells;
    d.score = r.score;
    int aln_len = r.pair.size();
    // getLengthNormalizedAlignmentScore() > 0 (:212, NW :119-126); 0/0 = NaN -> false
    bool aligned_ok = aln_len > 0 && ((double)r.score / (double)aln_len) > 0.0;
    if (!aligned_ok) return d;
    AlignmentPair a = r.pair.without_ending_query_inserts();  // :218
    int read_length = (int)rw.size() - 1;                     // :225
    if (!(read_length > LENGTH_THRESHOLD_FOR_EXTENDED_READ)) return d;
    std::string exp = expanded_cigar(a);
    d.nm = a.edit_distance();  // :222
    if (is_reversed) {         // :228-234
        std::string rexp = exp;
        std::reverse(rexp.begin(), rexp.end());
        d.cigar = standard_cigar(rexp);
        d.md = reverse_md_string(a);
        d.pos = adjacent_pos - cigar_length_of_reference(rexp);
    } else {
        d.cigar = standard_cigar(exp);
        d.md = md_string(a);
        d.pos = adjacent_pos + 1;  // :236
    }
    d.read_start = start_in_read;
    d.read_stop = start_in_read + read_length;  // ReadExtensionDetails.getReadStop (:404-406)
    d.ok = true;
    return d;
}

struct Probe {  // COMMON/utils/probeinfo/Probe.java:53-69
    bool reverse_strand = false;
    int ext_start = 0, ext_stop = 0, lig_start = 0, lig_stop = 0, target_start = 0, target_stop = 0;
    seq_t ext, lig, target;
};

struct PairExtension {
    ReadEndDetails one, two;
    std::string seq1, qual1, seq2, qual2;  // record read string / base qualities (createRecord :247-284)
    int tlen1 = 0, tlen2 = 0;
    bool neg1 = false, neg2 = false;
    bool pair_ok() const { return one.ok && two.ok; }
};

// extendReadPair (HSQ/process/ExtendReadsToPrimer.java:96-198)
inline PairExtension extend_read_pair(const Probe& p, const seq_t& seq1, const std::string& q1, const seq_t& seq2_raw,
                                      const std::string& q2, const Scorer& user, NwWorkspace* ws, int64_t* cells) {
    PairExtension out;
    seq_t seq2 = seq_complement(seq2_raw);  // :104
    bool one_rev = p.reverse_strand, two_rev = !one_rev;
    int adjacent = one_rev ? p.ext_start : p.ext_stop;  // :109-112
    seq_t capture, ext_primer, one_c;
    if (one_rev) { capture = seq_complement(p.target); ext_primer = seq_complement(p.ext); one_c = seq_complement(seq1); }
    else { capture = p.target; ext_primer = p.ext; one_c = seq1; }
    out.one = extend_read_end(ext_primer, adjacent, capture, one_c, one_rev, user, ws, cells);  // :128
    out.neg1 = one_rev;
    out.neg2 = two_rev;
    if (out.one.ok) {
        seq_t s = seq_sub(seq1, out.one.read_start, out.one.read_stop);
        std::string q = q1.substr((size_t)out.one.read_start, (size_t)(out.one.read_stop + 1 - out.one.read_start));
        out.tlen1 = (int)p.target.size();
        adjacent = p.lig_start;  // :137
        if (one_rev) {
            s = seq_revcomp(s);
            std::reverse(q.begin(), q.end());
            out.tlen1 = -out.tlen1;
            adjacent = p.lig_stop;  // :143
        }
        out.seq1 = decode_seq(s);
        out.qual1 = q;
    }
    seq_t lig_primer, two_c;
    if (one_rev) { lig_primer = seq_revcomp(p.lig); two_c = seq_complement(seq2); capture = seq_revcomp(p.target); }  // :155-157
    else { lig_primer = seq_reverse(p.lig); two_c = seq2; capture = seq_reverse(p.target); }                           // :159-161
    out.two = extend_read_end(lig_primer, adjacent, capture, two_c, two_rev, user, ws, cells);                        // :164
    if (out.two.ok) {
        seq_t s = seq_complement(seq_sub(seq2, out.two.read_start, out.two.read_stop));  // :170
        std::string q = q2.substr((size_t)out.two.read_start, (size_t)(out.two.read_stop + 1 - out.two.read_start));
        out.tlen2 = (int)p.target.size();
        if (two_rev) {
            s = seq_revcomp(s);
            std::reverse(q.begin(), q.end());
            out.tlen2 = -out.tlen2;
        }
        out.seq2 = decode_seq(s);
        out.qual2 = q;
    }
    return out;
}

// BamFileUtil.getQualityScore (HSQ/utils/BamFileUtil.java:74-85): Java short arithmetic, wraps.
inline int16_t quality_score(const char* q, int64_t n) {
    int16_t sum = 0;
    for (int64_t i = 0; i < n; i++) {
        int16_t base = (int16_t)((int16_t)(uint8_t)q[i] - 33);
        sum = (int16_t)(sum + base);
    }
    return sum;
}
// ReadPair.getTotalSequenceQualityScore (HSQ/objects/ReadPair.java:101-103)
inline int16_t total_quality_score(const char* q1, int64_t n1, const char* q2, int64_t n2) {
    return (int16_t)(quality_score(q1, n1) + quality_score(q2, n2));
}

// FastqReadTrimmer.trim (HSQ/process/FastqReadTrimmer.java:309-328).  Returns [first, last] inclusive kept range.
inline void trim_range(int read_len, int first_to_keep, int last_to_keep, bool three_prime, int* first, int* last) {
    if (first_to_keep >= read_len) throw std::invalid_argument("Unable to trim bases from the beginning of a sequence");
    int lb = read_len - 1;
    if (three_prime) lb = std::min(last_to_keep, read_len - 1);
    *first = first_to_keep;
    *last = lb;  // substring(first, last+1): Java throws if last+1 < first
    if (lb + 1 < first_to_keep) throw std::out_of_range("substring end < begin");
}

// Multi-probe tie-break metrics for one (read pair, candidate probe)
// (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:712-734)
inline void tie_break_metrics(const Probe& p, const seq_t& r1, const seq_t& r2, int ext_uid_len, int lig_uid_len,
                              NwWorkspace* ws, int64_t* cells, int* matches, int* edit) {
    seq_t probe_seq = p.ext;  // Probe.getProbeSequence (Probe.java:83-88)
    probe_seq.insert(probe_seq.end(), p.target.begin(), p.target.end());
    probe_seq.insert(probe_seq.end(), p.lig.begin(), p.lig.end());
    Scorer def;
    NwResult a1 = nw_align(seq_sub(r1, ext_uid_len, (int)r1.size() - 1), probe_seq, def, ws);
    NwResult a2 = nw_align(seq_sub(r2, lig_uid_len, (int)r2.size() - 1), seq_revcomp(probe_seq), def, ws);
    if (cells) *cells += a1.cells + a2.cells;
    *edit = a1.pair.first_q + a1.pair.number_of_mismatches() + a1.pair.number_of_non_terminal_gaps() + a2.pair.first_q +
            a2.pair.number_of_mismatches() + a2.pair.number_of_non_terminal_gaps();
    *matches = a1.pair.number_of_matches() + a2.pair.number_of_matches();
}

// mergeSequences (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:1362-1552).  Sequences are ASCII; every
// output base is the canonical letter of its code (IupacNucleotideCode.toString), as the Java builds the result from codes.
// Throws std::out_of_range where the Java throws (substring / subSequence bounds).
inline void merge_sequences(bool negative_strand, const std::string& up, const std::string& down, const std::string& upq,
                            const std::string& downq, int up_start, int down_start, std::string* out_seq, std::string* out_qual) {
    if (upq.size() != up.size() || downq.size() != down.size()) throw std::out_of_range("quality length differs from sequence length");
    const int ul = (int)up.size(), dl = (int)down.size();
    const int up_end = up_start + ul - 1, down_end = down_start + dl - 1;  // :1371-1374
    int lead_s = 0, lead_e = 0, lead_len = 0, ov_s = 0, ov_e = 0, ov_len = 0, tr_s = 0, tr_e = 0, tr_len = 0;
    const int min_end = std::min(up_end, down_end), max_end = std::max(up_end, down_end);  // :1388-1389
    const bool trail_from_down = (down_end == max_end);                                    // :1394-1402
    const std::string& tsrc = trail_from_down ? down : up;
    const std::string& tsrcq = trail_from_down ? downq : upq;
    const int tsrc_start = trail_from_down ? down_start : up_start;
    if (down_start <= up_end) {  // :1405-1421
        ov_s = down_start;
        ov_e = min_end;
        ov_len = ov_e - ov_s + 1;
        if (up_start < ov_s) { lead_s = up_start; lead_e = ov_s - 1; lead_len = lead_e - lead_s + 1; }
        if (max_end > ov_e) { tr_s = ov_e + 1; tr_e = max_end; tr_len = tr_e - tr_s + 1; }
    } else {  // :1422-1430
        lead_s = up_start; lead_e = lead_s + ul - 1; lead_len = lead_e - lead_s + 1;
        tr_s = down_start; tr_e = tr_s + dl - 1; tr_len = tr_e - tr_s + 1;
    }
    auto canon = [](const std::string& x) {
        std::string r(x.size(), '?');
        for (size_t k = 0; k < x.size(); k++) {
            int c = encode_char(x[k]);
            if (c < 0) throw std::invalid_argument("not an IUPAC code");
            r[k] = decode_code((code_t)c);
        }
        return r;
    };
    auto sub = [](const std::string& x, int a, int b_incl) {  // String.substring(a, b+1) / subSequence(a, b) bounds
        if (a < 0 || b_incl + 1 > (int)x.size() || a > b_incl + 1) throw std::out_of_range("substring out of bounds");
        return x.substr((size_t)a, (size_t)(b_incl + 1 - a));
    };
    std::string lead, leadq, ov, ovq, trail, trailq;
    if (ov_len == 0) {  // :1449-1468
        const int gs = lead_e + 1, ge = tr_s - 1;
        const int fill = ge >= gs ? ge - gs + 1 : 0;
        if (fill > 0) { ov.assign((size_t)fill, 'N'); ovq.assign((size_t)fill, '!'); ov_len = fill; }
        lead = canon(up); leadq = upq; trail = canon(down); trailq = downq;
    } else if (ov_len >= 1) {  // :1469-1521
        if (lead_len >= 1) { lead = canon(sub(up, lead_s - up_start, lead_e - up_start)); leadq = sub(upq, lead_s - up_start, lead_e - up_start); }
        const std::string uo = canon(sub(up, ov_s - up_start, ov_e - up_start)), uoq = sub(upq, ov_s - up_start, ov_e - up_start);
        const std::string dox = canon(sub(down, ov_s - down_start, ov_e - down_start)), doq = sub(downq, ov_s - down_start, ov_e - down_start);
        for (int i = 0; i < ov_len; i++) {
            if (i + 1 > (int)downq.size()) throw std::out_of_range("substring out of bounds");
            const int uqs = (int)(unsigned char)uoq[(size_t)i] - 33;
            const int dqs = (int)(unsigned char)downq[(size_t)i] - 33;  // :1506 indexes the un-sliced downstream quality
            if (uqs >= dqs) { ov += uo[(size_t)i]; ovq += uoq[(size_t)i]; }
            else { ov += dox[(size_t)i]; ovq += doq[(size_t)i]; }
        }
        if (tr_len >= 1) { trail = canon(sub(tsrc, tr_s - tsrc_start, tr_e - tsrc_start)); trailq = sub(tsrcq, tr_s - tsrc_start, tr_e - tsrc_start); }
    } else {
        throw std::out_of_range("negative overlap");
    }
    std::string ms = (lead_len > 0 ? lead : "") + (ov_len > 0 ? ov : "") + (tr_len > 0 ? trail : "");      // :1530-1542
    std::string mq = (lead_len > 0 ? leadq : "") + (ov_len > 0 ? ovq : "") + (tr_len > 0 ? trailq : "");
    if (negative_strand) {  // :1544-1548
        std::string rc(ms.rbegin(), ms.rend());
        for (auto& ch : rc) ch = decode_code(complement_code((code_t)encode_char(ch)));
        ms = rc;
        mq = std::string(mq.rbegin(), mq.rend());
    }
    *out_seq = ms;
    *out_qual = mq;
}

}  // namespace hsqo
