// hsq_oracle_align.hpp -- TEST INFRASTRUCTURE ONLY.  See hsq_oracle_core.hpp for the header note.
// Needleman-Wunsch restatement + CIGAR / MD derivation.
#pragma once
#include "hsq_oracle_core.hpp"

namespace hsqo {

// ---------------------------------------------------------------------------------------------
// NeedlemanWunschGlobalAlignment.  COMMON/alignment/NeedlemanWunschGlobalAlignment.java
//   matrix rows = query (+1), cols = reference (+1)                         :44-49
//   init   (getInitialScore / getInitialSourceCell)                         :342-376
//   fill   (completeCell, row-major)                                        :72-78, :261-328
//   traceback from (len(q),len(r)) until the source pointer is null         :85-114, :334-340
// With SimpleAlignmentScorer the boxed gap indices carry no information beyond the pointer:
// a continuation costs getGapScore(k+1)=gapExtend, a start costs gapStart+getGapScore(0)=gapStart.
// ---------------------------------------------------------------------------------------------
struct NwResult {
    int score = 0;
    AlignmentPair pair;
    int64_t cells = 0;
};

struct NwWorkspace {  // reused between calls so the timed CPU baseline has no per-call allocation
    std::vector<int> S;
    std::vector<uint8_t> P;
};

inline NwResult nw_align(const seq_t& ref, const seq_t& qry, const Scorer& sc, NwWorkspace* wsp = nullptr) {
    NwWorkspace local;
    NwWorkspace& ws = wsp ? *wsp : local;
    const int R = (int)ref.size(), Q = (int)qry.size();
    const int W = R + 1;
    ws.S.assign((size_t)(Q + 1) * W, 0);
    ws.P.assign((size_t)(Q + 1) * W, PTR_NONE);
    int* S = ws.S.data();
    uint8_t* P = ws.P.data();
    // :342-376 -- note (k-1)*(gapScore(0)-1), NOT gapExtend
    for (int j = 1; j <= R; j++) {
        S[j] = sc.penalize_start ? sc.gap_start + (j - 1) * (sc.gap_score(0) - 1) : 0;
        P[j] = PTR_LEFT;
    }
    for (int i = 1; i <= Q; i++) {
        S[(size_t)i * W] = sc.penalize_start ? sc.gap_start + (i - 1) * (sc.gap_score(0) - 1) : 0;
        P[(size_t)i * W] = PTR_UP;
    }
    for (int i = 1; i <= Q; i++) {
        const bool bottom = (i == Q);
        for (int j = 1; j <= R; j++) {
            const bool end_of_row = (j == R);
            const size_t up = (size_t)(i - 1) * W + j, left = (size_t)i * W + j - 1, diag = (size_t)(i - 1) * W + j - 1;
            // vertical: source of cellAbove is in the same column <=> its pointer is UP (:270-284)
            int v = S[up];
            if (!end_of_row || sc.penalize_end) v += (P[up] == PTR_UP) ? sc.gap_score(1) : sc.gap_start + sc.gap_score(0);
            // horizontal (:286-300)
            int h = S[left];
            if (!bottom || sc.penalize_end) h += (P[left] == PTR_LEFT) ? sc.gap_score(1) : sc.gap_start + sc.gap_score(0);
            int d = S[diag] + (code_matches(qry[i - 1], ref[j - 1]) ? sc.match : sc.mismatch);  // :302-303
            int m = std::max(h, std::max(v, d));                                             // :305
            size_t cur = (size_t)i * W + j;
            S[cur] = m;
            P[cur] = (m == h) ? PTR_LEFT : (m == v) ? PTR_UP : PTR_DIAG;  // :309-327  tie order H > V > D
        }
    }
    NwResult out;
    out.cells = (int64_t)Q * R;
    out.score = S[(size_t)Q * W + R];
    seq_t a1, a2;  // reference alignment, query alignment (built backwards, then reversed :110)
    int i = Q, j = R;
    while (P[(size_t)i * W + j] != PTR_NONE) {
        uint8_t p = P[(size_t)i * W + j];
        if (p == PTR_LEFT) { a2.push_back(GAP); a1.push_back(ref[j - 1]); j--; }
        else if (p == PTR_UP) { a2.push_back(qry[i - 1]); a1.push_back(GAP); i--; }
        else { a2.push_back(qry[i - 1]); a1.push_back(ref[j - 1]); i--; j--; }
    }
    out.pair = AlignmentPair(seq_reverse(a1), seq_reverse(a2));
    return out;
}

// getIndexOfFirstMatchInReference (:412-428): first non-GAP column of the QUERY alignment, -1 if none.
inline int index_of_first_match_in_reference(const AlignmentPair& p) {
    for (int i = 0; i < (int)p.qry_aln.size(); i++) if (p.qry_aln[i] != GAP) return i;
    return -1;
}

// ---------------------------------------------------------------------------------------------
// CigarStringUtil.getCigarString (COMMON/alignment/CigarStringUtil.java:61-83): expanded =/I/D/X
// ---------------------------------------------------------------------------------------------
inline std::string expanded_cigar(const seq_t& ref_aln, const seq_t& qry_aln) {
    std::string s;
    for (size_t i = 0; i < ref_aln.size(); i++) {
        code_t r = ref_aln[i], q = qry_aln[i];
        if (code_matches(r, q)) s.push_back('=');
        else if (r == GAP) s.push_back('I');
        else if (q == GAP) s.push_back('D');
        else s.push_back('X');
    }
    return s;
}
inline std::string expanded_cigar(const AlignmentPair& p) { return expanded_cigar(p.ref_aln, p.qry_aln); }

// summarizeCigarString (:206-229)
inline std::string summarize_cigar(const std::string& c) {
    std::string o;
    if (c.empty()) return o;
    int n = 1;
    char cur = c[0];
    for (size_t i = 1; i < c.size(); i++) {
        if (c[i] == cur) n++;
        else { o += std::to_string(n); o.push_back(cur); n = 1; cur = c[i]; }
    }
    o += std::to_string(n);
    o.push_back(cur);
    return o;
}
// CigarString.getStandardCigarString (CigarString.java:75-77): =,X -> M then summarize
inline std::string standard_cigar(const std::string& expanded) {
    std::string t = expanded;
    for (auto& ch : t) if (ch == 'X' || ch == '=') ch = 'M';
    return summarize_cigar(t);
}
inline int cigar_length_of_reference(const std::string& expanded) {  // CigarString.java:61-70
    int n = 0;
    for (char ch : expanded) if (ch != 'I') n++;
    return n;
}
inline int cigar_edit_distance(const std::string& expanded) {  // CigarStringUtil.java:231-247
    int e = 0;
    for (char ch : expanded) if (ch == 'I' || ch == 'X') e++;
    return e;
}

// getMismatchDetailsString(refAln, qryAln, cigar) (CigarStringUtil.java:98-160, entries :162-204).
// The cigar is a separate argument: the reference's own KATs pass an untrimmed cigar with a
// trimmed alignment (CigarStringUtilTest.java:75-96) and rely on subSequence's end clamp.
inline std::string md_string(const seq_t& ref_aln, const std::string& expanded) {
    struct Entry { char type; int count; seq_t seq; };
    std::vector<Entry> entries;
    std::string summarized = summarize_cigar(expanded);
    int aln_index = 0;
    std::string num;
    for (char ch : summarized) {
        if (ch >= '0' && ch <= '9') { num.push_back(ch); continue; }
        int number = std::stoi(num);
        num.clear();
        bool last_is_match = !entries.empty() && (entries.back().type == 'M' || entries.back().type == '=');
        if (ch == 'M' || ch == '=') {
            if (last_is_match) entries.back().count += number;
            else entries.push_back({ch, number, {}});
        } else if (ch == 'I') {
        } else if (ch == 'D' || ch == 'X') {
            entries.push_back({ch, number, seq_sub(ref_aln, aln_index, aln_index + number - 1)});
        } else {
            throw std::logic_error("Unrecognized character in cigar string");
        }
        aln_index += number;
    }
    std::string out;
    const Entry* prev = nullptr;
    for (const Entry& e : entries) {
        bool is_match = (e.type == 'M' || e.type == '=');
        if (!prev) { if (!is_match) out += "0"; }
        else {
            bool prev_match = (prev->type == 'M' || prev->type == '=');
            if (!prev_match && !is_match) out += "0";
        }
        if (is_match) out += std::to_string(e.count);
        else if (e.type == 'D') { out += "^"; out += decode_seq(e.seq); }
        else {  // 'X': a zero between consecutive mismatches (:182-191)
            for (size_t i = 0; i < e.seq.size(); i++) {
                out.push_back(decode_code(e.seq[i]));
                if (i + 1 != e.seq.size()) out += "0";
            }
        }
        prev = &e;
    }
    if (prev && !(prev->type == 'M' || prev->type == '=')) out += "0";
    return out;
}
inline std::string md_string(const AlignmentPair& p) { return md_string(p.ref_aln, expanded_cigar(p)); }
// AlignmentPair.getReverseMismatchDetailsString (AlignmentPair.java:157-162)
inline std::string reverse_md_string(const AlignmentPair& p) {
    std::string e = expanded_cigar(p);
    std::reverse(e.begin(), e.end());
    return md_string(seq_reverse(p.ref_aln), e);
}

}  // namespace hsqo
