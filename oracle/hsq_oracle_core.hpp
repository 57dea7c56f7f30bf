// hsq_oracle_core.hpp -- TEST INFRASTRUCTURE ONLY (parity oracle / timed CPU baseline).
//
// CPU restatement (C++17, no dependencies) of the HSQUtils hot path of NimbleGen/bioinformatics.
// Nothing in the product path (bioinformatics_b200/) may include, link or call this file; only
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
//
// Parity status: the alignment core (nw_align, cigar/MD/edit distance) is PINNED against the
// reference's own 9 alignment known-answer tests (CigarStringUtilTest, AlignmentTest) and the
// trim KATs (FastqReadTrimmerTest) -- see tests/test_oracle_kat.py.  primer_end_index_in_read,
// extend_read_end, dedup and the multi-probe tie-break are "parity unpinned" by the reference
// (its tests for them are absent or commented out); they are line-by-line
// restatements of the cited reference lines.  The JVM reference cannot run in this image.
//
// Reference roots (under /root/reference):
//   COMMON/ = bioinformatics_common/src/main/java/com/roche/sequencing/bioinformatics/common/
//   HSQ/    = nimblegen_heatseq/src/main/java/com/roche/heatseq/
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include <stdexcept>
#include <algorithm>
#include <cmath>

namespace hsqo {

// ---------------------------------------------------------------------------------------------
// Codes.  One byte per code: low 4 bits = nucleotide set (A=1,C=2,G=4,T=8), bit 4 marks 'U'
// (same nucleotide set as T but a distinct enum constant, so equals() differs while matches()
// agrees).  GAP = 0 (empty nucleotide set).   COMMON/sequence/IupacNucleotideCode.java:31-39
// ---------------------------------------------------------------------------------------------
typedef uint8_t code_t;
static const code_t GAP = 0;
static const code_t CODE_U = 0x18;

inline int encode_char(char ch) {  // IupacNucleotideCode.getCodesFromString: toUpperCase + lookup (:86-104)
    switch (ch) {
        case 'A': case 'a': return 1;
        case 'C': case 'c': return 2;
        case 'G': case 'g': return 4;
        case 'T': case 't': return 8;
        case 'U': case 'u': return CODE_U;
        case 'R': case 'r': return 1 | 4;
        case 'Y': case 'y': return 2 | 8;
        case 'S': case 's': return 4 | 2;
        case 'W': case 'w': return 1 | 8;
        case 'K': case 'k': return 4 | 8;
        case 'M': case 'm': return 1 | 2;
        case 'B': case 'b': return 2 | 4 | 8;
        case 'D': case 'd': return 1 | 4 | 8;
        case 'H': case 'h': return 1 | 2 | 8;
        case 'V': case 'v': return 1 | 2 | 4;
        case 'N': case 'n': return 15;
        case '_': case '.': case '-': return GAP;
        default: return -1;
    }
}

inline char decode_code(code_t c) {  // toString(): abbreviations[0]; GAP prints as "_" (:37,:108-114)
    static const char tbl[16] = {'_', 'A', 'C', 'M', 'G', 'R', 'S', 'V', 'T', 'W', 'Y', 'H', 'K', 'D', 'B', 'N'};
    if (c == CODE_U) return 'U';
    return tbl[c & 15];
}

// getComplimentCode (:208-211): A<->T, C<->G, U->A, R<->Y, K<->M, B<->V, D<->H, S,W,N,GAP fixed.
// On nucleotide sets that is the swap A<->T, C<->G, i.e. a reversal of the 4 mask bits.
inline code_t complement_code(code_t c) {
    code_t m = c & 15;
    return (code_t)(((m & 1) << 3) | ((m & 2) << 1) | ((m & 4) >> 1) | ((m & 8) >> 3));
}

// IupacNucleotideCode.matches (:185-200): equals, else any shared nucleotide.  GAP matches only GAP.
inline bool code_matches(code_t a, code_t b) { return a == b || ((a & b & 15) != 0); }

typedef std::vector<code_t> seq_t;

inline seq_t encode_string(const char* s, int64_t n) {
    seq_t out((size_t)n);
    for (int64_t i = 0; i < n; i++) {
        int c = encode_char(s[i]);
        if (c < 0) throw std::invalid_argument(std::string("Unable to translate character[") + s[i] + "]");
        out[(size_t)i] = (code_t)c;
    }
    return out;
}
inline seq_t encode_string(const std::string& s) { return encode_string(s.data(), (int64_t)s.size()); }
inline std::string decode_seq(const seq_t& s) {
    std::string o(s.size(), '?');
    for (size_t i = 0; i < s.size(); i++) o[i] = decode_code(s[i]);
    return o;
}
inline seq_t seq_reverse(const seq_t& s) { return seq_t(s.rbegin(), s.rend()); }
inline seq_t seq_complement(const seq_t& s) {
    seq_t o(s.size());
    for (size_t i = 0; i < s.size(); i++) o[i] = complement_code(s[i]);
    return o;
}
inline seq_t seq_revcomp(const seq_t& s) { return seq_complement(seq_reverse(s)); }
// ISequence.subSequence(start,end): inclusive, end clamped to size-1, empty when start>end
// (COMMON/sequence/IupacNucleotideCodeSequence.java:103-117)
inline seq_t seq_sub(const seq_t& s, int start, int end) {
    end = std::min(end, (int)s.size() - 1);
    seq_t o;
    for (int i = start; i <= end; i++) {
        if (i < 0) throw std::out_of_range("subSequence index < 0");
        o.push_back(s[(size_t)i]);
    }
    return o;
}

// ---------------------------------------------------------------------------------------------
// Scorer.  COMMON/alignment/SimpleAlignmentScorer.java:23-96.  Integer valued in every reachable
// configuration, so int arithmetic is exact.
// ---------------------------------------------------------------------------------------------
struct Scorer {
    int match = 1, mismatch = -4, gap_extend = -1, gap_start = -6;
    bool penalize_start = false, penalize_end = false;
    int gap_score(int index_in_gap) const { return index_in_gap == 0 ? 0 : gap_extend; }  // :73-81
};

enum : uint8_t { PTR_NONE = 3, PTR_DIAG = 0, PTR_UP = 1, PTR_LEFT = 2 };

// ---------------------------------------------------------------------------------------------
// AlignmentPair.  COMMON/alignment/AlignmentPair.java:39-64 (ctor computes the four indices).
// ---------------------------------------------------------------------------------------------
struct AlignmentPair {
    seq_t ref_aln, qry_aln;
    int first_q = 0, last_q = -1, first_r = 0, last_r = -1;
    AlignmentPair() {}
    AlignmentPair(const seq_t& r, const seq_t& q) : ref_aln(r), qry_aln(q) {
        int n = (int)qry_aln.size();
        first_q = 0;
        while (first_q < n && qry_aln[first_q] == GAP) first_q++;
        last_q = n - 1;
        while (last_q >= 0 && qry_aln[last_q] == GAP) last_q--;
        first_r = 0;
        while (first_r < (int)ref_aln.size() && ref_aln[first_r] == GAP) first_r++;
        last_r = n - 1;  // sic: starts from queryAlignment.size()-1 (:58)
        while (last_r >= 0 && ref_aln[last_r] == GAP) last_r--;
    }
    int size() const { return (int)ref_aln.size(); }
    AlignmentPair sub(int a, int b) const { return AlignmentPair(seq_sub(ref_aln, a, b), seq_sub(qry_aln, a, b)); }
    AlignmentPair without_ending_query_inserts() const { return sub(0, last_q); }                        // :118-121
    AlignmentPair without_ending_and_beginning_query_inserts() const { return sub(first_q, last_q); }    // :95-99
    AlignmentPair without_ending_and_beginning_reference_inserts() const { return sub(first_r, last_r); }  // :111-114
    int edit_distance() const {  // :144-152  (equals, not matches)
        int e = 0;
        for (int i = 0; i < size(); i++) if (ref_aln[i] != qry_aln[i]) e++;
        return e;
    }
    int number_of_mismatches() const {  // :246-256
        int c = 0;
        for (int i = 0; i < size(); i++)
            if (ref_aln[i] != GAP && qry_aln[i] != GAP && !code_matches(ref_aln[i], qry_aln[i])) c++;
        return c;
    }
    int number_of_matches() const {  // :280-290
        int c = 0;
        for (int i = 0; i < size(); i++)
            if (ref_aln[i] != GAP && qry_aln[i] != GAP && code_matches(ref_aln[i], qry_aln[i])) c++;
        return c;
    }
    int insertions_rel_ref() const { int c = 0; for (auto x : ref_aln) if (x == GAP) c++; return c; }  // :190-199
    int deletions_rel_ref() const { int c = 0; for (auto x : qry_aln) if (x == GAP) c++; return c; }   // :218-227
    int number_of_non_terminal_gaps() const {  // :274-278
        AlignmentPair t = without_ending_and_beginning_reference_inserts().without_ending_and_beginning_query_inserts();
        return t.insertions_rel_ref() + t.deletions_rel_ref();
    }
};

}  // namespace hsqo
