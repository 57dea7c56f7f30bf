// hsq_emit.cuh -- what the reference derives from an AlignmentPair, computed from the traceback columns.
//   AlignmentPair indices / counters    COMMON/alignment/AlignmentPair.java:39-64,144-152,190-290
//   expanded + standard CIGAR           COMMON/alignment/CigarStringUtil.java:61-89,206-229; CigarString.java:61-77
//   MD string                           COMMON/alignment/CigarStringUtil.java:98-204
// Columns are stored last-first by dp_traceback: forward column c of an n-column alignment is sc.col(n-1-c).
#pragma once
#include "hsq_dp.cuh"

namespace hsq {

struct SlotWriter {  // bounded output slot; len keeps counting past cap so overflow is visible
    uint8_t* p;
    int cap;
    int len;
    __device__ inline void put(uint8_t ch) {
        if (len < cap) p[len] = ch;
        len++;
    }
    __device__ inline void put_int(int v) {  // v >= 0
        char buf[12];
        int k = 0;
        do { buf[k++] = (char)('0' + v % 10); v /= 10; } while (v > 0);
        while (k > 0) put((uint8_t)buf[--k]);
    }
};

struct Columns {
    ColStore cs;
    int n;
    __device__ inline uint32_t fwd(int c) const { return cs.col(n - 1 - c); }
};

__device__ __forceinline__ uint32_t col_ref(uint32_t w) { return w & 0xFFu; }
__device__ __forceinline__ uint32_t col_qry(uint32_t w) { return w >> 8; }
__device__ __forceinline__ bool codes_match(uint32_t a, uint32_t b) { return (a & b & MATCH_MASK) != 0; }

// CigarStringUtil.getCigarString (:61-83)
__device__ __forceinline__ uint8_t column_op(uint32_t w) {
    uint32_t r = col_ref(w), q = col_qry(w);
    if (codes_match(r, q)) return '=';
    if (r == CODE_GAP) return 'I';
    if (q == CODE_GAP) return 'D';
    return 'X';
}

struct PairIndices { int first_q, last_q, first_r, last_r; };

// AlignmentPair constructor (:39-64)
__device__ inline PairIndices pair_indices(const Columns& cv) {
    PairIndices x;
    int n = cv.n;
    x.first_q = 0;
    while (x.first_q < n && col_qry(cv.fwd(x.first_q)) == CODE_GAP) x.first_q++;
    x.last_q = n - 1;
    while (x.last_q >= 0 && col_qry(cv.fwd(x.last_q)) == CODE_GAP) x.last_q--;
    x.first_r = 0;
    while (x.first_r < n && col_ref(cv.fwd(x.first_r)) == CODE_GAP) x.first_r++;
    x.last_r = n - 1;
    while (x.last_r >= 0 && col_ref(cv.fwd(x.last_r)) == CODE_GAP) x.last_r--;
    return x;
}

// Streams forward columns [a..b] (or b..a when reversed) and emits the standard CIGAR, the MD string, the
// equality-based edit distance (AlignmentPair.getEditDistance :144-152) and the reference length
// (CigarString.getLengthOfReference).  The MD state machine restates getMismatchDetailsString:
//   - '='/'M' runs merge across insertions (:116-121); insertions emit nothing (:122);
//   - every 'D' run and every 'X' run of the summarized cigar is its own entry (:123-128);
//   - "0" before a leading non-match, between two adjacent non-match entries, and after a trailing one (:139-157);
//   - mismatch runs print their reference letters joined by "0" (:182-191), deletions print "^" + letters.
__device__ inline void emit_cigar_md(const Columns& cv, int a, int b, bool reversed, SlotWriter* cigar, SlotWriter* md,
                                     int* nm_out, int* ref_len_out) {
    int nm = 0, ref_len = 0;
    uint8_t run_ch = 0;  // standard-cigar run
    int run_n = 0;
    int last_entry = 0;  // 0 none, 1 match, 2 X, 3 D
    int match_count = 0;
    uint8_t prev_op = 0;
    int count = (b >= a) ? b - a + 1 : 0;
    for (int t = 0; t < count; t++) {
        int c = reversed ? b - t : a + t;
        uint32_t w = cv.fwd(c);
        uint8_t op = column_op(w);
        if (col_ref(w) != col_qry(w)) nm++;
        if (op != 'I') ref_len++;
        if (cigar) {
            uint8_t sch = (op == '=' || op == 'X') ? (uint8_t)'M' : op;
            if (sch == run_ch) run_n++;
            else {
                if (run_n > 0) { cigar->put_int(run_n); cigar->put(run_ch); }
                run_ch = sch;
                run_n = 1;
            }
        }
        if (md) {
            if (op == '=') {
                if (last_entry == 1) match_count++;
                else { match_count = 1; last_entry = 1; }
            } else if (op == 'X') {
                if (last_entry == 1) md->put_int(match_count);
                else md->put('0');  // leading, or separator after a non-match entry / inside a mismatch run
                md->put(decode_code((uint8_t)col_ref(w)));
                last_entry = 2;
            } else if (op == 'D') {
                if (prev_op != 'D') {
                    if (last_entry == 1) md->put_int(match_count);
                    else md->put('0');
                    md->put('^');
                }
                md->put(decode_code((uint8_t)col_ref(w)));
                last_entry = 3;
            }
        }
        prev_op = op;
    }
    if (cigar && run_n > 0) { cigar->put_int(run_n); cigar->put(run_ch); }
    if (md) {
        if (last_entry == 1) md->put_int(match_count);
        else if (last_entry >= 2) md->put('0');
    }
    *nm_out = nm;
    *ref_len_out = ref_len;
}

// ---- streamed emission (no column store) -------------------------------------------------------------------------
// The traceback produces columns last-first.  For a read on the reverse strand that IS the emission order of the
// reference (CIGAR and MD are computed on the reversed alignment, ExtendReadsToPrimer.java:229-231): tokens are appended.
// For a forward read the emission order is the opposite, so the same state machine runs on the reversed column sequence
// and writes its tokens right to left ("back" mode: numbers least-significant digit first, a deletion's '^' when the
// run ends); finish() then moves the string to the start of the slot.  The separator rules of getMismatchDetailsString
// (a "0" before a leading non-match entry, between adjacent non-match entries, after a trailing one) are symmetric.
struct StreamWriter {
    uint8_t* p;
    int cap;
    int len;
    bool back;
    __device__ inline void put(uint8_t ch) {
        if (len < cap) p[back ? cap - 1 - len : len] = ch;
        len++;
    }
    __device__ inline void put_int(int v) {  // v >= 0
        if (back) {
            do { put((uint8_t)('0' + v % 10)); v /= 10; } while (v > 0);
        } else {
            char buf[12];
            int k = 0;
            do { buf[k++] = (char)('0' + v % 10); v /= 10; } while (v > 0);
            while (k > 0) put((uint8_t)buf[--k]);
        }
    }
    __device__ inline void finish() {  // on overflow (len > cap) the slot content is unspecified
        if (back && len > 0 && len < cap) {
            const int off = cap - len;
            for (int k = 0; k < len; k++) p[k] = p[off + k];
        }
    }
};

struct CigarMdStream {
    StreamWriter cw, mw;
    bool back;
    bool started;  // a query base has been seen: trailing query-gap columns are dropped (AlignmentPair.java:95-103)
    int nm, ref_len;
    uint8_t run_ch;
    int run_n;
    int last_entry;  // 0 none, 1 match, 2 X, 3 D
    int match_count;
    uint8_t prev_op;
    __device__ inline void init(uint8_t* cigar, int cigar_cap, uint8_t* md, int md_cap, bool back_) {
        back = back_;
        cw = StreamWriter{cigar, cigar_cap, 0, back_};
        mw = StreamWriter{md, md_cap, 0, back_};
        started = false;
        nm = 0; ref_len = 0; run_ch = 0; run_n = 0; last_entry = 0; match_count = 0; prev_op = 0;
    }
    // fast paths of the packed traceback.  A '=' column that extends an 'M' run (the previous column was '=' or 'X') only
    // moves counters; a column without a query base before the first query base is dropped (see `started`).
    static constexpr bool HAS_FAST = true;
    __device__ inline bool fast_ready() const { return run_ch == 'M'; }
    __device__ inline void fast_match() {
        ref_len++;
        run_n++;
        match_count = (last_entry == 1) ? match_count + 1 : 1;
        last_entry = 1;
        prev_op = '=';
    }
    __device__ inline void fast_match_n(int n) {  // n times fast_match
        ref_len += n;
        run_n += n;
        match_count = (last_entry == 1) ? match_count + n : n;
        last_entry = 1;
        prev_op = '=';
    }
    __device__ inline bool drops_query_gaps() const { return !started; }
    __device__ inline void flush_run() {
        if (run_n > 0) {
            if (back) { cw.put(run_ch); cw.put_int(run_n); }
            else { cw.put_int(run_n); cw.put(run_ch); }
        }
    }
    __device__ inline void sep() {
        if (last_entry == 1) mw.put_int(match_count);
        else mw.put('0');
    }
    __device__ inline void push(uint32_t w) {
        if (!started) {
            if (col_qry(w) == CODE_GAP) return;
            started = true;
        }
        const uint8_t op = column_op(w);
        if (col_ref(w) != col_qry(w)) nm++;
        if (op != 'I') ref_len++;
        const uint8_t sch = (op == '=' || op == 'X') ? (uint8_t)'M' : op;
        if (sch == run_ch) run_n++;
        else {
            flush_run();
            run_ch = sch;
            run_n = 1;
        }
        if (back && prev_op == 'D' && op != 'D') mw.put('^');
        if (op == '=') {
            if (last_entry == 1) match_count++;
            else { match_count = 1; last_entry = 1; }
        } else if (op == 'X') {
            sep();
            mw.put(decode_code((uint8_t)col_ref(w)));
            last_entry = 2;
        } else if (op == 'D') {
            if (prev_op != 'D') {
                sep();
                if (!back) mw.put('^');
            }
            mw.put(decode_code((uint8_t)col_ref(w)));
            last_entry = 3;
        }
        prev_op = op;
    }
    __device__ inline void finish() {
        flush_run();
        if (back && prev_op == 'D') mw.put('^');
        if (last_entry == 1) mw.put_int(match_count);
        else if (last_entry >= 2) mw.put('0');
        cw.finish();
        mw.finish();
    }
};

// counters over forward columns [a..b]
__device__ inline void count_columns(const Columns& cv, int a, int b, int* matches, int* mismatches, int* ref_gaps,
                                     int* qry_gaps) {
    int m = 0, x = 0, rg = 0, qg = 0;
    for (int c = a; c <= b; c++) {
        uint32_t w = cv.fwd(c);
        uint32_t r = col_ref(w), q = col_qry(w);
        if (r == CODE_GAP) rg++;
        if (q == CODE_GAP) qg++;
        if (r != CODE_GAP && q != CODE_GAP) {
            if (codes_match(r, q)) m++;
            else x++;
        }
    }
    *matches = m; *mismatches = x; *ref_gaps = rg; *qry_gaps = qg;
}

// AlignmentPair.getNumberOfNonTerminalGaps (:274-278): trim by reference gaps, then by query gaps of that
// sub-alignment, then count gaps of both kinds.
__device__ inline int non_terminal_gaps(const Columns& cv, const PairIndices& ix) {
    int a = ix.first_r, b = ix.last_r;  // subSequence(first_r, last_r): empty when first_r > last_r
    if (a > b) return 0;
    int fa = a;
    while (fa <= b && col_qry(cv.fwd(fa)) == CODE_GAP) fa++;
    int lb = b;
    while (lb >= a && col_qry(cv.fwd(lb)) == CODE_GAP) lb--;
    if (fa > lb) return 0;
    int m, x, rg, qg;
    count_columns(cv, fa, lb, &m, &x, &rg, &qg);
    return rg + qg;
}

}  // namespace hsq
