// hsq_dp2.cuh -- the DP fill of hsq_dp.cuh for TWO alignments per thread, packed as the two u16 halves of a register.
//
// Same recurrence, same cell encoding (X = 4*S + p, candidates 4*score + {2,1,0}, see hsq_dp.cuh), but each half
// carries its alignment's X as an unsigned 16-bit number:
//     X'(i,j) = 4*S(i,j) + p + 0x8000 + (i + j) * SB/2
//   - 0x8000 makes the half unsigned;
//   - the position term makes every substitution score non-negative: a diagonal step adds 4*sub + SB >= 0, a gap step
//     adds its penalty + SB/2 (folded into the constants kh, kv), so all three candidates of a cell carry the same
//     offset and the max / tie-break is unchanged.  Non-negative profile entries can be merged bit-wise (PRMT) from two
//     different profile columns -- alignment A's reference base selects the low half, alignment B's the high half.
//   - the two maxima of a cell are VIADDMNMX.U16x2 (per-half wrapping add, per-half unsigned max);
//   - every other step of the cell (mask, multiply by a small constant, add) is ordinary 32-bit integer arithmetic on the
//     packed word, which is exact per half as long as every half stays inside [0, 65535] -- the host checks the score
//     range of the scorer / lengths before choosing these kernels (dp2_score_range_ok), else the 32-bit kernels run.
// One instruction stream fills two matrices: 6 ALU-pipe + 5 FMA-pipe instructions per pair of cells.
//
// The two alignments of a thread have the SAME shape (Q rows, R columns): the bucket sorts of hsq_extend.cuh key the
// read ends by exact (Q, R) and pad every bucket to an even size.
// Pointer words: per strip column two words, U bits and L bits; row r of the strip sits at bit 15-r of its half.
// Traceback is streamed (no column store): pointer words are fetched 16 columns at a time into the thread's (by then
// dead) profile slots, and each column goes straight to a sink (CIGAR/MD state machine or the primer gate).
#pragma once
#include <algorithm>
#include "hsq_dp.cuh"

namespace hsq {

static constexpr int DP2_BLOCK = 128;              // threads per block of the packed kernels
static constexpr int PROF2_INT4_PER_THREAD = 16;   // 4 one-hot bases x 16 rows x (A | B << 16)
static constexpr int TB_WINDOW = 16;               // traceback window: columns of pointer words staged in shared memory
static constexpr uint32_t DP2_BIAS = 0x8000u;
static constexpr uint32_t DP2_REP = 0x10001u;
static constexpr uint32_t DP2_MASK_D4 = 0xFFFCFFFCu, DP2_MASK_B0 = 0x00010001u, DP2_MASK_B1X2 = 0x00020002u;

struct ScoreParams2 {
    uint32_t kh2, kv2;        // per-half wrapping addends: 4*gapStart + 2 + SB/2, 4*gapStart + 1 + SB/2
    uint32_t c2, c4;          // 2*(gapExtend-gapStart), 4*(gapExtend-gapStart) as 32-bit multipliers (mod 2^32)
    uint32_t hfree2, vfree2;  // (2 - kh) and (1 - kv) replicated: turn a D4 into the candidate-neutral free-gap output
    int ms4o, mm4o;           // 4*match + SB, 4*mismatch + SB  (both in [0, 65535])
    int sb_half;              // SB/2, a multiple of 4
    int gap_start;
    uint32_t neg2;            // banded fill only: the value of a cell that is not computed, replicated (0 = band not available)
    int match, gap_extend;    // raw scores (the acceptance test of the banded kernel)
};

inline ScoreParams2 make_score_params2(const ScoreParams& s) {
    ScoreParams2 p;
    const int lo = std::min(0, std::min(s.ms4, s.mm4));
    const int sb = ((-lo + 7) / 8) * 8;
    p.sb_half = sb / 2;
    p.kh2 = ((uint32_t)(s.kh + p.sb_half) & 0xFFFFu) * DP2_REP;
    p.kv2 = ((uint32_t)(s.kv + p.sb_half) & 0xFFFFu) * DP2_REP;
    p.c2 = (uint32_t)s.c2;
    p.c4 = (uint32_t)s.c4;
    p.hfree2 = (uint32_t)(2 - s.kh) * DP2_REP;
    p.vfree2 = (uint32_t)(1 - s.kv) * DP2_REP;
    p.ms4o = s.ms4 + sb;
    p.mm4o = s.mm4 + sb;
    p.gap_start = s.gap_start;
    p.neg2 = 0;
    p.match = s.ms4 / 4;
    p.gap_extend = s.gap_extend;
    return p;
}

// every value the packed fill can form (cells, candidates, the padding rows of the last strip, the position offset)
// stays inside an unsigned half
inline bool dp2_score_range_ok(int match, int mismatch, int gap_extend, int gap_start, int q_max, int r_max) {
    int64_t m = 1;
    for (int v : {match, mismatch, gap_extend, gap_start}) m = std::max<int64_t>(m, v < 0 ? -(int64_t)v : v);
    const int64_t span = (int64_t)q_max + KROWS + r_max + 4;
    const int64_t bound = span * (m + 1);       // |score| of any cell or candidate
    const int64_t offset = span * (2 * m + 4);  // (i + j) * SB/2 with SB <= 4*m + 7
    return 4 * (bound + 2 * m + 2) + 3 + offset < 32768;
}

// Banded fill (opt-in, hsq_config.band): a cell outside the computed staircase reads as NEG, one constant for the whole matrix.
// With m = the largest |score| of the scorer and span = Q + R + the padding rows:
//   - every computed cell has a computed predecessor and a step costs at most m, so its score is >= -m * span (free start gaps:
//     the borders are 0), i.e. its X' is >= 0x8000 + offset(i, j) - 4*m*span;
//   - a chain of steps from a NEG cell gains at most 4*m per unit of (i + j) on top of the position offset, and never loses (a
//     diagonal step adds 4*sub + SB >= 0), so a NEG-derived X' stays inside [NEG, NEG + offset(i, j) + 4*m*span];
//   so with NEG = 0x8000 - 8*m*span - margin a candidate derived from NEG loses against every candidate derived from computed cells,
//   and with NEG >= margin no candidate formed from it wraps below zero (a wrapped half would win every max).
// Returns 0 when the two cannot both hold, or when the scorer is not of the shape the acceptance test of the banded kernel assumes
// (match > 0 and gapStart <= gapExtend <= 0): the caller then fills full matrices.
inline uint32_t dp2_band_neg(int match, int mismatch, int gap_extend, int gap_start, int q_max, int r_max) {
    if (match <= 0 || gap_extend > 0 || gap_start > gap_extend) return 0;
    int64_t m = 1;
    for (int v : {match, mismatch, gap_extend, gap_start}) m = std::max<int64_t>(m, v < 0 ? -(int64_t)v : v);
    const int64_t span = (int64_t)q_max + KROWS + r_max + 4;
    const int64_t neg = 32768 - 8 * m * span - 16 * m - 16;
    if (neg < 16 * m + 16) return 0;
    return (uint32_t)neg & ~3u;  // pointer bits clear
}

// The staircase a banded fill computes: strip s (rows 16s+1 .. 16s+16) covers the columns [lo, hi]; the last strip runs to R because
// the last row's horizontal moves are free.  A traceback is accepted only while it stays on diagonals |j - i| < band, all of whose
// predecessors lie inside the staircase.
struct BandCols { int lo, hi; };
__device__ __forceinline__ BandCols band_cols(int s, int band, int R, bool last_strip) {
    BandCols c;
    c.lo = min(R, max(1, s * KROWS + 1 - band));  // a strip below the band still computes the last column (free vertical moves)
    c.hi = last_strip ? R : min(R, s * KROWS + KROWS + band);
    return c;
}

struct DpScratch2 {  // every pointer is already offset to this thread's element 0; consecutive elements are nthreads apart
    uint2* ptr_words;   // [n_strips_max * (r_max + 1)] per thread: (U word, L word)
    uint32_t* brow;     // [r_max + 1] per thread: packed X' of the row above the current strip
    uint32_t* colinfo;  // [(r_max + 4) / 4] per thread: byte j & 3 of word j >> 2 = profile column of alignment A (low nibble) and B
                        // (high nibble) at reference position j; 4 = not A/C/G/T
    uint32_t* slow;     // [KROWS] per thread: packed substitution scores of a column with a non-ACGT reference code
    int4* prof;         // shared memory: [PROF2_INT4_PER_THREAD][DP2_BLOCK]
    int64_t nthreads;   // a kernel parameter (ScratchPool2::nthreads), not gridDim * blockDim: the compiler re-derives either wherever
                        // it runs out of registers, and a constant-bank operand costs nothing
    int r_stride;
    __device__ inline uint2& word(int strip, int j) const { return ptr_words[((int64_t)strip * r_stride + j) * nthreads]; }
    __device__ inline uint32_t& b(int j) const { return brow[(int64_t)j * nthreads]; }
    __device__ inline uint32_t& ciw(int w) const { return colinfo[(int64_t)w * nthreads]; }
};

// packed, biased, offset 4*border_score of border cell k (row 0 or column 0)
__device__ __forceinline__ uint32_t border2(int k, int pen_start, const ScoreParams2& sp) {
    return (uint32_t)(4 * border_score(k, pen_start, sp.gap_start) + (int)DP2_BIAS + k * sp.sb_half) * DP2_REP;
}

// One column of one strip for both alignments.  State per row: D4[r], H0[r] as in dp_column.
template <bool HFREE, bool VFREE, bool SLOW, bool SAME>
__device__ __forceinline__ uint32_t dp_column2(uint32_t (&D4)[KROWS], uint32_t (&H0)[KROWS], const int4* __restrict__ pcolA,
                                               const int4* __restrict__ pcolB, const uint32_t* __restrict__ slow, int64_t nt,
                                               uint32_t up_v0, uint32_t diag_d4, const ScoreParams2& sp, int rstar, uint2* word_out) {
    uint32_t acc_l = 0, acc_u = 0;
    uint32_t vin = up_v0;
    uint32_t dg = diag_d4;
    uint32_t x = 0;
#pragma unroll
    for (int r = 0; r < KROWS; r++) {
        uint32_t sub;
        if (SLOW) {
            sub = slow[r * nt];
        } else {
            int4 sa, sb;
            if ((r & 3) == 0) {
                sa = pcolA[(r >> 2) * DP2_BLOCK];
                if (!SAME) sb = pcolB[(r >> 2) * DP2_BLOCK];
            }
            const int a = (r & 3) == 0 ? sa.x : (r & 3) == 1 ? sa.y : (r & 3) == 2 ? sa.z : sa.w;
            if (SAME) {
                sub = (uint32_t)a;  // both alignments run against the same reference: one profile column serves both halves
            } else {
                const int b = (r & 3) == 0 ? sb.x : (r & 3) == 1 ? sb.y : (r & 3) == 2 ? sb.z : sb.w;
                sub = __byte_perm((uint32_t)a, (uint32_t)b, 0x7610);  // low half from A's column, high half from B's column
            }
        }
        const uint32_t dsub = dg + sub;
        // the horizontal and diagonal candidates do not depend on the row above: only the outer max sits on the row-to-row chain
        // (max is associative on the tie-break encoded values, so the result is the one of max(h, max(v, d)))
        x = __viaddmax_u16x2(vin, sp.kv2, __viaddmax_u16x2(H0[r], sp.kh2, dsub));
        dg = D4[r];
        const uint32_t d4 = x & DP2_MASK_D4;
        const uint32_t b1x2 = x & DP2_MASK_B1X2;
        const uint32_t b0 = x & DP2_MASK_B0;
        uint32_t h0 = b1x2 * sp.c2 + d4;
        uint32_t v0 = b0 * sp.c4 + d4;
        acc_u = mad2(acc_u, b0);
        if (r < KROWS - 1) acc_l = mad2(acc_l, b1x2);  // rows 0..14 land at bits 15..1 of their half
        else acc_l = acc_l + (b1x2 >> 1);              // row 15 at bit 0
        if (HFREE) h0 = (r == rstar) ? d4 + sp.hfree2 : h0;
        if (VFREE) v0 = d4 + sp.vfree2;
        D4[r] = d4;
        H0[r] = h0;
        vin = v0;
    }
    *word_out = make_uint2(acc_u, acc_l);
    return x;
}

// Profile of one strip: entry [base][row] = A's score in the low half, B's score in the high half (both >= 0).
__device__ inline void build_profile2(const uint32_t (&qa)[KROWS], const uint32_t (&qb)[KROWS], const ScoreParams2& sp, int4* prof) {
#pragma unroll
    for (int b = 0; b < 4; b++) {
        const uint32_t code = 1u << b;
#pragma unroll
        for (int ch = 0; ch < 4; ch++) {
            int e[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int va = (qa[4 * ch + k] & code) ? sp.ms4o : sp.mm4o;
                const int vb = (qb[4 * ch + k] & code) ? sp.ms4o : sp.mm4o;
                e[k] = va + vb * 65536;
            }
            prof[(b * 4 + ch) * DP2_BLOCK + threadIdx.x] = make_int4(e[0], e[1], e[2], e[3]);
        }
    }
}

// A sequence view read element by element from its END towards its start through aligned 32-bit words: the word in use
// and the next one in walking direction sit in registers, so the byte loads of the traceback (which miss L1 while other
// warps stream their DP scratch through it) become one prefetched word load per four elements.
struct CodeWalker {
    const uint32_t* wp;  // aligned word holding the current element
    uint32_t cur, nxt;
    int sh;              // 8 * byte index of the current element inside cur
    int dsh;             // +8 / -8: memory direction of the walk (= -8 * view step)
    int comp;
    // elements of the view still to be read after the ones in `cur`: a word is fetched only when it holds one of them
    // (never touches memory before the first / after the last word of the view)
    __device__ __forceinline__ bool more_after_cur(int remaining) const {
        const int in_cur = (dsh > 0 ? (24 - sh) : sh) >> 3;  // further elements in cur, walking direction
        return remaining - 1 > in_cur;
    }
    __device__ __forceinline__ void init(const SeqView& v, int remaining) {
        comp = v.comp;
        dsh = -8 * v.step;
        cur = 0; nxt = 0; sh = 0; wp = nullptr;
        if (remaining <= 0) return;
        const uint8_t* p = v.base + v.start + (int64_t)(remaining - 1) * v.step;
        const uintptr_t a = reinterpret_cast<uintptr_t>(p);
        wp = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
        sh = (int)(a & 3) * 8;
        cur = *wp;
        if (more_after_cur(remaining)) nxt = wp[dsh >> 3];
    }
    __device__ __forceinline__ uint32_t peek_raw() const { return (cur >> sh) & 0xFFu; }  // before the complement
    __device__ __forceinline__ uint32_t peek() const {  // code of the current element (garbage when none remains: unused)
        const uint32_t c = (cur >> sh) & 0xFFu;
        return comp ? (uint32_t)complement_code((uint8_t)c) : c;
    }
    // `remaining` counts the elements before this step consumes one
    __device__ __forceinline__ void advance(int remaining) {
        sh += dsh;
        if ((uint32_t)sh > 24u) {  // left the word (sh == 32 or -8)
            sh &= 31;
            wp += dsh >> 3;
            cur = nxt;
            if (more_after_cur(remaining - 1)) nxt = wp[dsh >> 3];
        }
    }
};

// complement_code on the four codes of a word
__device__ __forceinline__ uint32_t complement_codes4(uint32_t x) {
    uint32_t t = x & 0x0F0F0F0Fu;
    t = ((t & 0x05050505u) << 1) | ((t >> 1) & 0x05050505u);
    t = ((t & 0x03030303u) << 2) | ((t >> 2) & 0x03030303u);
    return t | (x & (0x01010101u * CODE_GAP));
}

// Elements e0 .. e0+15 of a view (all inside the view), element e0+m in byte m of the four words: what SeqView::get returns, 16 at a
// time.  Reads whole aligned words around the elements (inside the slack of every code buffer).
__device__ __forceinline__ void load_codes16(const SeqView& v, int e0, uint32_t (&w)[4]) {
    if (v.step == 1) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(v.base + v.start + e0);
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
        const uint32_t sh = (uint32_t)(a & 3) * 8u;
        uint32_t x[5];
#pragma unroll
        for (int k = 0; k < 5; k++) x[k] = wp[k];
#pragma unroll
        for (int k = 0; k < 4; k++) w[k] = __funnelshift_r(x[k], x[k + 1], sh);
    } else {  // step == -1: descending addresses, element e0+15 at the lowest
        const uintptr_t a = reinterpret_cast<uintptr_t>(v.base + v.start - e0 - 15);
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
        const uint32_t sh = (uint32_t)(a & 3) * 8u;
        uint32_t x[5];
#pragma unroll
        for (int k = 0; k < 5; k++) x[k] = wp[k];
#pragma unroll
        for (int k = 0; k < 4; k++) w[3 - k] = __byte_perm(__funnelshift_r(x[k], x[k + 1], sh), 0u, 0x0123);
    }
    if (v.comp) {
#pragma unroll
        for (int k = 0; k < 4; k++) w[k] = complement_codes4(w[k]);
    }
}

// keeps a per-alignment invariant in ONE register (without it ptxas re-selects it from the two alignments at every step)
__device__ __forceinline__ void pin(int& x) { asm volatile("" : "+r"(x)); }
__device__ __forceinline__ void pin(uint32_t& x) { asm volatile("" : "+r"(x)); }

// The KROWS query codes of one strip (rows i0+1 .. i0+KROWS; 0 beyond the query).  A forward view is read through five
// aligned 32-bit words realigned with funnel shifts: sixteen byte loads per lane, each to a different line, kept the LSU
// busy for longer than the rest of the strip set-up (profiles/README_r1.md).  Reads up to 19 bytes past element i0, which
// stays inside the slack every code buffer is allocated with.
__device__ __forceinline__ void load_strip_codes(const SeqView& q, int i0, int Q, uint32_t (&out)[KROWS]) {
    static_assert(KROWS == 16, "four words per strip");
    if (q.step == 1) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(q.base + q.start + i0);
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
        const uint32_t sh = (uint32_t)(a & 3) * 8u;
        uint32_t w[5];
#pragma unroll
        for (int k = 0; k < 5; k++) w[k] = wp[k];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t v = __funnelshift_r(w[k], w[k + 1], sh);
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const int r = 4 * k + b;
                uint32_t c = (v >> (8 * b)) & 0xFFu;
                if (q.comp) c = complement_code((uint8_t)c);
                out[r] = (i0 + r < Q) ? c : 0u;
            }
        }
    } else if (q.step == -1 && q.start - i0 - (KROWS - 1) >= 0) {  // reversed view, strip inside the sequence: descending addresses
        const uintptr_t lo = reinterpret_cast<uintptr_t>(q.base + q.start - i0 - (KROWS - 1));
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(lo & ~(uintptr_t)3);
        const uint32_t sh = (uint32_t)(lo & 3) * 8u;
        uint32_t w[5], v[4];
#pragma unroll
        for (int k = 0; k < 5; k++) w[k] = wp[k];
#pragma unroll
        for (int k = 0; k < 4; k++) v[k] = __funnelshift_r(w[k], w[k + 1], sh);
#pragma unroll
        for (int r = 0; r < KROWS; r++) {
            const int byte = KROWS - 1 - r;
            uint32_t c = (v[byte >> 2] >> (8 * (byte & 3))) & 0xFFu;
            if (q.comp) c = complement_code((uint8_t)c);
            out[r] = (i0 + r < Q) ? c : 0u;
        }
    } else {
#pragma unroll
        for (int r = 0; r < KROWS; r++) out[r] = (i0 + r < Q) ? (uint32_t)q.get(i0 + r) : 0u;
    }
}

template <bool HFREE, bool SAME, bool BAND = false>
__device__ __forceinline__ void dp_strip2(const SeqView& refA, const SeqView& refB, const SeqView& qryA, const SeqView& qryB,
                                          const ScoreParams2& sp, int pen_start, bool free_end, const DpScratch2& sc, int s,
                                          bool last_strip, uint32_t* final_d4, int band = 0, int n_strips = 0) {
    const int R = refA.len, Q = qryA.len;
    const int i0 = s * KROWS;
    const int rstar = last_strip ? (Q - 1 - i0) : -1;
    uint32_t D4[KROWS], H0[KROWS];
    {
        uint32_t qa[KROWS], qb[KROWS];
        load_strip_codes(qryA, i0, Q, qa);
        load_strip_codes(qryB, i0, Q, qb);
#pragma unroll
        for (int r = 0; r < KROWS; r++) {
            const int i = i0 + 1 + r;
            const uint32_t s0 = border2(i, pen_start, sp);
            D4[r] = s0;
            H0[r] = (HFREE && r == rstar) ? s0 + sp.hfree2 : s0;
        }
        build_profile2(qa, qb, sp, sc.prof);
    }
    uint32_t diag_d4 = (i0 == 0) ? DP2_BIAS * DP2_REP : border2(i0, pen_start, sp);
    int jlo = 1, jhi = R;
    if (BAND) {
        const BandCols bc = band_cols(s, band, R, last_strip);
        jlo = bc.lo;
        jhi = bc.hi;
        if (jlo > 1) {  // the column to the left of the strip's first column is not computed
#pragma unroll
            for (int r = 0; r < KROWS; r++) D4[r] = H0[r] = sp.neg2;
            // the cell above-left: computed by the strip above unless that strip, too, was cut down to the last column
            const bool above = jlo - 1 >= max(1, (s - 1) * KROWS + 1 - band);
            diag_d4 = above ? (sc.b(jlo - 1) & DP2_MASK_D4) : sp.neg2;
        }
    }
    uint2* wordp = &sc.word(s, jlo);
    uint32_t* browp = &sc.b(jlo);
    const uint32_t* cwp = &sc.ciw(jlo >> 2);
    const int64_t nt = sc.nthreads;
    const int4* prof_t = sc.prof + threadIdx.x;
    // the row above arrives one column ahead of its use, the profile columns a word (four columns) ahead
    uint32_t up0 = 0, ciw = 0, ciw_next = 0;
    const int cw_last = jhi >> 2;
    if (!BAND || jlo <= jhi) {
        up0 = *browp;
        ciw = *cwp;
        if ((jlo >> 2) < cw_last) ciw_next = cwp[nt];
    }
#ifdef HSQ_EXPERIMENT_NO_COLUMNS  // (development experiment: what a strip costs without its columns)
    jhi = jlo - 1;
#endif
    for (int j = jlo; j <= jhi; j++) {
        const uint32_t ci = (ciw >> (8 * (j & 3))) & 0xFFu;
        const uint32_t upx = up0;
        if (j < jhi) up0 = browp[nt];  // (two columns ahead costs more in register moves than the wait it saves; so does dropping the test)
        if ((j & 3) == 3) {
            ciw = ciw_next;
            cwp += nt;
            if (((j + 1) >> 2) < cw_last) ciw_next = cwp[nt];
        }
        const uint32_t up_d4 = upx & DP2_MASK_D4;
        const bool vfree = free_end && (j == R);
        const uint32_t up_v0 = vfree ? up_d4 + sp.vfree2 : (upx & DP2_MASK_B0) * sp.c4 + up_d4;
        uint2 w;
        uint32_t xlast;
        if ((ci & 0x44u) == 0) {
            const int4* pa = prof_t + (ci & 0xFu) * (4 * DP2_BLOCK);
            const int4* pb = SAME ? pa : prof_t + (ci >> 4) * (4 * DP2_BLOCK);
            if (!vfree) xlast = dp_column2<HFREE, false, false, SAME>(D4, H0, pa, pb, nullptr, nt, up_v0, diag_d4, sp, rstar, &w);
            else xlast = dp_column2<HFREE, true, false, SAME>(D4, H0, pa, pb, nullptr, nt, up_v0, diag_d4, sp, rstar, &w);
        } else {  // ambiguity code, N, U or gap character in a reference: this column's scores go through global scratch
            const uint32_t ca = refA.get(j - 1), cb = refB.get(j - 1);
            uint32_t* sl = sc.slow;
#pragma unroll 1
            for (int r = 0; r < KROWS; r++) {
                const int i = i0 + 1 + r;
                const uint32_t qa = (i <= Q) ? (uint32_t)qryA.get(i - 1) : 0u;
                const uint32_t qb = (i <= Q) ? (uint32_t)qryB.get(i - 1) : 0u;
                const int va = ((qa & ca & MATCH_MASK) != 0) ? sp.ms4o : sp.mm4o;
                const int vb = ((qb & cb & MATCH_MASK) != 0) ? sp.ms4o : sp.mm4o;
                sl[r * nt] = (uint32_t)va + (uint32_t)vb * 65536u;
            }
            if (!vfree) xlast = dp_column2<HFREE, false, true, SAME>(D4, H0, nullptr, nullptr, sl, nt, up_v0, diag_d4, sp, rstar, &w);
            else xlast = dp_column2<HFREE, true, true, SAME>(D4, H0, nullptr, nullptr, sl, nt, up_v0, diag_d4, sp, rstar, &w);
        }
        __stcs(wordp, w);  // streamed: read back once by the traceback, must not evict the boundary rows from L2
        if (!last_strip) *browp = xlast;
        diag_d4 = up_d4;
        wordp += nt;
        browp += nt;
    }
    if (BAND && !last_strip) {
        // the row above the next strip, to the right of this strip's last column, is not computed (browp is at column jhi + 1)
        const int next_hi = (s + 2 == n_strips) ? R : min(R, jhi + KROWS);
        for (int j = max(jhi, jlo - 1) + 1; j <= next_hi; j++, browp += nt) *browp = sp.neg2;
    }
    if (last_strip) {
#pragma unroll
        for (int r = 0; r < KROWS; r++)
            if (r == rstar) *final_d4 = D4[r];
    }
}

// Fills both matrices (same shape: refA.len == refB.len, qryA.len == qryB.len).  Scores S[Q][R] go to score[0], score[1].
// SAME: refA and refB are the same sequence in the same orientation (two reads of one probe and mate): the profile column of a
// reference position serves both halves, so the cell needs no PRMT merge and one profile load instead of two.
// BAND (free terminal gaps on both ends only: pen_start == pen_end == 0): only the staircase of band_cols is computed.
template <bool SAME = false, bool BAND = false>
__device__ inline void dp_fill2(const SeqView& refA, const SeqView& refB, const SeqView& qryA, const SeqView& qryB,
                                const ScoreParams2& sp, int pen_start, int pen_end, const DpScratch2& sc, int* score, int band = 0) {
    const int R = refA.len, Q = qryA.len;
    if (Q == 0 || R == 0) {
        const int k = Q == 0 ? R : Q;
        score[0] = score[1] = (k == 0) ? 0 : border_score(k, pen_start, sp.gap_start);
        return;
    }
    const int n_strips = (Q + KROWS - 1) / KROWS;
    const bool free_end = !pen_end;
    {   // matrix row 0 (cells point LEFT: never an UP continuation) and the profile column of every reference position
        uint32_t* bp = &sc.b(R);
        uint32_t* cwp = &sc.ciw(R >> 2);
        uint32_t cw = 0;
        CodeWalker wa, wb;  // last column first: the walkers read from the end of a view
        wa.init(refA, R);
        wb.init(refB, R);
        for (int j = R; j >= 1; j--, bp -= sc.nthreads) {
            *bp = border2(j, pen_start, sp) + (uint32_t)P_LEFT * DP2_REP;
            const uint32_t ba = (uint32_t)prof_slot(wa.peek_raw(), refA.comp);
            const uint32_t bb = (uint32_t)prof_slot(wb.peek_raw(), refB.comp);
            cw |= (ba | (bb << 4)) << (8 * (j & 3));
            if ((j & 3) == 0) {
                *cwp = cw;
                cwp -= sc.nthreads;
                cw = 0;
            }
            wa.advance(j);
            wb.advance(j);
        }
        *cwp = cw;  // word 0: columns 1 .. 3
    }
    uint32_t final_d4 = 0;
    for (int s = 0; s < n_strips - 1; s++)
        dp_strip2<false, SAME, BAND>(refA, refB, qryA, qryB, sp, pen_start, free_end, sc, s, false, &final_d4, band, n_strips);
    if (free_end) dp_strip2<true, SAME, BAND>(refA, refB, qryA, qryB, sp, pen_start, true, sc, n_strips - 1, true, &final_d4, band, n_strips);
    else dp_strip2<false, SAME, BAND>(refA, refB, qryA, qryB, sp, pen_start, false, sc, n_strips - 1, true, &final_d4, band, n_strips);
    const int off = (int)DP2_BIAS + (Q + R) * sp.sb_half;
    score[0] = ((int)(final_d4 & 0xFFFFu) - off) >> 2;
    score[1] = ((int)(final_d4 >> 16) - off) >> 2;
}

// Streamed traceback of alignment `which` of the pair (NeedlemanWunschGlobalAlignment.java:85-114): walks from (Q,R) to
// (0,0) and hands every column, LAST column first, to sink.push(ref code | qry code << 8) (CODE_GAP where the pointer
// did not consume that sequence).  Pointer words come through a 16-column window in the thread's own profile slots.
// BAND: the walk stops and returns false as soon as it stands on a cell that is outside the diagonals |j - i| < band (cells of the
// last row and of the last column, where the free terminal gaps run, only have to lie inside the staircase); the caller then
// discards what the sink has seen and the alignment is redone with full matrices.
template <bool BAND = false, class Sink>
__device__ __forceinline__ bool dp_traceback2(const SeqView& ref, const SeqView& qry, const DpScratch2& sc, int which, Sink& sink, int band = 0) {
    int i = qry.len, j = ref.len;
    const int Q = qry.len, R = ref.len;
    int ws = -1, whi = 0;
    // window slot k of this thread: shared address win_s + (k >> 1) * 16 * DP2_BLOCK + (k & 1) * 8
    uint32_t win_s = (uint32_t)__cvta_generic_to_shared(sc.prof + threadIdx.x);
    pin(win_s);
    int shift = 16 * which;
    pin(shift);
    const int64_t nt = sc.nthreads;
    CodeWalker rw, qw;
    rw.init(ref, j);
    qw.init(qry, i);
    pin(rw.dsh); pin(rw.comp); pin(qw.dsh); pin(qw.comp);
    while (i > 0 || j > 0) {
        uint32_t p;
        if (i == 0) p = P_LEFT;
        else if (j == 0) p = P_UP;
        else {
            const int s = (int)((uint32_t)(i - 1) / KROWS), r = (int)((uint32_t)(i - 1) % KROWS);
            bool off_band = false;
            if (BAND) {
                off_band = (uint32_t)(j - i + band - 1) > (uint32_t)(2 * band - 2);
                if (off_band) {  // not on a diagonal strictly inside the band
                    if (i != Q && j != R) return false;
                    const BandCols bc = band_cols(s, band, R, s == (Q - 1) / KROWS);
                    if (j < bc.lo || j > bc.hi) return false;
                }
            }
            if (s != ws || j <= whi - TB_WINDOW) {
                ws = s;
                whi = j;
                const uint2* g = &sc.word(s, j);
                uint2 t[TB_WINDOW];
#pragma unroll
                for (int k = 0; k < TB_WINDOW; k++) {
                    t[k] = (j - k >= 1) ? __ldcs(g) : make_uint2(0u, 0u);
                    g -= nt;
                }
                uint32_t not_diag = 1u;
#ifdef HSQ_EXPERIMENT_STRIP_FAST  // (development experiment, slower as it stands: three dependent load latencies per strip instead of 16 short steps)
                if constexpr (Sink::HAS_FAST) {
                    // entering a strip at its last row: are the next 16 steps all diagonal?  (row 15-k of column j-k: bit k + shift)
                    if (r == KROWS - 1 && j >= KROWS && !off_band && sink.fast_ready()) {
                        uint32_t acc = 0;
#pragma unroll
                        for (int k = 0; k < TB_WINDOW; k++) acc |= (t[k].x | t[k].y) >> k;
                        not_diag = (acc >> shift) & 1u;
                    }
                }
#endif
#pragma unroll
                for (int k = 0; k < TB_WINDOW; k++)
                    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(win_s + (uint32_t)((k >> 1) * 16 * DP2_BLOCK + (k & 1) * 8)),
                                 "r"(t[k].x), "r"(t[k].y)
                                 : "memory");
#ifdef HSQ_EXPERIMENT_STRIP_FAST
                if constexpr (Sink::HAS_FAST) {
                    // ... over 16 equal, non-gap bases?  Then the whole strip is one extension of the sink's run of matches.
                    if (!not_diag) {
                        uint32_t rc[4], qc[4];
                        load_codes16(ref, j - KROWS, rc);
                        load_codes16(qry, i - KROWS, qc);
                        const uint32_t diff = (rc[0] ^ qc[0]) | (rc[1] ^ qc[1]) | (rc[2] ^ qc[2]) | (rc[3] ^ qc[3]);
                        const uint32_t any = rc[0] | rc[1] | rc[2] | rc[3];
                        if (diff == 0 && (any & (0x01010101u * CODE_GAP)) == 0) {
                            sink.fast_match_n(KROWS);
                            i -= KROWS;
                            j -= KROWS;
                            rw.init(ref, j);
                            qw.init(qry, i);
                            continue;  // (the window now belongs to the strip that was skipped: s != ws reloads it)
                        }
                    }
                }
#endif
                (void)not_diag;
            }
            const uint32_t k = (uint32_t)(whi - j);
            uint2 w;
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(w.x), "=r"(w.y) : "r"(win_s + ((k >> 1) * (16 * DP2_BLOCK)) + ((k & 1u) * 8u)) : "memory");
            const int bit = 15 - r + shift;
            p = (((w.y >> bit) & 1u) << 1) | ((w.x >> bit) & 1u);
        }
        const uint32_t rcc = rw.peek(), qcc = qw.peek();
        if constexpr (Sink::HAS_FAST) {
            // the common step: a diagonal move over two equal bases while the sink is inside a run of matches
            if (p == P_DIAG && rcc == qcc && rcc != CODE_GAP && sink.fast_ready()) {
                sink.fast_match();
                rw.advance(j);
                qw.advance(i);
                i--;
                j--;
                continue;
            }
            // columns after the last query base (they come first here) are not part of the reported alignment
            if (p == P_LEFT && sink.drops_query_gaps()) {
                rw.advance(j);
                j--;
                continue;
            }
        }
        uint32_t rc = CODE_GAP, qc = CODE_GAP;
        if (p != P_UP) {
            rc = rcc;
            rw.advance(j);
            j--;
        }
        if (p != P_LEFT) {
            qc = qcc;
            qw.advance(i);
            i--;
        }
        sink.push(rc | (qc << 8));
    }
    return true;
}

}  // namespace hsq
