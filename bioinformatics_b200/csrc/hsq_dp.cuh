// hsq_dp.cuh -- the DP fill of NeedlemanWunschGlobalAlignment, one alignment per thread.
//
// Recurrence (COMMON/alignment/NeedlemanWunschGlobalAlignment.java:261-328, init :342-376; rows = query,
// columns = reference):
//   v = S[i-1][j] + (j==R && !penalizeEnd ? 0 : (P[i-1][j]==UP   ? gapExtend : gapStart))
//   h = S[i][j-1] + (i==Q && !penalizeEnd ? 0 : (P[i][j-1]==LEFT ? gapExtend : gapStart))
//   d = S[i-1][j-1] + (matches(q[i-1], r[j-1]) ? match : mismatch)
//   S = max(h,v,d);  P = (S==h) ? LEFT : (S==v) ? UP : DIAG          -- tie order H > V > D
// This is NOT Gotoh: "continuation" is read off the neighbour's chosen pointer.
//
// Cell encoding: X = 4*S + p with p = 0 DIAG, 1 UP, 2 LEFT.  Candidates are formed as 4*score + {2,1,0}
// for h, v, d, so ONE signed 3-way max (VIMNMX3) yields both the score and the tie-broken pointer.
// What a cell hands to its neighbours:
//   D4   = X & ~3                           (to the diagonal neighbour; + 4*sub there)
//   Hout = D4 + 2 + 4*(p==LEFT ? ge : gs)   (to the right neighbour)   = X - hpen[p]
//   Vout = D4 + 1 + 4*(p==UP   ? ge : gs)   (to the lower neighbour)   = X - vpen[p]
// hpen/vpen are 3-entry byte tables held in one register each and indexed with PRMT.
//
// Work decomposition: a thread sweeps strips of K query rows left to right; the K cells of a column
// live in registers; the row above a strip comes from a per-thread boundary row in scratch; pointers of a
// strip column (K x 2 bit) are one 32-bit word written to scratch for the traceback.
#pragma once
#include <stdint.h>
#include "hsq_codes.cuh"

namespace hsq {

static constexpr int KROWS = 16;  // rows per strip: 16 x 2 bit = one pointer word

enum : uint32_t { P_DIAG = 0, P_UP = 1, P_LEFT = 2 };

struct ScoreParams {
    int ms4, mm4;          // 4*match, 4*mismatch
    int hs, he;            // 4*gapStart+2, 4*gapExtend+2
    int vs, ve;            // 4*gapStart+1, 4*gapExtend+1
    uint32_t tabH, tabV;   // byte LUTs hpen[p], vpen[p] (valid when lut_ok)
    int lut_ok;
    int gap_start, gap_extend;
};

inline ScoreParams make_score_params(int match, int mismatch, int gap_extend, int gap_start) {
    ScoreParams s;
    s.ms4 = 4 * match;
    s.mm4 = 4 * mismatch;
    s.hs = 4 * gap_start + 2;
    s.he = 4 * gap_extend + 2;
    s.vs = 4 * gap_start + 1;
    s.ve = 4 * gap_extend + 1;
    s.gap_start = gap_start;
    s.gap_extend = gap_extend;
    int h[3] = {0 - s.hs, 1 - s.hs, 2 - s.he};
    int v[3] = {0 - s.vs, 1 - s.ve, 2 - s.vs};
    s.lut_ok = 1;
    for (int k = 0; k < 3; k++)
        if (h[k] < 0 || h[k] > 255 || v[k] < 0 || v[k] > 255) s.lut_ok = 0;
    s.tabH = s.lut_ok ? ((uint32_t)h[0] | ((uint32_t)h[1] << 8) | ((uint32_t)h[2] << 16)) : 0;
    s.tabV = s.lut_ok ? ((uint32_t)v[0] | ((uint32_t)v[1] << 8) | ((uint32_t)v[2] << 16)) : 0;
    return s;
}

// Per-thread scratch, interleaved by thread so that a warp's accesses coalesce: element e of thread t
// sits at base[e * nthreads + t].
struct DpScratch {
    uint32_t* ptr_words;  // [n_strips_max * (r_max + 1)] per thread
    int32_t* brow;        // [r_max + 1] per thread: X of the row above the current strip
    uint16_t* cols;       // [q_max + r_max] per thread: traceback columns (ref code | qry code << 8), last column first
    int64_t nthreads;
    int64_t tid;
    int r_stride;         // r_max + 1
    __device__ inline uint32_t& word(int strip, int j) const {
        return ptr_words[((int64_t)strip * r_stride + j) * nthreads + tid];
    }
    __device__ inline int32_t& b(int j) const { return brow[(int64_t)j * nthreads + tid]; }
    __device__ inline uint16_t& col(int k) const { return cols[(int64_t)k * nthreads + tid]; }
};

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}

// (a & b) | c in one LOP3
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
// a * 4 + b on the FMA pipe (kept opaque so the compiler does not re-derive it as shifts and masks on the ALU pipe)
__device__ __forceinline__ uint32_t mad4(uint32_t a, uint32_t b) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, 4, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

// init score of a border cell k >= 1 (getInitialScore :342-376): gapStart + (k-1)*(gapScore(0)-1)
__device__ __forceinline__ int border_score(int k, int pen_start, int gap_start) {
    return pen_start ? gap_start - (k - 1) : 0;
}

// One column of one strip.  State per row r: D4[r], Hout[r] of cell (i0+1+r, j-1) on entry, of column j on exit.
//   up_d4 / up_vout: what cell (i0, j) hands down;  diag_d4: D4 of cell (i0, j-1).
// Returns the packed X of the strip's last row (for the boundary row); *acc_out gets the pointer word.
template <bool LUT, bool HFREE, bool VFREE>
__device__ __forceinline__ int dp_column(int (&D4)[KROWS], int (&Hout)[KROWS], const uint32_t (&q)[KROWS], uint32_t rcode,
                                         int up_d4, int up_vout, int diag_d4, const ScoreParams& sp, int rstar,
                                         uint32_t* acc_out) {
    uint32_t acc = 0;
    int vin = up_vout;
    int dg = diag_d4;
    int m = 0;
    (void)up_d4;
#pragma unroll
    for (int r = 0; r < KROWS; r++) {
        int sub = ((q[r] & rcode & MATCH_MASK) != 0) ? sp.ms4 : sp.mm4;
        int dc = dg + sub;
        m = __vimax3_s32(Hout[r], vin, dc);
        dg = D4[r];
        int d4 = m & ~3;
        D4[r] = d4;
        int hout, vout;
        if (LUT) {
            uint32_t sel = lop3_and_or((uint32_t)m, 3u, 0x4440u);
            hout = m - (int)prmt(sp.tabH, 0u, sel);
            vout = m - (int)prmt(sp.tabV, 0u, sel);
            acc = mad4(acc, sel);
        } else {
            uint32_t p = (uint32_t)m & 3u;
            hout = d4 + (p == P_LEFT ? sp.he : sp.hs);
            vout = d4 + (p == P_UP ? sp.ve : sp.vs);
            acc = acc * 4u + p;
        }
        if (HFREE) hout = (r == rstar) ? d4 + 2 : hout;  // bottom row, terminal gaps free (:286-300)
        if (VFREE) vout = d4 + 1;                         // last column, terminal gaps free (:270-284)
        Hout[r] = hout;
        vin = vout;
    }
    if (LUT) acc -= 0x4440u * 0x55555555u;  // the constant the sel bits 0x4440 accumulated (KROWS == 16)
    *acc_out = acc;
    return m;
}

// Fills the matrix for (ref R columns, query Q rows); pointer words go to scratch.  Returns the score S[Q][R].
template <bool LUT>
__device__ inline int dp_fill(const SeqView& ref, const SeqView& qry, const ScoreParams& sp, int pen_start, int pen_end,
                              const DpScratch& sc) {
    const int R = ref.len, Q = qry.len;
    if (Q == 0) return R == 0 ? 0 : border_score(R, pen_start, sp.gap_start);
    if (R == 0) return border_score(Q, pen_start, sp.gap_start);
    const int n_strips = (Q + KROWS - 1) / KROWS;
    const bool free_end = !pen_end;
    int final_d4 = 0;
    for (int s = 0; s < n_strips; s++) {
        const int i0 = s * KROWS;  // row above the strip
        const bool last_strip = (s == n_strips - 1);
        const int rstar = last_strip ? (Q - 1 - i0) : -1;  // strip-local index of row Q
        uint32_t q[KROWS];
        int D4[KROWS], Hout[KROWS];
#pragma unroll
        for (int r = 0; r < KROWS; r++) {
            int i = i0 + 1 + r;
            q[r] = (i <= Q) ? (uint32_t)qry.get(i - 1) : 0u;
            int s0 = 4 * border_score(i, pen_start, sp.gap_start);
            D4[r] = s0;
            Hout[r] = (free_end && r == rstar) ? s0 + 2 : s0 + sp.hs;  // column-0 cells point UP: never a LEFT continuation
        }
        int diag_d4 = (i0 == 0) ? 0 : 4 * border_score(i0, pen_start, sp.gap_start);
        const bool hfree = free_end && last_strip;
        for (int j = 1; j <= R; j++) {
            const bool vfree = free_end && (j == R);
            uint32_t rcode = (uint32_t)ref.get(j - 1);
            int up_d4, up_vout;
            if (s == 0) {  // row 0 cells point LEFT: never an UP continuation
                up_d4 = 4 * border_score(j, pen_start, sp.gap_start);
                up_vout = vfree ? up_d4 + 1 : up_d4 + sp.vs;
            } else {
                int x = sc.b(j);
                up_d4 = x & ~3;
                up_vout = vfree ? up_d4 + 1 : up_d4 + (((uint32_t)x & 3u) == P_UP ? sp.ve : sp.vs);
            }
            uint32_t acc;
            int mlast;
            if (!hfree) {
                if (!vfree) mlast = dp_column<LUT, false, false>(D4, Hout, q, rcode, up_d4, up_vout, diag_d4, sp, rstar, &acc);
                else mlast = dp_column<LUT, false, true>(D4, Hout, q, rcode, up_d4, up_vout, diag_d4, sp, rstar, &acc);
            } else {
                if (!vfree) mlast = dp_column<LUT, true, false>(D4, Hout, q, rcode, up_d4, up_vout, diag_d4, sp, rstar, &acc);
                else mlast = dp_column<LUT, true, true>(D4, Hout, q, rcode, up_d4, up_vout, diag_d4, sp, rstar, &acc);
            }
            sc.word(s, j) = acc;
            if (!last_strip) sc.b(j) = mlast;
            diag_d4 = up_d4;
        }
        if (last_strip) {
#pragma unroll
            for (int r = 0; r < KROWS; r++)
                if (r == rstar) final_d4 = D4[r];
        }
    }
    return final_d4 >> 2;
}

// Pointer of cell (i,j), i,j >= 0, not both 0 (getInitialSourceCell :342-352 for the borders).
__device__ __forceinline__ uint32_t dp_pointer(int i, int j, const DpScratch& sc) {
    if (i == 0) return P_LEFT;
    if (j == 0) return P_UP;
    int s = (i - 1) / KROWS, r = (i - 1) % KROWS;
    uint32_t w = sc.word(s, j);
    return (w >> (2 * (KROWS - 1 - r))) & 3u;
}

// Traceback (:85-114): walks from (Q,R) to (0,0) and stores the columns, LAST column first, as
// (ref code | qry code << 8) with CODE_GAP where the pointer did not consume that sequence.  Returns #columns.
__device__ inline int dp_traceback(const SeqView& ref, const SeqView& qry, const DpScratch& sc) {
    int i = qry.len, j = ref.len, n = 0;
    while (i > 0 || j > 0) {
        uint32_t p = dp_pointer(i, j, sc);
        uint32_t rc = CODE_GAP, qc = CODE_GAP;
        if (p != P_UP) { rc = ref.get(j - 1); }
        if (p != P_LEFT) { qc = qry.get(i - 1); }
        if (p != P_UP) j--;
        if (p != P_LEFT) i--;
        sc.col(n) = (uint16_t)(rc | (qc << 8));
        n++;
    }
    return n;
}

}  // namespace hsq
