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
// Cell encoding: X = 4*S + p with p = 0 DIAG, 1 UP (bit 0), 2 LEFT (bit 1).  Candidates are formed as
// 4*score + {2,1,0} for h, v, d, so a signed max yields both the score and the tie-broken pointer.
// What a cell hands to its neighbours (c = gapExtend - gapStart):
//   D4    = X & ~3                      to the diagonal neighbour (+ 4*sub there)
//   hout0 = D4 + (X & 2) * 2c           to the right neighbour; the candidate is hout0 + kh, kh = 4*gapStart + 2
//   vout0 = D4 + (X & 1) * 4c           to the lower neighbour; the candidate is vout0 + kv, kv = 4*gapStart + 1
// so one cell is  X = max(hout0_left + kh, max(vout0_up + kv, D4_diag + sub))  = two VIADDMNMX, and its outputs are
// three LOP3 and two IMAD; pointer bits are accumulated with two more IMAD (x*2 + bit).  No table, any integer scorer.
// 4*sub comes from a per-thread profile in shared memory (4 one-hot reference bases x 16 rows), so the match test
// costs no ALU work in the common case; reference codes that are not one of A,C,G,T take a slow column.
//
// Work decomposition: a thread sweeps strips of 16 query rows left to right; the 16 cells of a column live in
// registers; the row above a strip comes from a per-thread boundary row in scratch; the pointers of a strip column
// (16 x 2 bit) are one 32-bit word written to scratch for the traceback.
#pragma once
#include <stdint.h>
#include "hsq_codes.cuh"

namespace hsq {

static constexpr int KROWS = 16;       // rows per strip: 16 x 2 bit = one pointer word
static constexpr int DP_BLOCK = 128;   // threads per block of every DP kernel (sizes the shared profile)
static constexpr int PROF_INT4_PER_THREAD = 20;  // (4 one-hot bases + 1 scratch column) x 16 rows x int32

enum : uint32_t { P_DIAG = 0, P_UP = 1, P_LEFT = 2 };

struct ScoreParams {
    int ms4, mm4;   // 4*match, 4*mismatch
    int kh, kv;     // 4*gapStart + 2, 4*gapStart + 1
    int c2, c4;     // 2*(gapExtend-gapStart), 4*(gapExtend-gapStart)
    int gap_start, gap_extend;
};

inline ScoreParams make_score_params(int match, int mismatch, int gap_extend, int gap_start) {
    ScoreParams s;
    s.ms4 = 4 * match;
    s.mm4 = 4 * mismatch;
    s.kh = 4 * gap_start + 2;
    s.kv = 4 * gap_start + 1;
    s.c2 = 2 * (gap_extend - gap_start);
    s.c4 = 4 * (gap_extend - gap_start);
    s.gap_start = gap_start;
    s.gap_extend = gap_extend;
    return s;
}

// Traceback columns of one thread (the view hsq_emit.cuh works on).
struct ColStore {
    uint16_t* cols;
    int64_t nthreads;
    int64_t tid;
    __device__ inline uint16_t& col(int k) const { return cols[(int64_t)k * nthreads + tid]; }
};

// Per-thread scratch, interleaved by thread so that a warp's accesses coalesce: element e of thread t
// sits at base[e * nthreads + t].
struct DpScratch {
    uint32_t* ptr_words;  // [n_strips_max * (r_max + 1)] per thread
    int32_t* brow;        // [r_max + 1] per thread: X of the row above the current strip
    uint16_t* cols;       // [q_max + r_max] per thread: traceback columns (ref code | qry code << 8), last column first
    int4* prof;           // shared memory: [PROF_INT4_PER_THREAD][DP_BLOCK]
    int64_t nthreads;
    int64_t tid;
    int r_stride;         // r_max + 1
    __device__ inline uint32_t& word(int strip, int j) const {
        return ptr_words[((int64_t)strip * r_stride + j) * nthreads + tid];
    }
    __device__ inline int32_t& b(int j) const { return brow[(int64_t)j * nthreads + tid]; }
    __device__ inline uint16_t& col(int k) const { return cols[(int64_t)k * nthreads + tid]; }
    __device__ inline ColStore colstore() const { return ColStore{cols, nthreads, tid}; }
};

// init score of a border cell k >= 1 (getInitialScore :342-376): gapStart + (k-1)*(gapScore(0)-1)
__device__ __forceinline__ int border_score(int k, int pen_start, int gap_start) {
    return pen_start ? gap_start - (k - 1) : 0;
}

// acc * 2 + bit as one IMAD the compiler cannot re-associate (it otherwise rebuilds the bits with extra LOP3s)
__device__ __forceinline__ uint32_t mad2(uint32_t acc, uint32_t bit) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, 2, %2;" : "=r"(r) : "r"(acc), "r"(bit));
    return r;
}

// One column of one strip.  State per row r: D4[r], H0[r] (= hout0) of cell (i0+1+r, j-1) on entry, of column j on exit.
//   up_v0: vout0 of cell (i0, j);  diag_d4: D4 of cell (i0, j-1);  pcol: this thread's profile column (4*sub per row).
// Returns the packed X of the strip's last row (for the boundary row); *word_out gets the pointer word.
template <bool HFREE, bool VFREE>
__device__ __forceinline__ int dp_column(int (&D4)[KROWS], int (&H0)[KROWS], const int4* __restrict__ pcol, int up_v0, int diag_d4,
                                         const ScoreParams& sp, int rstar, uint32_t* word_out) {
    uint32_t acc_l = 0, acc_u = 0;
    int vin = up_v0;
    int dg = diag_d4;
    int x = 0;
#pragma unroll
    for (int r = 0; r < KROWS; r++) {
        int4 sv;
        if ((r & 3) == 0) sv = pcol[(r >> 2) * DP_BLOCK];  // 4 rows of 4*sub per LDS.128, loaded just before use
        const int sub = (r & 3) == 0 ? sv.x : (r & 3) == 1 ? sv.y : (r & 3) == 2 ? sv.z : sv.w;
        const int dsub = dg + sub;
        x = __viaddmax_s32(vin, sp.kv, __viaddmax_s32(H0[r], sp.kh, dsub));  // only the outer max is on the row-to-row chain
        dg = D4[r];
        const int d4 = x & ~3;
        const int b1x2 = x & 2;
        const int b0 = x & 1;
        int h0 = b1x2 * sp.c2 + d4;
        int v0 = b0 * sp.c4 + d4;
        acc_l = mad2(acc_l, (uint32_t)b1x2);
        acc_u = mad2(acc_u, (uint32_t)b0);
        if (HFREE) h0 = (r == rstar) ? d4 + (2 - sp.kh) : h0;  // bottom row, terminal gaps free (:286-300)
        if (VFREE) v0 = d4 + (1 - sp.kv);                       // last column, terminal gaps free (:270-284)
        D4[r] = d4;
        H0[r] = h0;
        vin = v0;
    }
    *word_out = (acc_l << 15) | acc_u;  // bit1 of row r at 31-r, bit0 at 15-r
    return x;
}

}  // namespace hsq

#include "hsq_dp_fill.cuh"
