// hsq_dp_fill.cuh -- strip sweep, shared-memory profile, pointer lookup and traceback (included by hsq_dp.cuh).
#pragma once

namespace hsq {

__device__ __forceinline__ int match4(uint32_t q, uint32_t r, const ScoreParams& sp) {
    return ((q & r & MATCH_MASK) != 0) ? sp.ms4 : sp.mm4;
}

// Profile of one strip: for each of the 4 one-hot reference bases, 4*sub of the strip's 16 query rows.
// Layout [base][row/4][thread] as int4, so a warp's LDS.128 is conflict free whatever base each lane wants.
__device__ inline void build_profile(const uint32_t (&q)[KROWS], const ScoreParams& sp, int4* prof) {
#pragma unroll
    for (int b = 0; b < 4; b++) {
        const uint32_t code = 1u << b;
#pragma unroll
        for (int ch = 0; ch < 4; ch++) {
            int4 v;
            v.x = match4(q[4 * ch + 0], code, sp);
            v.y = match4(q[4 * ch + 1], code, sp);
            v.z = match4(q[4 * ch + 2], code, sp);
            v.w = match4(q[4 * ch + 3], code, sp);
            prof[(b * 4 + ch) * DP_BLOCK + threadIdx.x] = v;
        }
    }
}

// Fills the matrix for (ref R columns, query Q rows); pointer words go to scratch.  Returns the score S[Q][R].
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
        int D4[KROWS], H0[KROWS];
        {
            uint32_t q[KROWS];
#pragma unroll
            for (int r = 0; r < KROWS; r++) {
                const int i = i0 + 1 + r;
                q[r] = (i <= Q) ? (uint32_t)qry.get(i - 1) : 0u;
                const int s0 = 4 * border_score(i, pen_start, sp.gap_start);
                D4[r] = s0;
                // column-0 cells point UP: never a LEFT continuation -> hout0 = D4
                H0[r] = (free_end && r == rstar) ? s0 + (2 - sp.kh) : s0;
            }
            build_profile(q, sp, sc.prof);
        }
        int diag_d4 = (i0 == 0) ? 0 : 4 * border_score(i0, pen_start, sp.gap_start);
        const bool hfree = free_end && last_strip;
        uint32_t* wordp = &sc.word(s, 1);
        int32_t* browp = &sc.b(1);
        for (int j = 1; j <= R; j++, wordp += sc.nthreads, browp += sc.nthreads) {
            const bool vfree = free_end && (j == R);
            const uint32_t rcode = (uint32_t)ref.get(j - 1);
            int up_d4, up_v0;
            if (s == 0) {  // row-0 cells point LEFT: never an UP continuation
                up_d4 = 4 * border_score(j, pen_start, sp.gap_start);
                up_v0 = up_d4;
            } else {
                const int x = *browp;
                up_d4 = x & ~3;
                up_v0 = (x & 1) * sp.c4 + up_d4;
            }
            if (vfree) up_v0 = up_d4 + (1 - sp.kv);
            int b;
            if (rcode == 1u || rcode == 2u || rcode == 4u || rcode == 8u) {
                b = 31 - __clz(rcode);
            } else {  // ambiguity code, N, U or gap character in the reference: build this column's scores in the scratch slot
                b = 4;
                int* slot = reinterpret_cast<int*>(sc.prof + 16 * DP_BLOCK + threadIdx.x);
#pragma unroll 1
                for (int r = 0; r < KROWS; r++) {
                    const int i = i0 + 1 + r;
                    slot[(r >> 2) * DP_BLOCK * 4 + (r & 3)] = match4((i <= Q) ? (uint32_t)qry.get(i - 1) : 0u, rcode, sp);
                }
            }
            const int4* pcol = sc.prof + b * 4 * DP_BLOCK + threadIdx.x;
            uint32_t w;
            int xlast;
            if (!hfree) {
                if (!vfree) xlast = dp_column<false, false>(D4, H0, pcol, up_v0, diag_d4, sp, rstar, &w);
                else xlast = dp_column<false, true>(D4, H0, pcol, up_v0, diag_d4, sp, rstar, &w);
            } else {
                if (!vfree) xlast = dp_column<true, false>(D4, H0, pcol, up_v0, diag_d4, sp, rstar, &w);
                else xlast = dp_column<true, true>(D4, H0, pcol, up_v0, diag_d4, sp, rstar, &w);
            }
            *wordp = w;
            if (!last_strip) *browp = xlast;
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
    const int s = (i - 1) / KROWS, r = (i - 1) % KROWS;
    const uint32_t w = sc.word(s, j);
    return (((w >> (31 - r)) & 1u) << 1) | ((w >> (15 - r)) & 1u);
}

// Traceback (:85-114): walks from (Q,R) to (0,0) and stores the columns, LAST column first, as
// (ref code | qry code << 8) with CODE_GAP where the pointer did not consume that sequence.  Returns #columns.
__device__ inline int dp_traceback(const SeqView& ref, const SeqView& qry, const DpScratch& sc) {
    int i = qry.len, j = ref.len, n = 0;
    while (i > 0 || j > 0) {
        const uint32_t p = dp_pointer(i, j, sc);
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
