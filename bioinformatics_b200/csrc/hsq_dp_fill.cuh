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

// profile slot of a raw reference code: 0..3 for A,C,G,T (after the view's complement), 4 = build the column on the fly
__device__ __forceinline__ int prof_slot(uint32_t raw, int comp) {
    const bool one_hot = (raw - 1u) < 8u && (raw & (raw - 1u)) == 0u;
    const int idx = 31 - __clz((int)raw);
    return one_hot ? (comp ? 3 - idx : idx) : 4;
}

// One strip (16 query rows starting below row i0) swept over all R columns.  The row above the strip is read from the
// boundary row in scratch (X per column) and replaced by the strip's last row.  Reference code and boundary value of
// column j+1 are fetched while column j is computed.
template <bool HFREE>
__device__ __forceinline__ void dp_strip(const SeqView& ref, const SeqView& qry, const ScoreParams& sp, int pen_start, bool free_end,
                                         const DpScratch& sc, int s, bool last_strip, int* final_d4) {
    const int R = ref.len, Q = qry.len;
    const int i0 = s * KROWS;
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
            H0[r] = (HFREE && r == rstar) ? s0 + (2 - sp.kh) : s0;
        }
        build_profile(q, sp, sc.prof);
    }
    int diag_d4 = (i0 == 0) ? 0 : 4 * border_score(i0, pen_start, sp.gap_start);
    uint32_t* wordp = &sc.word(s, 1);
    int32_t* browp = &sc.b(1);
    const uint8_t* rp = ref.base + ref.start;
    const int64_t rstep = ref.step;
    const int64_t nt = sc.nthreads;
    uint32_t next_raw = *rp;
    int next_up = *browp;
    for (int j = 1; j <= R; j++) {
        const uint32_t raw = next_raw;
        const int upx = next_up;
        if (j < R) {  // prefetch column j+1
            rp += rstep;
            next_raw = *rp;
            next_up = browp[nt];
        }
        const int up_d4 = upx & ~3;
        const bool vfree = free_end && (j == R);
        const int up_v0 = vfree ? up_d4 + (1 - sp.kv) : (upx & 1) * sp.c4 + up_d4;
        int b = prof_slot(raw, ref.comp);
        if (b == 4) {  // ambiguity code, N, U or gap character in the reference: build this column's scores in the scratch slot
            const uint32_t rcode = ref.comp ? complement_code((uint8_t)raw) : raw;
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
        if (!vfree) xlast = dp_column<HFREE, false>(D4, H0, pcol, up_v0, diag_d4, sp, rstar, &w);
        else xlast = dp_column<HFREE, true>(D4, H0, pcol, up_v0, diag_d4, sp, rstar, &w);
        *wordp = w;
        if (!last_strip) *browp = xlast;
        diag_d4 = up_d4;
        wordp += nt;
        browp += nt;
    }
    if (last_strip) {
#pragma unroll
        for (int r = 0; r < KROWS; r++)
            if (r == rstar) *final_d4 = D4[r];
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
    // boundary row for strip 0 = matrix row 0: cells point LEFT (never an UP continuation)
    {
        int32_t* bp = &sc.b(1);
        for (int j = 1; j <= R; j++, bp += sc.nthreads) *bp = 4 * border_score(j, pen_start, sp.gap_start) + (int)P_LEFT;
    }
    int final_d4 = 0;
    for (int s = 0; s < n_strips - 1; s++) dp_strip<false>(ref, qry, sp, pen_start, free_end, sc, s, false, &final_d4);
    if (free_end) dp_strip<true>(ref, qry, sp, pen_start, true, sc, n_strips - 1, true, &final_d4);
    else dp_strip<false>(ref, qry, sp, pen_start, false, sc, n_strips - 1, true, &final_d4);
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

// Streamed traceback: same walk, but every column (last column first) goes to sink.push(ref code | qry code << 8)
// instead of the column store.
template <class Sink>
__device__ __forceinline__ void dp_traceback_stream(const SeqView& ref, const SeqView& qry, const DpScratch& sc, Sink& sink) {
    int i = qry.len, j = ref.len;
    while (i > 0 || j > 0) {
        const uint32_t p = dp_pointer(i, j, sc);
        uint32_t rc = CODE_GAP, qc = CODE_GAP;
        if (p != P_UP) { rc = ref.get(j - 1); j--; }
        if (p != P_LEFT) { qc = qry.get(i - 1); i--; }
        sink.push(rc | (qc << 8));
    }
}

}  // namespace hsq
