// hsq_kernels.cuh -- sm_100a kernels of the HSQUtils hot path (one alignment task per thread, persistent grid).
#pragma once
#include <stdint.h>
#include "hsq_emit.cuh"
#include "../../include/hsq_b200.h"

namespace hsq {

static constexpr int POW95_N = 2048;
__constant__ double c_pow95[POW95_N];  // Math.pow(0.95, d), filled by the host with pow() (ExtendReadsToPrimer.java:326)

static constexpr int PRIMER_ALIGNMENT_BUFFER = 25;             // HSQ/process/ExtendReadsToPrimer.java:54
static constexpr int LENGTH_THRESHOLD_FOR_EXTENDED_READ = 10;  // :56
static constexpr double WEIGHTED_ALIGNMENT_SCORE_THRESHOLD = 4.0;  // :58

// ---------------------------------------------------------------------------------------------
// ASCII -> device codes; 16 bytes per thread.  HBM-bound: 1 B read + 1 B written per base.
// ---------------------------------------------------------------------------------------------
__global__ void encode_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int64_t n, int* __restrict__ status) {
    int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 16;
    bool bad = false;
    for (; i < n; i += stride) {
        if (i + 16 <= n) {
            uint4 v = *reinterpret_cast<const uint4*>(in + i);
            uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; k++) {
                uint32_t o = 0;
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    uint8_t c = encode_ascii((uint8_t)(w[k] >> (8 * b)));
                    bad |= (c == CODE_BAD);
                    o |= (uint32_t)c << (8 * b);
                }
                w[k] = o;
            }
            *reinterpret_cast<uint4*>(out + i) = make_uint4(w[0], w[1], w[2], w[3]);
        } else {
            for (int64_t k = i; k < n; k++) {
                uint8_t c = encode_ascii(in[k]);
                bad |= (c == CODE_BAD);
                out[k] = c;
            }
        }
    }
    if (bad) atomicMin(status, HSQ_ERR_BAD_CHARACTER);
}

// max element length of a ragged batch (offsets has n+1 entries)
__global__ void max_len_kernel(const int64_t* __restrict__ offsets, int64_t n, int* __restrict__ out_max) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int mx = 0;
    for (; i < n; i += stride) mx = max(mx, (int)(offsets[i + 1] - offsets[i]));
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0 && mx > 0) atomicMax(out_max, mx);
}

struct DeviceSeqs {
    const uint8_t* codes;
    const int64_t* offsets;
};

struct DeviceProbes {
    const uint8_t* reverse_strand;
    const int32_t *ext_start, *ext_stop, *lig_start, *lig_stop;
    DeviceSeqs ext, lig, target;
};

struct ExtendDeviceOut {
    int32_t cigar_cap, md_cap;
    uint8_t* ok;
    int32_t *primer_end, *score, *pos, *read_start, *read_stop, *nm, *tlen, *cigar_len, *md_len, *ref_len;
    uint8_t *cigar, *md, *overflow;
};

struct ScratchPool {
    uint32_t* ptr_words;
    int32_t* brow;
    uint16_t* cols;
    int r_stride;
};

__device__ inline DpScratch thread_scratch(const ScratchPool& pool, int4* prof) {
    DpScratch sc;
    sc.prof = prof;
    sc.ptr_words = pool.ptr_words;
    sc.brow = pool.brow;
    sc.cols = pool.cols;
    sc.nthreads = (int64_t)gridDim.x * blockDim.x;
    sc.tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    sc.r_stride = pool.r_stride;
    return sc;
}

// The scan of getPrimerEndIndexInRead (HSQ/process/ExtendReadsToPrimer.java:305-343) as a traceback sink: columns arrive
// last-first, which is the order of the reversed alignments the reference walks (i = 0 is the last column).
struct PrimerGateSink {
    static constexpr bool HAS_FAST = false;
    bool passed = false;
    int pe = 0, first = -1, i = 0;
    double w = 0.0;
    __device__ inline void push(uint32_t c) {
        const uint32_t rc = c & 0xFFu, pc = c >> 8;  // reference = read prefix, query = primer
        if (!passed && pc != CODE_GAP) { passed = true; first = i; }
        if (passed && rc != CODE_GAP) {
            pe++;
            if ((rc & pc & MATCH_MASK) != 0) w += c_pow95[i - first];  // Math.pow(0.95, i - firstNonGapIndex), summed in this order
        }
        i++;
    }
    __device__ inline int result() const { return (w < WEIGHTED_ALIGNMENT_SCORE_THRESHOLD) ? HSQ_PRIMER_NOT_FOUND : pe - 1; }
};

// read[0 .. min(L-1, p+25)] inclusive (:288)
__device__ inline SeqView primer_prefix(const SeqView& primer, const SeqView& read) {
    return sub_view(read, 0, min(read.len, primer.len + PRIMER_ALIGNMENT_BUFFER + 1));
}

// getPrimerEndIndexInRead (HSQ/process/ExtendReadsToPrimer.java:286-345), 32-bit cells
__device__ inline int primer_end_index(const SeqView& primer, const SeqView& read, const ScoreParams& sp_default,
                                       const DpScratch& sc, int* status, int64_t* cells) {
    SeqView prefix = primer_prefix(primer, read);
    if (primer.len == 0) {  // Java: IndexOutOfBounds at :299 (all-gap primer alignment)
        atomicMin(status, HSQ_ERR_OUT_OF_RANGE);
        return HSQ_PRIMER_NOT_FOUND;
    }
    dp_fill(prefix, primer, sp_default, /*pen_start=*/1, /*pen_end=*/0, sc);  // ref = read prefix, query = primer (:290-293)
    *cells += (int64_t)prefix.len * primer.len;
    PrimerGateSink gate;
    dp_traceback_stream(prefix, primer, sc, gate);
    return gate.result();
}

// The extend pipeline (primer stage, bucketing, capture stage) lives in hsq_extend.cuh.
#define HSQ_EXTEND_KERNELS_HERE

// Per pair: both-mates flag (ExtendReadsToPrimer.java:364), TLEN (:136-142,:168-178), and the reference's quirk that
// read two's adjacent primer position is only switched to the ligation primer when read one extended (:133-144).
__global__ void finalize_extend_kernel(DeviceProbes probes, const int32_t* __restrict__ probe_id, int64_t n_pairs,
                                       ExtendDeviceOut out, uint8_t* __restrict__ pair_ok, int* status) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_pairs) return;
    const int pid = probe_id[k];
    const bool one_rev = probes.reverse_strand[pid] != 0;
    const int tl = (int)(probes.target.offsets[pid + 1] - probes.target.offsets[pid]);
    const bool ok1 = out.ok[2 * k] != 0, ok2 = out.ok[2 * k + 1] != 0;
    out.tlen[2 * k] = ok1 ? (one_rev ? -tl : tl) : 0;
    out.tlen[2 * k + 1] = ok2 ? (!one_rev ? -tl : tl) : 0;
    if (!ok1 && ok2) {
        int adjacent = one_rev ? probes.ext_start[pid] : probes.ext_stop[pid];
        out.pos[2 * k + 1] = (!one_rev) ? adjacent - out.ref_len[2 * k + 1] : adjacent + 1;
    }
    pair_ok[k] = (ok1 && ok2) ? 1 : 0;
    if ((out.overflow[2 * k] | out.overflow[2 * k + 1]) != 0) atomicMin(status + 3, HSQ_ERR_OVERFLOW);  // word 3: not fatal
}

struct AlignDeviceOut {
    int32_t cols_cap, md_cap;
    int32_t *score, *n_cols, *first_q, *last_q, *first_r, *last_r, *matches, *mismatches, *non_terminal_gaps, *edit_distance,
        *md_len, *rmd_len;
    uint8_t *ref_aln, *qry_aln, *ops, *md, *rmd, *overflow;
    int32_t* find_counts;  // [n][5] or null
};

// The five counters FastqFindAndTrimTool filters on (FastqFindAndTrimTool.java:932-938) over the columns [first_q .. last_q] of
// getAlignmentWithoutEndingAndBeginningQueryInserts (AlignmentPair.java:95-99; empty when the query has no base: subSequence
// clamps): mismatches, insertion gaps / bases, deletion gaps / bases relative to the reference (AlignmentPair.java:190-256).
__device__ inline void find_counters(const Columns& cv, const PairIndices& ix, int32_t* fc) {
    int fx = 0, ig = 0, ib = 0, dg = 0, db = 0;
    bool in_i = false, in_d = false;
    for (int c = max(ix.first_q, 0); c <= ix.last_q && c < cv.n; c++) {
        const uint32_t w = cv.fwd(c);
        const uint32_t r = col_ref(w), q = col_qry(w);
        if (r == CODE_GAP) { ib++; if (!in_i) ig++; in_i = true; } else in_i = false;
        if (q == CODE_GAP) { db++; if (!in_d) dg++; in_d = true; } else in_d = false;
        if (r != CODE_GAP && q != CODE_GAP && !codes_match(r, q)) fx++;
    }
    fc[0] = fx; fc[1] = ig; fc[2] = ib; fc[3] = dg; fc[4] = db;
}

// Operator seam: NeedlemanWunschGlobalAlignment + AlignmentPair getters, one alignment per thread.
__global__ void __launch_bounds__(DP_BLOCK) align_kernel(DeviceSeqs refs, DeviceSeqs qrys, int64_t n, ScoreParams sp, int pen_start,
                                                    int pen_end, ScratchPool pool, AlignDeviceOut out, int* status,
                                                    unsigned long long* cells_total) {
    __shared__ int4 prof[PROF_INT4_PER_THREAD * DP_BLOCK];
    DpScratch sc = thread_scratch(pool, prof);
    int64_t cells = 0;
    for (int64_t k = sc.tid; k < n; k += sc.nthreads) {
        int64_t r0 = refs.offsets[k], q0 = qrys.offsets[k];
        SeqView ref = make_view(refs.codes + r0, (int)(refs.offsets[k + 1] - r0), false, false);
        SeqView qry = make_view(qrys.codes + q0, (int)(qrys.offsets[k + 1] - q0), false, false);
        int score = dp_fill(ref, qry, sp, pen_start, pen_end, sc);
        cells += (int64_t)ref.len * qry.len;
        Columns cv{sc.colstore(), dp_traceback(ref, qry, sc)};
        PairIndices ix = pair_indices(cv);
        out.score[k] = score;
        out.n_cols[k] = cv.n;
        out.first_q[k] = ix.first_q;
        out.last_q[k] = ix.last_q;
        out.first_r[k] = ix.first_r;
        out.last_r[k] = ix.last_r;
        int m, x, rg, qg;
        count_columns(cv, 0, cv.n - 1, &m, &x, &rg, &qg);
        out.matches[k] = m;
        out.mismatches[k] = x;
        out.non_terminal_gaps[k] = non_terminal_gaps(cv, ix);
        if (out.find_counts) find_counters(cv, ix, out.find_counts + 5 * k);
        uint8_t ovf = cv.n > out.cols_cap ? 1 : 0;
        for (int c = 0; c < cv.n && c < out.cols_cap; c++) {
            uint32_t w = cv.fwd(c);
            out.ref_aln[k * out.cols_cap + c] = decode_code((uint8_t)col_ref(w));
            out.qry_aln[k * out.cols_cap + c] = decode_code((uint8_t)col_qry(w));
            out.ops[k * out.cols_cap + c] = column_op(w);
        }
        SlotWriter mw{out.md + k * out.md_cap, out.md_cap, 0};
        SlotWriter rw{out.rmd + k * out.md_cap, out.md_cap, 0};
        int nm, rl;
        emit_cigar_md(cv, 0, cv.n - 1, false, nullptr, &mw, &nm, &rl);
        out.edit_distance[k] = nm;
        emit_cigar_md(cv, 0, cv.n - 1, true, nullptr, &rw, &nm, &rl);
        out.md_len[k] = mw.len;
        out.rmd_len[k] = rw.len;
        if (mw.len > mw.cap) ovf |= 2;
        if (rw.len > rw.cap) ovf |= 4;
        out.overflow[k] = ovf;
        if (ovf) atomicMin(status, HSQ_ERR_OVERFLOW);
    }
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
}

}  // namespace hsq
