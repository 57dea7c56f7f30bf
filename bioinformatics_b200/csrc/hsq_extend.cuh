// hsq_extend.cuh -- the ExtendReadsToPrimer pipeline on the device:
//   primer stage    DP#1 (primer x read prefix) + the 0.95^d gate           -> primer end index e, DP#2 bucket key
//   bucket sort     counting sort of the read ends by exact DP shape (rows, columns), buckets padded to even sizes
//   capture stage   DP#2 (capture target x read remainder) + streamed traceback + CIGAR / MD / NM / POS, in bucket order
// Each stage exists twice: packed (two read ends of the same DP shape per thread, u16x2 cells, hsq_dp2.cuh) -- the path
// HSQUtils' scorers take -- and 32-bit (one read end per thread, hsq_dp.cuh) for scorers whose scores do not fit 14 bits.
// Reference: HSQ/process/ExtendReadsToPrimer.java:96-245,286-345.
#pragma once
#include "hsq_kernels.cuh"
#include "hsq_dp2.cuh"

namespace hsq {

struct ReadEndViews {
    SeqView primer, capture, read;
    bool is_reversed;
    int adjacent;
};

// extendReadPair strand logic (:96-198): which sequences, in which orientation, a read end is aligned against.
__device__ inline ReadEndViews read_end_views(const DeviceProbes& probes, const int32_t* __restrict__ probe_id, const DeviceSeqs& r1,
                                              const DeviceSeqs& r2, int64_t m) {
    ReadEndViews v;
    const int64_t k = m >> 1;
    const int mate = (int)(m & 1);
    const int pid = probe_id[k];
    const bool one_rev = probes.reverse_strand[pid] != 0;
    const int64_t t0 = probes.target.offsets[pid];
    const int tl = (int)(probes.target.offsets[pid + 1] - t0);
    if (mate == 0) {  // :114-126
        const int64_t p0 = probes.ext.offsets[pid];
        v.primer = make_view(probes.ext.codes + p0, (int)(probes.ext.offsets[pid + 1] - p0), false, one_rev);
        v.capture = make_view(probes.target.codes + t0, tl, false, one_rev);
        const int64_t s0 = r1.offsets[k];
        v.read = make_view(r1.codes + s0, (int)(r1.offsets[k + 1] - s0), false, one_rev);
        v.is_reversed = one_rev;
        v.adjacent = one_rev ? probes.ext_start[pid] : probes.ext_stop[pid];  // :108-112
    } else {  // :150-162 (sequenceTwo was complemented at :104)
        const int64_t p0 = probes.lig.offsets[pid];
        v.primer = make_view(probes.lig.codes + p0, (int)(probes.lig.offsets[pid + 1] - p0), true, one_rev);
        v.capture = make_view(probes.target.codes + t0, tl, true, one_rev);
        const int64_t s0 = r2.offsets[k];
        v.read = make_view(r2.codes + s0, (int)(r2.offsets[k + 1] - s0), false, !one_rev);
        v.is_reversed = !one_rev;
        v.adjacent = one_rev ? probes.lig_stop[pid] : probes.lig_start[pid];  // :137,:143 (finalize restores the :133 quirk)
    }
    return v;
}

struct BucketKeys {
    uint32_t* key;      // [2n] bucket of each read end
    int key_stride;     // r_max + 1
    uint32_t skip_key;  // read ends that do not reach DP#2
};

// What follows DP#1 for one read end: primer end index, DP#2 bucket key; read ends that stop here get their (all-zero)
// results now.
__device__ __forceinline__ void primer_finish(const ReadEndViews& v, int e, const ExtendDeviceOut& out, const BucketKeys& keys, int64_t m) {
    const int L = v.read.len;
    out.primer_end[m] = e;
    const bool reaches = (e != HSQ_PRIMER_NOT_FOUND && e >= 0 && e < L);  // :206
    if (reaches) {
        const int q = L - (e + 1);
        keys.key[m] = (uint32_t)(q * keys.key_stride + v.capture.len);  // exact DP#2 shape (Q rows, R columns)
    } else {
        keys.key[m] = keys.skip_key;
        out.ok[m] = 0;
        out.score[m] = 0;
        out.pos[m] = 0;
        out.read_start[m] = 0;
        out.read_stop[m] = 0;
        out.nm[m] = 0;
        out.ref_len[m] = 0;
        out.cigar_len[m] = 0;
        out.md_len[m] = 0;
        out.overflow[m] = 0;
    }
}

// Stage 1 (32-bit cells, natural order): getPrimerEndIndexInRead for every read end.
__global__ void __launch_bounds__(DP_BLOCK) primer_kernel(DeviceProbes probes, const int32_t* __restrict__ probe_id, DeviceSeqs r1,
                                                          DeviceSeqs r2, int64_t n_pairs, ScoreParams sp_default, ScratchPool pool,
                                                          ExtendDeviceOut out, BucketKeys keys, int* status,
                                                          unsigned long long* cells_total) {
    __shared__ int4 prof[PROF_INT4_PER_THREAD * DP_BLOCK];
    DpScratch sc = thread_scratch(pool, prof);
    int64_t cells = 0;
    const int64_t n_tasks = 2 * n_pairs;
    for (int64_t m = sc.tid; m < n_tasks; m += sc.nthreads) {
        ReadEndViews v = read_end_views(probes, probe_id, r1, r2, m);
        const int e = primer_end_index(v.primer, v.read, sp_default, sc, status, &cells);
        primer_finish(v, e, out, keys, m);
    }
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
}

// Bucket key of DP#1 (rows = primer length, columns = read prefix length) for the packed primer stage.
__global__ void primer_key_kernel(DeviceProbes probes, const int32_t* __restrict__ probe_id, DeviceSeqs r1, DeviceSeqs r2, int64_t n_pairs,
                                  BucketKeys keys) {
    int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; m < 2 * n_pairs; m += stride) {
        const int64_t k = m >> 1;
        const int pid = probe_id[k];
        const DeviceSeqs& ps = (m & 1) ? probes.lig : probes.ext;
        const DeviceSeqs& rs = (m & 1) ? r2 : r1;
        const int p = (int)(ps.offsets[pid + 1] - ps.offsets[pid]);
        const int L = (int)(rs.offsets[k + 1] - rs.offsets[k]);
        keys.key[m] = (uint32_t)(p * keys.key_stride + min(L, p + PRIMER_ALIGNMENT_BUFFER + 1));
    }
}

// ---- counting sort of read ends by bucket key (order inside a bucket is irrelevant: results are written by id) ----
__global__ void bucket_hist_kernel(const uint32_t* __restrict__ key, int64_t n, uint32_t* __restrict__ hist) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i - threadIdx.x % 32 < n; i += stride) {  // whole warps stay in the loop for the match
        const bool valid = i < n;
        const uint32_t k = valid ? key[i] : 0xffffffffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, k);
        if (valid && (int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&hist[k], (uint32_t)__popc(peers));
    }
}

// exclusive scan of the histogram into bucket cursors, one block
// (every bucket is rounded up to an even size so that the packed kernel's pairs never straddle two shapes)
__global__ void bucket_scan_kernel(uint32_t* __restrict__ hist, int n_keys, uint32_t* __restrict__ cursor) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int base = 0; base < n_keys; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < n_keys ? ((hist[i] + 1u) & ~1u) : 0;
        uint32_t x = v;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_sums[wid] = x;
        __syncthreads();
        if (wid == 0) {
            uint32_t s = lane < nw ? warp_sums[lane] : 0;
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
                if (lane >= o) s += y;
            }
            warp_sums[lane] = s;
        }
        __syncthreads();
        const uint32_t before = carry + (wid > 0 ? warp_sums[wid - 1] : 0) + x - v;
        if (i < n_keys) cursor[i] = before;
        __syncthreads();
        if (threadIdx.x == 0) carry += warp_sums[nw - 1];
        __syncthreads();
    }
}

__global__ void bucket_scatter_kernel(const uint32_t* __restrict__ key, int64_t n, uint32_t* __restrict__ cursor,
                                      uint32_t* __restrict__ order) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    for (; i - lane < n; i += stride) {
        const bool valid = i < n;
        const uint32_t k = valid ? key[i] : 0xffffffffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, k);
        const int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if (valid && leader == lane) base = atomicAdd(&cursor[k], (uint32_t)__popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (valid) order[base + __popc(peers & ((1u << lane) - 1u))] = (uint32_t)i;
    }
}

// What follows DP#2 for one read end (:212-236): acceptance test, streamed traceback into the CIGAR / MD state machine,
// NM / POS, result stores.  `traceback(sink)` walks the alignment last column first.
template <class Traceback>
__device__ __forceinline__ int capture_finish(const ReadEndViews& v, const SeqView& rw, int start, int score, Traceback traceback,
                                               const ExtendDeviceOut& out, int64_t m) {
    int ok = 0, pos = 0, read_start = 0, read_stop = 0;
    CigarMdStream st;
    // a reverse-strand read emits in traceback order; a forward read fills its slots right to left
    st.init(out.cigar + m * out.cigar_cap, out.cigar_cap, out.md + m * out.md_cap, out.md_cap, !v.is_reversed);
    const int read_length = rw.len - 1;  // :225
    // score/alignmentLength > 0 (:212): the alignment has max(Q,R) >= 1 columns whenever score != 0
    if (score > 0 && read_length > LENGTH_THRESHOLD_FOR_EXTENDED_READ) {
#ifndef HSQ_EXPERIMENT_NO_TRACEBACK  // (development experiment: upper bound of what the traceback costs)
        traceback(st);  // trailing query-gap columns are dropped by the sink (getAlignmentWithoutEndingQueryInserts :218)
        st.finish();    // :220-231
#endif
        pos = v.is_reversed ? v.adjacent - st.ref_len : v.adjacent + 1;  // :234,:236
        read_start = start;
        read_stop = start + read_length;
        ok = 1;
    }
    out.ok[m] = (uint8_t)ok;
    out.score[m] = score;
    out.pos[m] = pos;
    out.read_start[m] = read_start;
    out.read_stop[m] = read_stop;
    out.nm[m] = st.nm;
    out.ref_len[m] = st.ref_len;
    out.cigar_len[m] = st.cw.len;
    out.md_len[m] = st.mw.len;
    out.overflow[m] = (uint8_t)((st.cw.len > st.cw.cap ? 1 : 0) | (st.mw.len > st.mw.cap ? 2 : 0));
    return ok;
}

static constexpr uint32_t ORDER_HOLE = 0xFFFFFFFFu;  // padding entry of an odd-sized bucket

// Dynamic work distribution of the packed kernels: a warp takes the next 32 pair slots from a device counter, so the
// last wave of a launch is shared out instead of leaving the threads of a static stride idle (matters for the short
// launches of the chunked host pipeline: ~3 slots per resident thread).
__device__ __forceinline__ int64_t next_slot(unsigned long long* counter) {
    unsigned long long base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(counter, 32ull);
    base = __shfl_sync(0xffffffffu, base, 0);
    return (int64_t)base + (threadIdx.x & 31);
}

// Stage 2: calculateDetailsForReadExtensionToPrimer (:200-245) for the read ends that reached DP#2, in bucket order.
// 32-bit cells, one read end per thread: the general kernel (any integer scorer, any length).
__global__ void __launch_bounds__(DP_BLOCK, 5) capture_kernel(DeviceProbes probes, const int32_t* __restrict__ probe_id, DeviceSeqs r1,
                                                           DeviceSeqs r2, const uint32_t* __restrict__ order,
                                                           const uint32_t* __restrict__ n_active_ptr, ScoreParams sp_user,
                                                           ScratchPool pool, ExtendDeviceOut out, unsigned long long* cells_total) {
    __shared__ int4 prof[PROF_INT4_PER_THREAD * DP_BLOCK];
    DpScratch sc = thread_scratch(pool, prof);
    int64_t cells = 0;
    const int64_t n_active = (int64_t)*n_active_ptr;
    for (int64_t t = sc.tid; t < n_active; t += sc.nthreads) {
        const uint32_t mo = order[t];
        if (mo == ORDER_HOLE) continue;
        const int64_t m = mo;
        ReadEndViews v = read_end_views(probes, probe_id, r1, r2, m);
        const int start = out.primer_end[m] + 1;
        SeqView rw = sub_view(v.read, start, v.read.len - start);  // read[e+1 .. L-1] (:210)
        const int score = dp_fill(v.capture, rw, sp_user, 0, 0, sc);  // :211
        cells += (int64_t)v.capture.len * rw.len;
        capture_finish(v, rw, start, score, [&](CigarMdStream& st) { dp_traceback_stream(v.capture, rw, sc, st); }, out, m);
    }
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
}

struct ScratchPool2 {
    uint2* ptr_words;
    uint32_t* brow;
    uint32_t* colinfo;
    uint32_t* slow;
    int r_stride;
    uint32_t nthreads;  // threads of the launch this copy of the pool is passed to (set at every launch site)
};

__device__ inline DpScratch2 thread_scratch2(const ScratchPool2& pool, int4* prof) {
    DpScratch2 sc;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    sc.prof = prof;
    sc.ptr_words = pool.ptr_words + tid;
    sc.brow = pool.brow + tid;
    sc.colinfo = pool.colinfo + tid;
    sc.slow = pool.slow + tid;
    sc.nthreads = (int64_t)pool.nthreads;
    sc.r_stride = pool.r_stride;
    return sc;
}

// Stage 1, packed: two read ends of the same DP#1 shape per thread (entries 2t, 2t+1 of the padded bucket order).
__global__ void __launch_bounds__(DP2_BLOCK, 5) primer2_kernel(DeviceProbes probes, const int32_t* __restrict__ probe_id, DeviceSeqs r1,
                                                            DeviceSeqs r2, const uint32_t* __restrict__ order,
                                                            const uint32_t* __restrict__ n_active_ptr, ScoreParams2 sp_default,
                                                            ScratchPool2 pool, ExtendDeviceOut out, BucketKeys keys, int* status,
                                                            unsigned long long* cells_total, unsigned long long* work_counter) {
    __shared__ int4 prof[PROF2_INT4_PER_THREAD * DP2_BLOCK];
    DpScratch2 sc = thread_scratch2(pool, prof);
    int64_t cells = 0;
    const int64_t n_slots = (int64_t)*n_active_ptr >> 1;  // bucket sizes are even
    for (;;) {
        const int64_t t = next_slot(work_counter);
        if (t - (threadIdx.x & 31) >= n_slots) break;  // whole warp leaves together
        if (t >= n_slots) continue;
        const uint2 mo = reinterpret_cast<const uint2*>(order)[t];
        const int n_here = (mo.y != ORDER_HOLE) ? 2 : 1;
        const int64_t ma = (int64_t)mo.x, mb = (int64_t)(n_here == 2 ? mo.y : mo.x);
        ReadEndViews va = read_end_views(probes, probe_id, r1, r2, ma);
        ReadEndViews vb = read_end_views(probes, probe_id, r1, r2, mb);
        if (va.primer.len == 0) {  // Java: IndexOutOfBounds at :299 (all-gap primer alignment); both have the same shape
            atomicMin(status, HSQ_ERR_OUT_OF_RANGE);
            primer_finish(va, HSQ_PRIMER_NOT_FOUND, out, keys, ma);
            if (n_here == 2) primer_finish(vb, HSQ_PRIMER_NOT_FOUND, out, keys, mb);
            continue;
        }
        SeqView pa = primer_prefix(va.primer, va.read), pb = primer_prefix(vb.primer, vb.read);
        int score[2];
        dp_fill2(pa, pb, va.primer, vb.primer, sp_default, /*pen_start=*/1, /*pen_end=*/0, sc, score);  // :290-293
        cells += (int64_t)pa.len * va.primer.len * n_here;
#pragma unroll 1
        for (int which = 0; which < n_here; which++) {
            const ReadEndViews v = which ? vb : va;  // by value: selected once, not at every use inside the column loop
            const SeqView px = which ? pb : pa;
            PrimerGateSink gate;
            dp_traceback2(px, v.primer, sc, which, gate);
            primer_finish(v, gate.result(), out, keys, which ? mb : ma);
        }
    }
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
}

// Stage 2, packed: two read ends of the same DP#2 shape per thread, cells as u16x2 halves (hsq_dp2.cuh).
// BAND (hsq_config.band > 0): the fill computes the staircase of band_cols only.  A read end keeps its banded result only if
//   (0) its matrix is much larger than the band: min(Q, R) >= 2 * band + 16,
//   (1) its score is above what any path that starts inside the band and reaches a diagonal |j - i| = band can score:
//       match * min(Q, R) + gapStart + (band - 1) * gapExtend (such a path carries a net run of >= band gaps), and
//   (2) its traceback stays on the diagonals |j - i| < band (terminal free-gap cells excepted);
// every other read end keeps its bucket key and is redone by the full kernel in a second pass (extend_capture), the accepted ones move
// to the skip bucket.  What neither test sees is an alignment that lies outside the band from its first to its last column (a
// read matching a shifted copy of its target better than the target itself); DESIGN.md 4.1 has the measured rate (0).
// band_stats: [0] read ends aligned in the band, [1] of them handed to the full pass.
#ifndef HSQ_EXP_CTAS
#define HSQ_EXP_CTAS 5
#endif
template <bool BAND>
__device__ __forceinline__ void capture2_body(const DeviceProbes& probes, const int32_t* __restrict__ probe_id, const DeviceSeqs& r1,
                                              const DeviceSeqs& r2, const uint32_t* __restrict__ order, const uint32_t* __restrict__ n_active_ptr,
                                              const ScoreParams2& sp_user, const ScratchPool2& pool, const ExtendDeviceOut& out,
                                              unsigned long long* cells_total, unsigned long long* work_counter, int band, BucketKeys keys,
                                              unsigned long long* band_stats, int4* prof) {
    DpScratch2 sc = thread_scratch2(pool, prof);
    int64_t cells = 0;
    int n_banded = 0, n_redo = 0;
    const int64_t n_slots = (int64_t)*n_active_ptr >> 1;  // bucket sizes are even
    for (;;) {
        const int64_t t = next_slot(work_counter);
        if (t - (threadIdx.x & 31) >= n_slots) break;  // whole warp leaves together
        if (t >= n_slots) continue;
        const uint2 mo = reinterpret_cast<const uint2*>(order)[t];
#ifdef HSQ_EXPERIMENT_DUPB  // (development experiment: both halves align the same read end, so the two references are identical)
        const int n_here = 2;
        const int64_t ma = (int64_t)mo.x, mb = ma;
#else
        const int n_here = (mo.y != ORDER_HOLE) ? 2 : 1;
        const int64_t ma = (int64_t)mo.x, mb = (int64_t)(n_here == 2 ? mo.y : mo.x);
#endif
        ReadEndViews va = read_end_views(probes, probe_id, r1, r2, ma);
        ReadEndViews vb = read_end_views(probes, probe_id, r1, r2, mb);
        const int start_a = out.primer_end[ma] + 1, start_b = out.primer_end[mb] + 1;
        SeqView rwa = sub_view(va.read, start_a, va.read.len - start_a);  // read[e+1 .. L-1] (:210)
        SeqView rwb = sub_view(vb.read, start_b, vb.read.len - start_b);
        if (BAND && min(rwa.len, va.capture.len) < 2 * band + KROWS) {
            // a matrix that is not much larger than the band: the band says nothing about where in it the alignment lies (a
            // three-base target matches anywhere along a read); both read ends keep their keys and go to the full pass
            n_banded += n_here;
            n_redo += n_here;
            continue;
        }
        int score[2];
#ifdef HSQ_EXPERIMENT_SAMEREF
        dp_fill2<true, BAND>(va.capture, vb.capture, rwa, rwb, sp_user, 0, 0, sc, score, band);  // :211
#else
        dp_fill2<false, BAND>(va.capture, vb.capture, rwa, rwb, sp_user, 0, 0, sc, score, band);  // :211
#endif
        const int score_a = score[0], score_b = score[1];
        if (!BAND) cells += (int64_t)va.capture.len * rwa.len * n_here;  // (a banded read end's cells are counted when it is accepted)
#pragma unroll 1
        for (int which = 0; which < n_here; which++) {
            const ReadEndViews v = which ? vb : va;  // by value: selected once, not at every use inside the column loop
            const SeqView rw = which ? rwb : rwa;
            const int64_t m = which ? mb : ma;
            const int score_x = which ? score_b : score_a;
            bool inside = true;
            int ok = 0;
            if (!BAND || score_x > sp_user.match * min(rw.len, v.capture.len) + sp_user.gap_start + (band - 1) * sp_user.gap_extend)
                ok = capture_finish(v, rw, which ? start_b : start_a, score_x,
                                    [&](CigarMdStream& st) { inside = dp_traceback2<BAND>(v.capture, rw, sc, which, st, band); }, out, m);
            if (BAND) {
                n_banded++;
                if (ok && inside) {
                    keys.key[m] = keys.skip_key;
                    cells += (int64_t)v.capture.len * rw.len;  // the cells the reference fills for this read end
                } else {
                    n_redo++;
                }
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
    if (BAND) {
        for (int o = 16; o > 0; o >>= 1) {
            n_banded += __shfl_xor_sync(0xffffffffu, n_banded, o);
            n_redo += __shfl_xor_sync(0xffffffffu, n_redo, o);
        }
        if ((threadIdx.x & 31) == 0 && n_banded) {
            atomicAdd(&band_stats[0], (unsigned long long)n_banded);
            if (n_redo) atomicAdd(&band_stats[1], (unsigned long long)n_redo);
        }
    }
}

__global__ void __launch_bounds__(DP2_BLOCK, HSQ_EXP_CTAS) capture2_kernel(DeviceProbes probes, const int32_t* __restrict__ probe_id, DeviceSeqs r1,
                                                             DeviceSeqs r2, const uint32_t* __restrict__ order,
                                                             const uint32_t* __restrict__ n_active_ptr, ScoreParams2 sp_user,
                                                             ScratchPool2 pool, ExtendDeviceOut out, unsigned long long* cells_total,
                                                             unsigned long long* work_counter) {
    __shared__ int4 prof[PROF2_INT4_PER_THREAD * DP2_BLOCK];
    capture2_body<false>(probes, probe_id, r1, r2, order, n_active_ptr, sp_user, pool, out, cells_total, work_counter, 0, BucketKeys{}, nullptr, prof);
}

__global__ void __launch_bounds__(DP2_BLOCK, HSQ_EXP_CTAS) capture2_band_kernel(DeviceProbes probes, const int32_t* __restrict__ probe_id, DeviceSeqs r1,
                                                                  DeviceSeqs r2, const uint32_t* __restrict__ order,
                                                                  const uint32_t* __restrict__ n_active_ptr, ScoreParams2 sp_user,
                                                                  ScratchPool2 pool, ExtendDeviceOut out, unsigned long long* cells_total,
                                                                  unsigned long long* work_counter, int band, BucketKeys keys,
                                                                  unsigned long long* band_stats) {
    __shared__ int4 prof[PROF2_INT4_PER_THREAD * DP2_BLOCK];
    capture2_body<true>(probes, probe_id, r1, r2, order, n_active_ptr, sp_user, pool, out, cells_total, work_counter, band, keys, band_stats, prof);
}

}  // namespace hsq
