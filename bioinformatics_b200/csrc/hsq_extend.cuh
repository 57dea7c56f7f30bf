// hsq_extend.cuh -- the ExtendReadsToPrimer pipeline on the device, one read end per thread:
//   primer_kernel   DP#1 (primer x read prefix) + the 0.95^d gate           -> primer end index e, bucket key
//   bucket sort     counting sort of the read ends by (strips of DP#2, capture-target length)
//   capture_kernel  DP#2 (capture target x read remainder) + traceback + CIGAR / MD / NM / POS, in bucket order,
//                   so the 32 lanes of a warp run loops of identical trip counts (no divergence in the fill)
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

// Stage 1: getPrimerEndIndexInRead for every read end; read ends that stop here get their (all-zero) results now.
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
        const int L = v.read.len;
        const int e = primer_end_index(v.primer, v.read, sp_default, sc, status, &cells);
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
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
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

// What follows DP#2 for one read end (:212-236): acceptance test, traceback, CIGAR / MD / NM / POS, result stores.
// `traceback` fills the thread's column store and returns the number of alignment columns.
template <class Traceback>
__device__ __forceinline__ void capture_finish(const ReadEndViews& v, const SeqView& rw, int start, int score, const ColStore& cs,
                                               Traceback traceback, const ExtendDeviceOut& out, int64_t m) {
    int ok = 0, pos = 0, read_start = 0, read_stop = 0, nm = 0, ref_len = 0;
    SlotWriter cw{out.cigar + m * out.cigar_cap, out.cigar_cap, 0};
    SlotWriter mw{out.md + m * out.md_cap, out.md_cap, 0};
    const int read_length = rw.len - 1;  // :225
    // score/alignmentLength > 0 (:212): the alignment has max(Q,R) >= 1 columns whenever score != 0
    if (score > 0 && read_length > LENGTH_THRESHOLD_FOR_EXTENDED_READ) {
        Columns cv{cs, traceback()};
        int last_q = cv.n - 1;  // getAlignmentWithoutEndingQueryInserts (:218)
        while (last_q >= 0 && col_qry(cv.fwd(last_q)) == CODE_GAP) last_q--;
        emit_cigar_md(cv, 0, last_q, v.is_reversed, &cw, &mw, &nm, &ref_len);  // :220-231
        pos = v.is_reversed ? v.adjacent - ref_len : v.adjacent + 1;           // :234,:236
        read_start = start;
        read_stop = start + read_length;
        ok = 1;
    }
    out.ok[m] = (uint8_t)ok;
    out.score[m] = score;
    out.pos[m] = pos;
    out.read_start[m] = read_start;
    out.read_stop[m] = read_stop;
    out.nm[m] = nm;
    out.ref_len[m] = ref_len;
    out.cigar_len[m] = cw.len;
    out.md_len[m] = mw.len;
    out.overflow[m] = (uint8_t)((cw.len > cw.cap ? 1 : 0) | (mw.len > mw.cap ? 2 : 0));
}

static constexpr uint32_t ORDER_HOLE = 0xFFFFFFFFu;  // padding entry of an odd-sized bucket

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
        capture_finish(v, rw, start, score, sc.colstore(), [&]() { return dp_traceback(v.capture, rw, sc); }, out, m);
    }
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
}

struct ScratchPool2 {
    uint2* ptr_words;
    uint32_t* brow;
    uint16_t* cols;
    uint32_t* slow;
    int r_stride;
};

__device__ inline DpScratch2 thread_scratch2(const ScratchPool2& pool, int4* prof) {
    DpScratch2 sc;
    sc.prof = prof;
    sc.ptr_words = pool.ptr_words;
    sc.brow = pool.brow;
    sc.cols = pool.cols;
    sc.slow = pool.slow;
    sc.nthreads = (int64_t)gridDim.x * blockDim.x;
    sc.tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    sc.r_stride = pool.r_stride;
    return sc;
}

// Stage 2, packed: two read ends of the same DP#2 shape per thread (entries 2t, 2t+1 of the padded bucket order),
// cells as u16x2 halves (hsq_dp2.cuh).  Chosen by the host when the scorer's score range fits 14 bits.
__global__ void __launch_bounds__(DP2_BLOCK, 7) capture2_kernel(DeviceProbes probes, const int32_t* __restrict__ probe_id, DeviceSeqs r1,
                                                             DeviceSeqs r2, const uint32_t* __restrict__ order,
                                                             const uint32_t* __restrict__ n_active_ptr, ScoreParams2 sp_user,
                                                             ScratchPool2 pool, ExtendDeviceOut out, unsigned long long* cells_total) {
    __shared__ int4 prof[PROF2_INT4_PER_THREAD * DP2_BLOCK];
    DpScratch2 sc = thread_scratch2(pool, prof);
    int64_t cells = 0;
    const int64_t n_slots = (int64_t)*n_active_ptr >> 1;  // bucket sizes are even
    for (int64_t t = sc.tid; t < n_slots; t += sc.nthreads) {
        const uint2 mo = reinterpret_cast<const uint2*>(order)[t];
        const bool has_b = mo.y != ORDER_HOLE;
        const int64_t ma = mo.x, mb = has_b ? mo.y : mo.x;
        ReadEndViews va = read_end_views(probes, probe_id, r1, r2, ma);
        ReadEndViews vb = read_end_views(probes, probe_id, r1, r2, mb);
        const int start_a = out.primer_end[ma] + 1, start_b = out.primer_end[mb] + 1;
        SeqView rwa = sub_view(va.read, start_a, va.read.len - start_a);
        SeqView rwb = sub_view(vb.read, start_b, vb.read.len - start_b);
        int score[2];
        dp_fill2(va.capture, vb.capture, rwa, rwb, sp_user, 0, 0, sc, score);
        cells += (int64_t)va.capture.len * rwa.len * (has_b ? 2 : 1);
        capture_finish(va, rwa, start_a, score[0], sc.colstore(), [&]() { return dp_traceback2(va.capture, rwa, sc, 0); }, out, ma);
        if (has_b)
            capture_finish(vb, rwb, start_b, score[1], sc.colstore(), [&]() { return dp_traceback2(vb.capture, rwb, sc, 1); }, out, mb);
    }
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
}

}  // namespace hsq
