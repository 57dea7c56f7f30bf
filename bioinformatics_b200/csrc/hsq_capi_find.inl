// hsq_capi_find.inl -- candidate stage of FastqFindAndTrimTool:
//   trimReadBasedOnQuality                     COMMON/fastqtool/FastqFindAndTrimTool.java:1188-1219
//   the N test on the trimmed read             :920
//   SimpleMapper (index, tallies, candidates)  COMMON/mapping/SimpleMapper.java:90-143,154-217
// Device plan.  Index: every (k-mer, sequence) occurrence becomes a 64-bit key kmer * R + row; radix sort, drop repeats
// (a k-mer counts once per sequence: the reference keeps a Set), count per k-mer, scan -> CSR; k-mers over the per-k-mer cap
// get length 0.  Query: one warp per read; tallies of all R sequences live in shared memory (global scratch when R is
// large), filled with one atomicAdd per (query position, sequence holding that k-mer); the candidate walk then takes the
// groups of equal tallies from the largest down, each group in ascending row order, exactly as the reference's loop does
// when equal tallies come out of its sort in that order.
namespace {

constexpr int FIND_MAX_K = 10;  // 4^10 CSR rows

__device__ __forceinline__ bool is_acgt(uint8_t c) {
    const uint8_t u = c & 0xDF;  // NucleotideCode.getNucleotidesFromString upper-cases first
    return u == 'A' || u == 'C' || u == 'G' || u == 'T';
}
__device__ __forceinline__ uint32_t base2(uint8_t c) { return (c >> 1) & 3u; }  // A/a 0, C/c 1, T/t 2, G/g 3

__device__ __forceinline__ uint32_t kmer_at(const uint8_t* s, int k) {
    uint32_t v = 0;
    for (int j = 0; j < k; j++) v = (v << 2) | base2(s[j]);
    return v;
}

// one thread per sequence to find: checks, size, number of indexed k-mers (start positions 0 .. size-k-1, SimpleMapper.java:93)
__global__ void find_seq_check_kernel(const uint8_t* __restrict__ data, const int64_t* __restrict__ off, int64_t R, int k, int32_t* __restrict__ size,
                                      int64_t* __restrict__ count, int* __restrict__ status) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < R; r += stride) {
        const int64_t a = off[r];
        const int len = (int)(off[r + 1] - a);
        bool ok = true;
        for (int i = 0; i < len; i++) ok &= is_acgt(data[a + i]);
        if (!ok) atomicCAS(status, 0, HSQ_ERR_BAD_CHARACTER);
        else if (len < k) atomicCAS(status, 0, HSQ_ERR_OUT_OF_RANGE);
        size[r] = len;
        count[r] = (ok && len > k) ? (len - k) : 0;
    }
}

__global__ void find_emit_keys_kernel(const uint8_t* __restrict__ data, const int64_t* __restrict__ off, const int64_t* __restrict__ pos, int64_t R,
                                      int k, uint64_t* __restrict__ keys) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < R; r += stride) {
        const int64_t a = off[r];
        const int m = (int)(pos[r + 1] - pos[r]);
        for (int i = 0; i < m; i++) keys[pos[r] + i] = (uint64_t)kmer_at(data + a + i, k) * (uint64_t)R + (uint64_t)r;
    }
}

// first occurrence of every (k-mer, sequence) key in the sorted list: flag + per-k-mer count
__global__ void find_heads_kernel(const uint64_t* __restrict__ keys, int64_t M, uint64_t R, int64_t* __restrict__ head, int64_t* __restrict__ per_kmer) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < M; i += stride) {
        const bool h = (i == 0) || keys[i] != keys[i - 1];
        head[i] = h ? 1 : 0;
        if (h) atomicAdd(reinterpret_cast<unsigned long long*>(per_kmer + keys[i] / R), 1ull);
    }
}

__global__ void find_compact_kernel(const uint64_t* __restrict__ keys, const int64_t* __restrict__ head, const int64_t* __restrict__ upos, int64_t M,
                                    uint64_t R, int32_t* __restrict__ refs) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < M; i += stride)
        if (head[i]) refs[upos[i]] = (int32_t)(keys[i] % R);
}

// length of every CSR row; a k-mer held by more than max_refs sequences is not in the index (SimpleMapper.java:135-138)
__global__ void find_row_len_kernel(const int64_t* __restrict__ per_kmer, int64_t rows, int max_refs, int32_t* __restrict__ len) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < rows; i += stride) len[i] = (max_refs > 0 && per_kmer[i] > max_refs) ? 0 : (int32_t)per_kmer[i];
}

}  // namespace

extern "C" int32_t hsq_set_find_sequences(hsq_ctx* ctx, const hsq_seqs* seqs, int32_t kmer, int32_t max_refs_per_kmer) {
    if (!ctx || !seqs) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (kmer < 1 || kmer > FIND_MAX_K) return fail(ctx, HSQ_ERR_UNSUPPORTED, "k-mer size must be 1..10");
    if (seqs->n < 1 || seqs->n >= (1ll << 31)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "need between 1 and 2^31-1 sequences to find");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    ctx->fi_n = -1;
    int rc;
    if ((rc = reset_status(ctx))) return rc;
    if ((rc = copy_seqs(ctx, seqs, &ctx->fs))) return rc;
    const int64_t R = seqs->n, rows = 1ll << (2 * kmer);
    HSQ_CUDA(ctx, ctx->fi_size.ensure((size_t)R * 4 + 16));
    HSQ_CUDA(ctx, ctx->cols.ensure((size_t)(R + 1) * 8));
    HSQ_CUDA(ctx, ctx->brow.ensure((size_t)(R + 1) * 8));
    HSQ_CUDA(ctx, ctx->misc.ensure((size_t)((std::max(R, rows) + 1 + SCAN_TILE) / SCAN_TILE + 1) * 8 + 64));
    int64_t* cnt = ctx->cols.as<int64_t>();
    int64_t* pos = ctx->brow.as<int64_t>();
    int64_t* tiles = ctx->misc.as<int64_t>();
    const uint8_t* data = ctx->fs.ascii.as<uint8_t>();
    const int64_t* off = ctx->fs.offsets.as<int64_t>();
    const int rblocks = (int)std::min<int64_t>((R + 127) / 128, (int64_t)ctx->sm_count * 16);
    find_seq_check_kernel<<<rblocks, 128, 0, ctx->stream>>>(data, off, R, kmer, ctx->fi_size.as<int32_t>(), cnt, ctx->status.as<int>());
    ctx->launches++;
    if ((rc = scan_i64(ctx, cnt, R, tiles, pos))) return rc;
    int64_t* h_m = reinterpret_cast<int64_t*>(ctx->h_cells) + 1;  // pinned
    HSQ_CUDA(ctx, cudaMemcpyAsync(h_m, pos + R, 8, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->h_status, ctx->status.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->h_status[0] != 0) return fail(ctx, ctx->h_status[0], ctx->h_status[0] == HSQ_ERR_BAD_CHARACTER
                                                                      ? "a sequence to find holds a character other than A, C, G, T"
                                                                      : "a sequence to find is shorter than the k-mer size");
    const int64_t M = *h_m;
    const size_t mm = (size_t)M;
    HSQ_CUDA(ctx, ctx->fi_start.ensure((size_t)(rows + 1) * 8));
    HSQ_CUDA(ctx, ctx->fi_len.ensure((size_t)rows * 4 + 16));
    HSQ_CUDA(ctx, ctx->fi_refs.ensure(mm * 4 + 16));
    HSQ_CUDA(ctx, ctx->sort_a.ensure(mm * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_b.ensure(mm * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_c.ensure(mm * 4 + 16));
    HSQ_CUDA(ctx, ctx->sort_d.ensure(mm * 4 + 16));
    HSQ_CUDA(ctx, ctx->flags.ensure((mm + 1) * 8 + 16));
    HSQ_CUDA(ctx, ctx->slow.ensure((std::max<size_t>(mm, (size_t)rows) + 1) * 8 + 16));
    HSQ_CUDA(ctx, ctx->colinfo.ensure((size_t)(rows + 1) * 8));
    int64_t* per_kmer = ctx->colinfo.as<int64_t>();
    HSQ_CUDA(ctx, cudaMemsetAsync(per_kmer, 0, (size_t)(rows + 1) * 8, ctx->stream));
    const int kblocks = (int)std::min<int64_t>((rows + 255) / 256, (int64_t)ctx->sm_count * 16);
    if (M > 0) {
        uint64_t* ka = ctx->sort_a.as<uint64_t>();
        uint64_t* kb = ctx->sort_b.as<uint64_t>();
        find_emit_keys_kernel<<<rblocks, 128, 0, ctx->stream>>>(data, off, pos, R, kmer, ka);
        ctx->launches++;
        if ((rc = sort_pass(ctx, ka, kb, ctx->sort_c.as<uint32_t>(), ctx->sort_d.as<uint32_t>(), M, 2 * kmer + bits_for((uint64_t)R)))) return rc;
        const int mblocks = (int)std::min<int64_t>((M + 255) / 256, (int64_t)ctx->sm_count * 16);
        int64_t* head = ctx->flags.as<int64_t>();
        int64_t* upos = ctx->slow.as<int64_t>();
        find_heads_kernel<<<mblocks, 256, 0, ctx->stream>>>(kb, M, (uint64_t)R, head, per_kmer);
        ctx->launches++;
        if ((rc = scan_i64(ctx, head, M, tiles, upos))) return rc;
        find_compact_kernel<<<mblocks, 256, 0, ctx->stream>>>(kb, head, upos, M, (uint64_t)R, ctx->fi_refs.as<int32_t>());
        ctx->launches++;
    }
    if ((rc = scan_i64(ctx, per_kmer, rows, tiles, ctx->fi_start.as<int64_t>()))) return rc;
    find_row_len_kernel<<<kblocks, 256, 0, ctx->stream>>>(per_kmer, rows, max_refs_per_kmer, ctx->fi_len.as<int32_t>());
    ctx->launches++;
    HSQ_CUDA(ctx, cudaGetLastError());
    if ((rc = encode_seqs(ctx, &ctx->fs, d_status(ctx), d_maxlen(ctx, 1)))) return rc;  // codes for hsq_find_matches' alignments
    int w[3];
    if ((rc = read_status(ctx, w, 3))) return rc;
    if ((rc = timer.finish())) return rc;
    ctx->fi_n = R;
    ctx->fi_k = kmer;
    ctx->fi_max_len = w[2];
    return HSQ_OK;
}

namespace {

struct FindIndex {
    const int64_t* start;  // [4^k + 1]
    const int32_t* len;    // [4^k] 0 = k-mer not indexed
    const int32_t* refs;
    const int32_t* size;   // [R] referenceAddressesToSizeMap
    int32_t R, k;
};

struct FindDeviceOut {
    int32_t K;
    int32_t *lead, *trail, *n_cand, *cand, *hits;
    uint8_t *has_n, *overflow, *order_dep;
};

// One warp per read.  Shared memory per warp: tallies of all R sequences, two 16-bit counters per word (or `tally_global`
// when R is too large; a tally never exceeds the read length, which the launcher keeps below 65536) and five
// arrays indexed by tally value (H = longest read - k + 1 entries): how many sequences hold that tally, how many of them
// fail the hit-ratio test, the smallest failing row, the output slot of the group's first entry, entries already placed.
__global__ void find_candidates_kernel(const uint8_t* __restrict__ reads, const uint8_t* __restrict__ quals, const int64_t* __restrict__ off, int64_t n,
                                       FindIndex ix, int lead_q, int trail_q, int limit, double min_ratio, FindDeviceOut o, int H,
                                       uint32_t* __restrict__ tally_global, int* __restrict__ status) {
    extern __shared__ uint32_t find_smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int R = ix.R, k = ix.k;
    const int W = (R + 1) >> 1;  // tally words
    const size_t per_warp = (size_t)(tally_global ? 0 : W) + 5 * (size_t)H;
    uint32_t* base_smem = find_smem + (size_t)wib * per_warp;
    uint32_t* tally = tally_global ? tally_global + warp0 * W : base_smem;
    uint32_t* hist = base_smem + (tally_global ? 0 : W);
    uint32_t* failcnt = hist + H;
    uint32_t* failmin = failcnt + H;
    uint32_t* slot0 = failmin + H;
    uint32_t* placed = slot0 + H;
    for (int64_t r = warp0; r < n; r += n_warps) {
        const int64_t a = off[r];
        const int len = (int)(off[r + 1] - a);
        const uint8_t* q = quals + a;
        // trimReadBasedOnQuality (:1194-1214): first position from each end whose quality reaches the threshold
        int lead = len;
        for (int b = 0; b < len; b += 32) {
            const int i = b + lane;
            const unsigned m = __ballot_sync(0xffffffffu, i < len && ((int)q[i] - 33) >= lead_q);
            if (m) { lead = b + __ffs(m) - 1; break; }
        }
        const int rest = len - lead;
        int trail = rest;
        for (int b = 0; b < rest; b += 32) {
            const int t = b + lane;  // candidate number of trailing bases already trimmed
            const unsigned m = __ballot_sync(0xffffffffu, t < rest && ((int)q[len - t - 1] - 33) >= trail_q);
            if (m) { trail = b + __ffs(m) - 1; break; }
        }
        const uint8_t* s = reads + a + lead;
        const int tlen = len - lead - trail;
        bool has_n = false, bad = false;
        for (int i = lane; i < tlen; i += 32) {
            const uint8_t c = s[i];
            has_n |= ((c | 0x20) == 'n');
            bad |= !is_acgt(c);
        }
        has_n = __any_sync(0xffffffffu, has_n);
        bad = __any_sync(0xffffffffu, bad);
        int count = 0;
        bool od = false;
        if (!has_n && bad && lane == 0) atomicCAS(status, 0, HSQ_ERR_BAD_CHARACTER);
        if (!has_n && !bad && tlen > k) {
            const int nlev = tlen - k;  // largest possible tally (< H)
            for (int i = lane; i < W; i += 32) tally[i] = 0;
            for (int i = lane; i <= nlev; i += 32) { hist[i] = 0; failcnt[i] = 0; failmin[i] = 0xffffffffu; placed[i] = 0; }
            __syncwarp();
            // getReferenceTallyMap (SimpleMapper.java:205-217): query positions 0 .. size-k-1
            for (int i = lane; i < nlev; i += 32) {
                const uint32_t code = kmer_at(s + i, k);
                const int32_t m = ix.len[code];
                const int32_t* row = ix.refs + ix.start[code];
                int j = 0;
                for (; j + 4 <= m; j += 4) {  // the four row loads are issued before the first atomic waits on one
                    const int32_t r0 = __ldg(row + j), r1 = __ldg(row + j + 1), r2 = __ldg(row + j + 2), r3 = __ldg(row + j + 3);
                    atomicAdd(&tally[r0 >> 1], (r0 & 1) ? 0x10000u : 1u);
                    atomicAdd(&tally[r1 >> 1], (r1 & 1) ? 0x10000u : 1u);
                    atomicAdd(&tally[r2 >> 1], (r2 & 1) ? 0x10000u : 1u);
                    atomicAdd(&tally[r3 >> 1], (r3 & 1) ? 0x10000u : 1u);
                }
                for (; j < m; j++) {
                    const int32_t r0 = __ldg(row + j);
                    atomicAdd(&tally[r0 >> 1], (r0 & 1) ? 0x10000u : 1u);
                }
            }
            __syncwarp();
            // one pass over the tallies: group sizes and, per group, the entries the ratio test rejects (:171)
            const double limit_q = min_ratio * (double)nlev;
            uint32_t top = 0;
            for (int ref = lane; ref < R; ref += 32) {
                const uint32_t t = (tally[ref >> 1] >> ((ref & 1) * 16)) & 0xffffu;
                if (t == 0) continue;
                top = max(top, t);
                atomicAdd(&hist[t], 1u);
                if ((double)t < limit_q && (double)t < min_ratio * (double)(ix.size[ref] - k)) {
                    atomicAdd(&failcnt[t], 1u);
                    atomicMin(&failmin[t], (uint32_t)ref);
                }
            }
            for (int d = 16; d > 0; d >>= 1) top = max(top, __shfl_xor_sync(0xffffffffu, top, d));
            __syncwarp();
            // getBestCandidateReferences (:154-187): groups of equal tallies from the largest down (every lane walks the same path)
            uint32_t cut = 0xffffffffu, stop_h = 0, stop_ref = 0;
            for (uint32_t h = top; h >= 1; h--) {
                const uint32_t c = hist[h];
                if (c == 0) continue;
                if (count > limit) break;  // candidate limit reached and this group is not tied with the last one added (:172)
                if (lane == 0) slot0[h] = (uint32_t)count;
                if (failcnt[h] > 0) {  // the walk ends at the first rejected entry: rows before it (all accepted) still get in
                    stop_h = h;
                    stop_ref = failmin[h];
                    od = c > failcnt[h];
                    break;
                }
                count += (int)c;
                cut = h;
            }
            __syncwarp();
            // placement: slot = first slot of the tally group + rank inside it in ascending row order
            for (int b = 0; b < R; b += 32) {
                const int ref = b + lane;
                const uint32_t t = ref < R ? (tally[ref >> 1] >> ((ref & 1) * 16)) & 0xffffu : 0;
                const bool part = stop_h != 0 && t == stop_h && (uint32_t)ref < stop_ref;
                const bool take = (t != 0 && t >= cut) || part;
                const unsigned m_take = __ballot_sync(0xffffffffu, take);
                if (m_take == 0) continue;
                count += __popc(__ballot_sync(0xffffffffu, part));
                if (take) {
                    const unsigned peers = __match_any_sync(m_take, t);
                    const int slot = (int)(slot0[t] + placed[t]) + __popc(peers & lt_mask);
                    if (slot < o.K) {
                        o.cand[r * o.K + slot] = ref;
                        o.hits[r * o.K + slot] = (int32_t)t;
                    }
                    __syncwarp(m_take);
                    if ((peers & lt_mask) == 0) placed[t] += (uint32_t)__popc(peers);
                    __syncwarp(m_take);
                }
            }
        }
        if (lane == 0) {
            o.lead[r] = lead;
            o.trail[r] = trail;
            o.has_n[r] = has_n ? 1 : 0;
            o.n_cand[r] = min(count, o.K);
            o.overflow[r] = count > o.K ? 1 : 0;
            o.order_dep[r] = od ? 1 : 0;
        }
        __syncwarp();
    }
}

// per item (read r, candidate slot j): NeedlemanWunschGlobalAlignment(trimmed read, candidate) with the default scorer, the five
// counters over the query-trimmed alignment and the limit test (FastqFindAndTrimTool.java:930-951)
__global__ void __launch_bounds__(DP_BLOCK) find_align_kernel(DeviceSeqs reads, DeviceSeqs seqs, const int32_t* __restrict__ item_read,
                                                         const int64_t* __restrict__ item_off, int64_t n_items, FindDeviceOut o, ScoreParams sp,
                                                         hsq_find_limits lim, ScratchPool pool, int32_t* __restrict__ counts,
                                                         uint8_t* __restrict__ accepted, int32_t* __restrict__ n_matches,
                                                         unsigned long long* cells_total) {
    __shared__ int4 prof[PROF_INT4_PER_THREAD * DP_BLOCK];
    DpScratch sc = thread_scratch(pool, prof);
    int64_t cells = 0;
    for (int64_t i = sc.tid; i < n_items; i += sc.nthreads) {
        const int64_t r = item_read[i];
        const int64_t slot = r * o.K + (i - item_off[r]);
        const int32_t cand = o.cand[slot];
        const int64_t r0 = reads.offsets[r] + o.lead[r], q0 = seqs.offsets[cand];
        SeqView ref = make_view(reads.codes + r0, (int)(reads.offsets[r + 1] - reads.offsets[r]) - o.lead[r] - o.trail[r], false, false);
        SeqView qry = make_view(seqs.codes + q0, (int)(seqs.offsets[cand + 1] - q0), false, false);
        dp_fill(ref, qry, sp, 0, 0, sc);
        cells += (int64_t)ref.len * qry.len;
        Columns cv{sc.colstore(), dp_traceback(ref, qry, sc)};
        PairIndices ix = pair_indices(cv);
        int32_t fc[5];
        find_counters(cv, ix, fc);
        for (int f = 0; f < 5; f++) counts[slot * 5 + f] = fc[f];
        const bool ok = fc[0] <= lim.mismatches && fc[1] <= lim.insertion_gaps && fc[2] <= lim.insertion_bases && fc[3] <= lim.deletion_gaps &&
                        fc[4] <= lim.deletion_bases;
        accepted[slot] = ok ? 1 : 0;
        if (ok) atomicAdd(&n_matches[r], 1);
    }
    for (int d = 16; d > 0; d >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, d);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
}

__global__ void find_item_counts_kernel(const int32_t* __restrict__ n_cand, int64_t n, int64_t* __restrict__ out) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) out[k] = n_cand[k];
}

__global__ void find_items_kernel(const int64_t* __restrict__ item_off, int64_t n, int32_t* __restrict__ item_read) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n; r += stride)
        for (int64_t i = item_off[r]; i < item_off[r + 1]; i++) item_read[i] = (int32_t)r;
}

// uploads the reads, runs the candidate kernel (results stay on the device in `o`) and the encoder: new
// IupacNucleotideCodeSequence(readString) (:906) throws on a byte that is no IUPAC code, wherever it sits in the read
int find_run_candidates(hsq_ctx* ctx, const hsq_seqs* reads, const uint8_t* quals, int lead_q, int trail_q, int limit, double ratio, int32_t K,
                        FindDeviceOut* o, int* r_max) {
    const int64_t n = reads->n;
    int rc;
    if ((rc = reset_status(ctx))) return rc;
    if ((rc = copy_seqs(ctx, reads, &ctx->r1))) return rc;
    const size_t total = (size_t)ctx->r1.total, nn = (size_t)n, kk = (size_t)K;
    if (total > 0 && !quals) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null qualities");
    HSQ_CUDA(ctx, ctx->q1.ensure(total + 16));
    if (total > 0) HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->q1.p, quals + reads->offsets[0], total, cudaMemcpyHostToDevice, ctx->stream));
    HSQ_CUDA(ctx, ctx->out_i32.ensure((3 * nn + 2 * nn * kk) * 4 + 16));
    HSQ_CUDA(ctx, ctx->out_u8.ensure(3 * nn + 16));
    o->K = K;
    o->lead = ctx->out_i32.as<int32_t>();
    o->trail = o->lead + nn;
    o->n_cand = o->trail + nn;
    o->cand = o->n_cand + nn;
    o->hits = o->cand + nn * kk;
    o->has_n = ctx->out_u8.as<uint8_t>();
    o->overflow = o->has_n + nn;
    o->order_dep = o->overflow + nn;
    if (n == 0) return HSQ_OK;
    if ((rc = encode_seqs(ctx, &ctx->r1, d_status(ctx), d_maxlen(ctx, 0)))) return rc;
    int w[2];
    if ((rc = read_status(ctx, w, 2))) return rc;
    if (w[0] != 0) return fail(ctx, w[0], "a read holds a byte that is not an IUPAC code");
    *r_max = w[1];
    const int H = std::max(1, w[1] - ctx->fi_k + 1);  // tallies run from 0 to read length - k
    FindIndex ix{ctx->fi_start.as<int64_t>(), ctx->fi_len.as<int32_t>(), ctx->fi_refs.as<int32_t>(), ctx->fi_size.as<int32_t>(), (int32_t)ctx->fi_n,
                 ctx->fi_k};
    if (H > 65535) return fail(ctx, HSQ_ERR_UNSUPPORTED, "reads of 65 535 bases or more are not supported");
    const size_t levels = 5 * (size_t)H * 4, words = ((size_t)ctx->fi_n + 1) / 2 * 4;
    size_t per_warp = words + levels;
    int warps = 4;
    while (warps > 1 && per_warp * warps > 100 * 1024) warps >>= 1;
    uint32_t* tally_global = nullptr;
    if (per_warp * warps > 200 * 1024) {  // tallies in global memory, one slice per resident warp
        warps = 4;
        while (warps > 1 && levels * warps > 100 * 1024) warps >>= 1;
        if (levels * warps > 200 * 1024) return fail(ctx, HSQ_ERR_UNSUPPORTED, "reads longer than 10 000 bases are not supported");
        per_warp = levels;
    }
    // persistent grid: as many blocks per SM as the shared memory lets in (the kernel waits on L2 loads of the index rows)
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(std::min<size_t>(64 / warps, 16), (220 * 1024) / (per_warp * warps + 1024)));
    const int blocks = (int)std::min<int64_t>((n + warps - 1) / warps, (int64_t)ctx->sm_count * per_sm);
    if (per_warp == levels) {
        HSQ_CUDA(ctx, ctx->slow.ensure(words * (size_t)blocks * warps + 16));
        tally_global = ctx->slow.as<uint32_t>();
    }
    const size_t smem = per_warp * warps;
    if (smem > 48 * 1024) HSQ_CUDA(ctx, cudaFuncSetAttribute(find_candidates_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    find_candidates_kernel<<<blocks, warps * 32, smem, ctx->stream>>>(ctx->r1.ascii.as<uint8_t>(), ctx->q1.as<uint8_t>(), ctx->r1.offsets.as<int64_t>(), n, ix,
                                                                     lead_q, trail_q, limit, ratio, *o, H, tally_global, ctx->status.as<int>());
    ctx->launches++;
    HSQ_CUDA(ctx, cudaGetLastError());
    return HSQ_OK;
}

int find_fetch_candidates(hsq_ctx* ctx, const FindDeviceOut& o, int64_t n, hsq_find_out* out) {
    const size_t nn = (size_t)n, kk = (size_t)o.K;
    int rc;
    if ((rc = d2h(ctx, out->lead_trimmed, o.lead, nn))) return rc;
    if ((rc = d2h(ctx, out->trail_trimmed, o.trail, nn))) return rc;
    if ((rc = d2h(ctx, out->n_candidates, o.n_cand, nn))) return rc;
    if ((rc = d2h(ctx, out->candidates, o.cand, nn * kk))) return rc;
    if ((rc = d2h(ctx, out->hits, o.hits, nn * kk))) return rc;
    if ((rc = d2h(ctx, out->has_n, o.has_n, nn))) return rc;
    if ((rc = d2h(ctx, out->overflow, o.overflow, nn))) return rc;
    if ((rc = d2h(ctx, out->order_dependent, o.order_dep, nn))) return rc;
    return HSQ_OK;
}

}  // namespace

extern "C" int32_t hsq_find_candidates(hsq_ctx* ctx, const hsq_seqs* reads, const uint8_t* quals, int32_t leading_quality,
                                       int32_t trailing_quality, int32_t candidate_limit, double min_hit_ratio, hsq_find_out* out) {
    if (!ctx || !reads || !out) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (ctx->fi_n < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_set_find_sequences first");
    if (out->max_candidates < 1) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "max_candidates must be positive");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    FindDeviceOut o;
    int rc, r_max = 0;
    if ((rc = find_run_candidates(ctx, reads, quals, leading_quality, trailing_quality, candidate_limit, min_hit_ratio, out->max_candidates, &o, &r_max)))
        return rc;
    int w[1] = {0};
    if (reads->n > 0) {
        if ((rc = find_fetch_candidates(ctx, o, reads->n, out))) return rc;
        if ((rc = read_status(ctx, w, 1))) return rc;
    }
    if ((rc = timer.finish())) return rc;
    if (w[0] != 0) return fail(ctx, w[0], "a read without N holds a character other than A, C, G, T");
    return HSQ_OK;
}

extern "C" int32_t hsq_find_matches(hsq_ctx* ctx, const hsq_seqs* reads, const uint8_t* quals, int32_t leading_quality, int32_t trailing_quality,
                                    int32_t candidate_limit, double min_hit_ratio, const hsq_find_limits* limits, hsq_find_match_out* out) {
    if (!ctx || !reads || !out || !limits) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (ctx->fi_n < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_set_find_sequences first");
    const int32_t K = out->cand.max_candidates;
    if (K < 1) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "max_candidates must be positive");
    const int64_t n = reads->n;
    const size_t nn = (size_t)n, kk = (size_t)K;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    FindDeviceOut o;
    int rc, r_max = 0;
    out->cells = 0;
    if ((rc = find_run_candidates(ctx, reads, quals, leading_quality, trailing_quality, candidate_limit, min_hit_ratio, K, &o, &r_max))) return rc;
    if (n == 0) return timer.finish();
    // item list: candidate slots of all reads in read order
    const int64_t n_tiles = (n + 1 + SCAN_TILE - 1) / SCAN_TILE;
    HSQ_CUDA(ctx, ctx->sort_a.ensure((nn + 1) * 8));
    HSQ_CUDA(ctx, ctx->sort_b.ensure((nn + 1) * 8));
    HSQ_CUDA(ctx, ctx->misc.ensure((size_t)n_tiles * 8 + 64));
    HSQ_CUDA(ctx, ctx->dd_i32.ensure((nn * kk * 5 + nn) * 4 + 16));
    HSQ_CUDA(ctx, ctx->dd_u8.ensure(nn * kk + 16));
    int64_t* cnt = ctx->sort_a.as<int64_t>();
    int64_t* item_off = ctx->sort_b.as<int64_t>();
    int32_t* counts = ctx->dd_i32.as<int32_t>();
    int32_t* n_matches = counts + nn * kk * 5;
    uint8_t* accepted = ctx->dd_u8.as<uint8_t>();
    const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
    find_item_counts_kernel<<<gblocks, 256, 0, ctx->stream>>>(o.n_cand, n, cnt);
    ctx->launches++;
    if ((rc = scan_i64(ctx, cnt, n, ctx->misc.as<int64_t>(), item_off))) return rc;
    HSQ_CUDA(ctx, cudaMemsetAsync(counts, 0, (nn * kk * 5 + nn) * 4, ctx->stream));
    HSQ_CUDA(ctx, cudaMemsetAsync(accepted, 0, nn * kk, ctx->stream));
    int64_t* h_m = reinterpret_cast<int64_t*>(ctx->h_cells) + 1;  // pinned
    HSQ_CUDA(ctx, cudaMemcpyAsync(h_m, item_off + n, 8, cudaMemcpyDeviceToHost, ctx->stream));
    int w[1];
    if ((rc = read_status(ctx, w, 1))) return rc;
    if (w[0] != 0) return fail(ctx, w[0], "a read without N holds a character other than A, C, G, T");
    const int64_t n_items = *h_m;
    if (n_items > 0) {
        HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)n_items * 4 + 16));
        int32_t* item_read = ctx->sort_c.as<int32_t>();
        find_items_kernel<<<gblocks, 256, 0, ctx->stream>>>(item_off, n, item_read);
        ctx->launches++;
        int grid = 1;
        if ((rc = persistent_grid(ctx, find_align_kernel, DP_BLOCK, n_items, &grid))) return rc;
        ScratchPool pool;
        if ((rc = alloc_scratch(ctx, (int64_t)grid * DP_BLOCK, 0, ctx->fi_max_len, r_max, &pool, nullptr))) return rc;
        ScoreParams sp = make_score_params(1, -4, -1, -6);  // new SimpleAlignmentScorer() (SimpleAlignmentScorer.java:23-26,35-43)
        find_align_kernel<<<grid, DP_BLOCK, 0, ctx->stream>>>(ctx->r1.view(), ctx->fs.view(), item_read, item_off, n_items, o, sp, *limits, pool, counts,
                                                             accepted, n_matches, ctx->cells.as<unsigned long long>());
        ctx->launches++;
        HSQ_CUDA(ctx, cudaGetLastError());
    }
    if ((rc = find_fetch_candidates(ctx, o, n, &out->cand))) return rc;
    if ((rc = d2h(ctx, out->counts, counts, nn * kk * 5))) return rc;
    if ((rc = d2h(ctx, out->accepted, accepted, nn * kk))) return rc;
    if ((rc = d2h(ctx, out->n_matches, n_matches, nn))) return rc;
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->h_cells, ctx->cells.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    if ((rc = timer.finish())) return rc;
    out->cells = (int64_t)ctx->h_cells[0];
    for (size_t k = 0; k < nn; k++)
        if (out->cand.overflow[k]) return fail(ctx, HSQ_ERR_OVERFLOW, "a read has more candidates than slots; see the overflow flags");
    return HSQ_OK;
}
