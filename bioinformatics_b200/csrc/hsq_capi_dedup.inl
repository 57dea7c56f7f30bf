// hsq_capi_dedup.inl -- hsq_dedup_batch: FilterByUid.reduceReadsByProbeAndUid + findBestData for all probes of a batch.
//   UID         = R1[0:u] ++ R2[0:v]                       HSQ/process/FilterByUid.java:100-101,136-138
//   group key   = (probe, UID)                              :134-147 (one HashMap per probe task)
//   best pair   = max (short)(sum(q1-33)+sum(q2-33)), tie -> smaller Integer.parseInt(readName)   :281-297,
//                 HSQ/utils/BamFileUtil.java:74-85, HSQ/objects/ReadPair.java:101-103 (Java short wrap-around)
// Device plan: key generation (HBM-bound, one warp per pair) -> counting sort of the pair indexes by probe -> one block per probe
// sorts its pairs by (UID, rank, pair) in shared memory and reduces them in the same kernel (heads = best pairs, group sizes,
// per-probe counters).  A batch in which one probe holds more pairs than a block can sort takes the radix path instead: stable
// LSD radix sorts by rank, UID, probe (cub::DeviceRadixSort) -> head flags -> group walk.
#include <cub/device/device_radix_sort.cuh>
#include "hsq_scan.cuh"
namespace hsq {

__global__ void dedup_keygen_kernel(const uint8_t* __restrict__ r1, const int64_t* __restrict__ off1, const uint8_t* __restrict__ q1,
                                    const uint8_t* __restrict__ r2, const int64_t* __restrict__ off2, const uint8_t* __restrict__ q2,
                                    const int32_t* __restrict__ read_index, int64_t n, int u, int v, uint64_t* __restrict__ uid_out,
                                    uint64_t* __restrict__ rank_out, uint32_t* __restrict__ idx_out, int* status) {
    const int lane = threadIdx.x & 31;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; w < n; w += stride) {
        const int64_t b1 = off1[w], b2 = off2[w];
        const int l1 = (int)(off1[w + 1] - b1), l2 = (int)(off2[w + 1] - b2);
        int sum = 0;  // byte loads: an aligned-word + __vsadu4 variant of these two loops measured 300 us, prefetching the next pair's offsets 216 us,
        // against 178 us for this form at 500 k pairs (profiles/README_r1.md, r1_o)
        for (int i = lane; i < l1; i += 32) sum += (int)q1[b1 + i] - 33;
        for (int i = lane; i < l2; i += 32) sum += (int)q2[b2 + i] - 33;
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        // UID bases: lane i < u reads R1[i], lane u <= i < u+v reads R2[i-u]   (u + v <= 32)
        uint64_t code = 0;
        bool bad = false, short_read = false;
        if (lane < u + v) {
            uint8_t ch;
            if (lane < u) { short_read = lane >= l1; ch = short_read ? 'A' : r1[b1 + lane]; }
            else { short_read = (lane - u) >= l2; ch = short_read ? 'A' : r2[b2 + lane - u]; }
            switch (ch) {
                case 'A': code = 0; break;
                case 'C': code = 1; break;
                case 'G': code = 2; break;
                case 'T': code = 3; break;
                default: bad = true;
            }
            code <<= 2 * (u + v - 1 - lane);
        }
        for (int o = 16; o > 0; o >>= 1) code |= __shfl_xor_sync(0xffffffffu, code, o);
        if (__any_sync(0xffffffffu, short_read)) { if (lane == 0) atomicMin(status, HSQ_ERR_OUT_OF_RANGE); }
        else if (__any_sync(0xffffffffu, bad)) { if (lane == 0) atomicMin(status, HSQ_ERR_UNSUPPORTED); }
        if (lane == 0) {
            const int16_t qsum = (int16_t)sum;  // Java short arithmetic wraps mod 2^16
            uid_out[w] = code;
            rank_out[w] = ((uint64_t)(uint16_t)(32767 - (int)qsum) << 32) | (uint64_t)((uint32_t)read_index[w] ^ 0x80000000u);
            idx_out[w] = (uint32_t)w;
        }
    }
}

template <class T>
__global__ void gather_u64_kernel(const T* __restrict__ src, const uint32_t* __restrict__ idx, int64_t n, uint64_t* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = (uint64_t)src[idx[i]];
}

// sorted position i holds pair idx[i]; a head starts a (probe, UID) group and is its best pair
__global__ void dedup_heads_kernel(const uint32_t* __restrict__ idx, const int32_t* __restrict__ probe, const uint64_t* __restrict__ uid,
                                   int64_t n, uint8_t* __restrict__ head, uint8_t* __restrict__ is_best) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const uint32_t p = idx[i];
        bool h = true;
        if (i > 0) {
            const uint32_t pp = idx[i - 1];
            h = probe[p] != probe[pp] || uid[p] != uid[pp];
        }
        head[i] = h ? 1 : 0;
        is_best[p] = h ? 1 : 0;
    }
}

// each head walks its group: size -> every member, per-probe counters (FilterByUid.java:149-171,251)
__global__ void dedup_groups_kernel(const uint32_t* __restrict__ idx, const int32_t* __restrict__ probe, const uint8_t* __restrict__ head,
                                    int64_t n, int32_t* __restrict__ group_size, int32_t* __restrict__ unique_pairs,
                                    int32_t* __restrict__ duplicate_pairs, int32_t* __restrict__ max_pairs) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        if (!head[i]) continue;
        int64_t e = i + 1;
        while (e < n && !head[e]) e++;
        const int size = (int)(e - i);
        for (int64_t k = i; k < e; k++) group_size[idx[k]] = size;
        const int pr = probe[idx[i]];
        atomicAdd(&unique_pairs[pr], 1);
        if (size > 1) atomicAdd(&duplicate_pairs[pr], size - 1);
        atomicMax(&max_pairs[pr], size);
    }
}

// ---- segment path: counting sort of the pair indexes by probe, then one block per probe groups its pairs by UID in a shared-memory
// hash table and reduces them in the same kernel (best pair = smallest rank, then smallest pair index; group sizes; the probe's
// counters).  No sort: FilterByUid groups with a HashMap per probe too (FilterByUid.java:134-147), and the result does not depend on
// any order.  Keys are read from HBM once and never move; the radix path below stays for batches in which one probe holds more
// pairs than a block's table takes.
static constexpr int DEDUP_SEG_BLOCK = 256;
static constexpr int DEDUP_SEG_MAX = 4096;      // pairs per probe the group kernel takes (52 B of shared memory each)

// cnt[p] = pairs of probe p; within[i] = arrival number of pair i inside its probe (any order: the group kernel is order free)
__global__ void dedup_probe_hist_kernel(const int32_t* __restrict__ probe, int64_t n, int32_t* __restrict__ cnt, uint32_t* __restrict__ within) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    for (; i < n; i += stride) {
        const int p = probe[i];
        const unsigned peers = __match_any_sync(__activemask(), p);  // one atomic per distinct probe in the warp
        const int leader = __ffs(peers) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(&cnt[p], __popc(peers));
        base = __shfl_sync(peers, base, leader);
        within[i] = (uint32_t)(base + __popc(peers & ((1u << lane) - 1u)));
    }
}

__global__ void dedup_seg_prepare_kernel(const int32_t* __restrict__ cnt, int64_t np, int64_t* __restrict__ cnt64, int* __restrict__ max_count) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int mx = 0;
    for (; p < np; p += stride) {
        cnt64[p] = cnt[p];
        mx = max(mx, cnt[p]);
    }
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0 && mx > 0) atomicMax(max_count, mx);
}

__global__ void dedup_scatter_kernel(const int32_t* __restrict__ probe, int64_t n, const int64_t* __restrict__ start, const uint32_t* __restrict__ within,
                                     uint32_t* __restrict__ seg) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) seg[start[probe[i]] + within[i]] = (uint32_t)i;
}

__global__ void dedup_iota_kernel(uint32_t* __restrict__ out, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(DEDUP_SEG_BLOCK) dedup_group_kernel(const uint32_t* __restrict__ seg, const int64_t* __restrict__ start, int64_t np, int cap,
                                                                      const uint64_t* __restrict__ uid, const uint64_t* __restrict__ rank,
                                                                      uint8_t* __restrict__ is_best, int32_t* __restrict__ group_size,
                                                                      int32_t* __restrict__ unique_pairs, int32_t* __restrict__ duplicate_pairs,
                                                                      int32_t* __restrict__ max_pairs, int skip_up_to,
                                                                      const uint32_t* __restrict__ tie) {
    extern __shared__ __align__(16) unsigned char dedup_smem[];
    // per pair: its UID and its table slot; per table slot (2 x cap of them): claiming pair + 1, best rank, best pair, pairs
    uint64_t* s_uid = reinterpret_cast<uint64_t*>(dedup_smem);
    unsigned long long* t_best = reinterpret_cast<unsigned long long*>(s_uid + cap);
    int* s_slot = reinterpret_cast<int*>(t_best + 2 * cap);
    int* t_tag = s_slot + cap;
    unsigned* t_bidx = reinterpret_cast<unsigned*>(t_tag + 2 * cap);
    int* t_cnt = reinterpret_cast<int*>(t_bidx + 2 * cap);
    __shared__ int s_groups, s_max;
    const int tid = threadIdx.x;
    for (int64_t p = blockIdx.x; p < np; p += gridDim.x) {
        const int64_t b = start[p];
        const int m = (int)(start[p + 1] - b);
        if (m <= skip_up_to) continue;  // block-uniform: empty probes, and the small ones dedup_group_small_kernel takes
        int table = 64;
        while (table < 2 * m) table <<= 1;
        const unsigned mask = (unsigned)table - 1u;
        for (int h = tid; h < table; h += DEDUP_SEG_BLOCK) {
            t_tag[h] = 0;
            t_best[h] = ~0ull;
            t_bidx[h] = ~0u;
            t_cnt[h] = 0;
        }
        for (int j = tid; j < m; j += DEDUP_SEG_BLOCK) s_uid[j] = uid[seg[b + j]];
        if (tid == 0) { s_groups = 0; s_max = 0; }
        __syncthreads();
        int claimed = 0;
        for (int j = tid; j < m; j += DEDUP_SEG_BLOCK) {  // find or claim the slot of this pair's UID
            const uint64_t u = s_uid[j];
            unsigned h = (unsigned)((u ^ (u >> 29)) * 0x9E3779B97F4A7C15ull >> 40) & mask;
            int slot;
            while (true) {
                int tag = t_tag[h];
                if (tag == 0) {
                    tag = atomicCAS(&t_tag[h], 0, j + 1);
                    if (tag == 0) { slot = (int)h; claimed++; break; }
                }
                if (s_uid[tag - 1] == u) { slot = (int)h; break; }
                h = (h + 1u) & mask;
            }
            s_slot[j] = slot;
            atomicMin(&t_best[slot], (unsigned long long)rank[seg[b + j]]);
            atomicAdd(&t_cnt[slot], 1);
        }
        __syncthreads();
        for (int j = tid; j < m; j += DEDUP_SEG_BLOCK) {  // among the pairs with the best rank the earliest pair wins
            const uint32_t i = seg[b + j];
            const int slot = s_slot[j];
            if ((unsigned long long)rank[i] == t_best[slot]) atomicMin(&t_bidx[slot], tie ? tie[i] : i);
        }
        __syncthreads();
        int mx = 0;
        for (int j = tid; j < m; j += DEDUP_SEG_BLOCK) {
            const uint32_t i = seg[b + j];
            const int slot = s_slot[j];
            const int size = t_cnt[slot];
            is_best[i] = t_bidx[slot] == (tie ? tie[i] : i) ? 1 : 0;
            group_size[i] = size;
            mx = max(mx, size);
        }
        for (int o = 16; o > 0; o >>= 1) {
            mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            claimed += __shfl_xor_sync(0xffffffffu, claimed, o);
        }
        if ((tid & 31) == 0) {
            atomicMax(&s_max, mx);
            atomicAdd(&s_groups, claimed);
        }
        __syncthreads();
        if (tid == 0) {  // FilterByUid.java:149-171,251
            unique_pairs[p] = s_groups;
            duplicate_pairs[p] = m - s_groups;
            max_pairs[p] = s_max;
        }
        __syncthreads();
    }
}

// probes with at most 32 pairs: one warp per probe, one pair per lane; the pairs of a UID find each other with __match_any_sync, the
// (rank, pair) columns of the warp sit in shared memory so every lane can look at its peers
static constexpr int DEDUP_SMALL_MAX = 32;
__global__ void __launch_bounds__(DEDUP_SEG_BLOCK) dedup_group_small_kernel(const uint32_t* __restrict__ seg, const int64_t* __restrict__ start, int64_t np,
                                                                            const uint64_t* __restrict__ uid, const uint64_t* __restrict__ rank,
                                                                            uint8_t* __restrict__ is_best, int32_t* __restrict__ group_size,
                                                                            int32_t* __restrict__ unique_pairs, int32_t* __restrict__ duplicate_pairs,
                                                                            int32_t* __restrict__ max_pairs, const uint32_t* __restrict__ tie) {
    __shared__ uint64_t s_rank[DEDUP_SEG_BLOCK / 32][32];
    __shared__ uint32_t s_idx[DEDUP_SEG_BLOCK / 32][32];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    int64_t p = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; p < np; p += stride) {
        const int64_t b = start[p];
        const int m = (int)(start[p + 1] - b);
        if (m == 0 || m > DEDUP_SMALL_MAX) continue;  // warp-uniform; larger probes belong to dedup_group_kernel
        const bool active = lane < m;
        const unsigned amask = __ballot_sync(0xffffffffu, active);
        uint32_t i = 0, t = 0;
        uint64_t u = 0, r = 0;
        if (active) {
            i = seg[b + lane];
            u = uid[i];
            r = rank[i];
            t = tie ? tie[i] : i;  // last tie-break: the pair's place in the whole (possibly sharded) batch
            s_rank[wib][lane] = r;
            s_idx[wib][lane] = t;
        }
        __syncwarp();
        int size = 0;
        bool leader = false;
        if (active) {
            const unsigned peers = __match_any_sync(amask, u);
            size = __popc(peers);
            leader = lane == __ffs(peers) - 1;
            bool best = true;  // smallest rank, then smallest pair index, among the pairs of this UID
            for (unsigned q = peers & ~(1u << lane); q; q &= q - 1) {
                const int o = __ffs(q) - 1;
                const uint64_t ro = s_rank[wib][o];
                if (ro < r || (ro == r && s_idx[wib][o] < t)) best = false;
            }
            is_best[i] = best ? 1 : 0;
            group_size[i] = size;
        }
        const int groups = __popc(__ballot_sync(0xffffffffu, leader));
        for (int o = 16; o > 0; o >>= 1) size = max(size, __shfl_xor_sync(0xffffffffu, size, o));
        if (lane == 0) {  // FilterByUid.java:149-171,251
            unique_pairs[p] = groups;
            duplicate_pairs[p] = m - groups;
            max_pairs[p] = size;
        }
        __syncwarp();  // the shared rows are rewritten by the next probe
    }
}

// key columns <-> 24-byte key records (hsq_dedup_key) for the multi-GPU exchange
__global__ void dedup_pack_keys_kernel(const uint64_t* __restrict__ uid, const uint64_t* __restrict__ rank, const int32_t* __restrict__ probe,
                                       int64_t n, uint32_t origin_base, hsq_dedup_key* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        hsq_dedup_key k;
        k.uid = uid[i];
        k.rank = rank[i];
        k.probe = probe[i];
        k.origin = origin_base + (uint32_t)i;
        out[i] = k;
    }
}

__global__ void dedup_unpack_keys_kernel(const hsq_dedup_key* __restrict__ in, int64_t n, int64_t n_probes, uint64_t* __restrict__ uid,
                                         uint64_t* __restrict__ rank, int32_t* __restrict__ probe, uint32_t* __restrict__ idx,
                                         uint32_t* __restrict__ origin, int* status) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const hsq_dedup_key k = in[i];
        uid[i] = k.uid;
        rank[i] = k.rank;
        probe[i] = k.probe;
        idx[i] = (uint32_t)i;
        origin[i] = k.origin;
        if (k.probe < 0 || k.probe >= n_probes) atomicMin(status, HSQ_ERR_INVALID_ARGUMENT);
    }
}

inline int bits_for(uint64_t max_value) {
    int b = 1;
    while (b < 64 && (max_value >> b) != 0) b++;
    return b;
}

}  // namespace hsq

namespace {

int sort_pass(hsq_ctx* ctx, const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out, int64_t n, int end_bit) {
    size_t tmp_bytes = 0;
    HSQ_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys_in, keys_out, vals_in, vals_out, n, 0, end_bit, ctx->stream));
    HSQ_CUDA(ctx, ctx->sort_tmp.ensure(tmp_bytes + 16));
    HSQ_CUDA(ctx, cub::DeviceRadixSort::SortPairs(ctx->sort_tmp.p, tmp_bytes, keys_in, keys_out, vals_in, vals_out, n, 0, end_bit, ctx->stream));
    ctx->launches += (end_bit + 7) / 8 + 2;  // library kernels (approximate; they are not counted as ours in bench.py)
    return HSQ_OK;
}

// exclusive scan of in[0..n) into out[0..n] (tile sums in `tiles`)
int scan_i64(hsq_ctx* ctx, const int64_t* in, int64_t n, int64_t* tiles, int64_t* out) {
    const int64_t n_tiles = (n + 1 + SCAN_TILE - 1) / SCAN_TILE;
    scan_tile_sums_kernel<<<(unsigned)n_tiles, SCAN_BLOCK, 0, ctx->stream>>>(in, n, tiles);
    scan_sums_kernel<<<1, SCAN_BLOCK, 0, ctx->stream>>>(tiles, n_tiles);
    scan_apply_kernel<<<(unsigned)n_tiles, SCAN_BLOCK, 0, ctx->stream>>>(in, n, tiles, out);
    ctx->launches += 3;
    HSQ_CUDA(ctx, cudaGetLastError());
    return HSQ_OK;
}

// host -> device copies of one pair batch + key generation: uid -> sort_a, rank -> sort_b, pair index -> sort_d[0..n)
int dedup_upload_keygen(hsq_ctx* ctx, const hsq_pair_batch* pb, int u, int v, int64_t np) {
    const int64_t n = pb->n;
    const size_t nn = (size_t)n;
    int rc = reset_status(ctx);
    if (rc) return rc;
    // raw bytes: UID grouping is on the read strings themselves, qualities are summed as chars
    if ((rc = upload_pairs_raw(ctx, pb))) return rc;  // probe ids + reads (ASCII, offsets rebased to 0)
    struct { hsq::DevBuf* d; const void* h; size_t bytes; } up[] = {
        {&ctx->q1, pb->q1 ? pb->q1 + pb->r1.offsets[0] : nullptr, (size_t)ctx->r1.total},
        {&ctx->q2, pb->q2 ? pb->q2 + pb->r2.offsets[0] : nullptr, (size_t)ctx->r2.total},
        {&ctx->read_index, pb->read_index, nn * 4}};
    for (auto& c : up) {
        if (c.bytes && !c.h) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_dedup: null input array");
        HSQ_CUDA(ctx, c.d->ensure(c.bytes + 16));
        if (c.bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(c.d->p, c.h, c.bytes, cudaMemcpyHostToDevice, ctx->stream));
    }
    HSQ_CUDA(ctx, ctx->sort_a.ensure(nn * 8 + 16));   // uid per pair
    HSQ_CUDA(ctx, ctx->sort_b.ensure(nn * 8 + 16));   // keys ping
    HSQ_CUDA(ctx, ctx->sort_d.ensure(nn * 8 + 16));   // idx ping | idx pong (2 x u32)
    if (n > 0) {
        const int blocks = (int)std::min<int64_t>((n * 32 + 255) / 256, (int64_t)ctx->sm_count * 16);
        const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
        span_begin(ctx, 0);
        check_probe_ids_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), n, np > 0 ? np : INT32_MAX, d_status(ctx));
        dedup_keygen_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->r1.ascii.as<uint8_t>(), ctx->r1.offsets.as<int64_t>(), ctx->q1.as<uint8_t>(),
                                                            ctx->r2.ascii.as<uint8_t>(), ctx->r2.offsets.as<int64_t>(), ctx->q2.as<uint8_t>(),
                                                            ctx->read_index.as<int32_t>(), n, u, v, ctx->sort_a.as<uint64_t>(),
                                                            ctx->sort_b.as<uint64_t>(), ctx->sort_d.as<uint32_t>(), d_status(ctx));
        span_end(ctx, 0);
        ctx->launches += 2;
        int w[1];
        if ((rc = read_status(ctx, w, 1))) return rc;
        if (w[0] != 0)
            return fail(ctx, w[0], w[0] == HSQ_ERR_UNSUPPORTED ? "hsq_dedup: a UID holds a character other than A,C,G,T (filter N reads first, "
                                                                 "PrimerReadExtensionAndPcrDuplicateIdentification.java:815-821)"
                                   : w[0] == HSQ_ERR_INVALID_ARGUMENT ? "hsq_dedup: probe_id out of range"
                                                                      : "hsq_dedup: a read is shorter than its UID");
    }
    return HSQ_OK;
}

// Sort + segmented reduce over key columns that are already on the device: probe -> ctx->probe_id, uid -> sort_a,
// rank -> sort_b, pair index -> sort_d[0..n).  Results (per pair, in input order) go to d_best / d_gs (device) and the
// per-probe counters to d_stats[3 * np] (device).
// tie: optional last tie-break per pair (smaller wins) among pairs of equal (quality sum, read index); null = the pair index.  The
// sharded path passes the pair's global position so that the answer does not depend on how the batch was split over ranks.
int dedup_reduce_device_impl(hsq_ctx* ctx, int64_t n, int64_t np, int uid_len, uint8_t* d_best, int32_t* d_gs, int32_t* d_stats,
                             const uint32_t* tie);

int dedup_reduce_device(hsq_ctx* ctx, int64_t n, int64_t np, int uid_len, uint8_t* d_best, int32_t* d_gs, int32_t* d_stats,
                        const uint32_t* tie = nullptr) {
    const bool own = !ctx->span_open[1];  // the sharded path opens span 1 around the whole exchange
    if (own) span_begin(ctx, 1);
    const int rc = dedup_reduce_device_impl(ctx, n, np, uid_len, d_best, d_gs, d_stats, tie);
    if (own) span_end(ctx, 1);
    return rc;
}

int dedup_reduce_device_impl(hsq_ctx* ctx, int64_t n, int64_t np, int uid_len, uint8_t* d_best, int32_t* d_gs, int32_t* d_stats,
                             const uint32_t* tie) {
    const size_t nn = (size_t)n;
    int rc;
    HSQ_CUDA(ctx, ctx->sort_c.ensure(nn * 8 + 16));   // keys pong
    HSQ_CUDA(ctx, ctx->flags.ensure(nn + 32));        // head flags
    uint64_t* uid = ctx->sort_a.as<uint64_t>();
    uint64_t* ka = ctx->sort_b.as<uint64_t>();
    uint64_t* kb = ctx->sort_c.as<uint64_t>();
    uint32_t* ia = ctx->sort_d.as<uint32_t>();
    uint32_t* ib = ia + nn;
    uint8_t* head = ctx->flags.as<uint8_t>();
    if (np) HSQ_CUDA(ctx, cudaMemsetAsync(d_stats, 0, 3 * (size_t)np * 4, ctx->stream));
    if (n > 0 && np > 0 && !ctx->cfg.force_dedup_radix) {
        // segment path: counting sort by probe + one block per probe (hash group-by and reduce fused)
        const size_t nps = (size_t)np;
        const int64_t n_tiles = (np + 1 + SCAN_TILE - 1) / SCAN_TILE;
        HSQ_CUDA(ctx, ctx->dd_seg.ensure(nps * 8 + (nps + 1) * 16 + (size_t)n_tiles * 8 + 64));
        int64_t* cnt64 = ctx->dd_seg.as<int64_t>();
        int64_t* start = cnt64 + nps + 1;
        int64_t* tiles = start + nps + 1;
        int32_t* cnt = reinterpret_cast<int32_t*>(tiles + n_tiles);
        int32_t* cursor = cnt + nps;
        HSQ_CUDA(ctx, cudaMemsetAsync(cnt, 0, nps * 8, ctx->stream));
        int* d_max = d_status(ctx) + 8;  // a status word nothing else uses
        const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
        const int pblocks = (int)std::max<int64_t>(1, std::min<int64_t>((np + 255) / 256, (int64_t)ctx->sm_count * 8));
        HSQ_CUDA(ctx, cudaMemsetAsync(d_max, 0, 4, ctx->stream));
        dedup_probe_hist_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), n, cnt, ia);  // ia: arrival numbers from here on
        dedup_seg_prepare_kernel<<<pblocks, 256, 0, ctx->stream>>>(cnt, np, cnt64, d_max);
        ctx->launches += 2;
        if ((rc = scan_i64(ctx, cnt64, np, tiles, start))) return rc;
        int w[9];
        if ((rc = read_status(ctx, w, 9))) return rc;
        const int max_count = w[8];
        if (max_count <= DEDUP_SEG_MAX) {
            int cap = 32;
            while (cap < max_count) cap <<= 1;
            const size_t smem = (size_t)cap * 52 + 16;
            HSQ_CUDA(ctx, cudaFuncSetAttribute(dedup_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            int per_sm = 1;
            HSQ_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dedup_group_kernel, DEDUP_SEG_BLOCK, smem));
            const int sblocks = (int)std::max<int64_t>(1, std::min<int64_t>(np, (int64_t)ctx->sm_count * std::max(per_sm, 1)));
            dedup_scatter_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), n, start, ia, ib);
            const int wblocks = (int)std::max<int64_t>(1, std::min<int64_t>((np + 7) / 8, (int64_t)ctx->sm_count * 8));
            dedup_group_small_kernel<<<wblocks, DEDUP_SEG_BLOCK, 0, ctx->stream>>>(ib, start, np, uid, ka, d_best, d_gs, d_stats, d_stats + np, d_stats + 2 * np, tie);
            ctx->launches += 2;
            if (max_count > DEDUP_SMALL_MAX) {
                dedup_group_kernel<<<sblocks, DEDUP_SEG_BLOCK, smem, ctx->stream>>>(ib, start, np, cap, uid, ka, d_best, d_gs, d_stats, d_stats + np, d_stats + 2 * np,
                                                                                   DEDUP_SMALL_MAX, tie);
                ctx->launches++;
            }
            HSQ_CUDA(ctx, cudaGetLastError());
            return HSQ_OK;
        }
        dedup_iota_kernel<<<gblocks, 256, 0, ctx->stream>>>(ia, n);  // the radix path wants the pair indexes back in ia
        ctx->launches++;
        // a probe with more pairs than a block's table takes: the whole batch goes down the radix path below
    }
    if (n > 0) {
        const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
        // stable LSD passes, least significant key first: [tie] -> rank (48 bits) -> UID -> probe.  cur holds the current order.
        uint32_t *cur = ia, *nxt = ib;
        if (tie) {
            HSQ_CUDA(ctx, ctx->misc.ensure(nn * 8 + 16));
            uint64_t* rank_copy = ctx->misc.as<uint64_t>();
            HSQ_CUDA(ctx, cudaMemcpyAsync(rank_copy, ka, nn * 8, cudaMemcpyDeviceToDevice, ctx->stream));
            gather_u64_kernel<uint32_t><<<gblocks, 256, 0, ctx->stream>>>(tie, cur, n, ka);
            ctx->launches++;
            if ((rc = sort_pass(ctx, ka, kb, cur, nxt, n, 32))) return rc;
            std::swap(cur, nxt);
            gather_u64_kernel<uint64_t><<<gblocks, 256, 0, ctx->stream>>>(rank_copy, cur, n, ka);
            ctx->launches++;
        }
        if ((rc = sort_pass(ctx, ka, kb, cur, nxt, n, 48))) return rc;
        std::swap(cur, nxt);
        if (uid_len > 0) {
            gather_u64_kernel<uint64_t><<<gblocks, 256, 0, ctx->stream>>>(uid, cur, n, ka);
            ctx->launches++;
            if ((rc = sort_pass(ctx, ka, kb, cur, nxt, n, 2 * uid_len))) return rc;
            std::swap(cur, nxt);
        }
        gather_u64_kernel<int32_t><<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), cur, n, ka);
        ctx->launches++;
        if ((rc = sort_pass(ctx, ka, kb, cur, nxt, n, bits_for((uint64_t)std::max<int64_t>(np, 1))))) return rc;
        std::swap(cur, nxt);
        dedup_heads_kernel<<<gblocks, 256, 0, ctx->stream>>>(cur, ctx->probe_id.as<int32_t>(), uid, n, head, d_best);
        dedup_groups_kernel<<<gblocks, 256, 0, ctx->stream>>>(cur, ctx->probe_id.as<int32_t>(), head, n, d_gs, d_stats, d_stats + np, d_stats + 2 * np);
        ctx->launches += 2;
        HSQ_CUDA(ctx, cudaGetLastError());
    }
    return HSQ_OK;
}

int fetch_stats(hsq_ctx* ctx, hsq_probe_stats* stats, const int32_t* d_stats) {
    int rc;
    const int64_t np = stats ? stats->n_probes : 0;
    if (stats && np) {
        if ((rc = d2h(ctx, stats->unique_pairs, d_stats, (size_t)np))) return rc;
        if ((rc = d2h(ctx, stats->duplicate_pairs, d_stats + np, (size_t)np))) return rc;
        if ((rc = d2h(ctx, stats->max_pairs_per_uid, d_stats + 2 * np, (size_t)np))) return rc;
    }
    return HSQ_OK;
}

int run_dedup(hsq_ctx* ctx, const hsq_pair_batch* pb, int u, int v, uint8_t* is_best, int32_t* group_size, hsq_probe_stats* stats) {
    const size_t nn = (size_t)pb->n;
    const int64_t np = stats ? stats->n_probes : 0;
    int rc = dedup_upload_keygen(ctx, pb, u, v, np);
    if (rc) return rc;
    HSQ_CUDA(ctx, ctx->out_u8.ensure(nn + 32));
    HSQ_CUDA(ctx, ctx->out_i32.ensure((nn + 3 * (size_t)np) * 4 + 16));
    uint8_t* best = ctx->out_u8.as<uint8_t>();
    int32_t* d_gs = ctx->out_i32.as<int32_t>();
    int32_t* d_stats = d_gs + nn;
    if ((rc = dedup_reduce_device(ctx, pb->n, np, u + v, best, d_gs, d_stats))) return rc;
    if ((rc = d2h(ctx, is_best, best, nn))) return rc;
    if ((rc = d2h(ctx, group_size, d_gs, nn))) return rc;
    return fetch_stats(ctx, stats, d_stats);
}

}  // namespace

extern "C" int32_t hsq_dedup_batch(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t ext_uid_len, int32_t lig_uid_len, uint8_t* is_best,
                                   int32_t* group_size, hsq_probe_stats* stats) {
    if (!ctx || !pairs || !is_best) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (pairs->n < 0 || pairs->r1.n != pairs->n || pairs->r2.n != pairs->n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "pair batch arrays disagree on n");
    if (ext_uid_len < 0 || lig_uid_len < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative UID length");
    if (ext_uid_len + lig_uid_len > 32) return fail(ctx, HSQ_ERR_UNSUPPORTED, "UID longer than 32 bases in total is outside the accelerated domain");
    if (!stats || stats->n_probes <= 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "probe stats (with n_probes) are required");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = run_dedup(ctx, pairs, ext_uid_len, lig_uid_len, is_best, group_size, stats);
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}

// ---- sharded dedup: the two halves of hsq_dedup_batch around the caller's all-to-all -----------------
extern "C" int32_t hsq_dedup_keys(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t ext_uid_len, int32_t lig_uid_len, int64_t n_probes,
                                  uint32_t origin_base, hsq_dedup_key* d_keys_out) {
    if (!ctx || !pairs) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (pairs->n < 0 || pairs->r1.n != pairs->n || pairs->r2.n != pairs->n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "pair batch arrays disagree on n");
    if (pairs->n > 0 && !d_keys_out) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null key buffer");
    if (ext_uid_len < 0 || lig_uid_len < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative UID length");
    if (ext_uid_len + lig_uid_len > 32) return fail(ctx, HSQ_ERR_UNSUPPORTED, "UID longer than 32 bases in total is outside the accelerated domain");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = dedup_upload_keygen(ctx, pairs, ext_uid_len, lig_uid_len, n_probes);
    if (!rc && pairs->n > 0) {
        const int gblocks = (int)std::min<int64_t>((pairs->n + 255) / 256, (int64_t)ctx->sm_count * 16);
        dedup_pack_keys_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->sort_a.as<uint64_t>(), ctx->sort_b.as<uint64_t>(),
                                                                ctx->probe_id.as<int32_t>(), pairs->n, origin_base, d_keys_out);
        ctx->launches++;
    }
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}

extern "C" int32_t hsq_dedup_reduce_keys(hsq_ctx* ctx, const hsq_dedup_key* d_keys, int64_t n, int32_t uid_len, uint8_t* d_is_best,
                                         int32_t* d_group_size, hsq_probe_stats* stats) {
    if (!ctx || n < 0 || (n > 0 && (!d_keys || !d_is_best || !d_group_size))) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (uid_len < 0 || uid_len > 32) return fail(ctx, HSQ_ERR_UNSUPPORTED, "UID longer than 32 bases in total is outside the accelerated domain");
    if (!stats || stats->n_probes <= 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "probe stats (with n_probes) are required");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    const size_t nn = (size_t)n;
    const int64_t np = stats->n_probes;
    int rc = reset_status(ctx);
    if (rc) return rc;
    HSQ_CUDA(ctx, ctx->sort_a.ensure(nn * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_b.ensure(nn * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_d.ensure(nn * 8 + 16));
    HSQ_CUDA(ctx, ctx->probe_id.ensure(nn * 4 + 16));
    HSQ_CUDA(ctx, ctx->out_i32.ensure(3 * (size_t)np * 4 + 16));
    HSQ_CUDA(ctx, ctx->cm_tie.ensure(nn * 4 + 16));
    if (n > 0) {
        const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
        dedup_unpack_keys_kernel<<<gblocks, 256, 0, ctx->stream>>>(d_keys, n, np, ctx->sort_a.as<uint64_t>(), ctx->sort_b.as<uint64_t>(),
                                                                  ctx->probe_id.as<int32_t>(), ctx->sort_d.as<uint32_t>(), ctx->cm_tie.as<uint32_t>(), d_status(ctx));
        ctx->launches++;
        int w[1];
        if ((rc = read_status(ctx, w, 1))) return rc;
        if (w[0] != 0) return fail(ctx, w[0], "hsq_dedup_reduce_keys: probe out of range in a key record");
    }
    int32_t* d_stats = ctx->out_i32.as<int32_t>();
    // the origin field is the last tie-break: with globally unique origins the answer does not depend on the arrival order of the records
    rc = dedup_reduce_device(ctx, n, np, uid_len, d_is_best, d_group_size, d_stats, ctx->cm_tie.as<uint32_t>());
    if (!rc) rc = fetch_stats(ctx, stats, d_stats);
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}
