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

// bytes [lo, hi) of a 16-byte chunk kept, the others zeroed (0 <= lo <= hi <= 16)
__device__ __forceinline__ uint32_t keep_bytes(uint32_t w, int word, int lo, int hi) {
    const int l = min(max(lo - 4 * word, 0), 4), h = min(max(hi - 4 * word, 0), 4);
    const uint32_t m = (h <= l) ? 0u : ((h == 4 ? 0xFFFFFFFFu : ((1u << (8 * h)) - 1u)) & (l == 4 ? 0u : (0xFFFFFFFFu << (8 * l))));
    return w & m;
}

// sum of the bytes q[b .. e) read as aligned 16-byte chunks (q is 16-byte aligned and has 16 bytes of slack after its last byte): the
// first and the last chunk are masked, the others go through four __vsadu4 each
__device__ __forceinline__ int byte_sum(const uint8_t* __restrict__ q, int64_t b, int64_t e) {
    if (e <= b) return 0;
    const int64_t p0 = b & ~(int64_t)15;
    int total = 0;
    for (int64_t p = p0; p < e; p += 16) {
        uint4 w = __ldg(reinterpret_cast<const uint4*>(q + p));
        if (p < b || p + 16 > e) {
            const int lo = (int)max((int64_t)0, b - p), hi = (int)min((int64_t)16, e - p);
            w.x = keep_bytes(w.x, 0, lo, hi);
            w.y = keep_bytes(w.y, 1, lo, hi);
            w.z = keep_bytes(w.z, 2, lo, hi);
            w.w = keep_bytes(w.w, 3, lo, hi);
        }
        total += (int)(__vsadu4(w.x, 0u) + __vsadu4(w.y, 0u) + __vsadu4(w.z, 0u) + __vsadu4(w.w, 0u));
    }
    return total;
}

// Key generation, one THREAD per pair.  (Up to r2_b a warp took one pair with byte loads: 240 warp instructions per pair, issue bound at
// 0.13 of the HBM rate.  A thread that walks its own two quality strings in 16-byte chunks issues ~9 warp instructions per pair; consecutive
// threads take consecutive pairs, so a warp still covers one contiguous stretch of each buffer.)
//   CODES: the reads are device codes (A=1 C=2 G=4 T=8; packed batches) instead of ASCII.
//   qs1 / qs2 non-null: the per-read quality sums come from the host (hsq_quality_sums) and the quality bytes are not touched.
template <bool CODES>
__global__ void dedup_keygen_kernel(const uint8_t* __restrict__ r1, const int64_t* __restrict__ off1, const uint8_t* __restrict__ q1,
                                    const uint8_t* __restrict__ r2, const int64_t* __restrict__ off2, const uint8_t* __restrict__ q2,
                                    const int16_t* __restrict__ qs1, const int16_t* __restrict__ qs2, const int32_t* __restrict__ read_index,
                                    int64_t n, int u, int v, uint64_t* __restrict__ uid_out, uint64_t* __restrict__ rank_out,
                                    uint32_t* __restrict__ idx_out, int* status) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) {
        const int64_t b1 = off1[k], e1 = off1[k + 1], b2 = off2[k], e2 = off2[k + 1];
        int sum;
        if (qs1) sum = (int)qs1[k] + (int)qs2[k];
        else sum = byte_sum(q1, b1, e1) + byte_sum(q2, b2, e2) - 33 * (int)((e1 - b1) + (e2 - b2));
        // UID = R1[0:u] ++ R2[0:v], 2 bit per base, first base most significant (u + v <= 32)
        uint64_t code = 0;
        bool bad = false;
        const bool short_read = (e1 - b1) < u || (e2 - b2) < v;
        if (!short_read) {
#pragma unroll 1
            for (int i = 0; i < u + v; i++) {
                const uint8_t ch = i < u ? r1[b1 + i] : r2[b2 + i - u];
                uint32_t c;
                if (CODES) {  // 1, 2, 4, 8 -> 0, 1, 2, 3
                    bad |= (ch == 0) | (ch > 8) | ((ch & (ch - 1)) != 0);
                    c = 31u - (uint32_t)__clz((int)(ch | 1u));
                } else {      // 'A' 'C' 'G' 'T' -> 0, 1, 2, 3 (bits 1-2 of the letter give A C T G)
                    bad |= !(ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T');
                    const uint32_t x = (ch >> 1) & 3u;
                    c = x ^ (x >> 1);
                }
                code = (code << 2) | (c & 3u);
            }
        }
        if (short_read) atomicMin(status, HSQ_ERR_OUT_OF_RANGE);
        else if (bad) atomicMin(status, HSQ_ERR_UNSUPPORTED);
        const int16_t qsum = (int16_t)sum;  // Java short arithmetic wraps mod 2^16
        uid_out[k] = code;
        rank_out[k] = ((uint64_t)(uint16_t)(32767 - (int)qsum) << 32) | (uint64_t)((uint32_t)read_index[k] ^ 0x80000000u);
        idx_out[k] = (uint32_t)k;
    }
}

// Per-probe nucleotide counts of the UIDs by position (the inputs of NucleotideCompositionUtil.getNucleotideComposition[ByPosition],
// FilterByUid.java:245-249): `weighted` counts every pair's UID, `unique` one UID per group (its best pair).  Tables are
// [probe][position][A, C, G, T]; uid holds the first base in its most significant used bit pair.
__global__ void uid_composition_kernel(const uint64_t* __restrict__ uid, const int32_t* __restrict__ probe, const uint8_t* __restrict__ is_best,
                                       int64_t n, int uid_len, int32_t* __restrict__ unique, int32_t* __restrict__ weighted) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) {
        const uint64_t code = uid[k];
        const int64_t row = (int64_t)probe[k] * uid_len;
        const bool best = is_best[k] != 0;
        for (int i = 0; i < uid_len; i++) {
            const int c = (int)((code >> (2 * (uid_len - 1 - i))) & 3u);
            atomicAdd(&weighted[(row + i) * 4 + c], 1);
            if (best) atomicAdd(&unique[(row + i) * 4 + c], 1);
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
                                                                      const uint32_t* __restrict__ tie, const int32_t* __restrict__ seg_probe) {
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
            if (seg_probe) {  // this segment is one UID-hash bucket of a split probe: the probe's counters add up over its buckets
                const int pr = seg_probe[p];
                atomicAdd(&unique_pairs[pr], s_groups);
                atomicAdd(&duplicate_pairs[pr], m - s_groups);
                atomicMax(&max_pairs[pr], s_max);
            } else {
                unique_pairs[p] = s_groups;
                duplicate_pairs[p] = m - s_groups;
                max_pairs[p] = s_max;
            }
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
                                                                            int32_t* __restrict__ max_pairs, const uint32_t* __restrict__ tie,
                                                                            const int32_t* __restrict__ seg_probe) {
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
            if (seg_probe) {
                const int pr = seg_probe[p];
                atomicAdd(&unique_pairs[pr], groups);
                atomicAdd(&duplicate_pairs[pr], m - groups);
                atomicMax(&max_pairs[pr], size);
            } else {
                unique_pairs[p] = groups;
                duplicate_pairs[p] = m - groups;
                max_pairs[p] = size;
            }
        }
        __syncwarp();  // the shared rows are rewritten by the next probe
    }
}

// ---- probes with more pairs than a block's table takes (a hot amplicon, or a small panel: 5 M pairs on 1 k probes is 5 k pairs per
// probe) are split into UID-hash buckets: all pairs of one UID land in one bucket, so every bucket is a segment the two group kernels
// above reduce on their own, and the probe's counters add up over its buckets (unique = sum, largest group = max).
static constexpr int DEDUP_BUCKET_TARGET = 1024;  // expected pairs per bucket (the table takes 4096: room for an uneven hash)

// buckets[p] = 1 for a probe a block can take, else the power of two >= pairs / DEDUP_BUCKET_TARGET
__global__ void dedup_bucket_count_kernel(const int32_t* __restrict__ cnt, int64_t np, int64_t* __restrict__ buckets) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; p < np; p += stride) {
        int64_t b = 1;
        if (cnt[p] > DEDUP_SEG_MAX)
            while (b * DEDUP_BUCKET_TARGET < cnt[p]) b <<= 1;
        buckets[p] = b;
    }
}

// virtual segment of every pair = first bucket of its probe + hash(UID) mod buckets; seg_probe[] maps a virtual segment back to its probe
__global__ void dedup_bucket_key_kernel(const int32_t* __restrict__ probe, const uint64_t* __restrict__ uid, int64_t n, const int64_t* __restrict__ vbase,
                                        int32_t* __restrict__ vseg) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const int p = probe[i];
        const int64_t b0 = vbase[p], nb = vbase[p + 1] - b0;
        const uint64_t h = (uid[i] ^ (uid[i] >> 31)) * 0xD6E8FEB86659FD93ull;  // not the multiplier of the in-block table
        vseg[i] = (int32_t)(b0 + (int64_t)((h >> 40) & (uint64_t)(nb - 1)));
    }
}

__global__ void dedup_bucket_probe_kernel(const int64_t* __restrict__ vbase, int64_t np, int32_t* __restrict__ seg_probe) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; p < np; p += stride)
        for (int64_t b = vbase[p]; b < vbase[p + 1]; b++) seg_probe[b] = (int32_t)p;
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
        const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
        const int kblocks = (int)std::min<int64_t>((n + 127) / 128, (int64_t)ctx->sm_count * 16);
        span_begin(ctx, 0);
        check_probe_ids_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), n, np > 0 ? np : INT32_MAX, d_status(ctx));
        dedup_keygen_kernel<false><<<kblocks, 128, 0, ctx->stream>>>(ctx->r1.ascii.as<uint8_t>(), ctx->r1.offsets.as<int64_t>(), ctx->q1.as<uint8_t>(),
                                                                   ctx->r2.ascii.as<uint8_t>(), ctx->r2.offsets.as<int64_t>(), ctx->q2.as<uint8_t>(), nullptr,
                                                                   nullptr, ctx->read_index.as<int32_t>(), n, u, v, ctx->sort_a.as<uint64_t>(),
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
        // scatter the pair indexes into their segments and reduce every segment: one warp per small segment, one block per larger one
        auto reduce_segments = [&](const int32_t* seg_of_pair, const int64_t* seg_start, int64_t n_seg, int seg_max, const int32_t* seg_probe) -> int {
            int cap = 32;
            while (cap < seg_max) cap <<= 1;
            const size_t smem = (size_t)cap * 52 + 16;
            HSQ_CUDA(ctx, cudaFuncSetAttribute(dedup_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            int per_sm = 1;
            HSQ_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dedup_group_kernel, DEDUP_SEG_BLOCK, smem));
            const int sblocks = (int)std::max<int64_t>(1, std::min<int64_t>(n_seg, (int64_t)ctx->sm_count * std::max(per_sm, 1)));
            dedup_scatter_kernel<<<gblocks, 256, 0, ctx->stream>>>(seg_of_pair, n, seg_start, ia, ib);
            const int wblocks = (int)std::max<int64_t>(1, std::min<int64_t>((n_seg + 7) / 8, (int64_t)ctx->sm_count * 8));
            dedup_group_small_kernel<<<wblocks, DEDUP_SEG_BLOCK, 0, ctx->stream>>>(ib, seg_start, n_seg, uid, ka, d_best, d_gs, d_stats, d_stats + np,
                                                                                   d_stats + 2 * np, tie, seg_probe);
            ctx->launches += 2;
            if (seg_max > DEDUP_SMALL_MAX) {
                dedup_group_kernel<<<sblocks, DEDUP_SEG_BLOCK, smem, ctx->stream>>>(ib, seg_start, n_seg, cap, uid, ka, d_best, d_gs, d_stats, d_stats + np,
                                                                                   d_stats + 2 * np, DEDUP_SMALL_MAX, tie, seg_probe);
                ctx->launches++;
            }
            HSQ_CUDA(ctx, cudaGetLastError());
            return HSQ_OK;
        };
        if (max_count <= DEDUP_SEG_MAX) return reduce_segments(ctx->probe_id.as<int32_t>(), start, np, max_count, nullptr);
        if (uid_len > 0) {
            // some probe holds more pairs than a block takes: split those probes into UID-hash buckets and reduce the buckets
            dedup_bucket_count_kernel<<<pblocks, 256, 0, ctx->stream>>>(cnt, np, cnt64);
            ctx->launches++;
            if ((rc = scan_i64(ctx, cnt64, np, tiles, start))) return rc;  // start: first bucket of every probe from here on
            int64_t* h_nv = reinterpret_cast<int64_t*>(ctx->h_cells) + 5;
            HSQ_CUDA(ctx, cudaMemcpyAsync(h_nv, start + np, 8, cudaMemcpyDeviceToHost, ctx->stream));
            HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            const int64_t nv = *h_nv;
            const size_t nvs = (size_t)nv;
            const int64_t v_tiles = (nv + 1 + SCAN_TILE - 1) / SCAN_TILE;
            HSQ_CUDA(ctx, ctx->dd_vseg.ensure(nvs * 8 + (nvs + 1) * 16 + (size_t)v_tiles * 8 + 64));
            int64_t* vcnt64 = ctx->dd_vseg.as<int64_t>();
            int64_t* vstart = vcnt64 + nvs + 1;
            int64_t* vtiles = vstart + nvs + 1;
            int32_t* vcnt = reinterpret_cast<int32_t*>(vtiles + v_tiles);
            int32_t* seg_probe = vcnt + nvs;
            int32_t* vseg = reinterpret_cast<int32_t*>(kb);  // the radix path's second key buffer is free here
            HSQ_CUDA(ctx, cudaMemsetAsync(vcnt, 0, nvs * 4, ctx->stream));
            HSQ_CUDA(ctx, cudaMemsetAsync(d_max, 0, 4, ctx->stream));
            const int vblocks = (int)std::max<int64_t>(1, std::min<int64_t>((nv + 255) / 256, (int64_t)ctx->sm_count * 8));
            dedup_bucket_key_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), uid, n, start, vseg);
            dedup_bucket_probe_kernel<<<pblocks, 256, 0, ctx->stream>>>(start, np, seg_probe);
            dedup_probe_hist_kernel<<<gblocks, 256, 0, ctx->stream>>>(vseg, n, vcnt, ia);
            dedup_seg_prepare_kernel<<<vblocks, 256, 0, ctx->stream>>>(vcnt, nv, vcnt64, d_max);
            ctx->launches += 4;
            if ((rc = scan_i64(ctx, vcnt64, nv, vtiles, vstart))) return rc;
            if ((rc = read_status(ctx, w, 9))) return rc;
            if (w[8] <= DEDUP_SEG_MAX) return reduce_segments(vseg, vstart, nv, w[8], seg_probe);
            // one UID group alone (or a very uneven hash) exceeds a block's table: the radix path takes the batch
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

extern "C" int32_t hsq_dedup_batch_report(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t ext_uid_len, int32_t lig_uid_len, uint8_t* is_best,
                                          int32_t* group_size, hsq_probe_stats* stats, hsq_uid_report* report) {
    if (!report || !report->unique_by_position || !report->weighted_by_position) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null report");
    int rc = hsq_dedup_batch(ctx, pairs, ext_uid_len, lig_uid_len, is_best, group_size, stats);
    if (rc) return rc;
    // the UID codes (sort_a), the probe ids and the answers of the reduction are still on the device
    const int L = ext_uid_len + lig_uid_len;
    report->uid_len = L;
    const size_t cells = (size_t)stats->n_probes * (size_t)L * 4;
    if (cells == 0) return HSQ_OK;
    // (no CallTimer: hsq_last_timing / hsq_last_spans keep describing the dedup call)
    HSQ_CUDA(ctx, ctx->misc.ensure(2 * cells * 4 + 16));
    int32_t* d_unique = ctx->misc.as<int32_t>();
    int32_t* d_weighted = d_unique + cells;
    HSQ_CUDA(ctx, cudaMemsetAsync(d_unique, 0, 2 * cells * 4, ctx->stream));
    if (pairs->n > 0) {
        const int blocks = (int)std::min<int64_t>((pairs->n + 255) / 256, (int64_t)ctx->sm_count * 16);
        uid_composition_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->sort_a.as<uint64_t>(), ctx->probe_id.as<int32_t>(), ctx->out_u8.as<uint8_t>(),
                                                                 pairs->n, L, d_unique, d_weighted);
        ctx->last_launches++;
        HSQ_CUDA(ctx, cudaGetLastError());
    }
    if ((rc = d2h(ctx, report->unique_by_position, d_unique, cells))) return rc;
    if ((rc = d2h(ctx, report->weighted_by_position, d_weighted, cells))) return rc;
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return HSQ_OK;
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
