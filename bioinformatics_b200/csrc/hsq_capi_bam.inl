// hsq_capi_bam.inl -- hsq_index_bam: record index of a BAM file whose BGZF blocks were written at compression level 0.
// The reference writes its intermediate BAMs that way (HSQ/process/FastqAndBamFileMerger.java:131-133: setCompressionLevel(0) before
// makeBAMWriter; records added at :386-396) and reads them back record by record through htsjdk (...Identification.java:439-474).  At
// level 0 a BGZF block holds `stored` deflate blocks: the record bytes sit in the file as they are, so an index needs no inflate.
//   host   : walk of the BGZF block headers (18 + 8 bytes per 64 KiB block) -> table of stored segments; BAM header (magic, text,
//            reference table) read through that table -> offset of the first alignment record
//   device : segments gathered into one contiguous stream; EVERY byte position is tested for "a record could start here" (sizes
//            consistent, reference ids in range, read name NUL terminated); the survivors, in order, are the records if each one ends
//            where the next begins -- checked for all of them at once.  Where that check fails (bytes inside a record that look like
//            a record header: it happens) the records are the candidates on the path first -> next -> ... -> end, marked by binary
//            lifting over the candidate list (log2(n) rounds of n-element kernels): exact, and still parallel.
//   device : one thread per record: offsets / lengths of the variable fields, the fixed columns, reference length of the CIGAR.
// BAM layout: SAM/BAM format specification 4.2 (htsjdk BAMRecordCodec writes it).
#include <vector>
namespace hsq {

__device__ __forceinline__ uint32_t ld_u32(const uint8_t* __restrict__ p) {  // little-endian, any alignment
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
__device__ __forceinline__ uint32_t ld_u16(const uint8_t* __restrict__ p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }

// one block per segment (grid stride): dst-aligned 16-byte chunks through load16_unaligned, the edges byte by byte
__global__ void bgzf_gather_kernel(const uint8_t* __restrict__ file, const int64_t* __restrict__ seg, int64_t n_seg, uint8_t* __restrict__ out) {
    for (int64_t s = blockIdx.x; s < n_seg; s += gridDim.x) {
        const int64_t src = seg[3 * s], dst = seg[3 * s + 1], len = seg[3 * s + 2];
        const int64_t end = dst + len;
        const int64_t lo = min(end, (dst + 15) & ~(int64_t)15), hi = max(lo, end & ~(int64_t)15);
        for (int64_t x = dst + threadIdx.x; x < lo; x += blockDim.x) out[x] = file[src + (x - dst)];
        for (int64_t c = lo + 16 * (int64_t)threadIdx.x; c < hi; c += 16 * (int64_t)blockDim.x)
            *reinterpret_cast<uint4*>(out + c) = load16_unaligned(file + src + (c - dst));
        for (int64_t x = hi + threadIdx.x; x < end; x += blockDim.x) out[x] = file[src + (x - dst)];
    }
}

static constexpr int BAM_FIXED = 36;  // block_size + the 32 fixed bytes of an alignment record

// could an alignment record start at p?  (every true start passes; what else passes is sorted out by the chain check)
__device__ __forceinline__ bool bam_start_plausible(const uint8_t* __restrict__ d, int64_t p, int64_t end, int n_ref) {
    if (p + BAM_FIXED > end) return false;
    const uint32_t bs = ld_u32(d + p);
    if (bs < 32u + 1u || bs > (1u << 28) || p + 4 + (int64_t)bs > end) return false;
    const int ref_id = (int)ld_u32(d + p + 4), next_ref = (int)ld_u32(d + p + 24);
    if (ref_id < -1 || ref_id >= n_ref || next_ref < -1 || next_ref >= n_ref) return false;
    if ((int)ld_u32(d + p + 8) < -1 || (int)ld_u32(d + p + 28) < -1) return false;
    const uint32_t l_name = d[p + 12], n_cigar = ld_u16(d + p + 16), l_seq = ld_u32(d + p + 20);
    if (l_name == 0 || l_seq > (1u << 28)) return false;
    const uint64_t need = 32ull + l_name + 4ull * n_cigar + (l_seq + 1ull) / 2 + l_seq;
    if (need > bs) return false;
    return d[p + BAM_FIXED + l_name - 1] == 0;
}

__device__ __forceinline__ uint32_t bam_mask16(const uint8_t* __restrict__ d, int64_t base, int64_t first, int64_t end, int n_ref) {
    uint32_t m = 0;
    for (int k = 0; k < NL_BYTES; k++) {
        const int64_t p = base + k;
        if (p >= first && bam_start_plausible(d, p, end, n_ref)) m |= 1u << k;
    }
    return m;
}

__global__ void bam_count_kernel(const uint8_t* __restrict__ d, int64_t first, int64_t end, int n_ref, int64_t* __restrict__ tile_counts) {
    const int64_t base = (int64_t)blockIdx.x * NL_TILE + (int64_t)threadIdx.x * NL_BYTES;
    const uint32_t m = base < end ? bam_mask16(d, base, first, end, n_ref) : 0u;
    int64_t total;
    block_exclusive_scan((int64_t)__popc(m), &total);
    if (threadIdx.x == 0) tile_counts[blockIdx.x] = total;
}

__global__ void bam_write_kernel(const uint8_t* __restrict__ d, int64_t first, int64_t end, int n_ref, const int64_t* __restrict__ tile_excl,
                                 int64_t* __restrict__ cand) {
    const int64_t base = (int64_t)blockIdx.x * NL_TILE + (int64_t)threadIdx.x * NL_BYTES;
    uint32_t m = base < end ? bam_mask16(d, base, first, end, n_ref) : 0u;
    int64_t total;
    int64_t rank = tile_excl[blockIdx.x] + block_exclusive_scan((int64_t)__popc(m), &total);
    while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        cand[rank++] = base + k;
    }
}

// flag[0] = 1 unless the candidates are exactly the chain first -> ... -> end
__global__ void bam_chain_check_kernel(const uint8_t* __restrict__ d, const int64_t* __restrict__ cand, int64_t n, int64_t first, int64_t end, int* flag) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    if (r == 0 && (n == 0 ? first != end : cand[0] != first)) *flag = 1;
    for (; r < n; r += stride) {
        const int64_t nxt = cand[r] + 4 + (int64_t)ld_u32(d + cand[r]);
        if (nxt != (r + 1 < n ? cand[r + 1] : end)) *flag = 1;
    }
}

// ---- the exact chain when some candidate is not a record: path marking by binary lifting over the candidate list -------------
// Node r < n is candidate r; node n = END (the end of the stream), node n + 1 = BAD (a successor that is not a candidate).
// succ[r] = node of cand[r] + 4 + block_size; END and BAD point to themselves.
__global__ void bam_succ_kernel(const uint8_t* __restrict__ d, const int64_t* __restrict__ cand, int64_t n, int64_t end, uint32_t* __restrict__ succ) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n + 2; r += stride) {
        if (r >= n) {
            succ[r] = (uint32_t)r;
            continue;
        }
        const int64_t nxt = cand[r] + 4 + (int64_t)ld_u32(d + cand[r]);
        uint32_t to = (uint32_t)(n + 1);
        if (nxt == end) to = (uint32_t)n;
        else {
            int64_t lo = r + 1, hi = n;  // first candidate at or after nxt
            while (lo < hi) {
                const int64_t mid = (lo + hi) >> 1;
                if (cand[mid] < nxt) lo = mid + 1;
                else hi = mid;
            }
            if (lo < n && cand[lo] == nxt) to = (uint32_t)lo;
        }
        succ[r] = to;
    }
}
__global__ void bam_lift_kernel(const uint32_t* __restrict__ prev, int64_t n_nodes, uint32_t* __restrict__ next) {  // next = prev o prev
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n_nodes; r += stride) next[r] = prev[prev[r]];
}
// every marked node marks its 2^k-th successor (marked nodes are on the path from the first record, so what they mark is too)
__global__ void bam_mark_kernel(const uint32_t* __restrict__ jump, int64_t n_nodes, uint8_t* __restrict__ mark) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n_nodes; r += stride)
        if (mark[r]) mark[jump[r]] = 1;
}
__global__ void bam_mark_flags_kernel(const uint8_t* __restrict__ mark, const uint32_t* __restrict__ succ, const int64_t* __restrict__ cand, int64_t n,
                                      int64_t* __restrict__ flags, unsigned long long* first_break) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n; r += stride) {
        flags[r] = mark[r] ? 1 : 0;
        if (mark[r] && succ[r] == (uint32_t)(n + 1)) atomicMin(first_break, (unsigned long long)cand[r]);
    }
}
__global__ void bam_keep_marked_kernel(const int64_t* __restrict__ cand, const int64_t* __restrict__ flags, const int64_t* __restrict__ pos, int64_t n,
                                       int64_t* __restrict__ rec) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n; r += stride)
        if (flags[r]) rec[pos[r]] = cand[r];
}

// reference bases a CIGAR consumes: M D N = X (ops 0 2 3 7 8)
__global__ void bam_fields_kernel(const uint8_t* __restrict__ d, const int64_t* __restrict__ rec, int64_t n, int64_t cap, int64_t* __restrict__ off,
                                  int32_t* __restrict__ val) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) {
        const int64_t p = rec[k];
        const uint8_t* r = d + p;
        const int64_t bs = (int64_t)ld_u32(r);
        const int l_name = r[12], n_cigar = (int)ld_u16(r + 16), l_seq = (int)ld_u32(r + 20);
        const int64_t name = p + BAM_FIXED, cigar = name + l_name, seq = cigar + 4ll * n_cigar, qual = seq + (l_seq + 1) / 2, tags = qual + l_seq;
        int ref_len = 0;
        for (int c = 0; c < n_cigar; c++) {
            const uint32_t op = ld_u32(d + cigar + 4ll * c);
            const uint32_t kind = op & 15u;
            if (kind == 0 || kind == 2 || kind == 3 || kind == 7 || kind == 8) ref_len += (int)(op >> 4);
        }
        off[HSQ_BAM_RECORD * cap + k] = p;
        off[HSQ_BAM_QNAME * cap + k] = name;
        off[HSQ_BAM_CIGAR * cap + k] = cigar;
        off[HSQ_BAM_SEQ * cap + k] = seq;
        off[HSQ_BAM_QUAL * cap + k] = qual;
        off[HSQ_BAM_TAGS * cap + k] = tags;
        val[HSQ_BAM_RECORD * cap + k] = (int32_t)(bs + 4);
        val[HSQ_BAM_QNAME * cap + k] = l_name - 1;
        val[HSQ_BAM_CIGAR * cap + k] = n_cigar;
        val[HSQ_BAM_SEQ * cap + k] = l_seq;
        val[HSQ_BAM_QUAL * cap + k] = l_seq;
        val[HSQ_BAM_TAGS * cap + k] = (int32_t)(p + 4 + bs - tags);
        val[HSQ_BAM_REFID * cap + k] = (int32_t)ld_u32(r + 4);
        val[HSQ_BAM_POS * cap + k] = (int32_t)ld_u32(r + 8) + 1;
        val[HSQ_BAM_FLAG * cap + k] = (int32_t)ld_u16(r + 18);
        val[HSQ_BAM_MAPQ * cap + k] = r[13];
        val[HSQ_BAM_NEXT_REFID * cap + k] = (int32_t)ld_u32(r + 24);
        val[HSQ_BAM_PNEXT * cap + k] = (int32_t)ld_u32(r + 28) + 1;
        val[HSQ_BAM_TLEN * cap + k] = (int32_t)ld_u32(r + 32);
        val[HSQ_BAM_REF_LEN * cap + k] = ref_len;
    }
}

}  // namespace hsq

namespace {

inline uint32_t h_u16(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }
inline uint32_t h_u32(const uint8_t* p) { return h_u16(p) | (h_u16(p + 2) << 16); }

// BGZF block walk (SAM/BAM specification 4.1): [src, dst, len] of every stored deflate block that carries data
int bgzf_stored_segments(hsq_ctx* ctx, const uint8_t* f, int64_t n, std::vector<int64_t>* seg, int64_t* payload_bytes) {
    int64_t pos = 0, out = 0;
    while (pos < n) {
        if (pos + 18 > n || f[pos] != 31 || f[pos + 1] != 139 || f[pos + 2] != 8 || !(f[pos + 3] & 4))
            return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: not a BGZF block header at byte " + std::to_string(pos));
        const int64_t xlen = h_u16(f + pos + 10);
        int64_t bsize = -1;
        for (int64_t x = pos + 12; x + 4 <= pos + 12 + xlen && x + 4 <= n;) {
            const int64_t slen = h_u16(f + x + 2);
            if (f[x] == 'B' && f[x + 1] == 'C' && slen == 2 && x + 6 <= n) bsize = h_u16(f + x + 4);
            x += 4 + slen;
        }
        if (bsize < 0) return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: gzip member without the BGZF 'BC' field at byte " + std::to_string(pos));
        const int64_t total = bsize + 1, c0 = pos + 12 + xlen, c1 = pos + total - 8;
        if (pos + total > n || c1 < c0) return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: truncated BGZF block at byte " + std::to_string(pos));
        const int64_t isize = h_u32(f + pos + total - 4);
        if (isize > 0) {  // (the end-of-file marker is an empty block in fixed-Huffman coding: nothing to take from it)
            int64_t q = c0, sum = 0;
            for (;;) {
                if (q >= c1) return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: deflate stream ends without a final block at byte " + std::to_string(pos));
                const int hdr = f[q];
                if (((hdr >> 1) & 3) != 0)
                    return fail(ctx, HSQ_ERR_UNSUPPORTED, "hsq_index_bam: BGZF block at byte " + std::to_string(pos) +
                                                              " is deflate-compressed; only files written at compression level 0 are indexed");
                if (q + 5 > c1) return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: truncated stored block at byte " + std::to_string(q));
                const int64_t len = h_u16(f + q + 1), nlen = h_u16(f + q + 3);
                if ((len ^ nlen) != 0xFFFF || q + 5 + len > c1) return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: bad stored block at byte " + std::to_string(q));
                if (len) {
                    seg->push_back(q + 5);
                    seg->push_back(out + sum);
                    seg->push_back(len);
                }
                sum += len;
                q += 5 + len;
                if (hdr & 1) break;
            }
            if (sum != isize || q != c1) return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: BGZF block at byte " + std::to_string(pos) + " does not hold ISIZE bytes");
            out += sum;
        }
        pos += total;
    }
    *payload_bytes = out;
    return HSQ_OK;
}

// sequential reads of the uncompressed stream through the segment table (the BAM header only)
struct SegmentReader {
    const uint8_t* f;
    const std::vector<int64_t>& seg;
    size_t s = 0;
    int64_t pos = 0;  // offset in the uncompressed stream
    bool read(uint8_t* out, int64_t n) {  // out may be null: skip
        while (n > 0) {
            while (s * 3 < seg.size() && seg[3 * s + 1] + seg[3 * s + 2] <= pos) s++;
            if (s * 3 >= seg.size()) return false;
            const int64_t in = pos - seg[3 * s + 1], take = std::min(n, seg[3 * s + 2] - in);
            if (out) {
                std::memcpy(out, f + seg[3 * s] + in, (size_t)take);
                out += take;
            }
            pos += take;
            n -= take;
        }
        return true;
    }
    bool i32(int32_t* v) {
        uint8_t b[4];
        if (!read(b, 4)) return false;
        *v = (int32_t)h_u32(b);
        return true;
    }
};

}  // namespace

extern "C" int32_t hsq_index_bam(hsq_ctx* ctx, const uint8_t* file, int64_t n_bytes, hsq_bam_index* out) {
    if (!ctx || !out || n_bytes < 0 || (n_bytes > 0 && !file)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_index_bam: null argument");
    if (out->cap < 0 || (out->cap > 0 && (!out->off || !out->val))) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_index_bam: null planes");
    out->n = 0;
    out->n_ref = 0;
    out->text_bytes = 0;
    out->first_record = 0;
    out->payload_bytes = 0;
    out->exact_walk = 0;
    std::vector<int64_t> seg;
    int64_t P = 0;
    int rc = bgzf_stored_segments(ctx, file, n_bytes, &seg, &P);
    if (rc) return rc;
    out->payload_bytes = P;
    // BAM header: magic, l_text, text, n_ref, { l_name, name, l_ref }
    SegmentReader rd{file, seg};
    uint8_t magic[4];
    int32_t l_text = 0, n_ref = 0;
    if (!rd.read(magic, 4) || std::memcmp(magic, "BAM\1", 4) != 0) return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: the stream does not start with BAM\\1");
    if (!rd.i32(&l_text) || l_text < 0 || !rd.read(nullptr, l_text) || !rd.i32(&n_ref) || n_ref < 0)
        return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: truncated BAM header");
    for (int32_t r = 0; r < n_ref; r++) {
        int32_t l_name = 0, l_ref = 0;
        if (!rd.i32(&l_name) || l_name < 0 || !rd.read(nullptr, l_name) || !rd.i32(&l_ref))
            return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: truncated reference table");
    }
    const int64_t first = rd.pos;
    out->text_bytes = l_text;
    out->n_ref = n_ref;
    out->first_record = first;
    if (out->payload && out->payload_cap < P) return fail(ctx, HSQ_ERR_OVERFLOW, "hsq_index_bam: payload_cap too small; payload_bytes holds the size");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    if ((rc = reset_status(ctx))) return rc;
    const int64_t n_seg = (int64_t)seg.size() / 3;
    const int64_t n_tiles = (P + NL_TILE - 1) / NL_TILE;
    const int64_t max_records = (P - first) / (BAM_FIXED + 1) + 2;
    HSQ_CUDA(ctx, ctx->r1.ascii.ensure((size_t)n_bytes + 64));
    HSQ_CUDA(ctx, ctx->r2.ascii.ensure((size_t)P + 64));
    HSQ_CUDA(ctx, ctx->sort_b.ensure((size_t)(3 * n_seg + 1) * 8));
    HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)(n_tiles + 2) * 8 + 16));
    uint8_t* d_file = ctx->r1.ascii.as<uint8_t>();
    uint8_t* d = ctx->r2.ascii.as<uint8_t>();
    int64_t* d_seg = ctx->sort_b.as<int64_t>();
    int64_t* d_tiles = ctx->sort_c.as<int64_t>();
    if (n_bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(d_file, file, (size_t)n_bytes, cudaMemcpyHostToDevice, ctx->stream));
    if (n_seg) {
        HSQ_CUDA(ctx, cudaMemcpyAsync(d_seg, seg.data(), (size_t)n_seg * 24, cudaMemcpyHostToDevice, ctx->stream));
        span_begin(ctx, 0);
        bgzf_gather_kernel<<<(unsigned)std::min<int64_t>(n_seg, (int64_t)ctx->sm_count * 8), 256, 0, ctx->stream>>>(d_file, d_seg, n_seg, d);
        ctx->launches++;
    }
    int64_t n_rec = 0;
    if (P > first) {
        HSQ_CUDA(ctx, cudaMemsetAsync(d_tiles + n_tiles, 0, 8, ctx->stream));
        bam_count_kernel<<<(unsigned)n_tiles, NL_BLOCK, 0, ctx->stream>>>(d, first, P, n_ref, d_tiles);
        scan_sums_kernel<<<1, SCAN_BLOCK, 0, ctx->stream>>>(d_tiles, n_tiles + 1);  // in place: exclusive; entry n_tiles = total
        ctx->launches += 2;
        int64_t n_cand = 0;
        HSQ_CUDA(ctx, cudaMemcpyAsync(&n_cand, d_tiles + n_tiles, 8, cudaMemcpyDeviceToHost, ctx->stream));
        HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        HSQ_CUDA(ctx, ctx->sort_a.ensure((size_t)std::max(max_records, n_cand + 2) * 8 + 16));
        int64_t* d_cand = ctx->sort_a.as<int64_t>();
        int* d_flag = d_status(ctx) + 4;
        int64_t* d_walk = reinterpret_cast<int64_t*>(d_status(ctx) + 6);
        bam_write_kernel<<<(unsigned)n_tiles, NL_BLOCK, 0, ctx->stream>>>(d, first, P, n_ref, d_tiles, d_cand);
        bam_chain_check_kernel<<<(unsigned)std::min<int64_t>((n_cand + 255) / 256 + 1, (int64_t)ctx->sm_count * 8), 256, 0, ctx->stream>>>(d, d_cand, n_cand, first,
                                                                                                                                       P, d_flag);
        ctx->launches += 2;
        int flag = 0;
        HSQ_CUDA(ctx, cudaMemcpyAsync(&flag, d_flag, 4, cudaMemcpyDeviceToHost, ctx->stream));
        HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        n_rec = n_cand;
        int64_t* d_rec = d_cand;
        if (flag) {  // some candidate is not a record (or the file is broken): mark the path first -> ... -> end among the candidates
            out->exact_walk = 1;
            if (n_cand + 2 >= (1ll << 32)) return fail(ctx, HSQ_ERR_UNSUPPORTED, "hsq_index_bam: more than 2^32 record candidates");
            if (n_cand == 0) {
                timer.finish();
                return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: no alignment record can start at byte " + std::to_string(first) + " of the uncompressed stream");
            }
            const int64_t nodes = n_cand + 2;
            int levels = 1;
            while ((1ll << levels) < nodes) levels++;
            // two jump tables (ping-pong: level k is needed only to build level k + 1 ... but the marking runs high to low, so all
            // levels are kept), the marks, the flags and their scan
            HSQ_CUDA(ctx, ctx->misc.ensure((size_t)levels * (size_t)nodes * 4 + (size_t)nodes + 64));
            HSQ_CUDA(ctx, ctx->sort_d.ensure(std::max<size_t>((size_t)(3 * nodes + 8) * 8, 64)));
            uint32_t* jump = ctx->misc.as<uint32_t>();
            uint8_t* mark = reinterpret_cast<uint8_t*>(jump + (size_t)levels * (size_t)nodes);
            int64_t* flags = ctx->sort_d.as<int64_t>();
            int64_t* pos = flags + nodes;
            int64_t* kept = pos + nodes + 1;
            const unsigned g = (unsigned)std::min<int64_t>((nodes + 255) / 256, (int64_t)ctx->sm_count * 16);
            bam_succ_kernel<<<g, 256, 0, ctx->stream>>>(d, d_cand, n_cand, P, jump);
            for (int k = 1; k < levels; k++) bam_lift_kernel<<<g, 256, 0, ctx->stream>>>(jump + (size_t)(k - 1) * nodes, nodes, jump + (size_t)k * nodes);
            HSQ_CUDA(ctx, cudaMemsetAsync(mark, 0, (size_t)nodes, ctx->stream));
            HSQ_CUDA(ctx, cudaMemsetAsync(mark, 1, 1, ctx->stream));  // cand[0] is the first record (checked below), or nothing is
            for (int k = levels - 1; k >= 0; k--) bam_mark_kernel<<<g, 256, 0, ctx->stream>>>(jump + (size_t)k * nodes, nodes, mark);
            unsigned long long* d_break = reinterpret_cast<unsigned long long*>(d_walk);
            HSQ_CUDA(ctx, cudaMemsetAsync(d_break, 0xff, 8, ctx->stream));
            bam_mark_flags_kernel<<<g, 256, 0, ctx->stream>>>(mark, jump, d_cand, n_cand, flags, d_break);
            ctx->launches += 2 * levels + 1;
            HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)((nodes + SCAN_TILE) / SCAN_TILE + 2) * 8 + 16));
            if ((rc = scan_i64(ctx, flags, n_cand, ctx->sort_c.as<int64_t>(), pos))) return rc;
            bam_keep_marked_kernel<<<g, 256, 0, ctx->stream>>>(d_cand, flags, pos, n_cand, kept);
            ctx->launches++;
            struct { int64_t first_cand; unsigned long long brk; int64_t n_kept; uint8_t end_marked; } h{};
            HSQ_CUDA(ctx, cudaMemcpyAsync(&h.first_cand, d_cand, 8, cudaMemcpyDeviceToHost, ctx->stream));
            HSQ_CUDA(ctx, cudaMemcpyAsync(&h.brk, d_break, 8, cudaMemcpyDeviceToHost, ctx->stream));
            HSQ_CUDA(ctx, cudaMemcpyAsync(&h.n_kept, pos + n_cand, 8, cudaMemcpyDeviceToHost, ctx->stream));
            HSQ_CUDA(ctx, cudaMemcpyAsync(&h.end_marked, mark + n_cand, 1, cudaMemcpyDeviceToHost, ctx->stream));
            HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            if (h.first_cand != first || h.brk != ~0ull || !h.end_marked) {
                timer.finish();
                const int64_t at = h.first_cand != first ? first : (h.brk != ~0ull ? (int64_t)h.brk : first);
                return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_bam: the chain of alignment records breaks at (or after) byte " + std::to_string(at) +
                                                        " of the uncompressed stream");
            }
            n_rec = h.n_kept;
            d_rec = kept;
        }
        out->n = n_rec;
        if (n_rec > out->cap) {
            timer.finish();
            return fail(ctx, HSQ_ERR_OVERFLOW, "hsq_index_bam: more records than cap; n holds the count");
        }
        if (n_rec) {
            const size_t cap = (size_t)out->cap;
            HSQ_CUDA(ctx, ctx->flags.ensure(cap * HSQ_BAM_N_FIELDS * 8 + 16));
            HSQ_CUDA(ctx, ctx->out_i32.ensure(cap * HSQ_BAM_N_VALUES * 4 + 16));
            bam_fields_kernel<<<(unsigned)std::min<int64_t>((n_rec + 255) / 256, (int64_t)ctx->sm_count * 16), 256, 0, ctx->stream>>>(
                d, d_rec, n_rec, out->cap, ctx->flags.as<int64_t>(), ctx->out_i32.as<int32_t>());
            ctx->launches++;
            span_end(ctx, 0);  // kernels only: the copies below are not part of the span
            HSQ_CUDA(ctx, cudaGetLastError());
            for (int f = 0; f < HSQ_BAM_N_FIELDS; f++)
                if ((rc = d2h(ctx, out->off + (size_t)f * cap, ctx->flags.as<int64_t>() + (size_t)f * cap, (size_t)n_rec))) return rc;
            for (int v = 0; v < HSQ_BAM_N_VALUES; v++)
                if ((rc = d2h(ctx, out->val + (size_t)v * cap, ctx->out_i32.as<int32_t>() + (size_t)v * cap, (size_t)n_rec))) return rc;
        }
    }
    if (!ctx->span_done[0]) span_end(ctx, 0);
    HSQ_CUDA(ctx, cudaGetLastError());
    if (out->payload && P) HSQ_CUDA(ctx, cudaMemcpyAsync(out->payload, d, (size_t)P, cudaMemcpyDeviceToHost, ctx->stream));
    return timer.finish();
}
