// hsq_capi_merge.inl -- hsq_merge_pairs: mergeSequences per item
// (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:1362-1552).  HBM-bound byte work: a length kernel,
// the exclusive scan of hsq_scan.cuh, and a write kernel (one warp per item).
namespace hsq {

// Section layout of one merge, restating the branch structure of :1371-1468 (quirks included: without an overlap the WHOLE
// upstream / downstream strings are the lead / trail whatever lead_len / tr_len say).
struct MergePlan {
    int lead_off, lead_n;   // from the upstream read
    int ov_off_up, ov_n;    // overlap: upstream offset (downstream offset is 0), or the length of the N fill
    int fill;               // 1: the "overlap" is an artificial gap of N / '!'
    int trail_from_down, trail_off, trail_n;
    int error;              // the Java throws (negative index into the upstream read)
    __device__ inline int64_t total() const { return (int64_t)lead_n + ov_n + trail_n; }
};

__device__ inline MergePlan merge_plan(int ul, int dl, int up_start, int down_start) {
    MergePlan m{};
    const int up_end = up_start + ul - 1, down_end = down_start + dl - 1;
    const int min_end = min(up_end, down_end), max_end = max(up_end, down_end);
    m.trail_from_down = (down_end == max_end) ? 1 : 0;
    const int tsrc_start = m.trail_from_down ? down_start : up_start;
    int lead_s = 0, lead_e = 0, lead_len = 0, ov_s = 0, ov_e = 0, ov_len = 0, tr_s = 0, tr_e = 0, tr_len = 0;
    if (down_start <= up_end) {
        ov_s = down_start; ov_e = min_end; ov_len = ov_e - ov_s + 1;
        if (up_start < ov_s) { lead_s = up_start; lead_e = ov_s - 1; lead_len = lead_e - lead_s + 1; }
        if (max_end > ov_e) { tr_s = ov_e + 1; tr_e = max_end; tr_len = tr_e - tr_s + 1; }
    } else {
        lead_s = up_start; lead_e = lead_s + ul - 1; lead_len = lead_e - lead_s + 1;
        tr_s = down_start; tr_e = tr_s + dl - 1; tr_len = tr_e - tr_s + 1;
    }
    if (ov_len == 0) {
        const int gs = lead_e + 1, ge = tr_s - 1;
        const int fill = ge >= gs ? ge - gs + 1 : 0;
        m.fill = 1;
        m.ov_n = fill;
        m.lead_off = 0; m.lead_n = lead_len > 0 ? ul : 0;
        m.trail_from_down = 1; m.trail_off = 0; m.trail_n = tr_len > 0 ? dl : 0;
    } else {
        m.fill = 0;
        m.ov_off_up = ov_s - up_start;
        m.ov_n = ov_len;
        if (lead_len >= 1) { m.lead_off = lead_s - up_start; m.lead_n = lead_len; }
        if (tr_len >= 1) { m.trail_off = tr_s - tsrc_start; m.trail_n = tr_len; }
        if (m.ov_off_up < 0 || m.lead_off < 0 || m.trail_off < 0) m.error = 1;  // String.substring would throw
    }
    return m;
}

__global__ void merge_len_kernel(const int64_t* __restrict__ up_off, const int64_t* __restrict__ down_off, const int32_t* __restrict__ up_start,
                                 const int32_t* __restrict__ down_start, int64_t n, int64_t* __restrict__ keep, int* status) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) {
        const MergePlan m = merge_plan((int)(up_off[k + 1] - up_off[k]), (int)(down_off[k + 1] - down_off[k]), up_start[k], down_start[k]);
        if (m.error) atomicMin(status, HSQ_ERR_OUT_OF_RANGE);
        keep[k] = m.error ? 0 : m.total();
    }
}

__global__ void merge_write_kernel(const uint8_t* __restrict__ up, const uint8_t* __restrict__ upq, const int64_t* __restrict__ up_off,
                                   const uint8_t* __restrict__ down, const uint8_t* __restrict__ downq, const int64_t* __restrict__ down_off,
                                   const int32_t* __restrict__ up_start, const int32_t* __restrict__ down_start,
                                   const uint8_t* __restrict__ negative, const int64_t* __restrict__ out_off, int64_t n,
                                   uint8_t* __restrict__ out_seq, uint8_t* __restrict__ out_qual, int* status) {
    const int lane = threadIdx.x & 31;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    bool bad = false;
    for (; w < n; w += stride) {
        const int64_t u0 = up_off[w], d0 = down_off[w];
        const MergePlan m = merge_plan((int)(up_off[w + 1] - u0), (int)(down_off[w + 1] - d0), up_start[w], down_start[w]);
        if (m.error) continue;
        const int64_t dst = out_off[w];
        const int total = (int)(out_off[w + 1] - dst);
        const bool neg = negative[w] != 0;
        const int64_t t0 = m.trail_from_down ? d0 : u0;
        const uint8_t* tsrc = m.trail_from_down ? down : up;
        const uint8_t* tsrcq = m.trail_from_down ? downq : upq;
        for (int p = lane; p < total; p += 32) {
            uint8_t base, q;
            if (p < m.lead_n) {
                base = up[u0 + m.lead_off + p];
                q = upq[u0 + m.lead_off + p];
            } else if (p < m.lead_n + m.ov_n) {
                const int i = p - m.lead_n;
                if (m.fill) { base = 'N'; q = '!'; }
                else {
                    const uint8_t uq = upq[u0 + m.ov_off_up + i], dq = downq[d0 + i];  // (:1505-1506)
                    const bool take_up = uq >= dq;                                     // higher quality wins, ties upstream (:1508)
                    base = take_up ? up[u0 + m.ov_off_up + i] : down[d0 + i];
                    q = take_up ? uq : dq;
                }
            } else {
                const int i = p - m.lead_n - m.ov_n;
                base = tsrc[t0 + m.trail_off + i];
                q = tsrcq[t0 + m.trail_off + i];
            }
            uint8_t code = encode_ascii(base);
            bad |= (code == CODE_BAD);
            if (neg) code = complement_code(code);               // getReverseCompliment + reversed qualities (:1544-1548)
            const int o = neg ? total - 1 - p : p;
            out_seq[dst + o] = decode_code(code);
            out_qual[dst + o] = q;
        }
    }
    if (bad) atomicMin(status, HSQ_ERR_BAD_CHARACTER);
}

}  // namespace hsq

extern "C" int32_t hsq_merge_pairs(hsq_ctx* ctx, const hsq_merge_batch* mb, uint8_t* out_seq, uint8_t* out_qual, int64_t* out_offsets,
                                   int64_t out_cap) {
    if (!ctx || !mb || !out_offsets || out_cap < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    const int64_t n = mb->n;
    if (n < 0 || mb->up.n != n || mb->down.n != n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "merge batch arrays disagree on n");
    if (n > 0 && (!mb->negative_strand || !mb->up_start || !mb->down_start)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null array");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = reset_status(ctx);
    if (rc) return rc;
    if ((rc = copy_seqs(ctx, &mb->up, &ctx->sa))) return rc;
    if ((rc = copy_seqs(ctx, &mb->down, &ctx->sb))) return rc;
    const size_t nn = (size_t)n;
    struct { hsq::DevBuf* d; const void* h; size_t bytes; } up[] = {
        {&ctx->q1, mb->up_qual ? mb->up_qual + mb->up.offsets[0] : nullptr, (size_t)ctx->sa.total},
        {&ctx->q2, mb->down_qual ? mb->down_qual + mb->down.offsets[0] : nullptr, (size_t)ctx->sb.total},
        {&ctx->probe_id, mb->up_start, nn * 4}, {&ctx->read_index, mb->down_start, nn * 4}, {&ctx->flags, mb->negative_strand, nn}};
    for (auto& c : up) {
        if (c.bytes && !c.h) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_merge_pairs: null input array");
        HSQ_CUDA(ctx, c.d->ensure(c.bytes + 16));
        if (c.bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(c.d->p, c.h, c.bytes, cudaMemcpyHostToDevice, ctx->stream));
    }
    HSQ_CUDA(ctx, ctx->sort_a.ensure((nn + 1) * 8));  // merged lengths
    HSQ_CUDA(ctx, ctx->sort_b.ensure((nn + 1) * 8));  // out offsets
    const int64_t n_tiles = (n + 1 + SCAN_TILE - 1) / SCAN_TILE;
    HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)n_tiles * 8 + 16));
    int64_t* keep = ctx->sort_a.as<int64_t>();
    int64_t* d_out_off = ctx->sort_b.as<int64_t>();
    if (n > 0) {
        const int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
        merge_len_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->sa.offsets.as<int64_t>(), ctx->sb.offsets.as<int64_t>(), ctx->probe_id.as<int32_t>(),
                                                         ctx->read_index.as<int32_t>(), n, keep, d_status(ctx));
        ctx->launches++;
    }
    scan_tile_sums_kernel<<<(unsigned)n_tiles, SCAN_BLOCK, 0, ctx->stream>>>(keep, n, ctx->sort_c.as<int64_t>());
    scan_sums_kernel<<<1, SCAN_BLOCK, 0, ctx->stream>>>(ctx->sort_c.as<int64_t>(), n_tiles);
    scan_apply_kernel<<<(unsigned)n_tiles, SCAN_BLOCK, 0, ctx->stream>>>(keep, n, ctx->sort_c.as<int64_t>(), d_out_off);
    ctx->launches += 3;
    HSQ_CUDA(ctx, cudaMemcpyAsync(out_offsets, d_out_off, (nn + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    int w[1];
    if ((rc = read_status(ctx, w, 1))) return rc;
    if (w[0] != 0) {
        timer.finish();
        return fail(ctx, w[0], "hsq_merge_pairs: the downstream read starts before the upstream read (the reference's substring would throw)");
    }
    const int64_t total = out_offsets[n];
    if (total > out_cap) {
        timer.finish();
        return fail(ctx, HSQ_ERR_OVERFLOW, "hsq_merge_pairs: output buffers too small; out_offsets holds the needed sizes");
    }
    if (total > 0) {
        if (!out_seq || !out_qual) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null output buffer");
        HSQ_CUDA(ctx, ctx->out_slots.ensure(2 * (size_t)total + 32));
        uint8_t* d_seq = ctx->out_slots.as<uint8_t>();
        uint8_t* d_qual = d_seq + total + 16;
        const int blocks = (int)std::min<int64_t>((n * 32 + 255) / 256, (int64_t)ctx->sm_count * 16);
        merge_write_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->sa.ascii.as<uint8_t>(), ctx->q1.as<uint8_t>(), ctx->sa.offsets.as<int64_t>(),
                                                           ctx->sb.ascii.as<uint8_t>(), ctx->q2.as<uint8_t>(), ctx->sb.offsets.as<int64_t>(),
                                                           ctx->probe_id.as<int32_t>(), ctx->read_index.as<int32_t>(), ctx->flags.as<uint8_t>(),
                                                           d_out_off, n, d_seq, d_qual, d_status(ctx));
        ctx->launches++;
        HSQ_CUDA(ctx, cudaGetLastError());
        HSQ_CUDA(ctx, cudaMemcpyAsync(out_seq, d_seq, (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(out_qual, d_qual, (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
        if ((rc = read_status(ctx, w, 1))) return rc;
        if (w[0] != 0) {
            timer.finish();
            return fail(ctx, w[0], std::string("hsq_merge_pairs: ") + status_message(w[0]));
        }
    }
    return timer.finish();
}
