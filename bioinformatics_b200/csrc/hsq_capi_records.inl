// hsq_capi_records.inl -- hsq_extend_records: the read string, base qualities, FLAG and PNEXT of the two SAM records of each
// extended pair (HSQ/process/ExtendReadsToPrimer.java:131-181 extendReadPair, :247-284 createRecord;
// HSQ/utils/SAMRecordUtil.java:270-289 setSAMRecordsAsPair).  HBM-bound byte work: length + flag kernel, the exclusive scan
// of hsq_scan.cuh, and a write kernel (one warp per record).
namespace hsq {

__global__ void record_len_kernel(DeviceProbes probes, const int32_t* __restrict__ probe_id, const uint8_t* __restrict__ pair_ok,
                                  const int32_t* __restrict__ read_start, const int32_t* __restrict__ read_stop, const int32_t* __restrict__ pos,
                                  const uint8_t* __restrict__ duplicate, const int64_t* __restrict__ off1, const int64_t* __restrict__ off2,
                                  int64_t n_pairs, int64_t* __restrict__ len, uint16_t* __restrict__ flag, int32_t* __restrict__ mate_pos, int* status) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; e < 2 * n_pairs; e += stride) {
        const int64_t k = e >> 1;
        const int m = (int)(e & 1);
        int64_t ln = 0;
        uint16_t fl = 0;
        int32_t mp = 0;
        if (pair_ok[k]) {
            const int L = (int)(m ? off2[k + 1] - off2[k] : off1[k + 1] - off1[k]);
            const int a = read_start[e], b = read_stop[e];
            if (a < 0 || b < a - 1 || b >= L) atomicMin(status, HSQ_ERR_OUT_OF_RANGE);  // String.substring would throw (:132,:171)
            else ln = b - a + 1;
            const bool one_rev = probes.reverse_strand[probe_id[k]] != 0;
            const bool neg = m ? !one_rev : one_rev;  // :106-107
            // paired + proper pair + first/second (SAMRecordUtil.java:271-276); strand of the read (:257) and of the mate (:283)
            fl = (uint16_t)(0x1 | 0x2 | (m ? 0x80 : 0x40) | (neg ? 0x10 : 0x20) | ((duplicate && duplicate[k]) ? 0x400 : 0));
            mp = pos[e ^ 1];  // setMateAlignmentStart (:281)
        }
        len[e] = ln;
        flag[e] = fl;
        mate_pos[e] = mp;
    }
}

__global__ void record_write_kernel(DeviceProbes probes, const int32_t* __restrict__ probe_id, const int32_t* __restrict__ read_start,
                                    const uint8_t* __restrict__ r1, const uint8_t* __restrict__ q1, const int64_t* __restrict__ off1,
                                    const uint8_t* __restrict__ r2, const uint8_t* __restrict__ q2, const int64_t* __restrict__ off2,
                                    const int64_t* __restrict__ out_off, int64_t n_pairs, uint8_t* __restrict__ out_seq,
                                    uint8_t* __restrict__ out_qual, int* status) {
    const int lane = threadIdx.x & 31;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    bool bad = false;
    for (; w < 2 * n_pairs; w += stride) {
        const int64_t dst = out_off[w];
        const int total = (int)(out_off[w + 1] - dst);
        if (total == 0) continue;
        const int64_t k = w >> 1;
        const int m = (int)(w & 1);
        const bool one_rev = probes.reverse_strand[probe_id[k]] != 0;
        const bool neg = m ? !one_rev : one_rev;
        const uint8_t* src = m ? r2 : r1;
        const uint8_t* srcq = m ? q2 : q1;
        const int64_t s0 = (m ? off2[k] : off1[k]) + read_start[w];
        for (int p = lane; p < total; p += 32) {
            uint8_t code = encode_ascii(src[s0 + p]);
            bad |= (code == CODE_BAD);
            // read two is complemented on entry and complemented back after the slice (:104,:170): identity except U -> A -> T
            if (m) code = complement_code(complement_code(code));
            if (neg) code = complement_code(code);  // getReverseCompliment (:139,:175) + reversed qualities (:140,:176)
            const int o = neg ? total - 1 - p : p;
            out_seq[dst + o] = decode_code(code);
            out_qual[dst + o] = srcq[s0 + p];
        }
    }
    if (bad) atomicMin(status, HSQ_ERR_BAD_CHARACTER);
}

}  // namespace hsq

extern "C" int32_t hsq_extend_records(hsq_ctx* ctx, const hsq_pair_batch* pb, const hsq_extend_out* ext, const uint8_t* duplicate,
                                      hsq_record_out* out, int64_t out_cap) {
    if (!ctx || !pb || !ext || !out || out_cap < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (!ctx->have_probes) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_set_probes has not been called");
    const int64_t n = pb->n;
    if (n < 0 || pb->r1.n != n || pb->r2.n != n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "pair batch arrays disagree on n");
    if (!out->flag || !out->mate_pos || !out->offsets) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null output array");
    if (n > 0 && (!ext->pair_ok || !ext->read_start || !ext->read_stop || !ext->pos)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null extend result array");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = reset_status(ctx);
    if (rc) return rc;
    if ((rc = upload_pairs_raw(ctx, pb))) return rc;
    const size_t nn = (size_t)n, n2 = 2 * nn;
    auto al = [](size_t b) { return (b + 15) & ~(size_t)15; };
    const size_t o_start = 0, o_stop = o_start + al(n2 * 4), o_pos = o_stop + al(n2 * 4), o_ok = o_pos + al(n2 * 4), o_dup = o_ok + al(nn);
    HSQ_CUDA(ctx, ctx->misc.ensure(o_dup + al(nn) + 16));
    uint8_t* in = ctx->misc.as<uint8_t>();
    struct { hsq::DevBuf* d; size_t off; const void* h; size_t bytes; } up[] = {
        {&ctx->q1, 0, pb->q1 ? pb->q1 + pb->r1.offsets[0] : nullptr, (size_t)ctx->r1.total},
        {&ctx->q2, 0, pb->q2 ? pb->q2 + pb->r2.offsets[0] : nullptr, (size_t)ctx->r2.total},
        {nullptr, o_start, ext->read_start, n2 * 4}, {nullptr, o_stop, ext->read_stop, n2 * 4}, {nullptr, o_pos, ext->pos, n2 * 4},
        {nullptr, o_ok, ext->pair_ok, nn}, {nullptr, o_dup, duplicate, duplicate ? nn : 0}};
    for (auto& c : up) {
        if (c.bytes && !c.h) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_extend_records: null input array (qualities are required)");
        if (c.d) HSQ_CUDA(ctx, c.d->ensure(c.bytes + 16));
        uint8_t* dst = c.d ? c.d->as<uint8_t>() : in + c.off;
        if (c.bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(dst, c.h, c.bytes, cudaMemcpyHostToDevice, ctx->stream));
    }
    const int64_t n_tiles = (2 * n + 1 + SCAN_TILE - 1) / SCAN_TILE;
    HSQ_CUDA(ctx, ctx->sort_a.ensure((n2 + 1) * 8));   // record lengths
    HSQ_CUDA(ctx, ctx->sort_b.ensure((n2 + 1) * 8));   // out offsets
    HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)n_tiles * 8 + 16));
    HSQ_CUDA(ctx, ctx->out_i32.ensure(n2 * 4 + 16));   // mate_pos
    HSQ_CUDA(ctx, ctx->out_u8.ensure(n2 * 2 + 16));    // flag
    int64_t* len = ctx->sort_a.as<int64_t>();
    int64_t* d_off = ctx->sort_b.as<int64_t>();
    DeviceProbes pv = probe_view(ctx);
    if (n > 0) {
        const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
        check_probe_ids_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), n, ctx->n_probes, d_status(ctx));
        ctx->launches++;
        int w[1];
        if ((rc = read_status(ctx, w, 1))) return rc;
        if (w[0] != 0) { timer.finish(); return fail(ctx, w[0], "hsq_extend_records: probe_id out of range"); }
        const int blocks = (int)std::min<int64_t>((2 * n + 255) / 256, (int64_t)ctx->sm_count * 16);
        record_len_kernel<<<blocks, 256, 0, ctx->stream>>>(pv, ctx->probe_id.as<int32_t>(), in + o_ok, (const int32_t*)(in + o_start),
                                                          (const int32_t*)(in + o_stop), (const int32_t*)(in + o_pos), duplicate ? in + o_dup : nullptr,
                                                          ctx->r1.offsets.as<int64_t>(), ctx->r2.offsets.as<int64_t>(), n, len,
                                                          ctx->out_u8.as<uint16_t>(), ctx->out_i32.as<int32_t>(), d_status(ctx));
        ctx->launches++;
    }
    scan_tile_sums_kernel<<<(unsigned)n_tiles, SCAN_BLOCK, 0, ctx->stream>>>(len, 2 * n, ctx->sort_c.as<int64_t>());
    scan_sums_kernel<<<1, SCAN_BLOCK, 0, ctx->stream>>>(ctx->sort_c.as<int64_t>(), n_tiles);
    scan_apply_kernel<<<(unsigned)n_tiles, SCAN_BLOCK, 0, ctx->stream>>>(len, 2 * n, ctx->sort_c.as<int64_t>(), d_off);
    ctx->launches += 3;
    HSQ_CUDA(ctx, cudaGetLastError());
    HSQ_CUDA(ctx, cudaMemcpyAsync(out->offsets, d_off, (n2 + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if ((rc = d2h(ctx, out->flag, ctx->out_u8.as<uint16_t>(), n2))) return rc;
    if ((rc = d2h(ctx, out->mate_pos, ctx->out_i32.as<int32_t>(), n2))) return rc;
    int w[1];
    if ((rc = read_status(ctx, w, 1))) return rc;
    if (w[0] != 0) { timer.finish(); return fail(ctx, w[0], "hsq_extend_records: a read range lies outside its read"); }
    const int64_t total = out->offsets[2 * n];
    if (total > out_cap) { timer.finish(); return fail(ctx, HSQ_ERR_OVERFLOW, "hsq_extend_records: output buffers too small; offsets holds the needed sizes"); }
    if (total > 0) {
        if (!out->seq || !out->qual) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null output buffer");
        HSQ_CUDA(ctx, ctx->out_slots.ensure(2 * (size_t)total + 32));
        uint8_t* d_seq = ctx->out_slots.as<uint8_t>();
        uint8_t* d_qual = d_seq + total + 16;
        const int blocks = (int)std::min<int64_t>((2 * n * 32 + 255) / 256, (int64_t)ctx->sm_count * 16);
        record_write_kernel<<<blocks, 256, 0, ctx->stream>>>(pv, ctx->probe_id.as<int32_t>(), (const int32_t*)(in + o_start), ctx->r1.ascii.as<uint8_t>(),
                                                            ctx->q1.as<uint8_t>(), ctx->r1.offsets.as<int64_t>(), ctx->r2.ascii.as<uint8_t>(),
                                                            ctx->q2.as<uint8_t>(), ctx->r2.offsets.as<int64_t>(), d_off, n, d_seq, d_qual, d_status(ctx));
        ctx->launches++;
        HSQ_CUDA(ctx, cudaGetLastError());
        HSQ_CUDA(ctx, cudaMemcpyAsync(out->seq, d_seq, (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(out->qual, d_qual, (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
        if ((rc = read_status(ctx, w, 1))) return rc;
        if (w[0] != 0) { timer.finish(); return fail(ctx, w[0], std::string("hsq_extend_records: ") + status_message(w[0])); }
    }
    return timer.finish();
}
