// hsq_capi_trim.inl -- hsq_trim_batch: FastqReadTrimmer.trim (HSQ/process/FastqReadTrimmer.java:309-328) per read.
#include "hsq_scan.cuh"
namespace hsq {

// kept length per read; flags the reads for which Java throws (first >= length, or substring end < begin)
__global__ void trim_len_kernel(const int64_t* __restrict__ offsets, int64_t n, int first, int last, int three_prime,
                                int64_t* __restrict__ keep, int* status) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) {
        int len = (int)(offsets[k + 1] - offsets[k]);
        int lb = len - 1;
        if (three_prime) lb = min(last, len - 1);  // :319-322
        int kp = lb + 1 - first;
        if (first >= len || kp < 0) {  // :310-312 IllegalArgumentException / String.substring bounds
            atomicMin(status, HSQ_ERR_OUT_OF_RANGE);
            kp = 0;
        }
        keep[k] = kp;
    }
}

// 16 bytes of src starting at byte address s (any alignment) as one uint4: the two aligned 16-byte vectors around it, realigned
// with funnel shifts.  The word offset (s & 12) is the same for all lanes of a warp here, so the switch does not diverge.
__device__ __forceinline__ uint4 load16_unaligned(const uint8_t* s) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(s);
    const uint4* vp = reinterpret_cast<const uint4*>(a & ~(uintptr_t)15);
    const uint4 lo = __ldg(vp);
    if ((a & 15) == 0) return lo;
    const uint4 hi = __ldg(vp + 1);
    const uint32_t sh = (uint32_t)(a & 3) * 8u;
    switch ((a >> 2) & 3) {
        case 0: return make_uint4(__funnelshift_r(lo.x, lo.y, sh), __funnelshift_r(lo.y, lo.z, sh), __funnelshift_r(lo.z, lo.w, sh), __funnelshift_r(lo.w, hi.x, sh));
        case 1: return make_uint4(__funnelshift_r(lo.y, lo.z, sh), __funnelshift_r(lo.z, lo.w, sh), __funnelshift_r(lo.w, hi.x, sh), __funnelshift_r(hi.x, hi.y, sh));
        case 2: return make_uint4(__funnelshift_r(lo.z, lo.w, sh), __funnelshift_r(lo.w, hi.x, sh), __funnelshift_r(hi.x, hi.y, sh), __funnelshift_r(hi.y, hi.z, sh));
        default: return make_uint4(__funnelshift_r(lo.w, hi.x, sh), __funnelshift_r(hi.x, hi.y, sh), __funnelshift_r(hi.y, hi.z, sh), __funnelshift_r(hi.z, hi.w, sh));
    }
}

// One warp per read: lanes 0-15 copy the bases, lanes 16-31 the qualities of [first, first+keep) to the packed outputs.  The 16-byte
// aligned chunks of the destination that lie inside the read are one 16-byte store each, assembled from the two aligned source
// vectors around them (one chunk per lane and round; a 150-base read takes one round); the up to 15 bytes before the first and
// after the last such chunk are one byte per lane.
// (Up to r2_c every lane moved one byte per step: 20 load / store instructions per read.)
__global__ void trim_copy_kernel(const uint8_t* __restrict__ seq, const uint8_t* __restrict__ qual, const int64_t* __restrict__ offsets,
                                 const int64_t* __restrict__ out_offsets, int64_t n, int first, uint8_t* __restrict__ out_seq,
                                 uint8_t* __restrict__ out_qual) {
    const int lane = threadIdx.x & 31, half = lane >> 4, sub = lane & 15;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const uint8_t* __restrict__ in = half ? qual : seq;
    uint8_t* __restrict__ out = half ? out_qual : out_seq;
    for (; w < n; w += stride) {
        const int64_t src = offsets[w] + first, dst = out_offsets[w];
        const int64_t end = out_offsets[w + 1];
        // destination chunks are aligned in the ADDRESS space of `out` (both output arrays sit at 16-byte aligned addresses)
        const int64_t body_lo = (dst + 15) & ~(int64_t)15, body_hi = end & ~(int64_t)15;
        if (body_lo <= body_hi) {
            for (int64_t c = body_lo + 16 * sub; c < body_hi; c += 256)
                *reinterpret_cast<uint4*>(out + c) = load16_unaligned(in + src + (c - dst));
            const int64_t h = dst + sub, t = body_hi + sub;
            if (h < body_lo) out[h] = in[src + sub];
            if (t < end) out[t] = in[src + (t - dst)];
        } else {  // fewer than 16 bytes, none of them on a chunk boundary
            const int64_t h = dst + sub;
            if (h < end) out[h] = in[src + sub];
        }
    }
}

}  // namespace hsq

extern "C" int32_t hsq_trim_batch(hsq_ctx* ctx, const hsq_seqs* reads, const uint8_t* quals, int32_t first, int32_t last,
                                  int32_t three_prime, uint8_t* out_seq, uint8_t* out_qual, int64_t* out_offsets) {
    if (!ctx || !reads || !reads->offsets || !out_offsets) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (first < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "first base to keep must be >= 0");  // FastqReadTrimmer.java:263-265
    const int64_t n = reads->n;
    const int64_t total = reads->offsets[n];
    if (total > 0 && (!reads->data || !quals || !out_seq || !out_qual)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null buffer");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = reset_status(ctx);
    if (rc) return rc;
    // raw bytes only: trim never decodes bases
    HSQ_CUDA(ctx, ctx->r1.ascii.ensure((size_t)total + 32));  // load16_unaligned reads whole words around a chunk
    HSQ_CUDA(ctx, ctx->q1.ensure((size_t)total + 32));
    HSQ_CUDA(ctx, ctx->r1.offsets.ensure((size_t)(n + 1) * 8));
    HSQ_CUDA(ctx, ctx->sort_a.ensure((size_t)(n + 1) * 8));  // keep lengths
    HSQ_CUDA(ctx, ctx->sort_b.ensure((size_t)(n + 1) * 8));  // out offsets
    const int64_t n_tiles = (n + 1 + SCAN_TILE - 1) / SCAN_TILE;
    HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)n_tiles * 8 + 16));
    HSQ_CUDA(ctx, ctx->out_slots.ensure(2 * (size_t)total + 64));
    // Phase 1: the offsets alone give every kept length and output offset (8 B in, 8 B out per read against 300 B of bases and qualities).
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->r1.offsets.p, reads->offsets, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
    int64_t* keep = ctx->sort_a.as<int64_t>();
    int64_t* d_out_off = ctx->sort_b.as<int64_t>();
    uint8_t* d_out_seq = ctx->out_slots.as<uint8_t>();
    uint8_t* d_out_qual = d_out_seq + ((total + 31) & ~(int64_t)15);  // 16-byte aligned like d_out_seq
    span_begin(ctx, 0);
    if (n > 0) {
        int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
        trim_len_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->r1.offsets.as<int64_t>(), n, first, last, three_prime, keep, d_status(ctx));
        ctx->launches++;
    }
    scan_tile_sums_kernel<<<(unsigned)n_tiles, SCAN_BLOCK, 0, ctx->stream>>>(keep, n, ctx->sort_c.as<int64_t>());
    scan_sums_kernel<<<1, SCAN_BLOCK, 0, ctx->stream>>>(ctx->sort_c.as<int64_t>(), n_tiles);
    scan_apply_kernel<<<(unsigned)n_tiles, SCAN_BLOCK, 0, ctx->stream>>>(keep, n, ctx->sort_c.as<int64_t>(), d_out_off);
    ctx->launches += 3;
    span_end(ctx, 0);
    HSQ_CUDA(ctx, cudaGetLastError());
    HSQ_CUDA(ctx, cudaMemcpyAsync(out_offsets, d_out_off, (size_t)(n + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    int w[1];
    if ((rc = read_status(ctx, w, 1))) return rc;
    if (w[0] != 0) {
        timer.finish();
        return fail(ctx, w[0], "hsq_trim_batch: unable to trim: first base to keep is beyond the end of a read");
    }
    // Phase 2: the bytes, in chunks of reads: H2D of chunk c + 1 | copy kernel of chunk c | D2H of chunk c - 1 on three streams (the two
    // PCIe directions run at the same time; with one chunk this is copy in, kernel, copy out as before).
    int n_chunks = ctx->cfg.trim_chunks > 0 ? ctx->cfg.trim_chunks : (int)std::min<int64_t>(8, std::max<int64_t>(1, total / (48ll << 20)));
    n_chunks = (int)std::max<int64_t>(1, std::min<int64_t>(n_chunks, std::min<int64_t>(n, ((int64_t)ctx->trim_events.size() - 1) / 2)));
    if (n_chunks > 1) HSQ_CUDA(ctx, cudaEventRecord(ctx->trim_events[0], ctx->stream));  // the copy streams start after phase 1
    for (int c = 0; c < n_chunks && n > 0; c++) {
        const int64_t a = n * c / n_chunks, b = n * (c + 1) / n_chunks;
        const int64_t in0 = reads->offsets[a], in1 = reads->offsets[b], o0 = out_offsets[a], o1 = out_offsets[b];
        cudaStream_t sin = n_chunks > 1 ? ctx->s_in : ctx->stream, sout = n_chunks > 1 ? ctx->s_out : ctx->stream;
        cudaEvent_t ev_in = ctx->trim_events[(size_t)(1 + 2 * c)], ev_k = ctx->trim_events[(size_t)(2 + 2 * c)];
        if (n_chunks > 1 && c == 0) {
            HSQ_CUDA(ctx, cudaStreamWaitEvent(sin, ctx->trim_events[0], 0));
            HSQ_CUDA(ctx, cudaStreamWaitEvent(sout, ctx->trim_events[0], 0));
        }
        if (in1 > in0) {
            HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->r1.ascii.as<uint8_t>() + in0, reads->data + in0, (size_t)(in1 - in0), cudaMemcpyHostToDevice, sin));
            HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->q1.as<uint8_t>() + in0, quals + in0, (size_t)(in1 - in0), cudaMemcpyHostToDevice, sin));
        }
        if (n_chunks > 1) {
            HSQ_CUDA(ctx, cudaEventRecord(ev_in, sin));
            HSQ_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ev_in, 0));
        }
        {
            const int64_t m = b - a;
            if (n_chunks == 1) span_begin(ctx, 1);  // one chunk: the copy kernel alone (the stream reaches this after the copies in)
            int blocks = (int)std::min<int64_t>((m * 32 + 255) / 256, (int64_t)ctx->sm_count * 16);
            trim_copy_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->r1.ascii.as<uint8_t>(), ctx->q1.as<uint8_t>(), ctx->r1.offsets.as<int64_t>() + a,
                                                             d_out_off + a, m, first, d_out_seq, d_out_qual);
            ctx->launches++;
            if (n_chunks == 1) span_end(ctx, 1);
        }
        if (n_chunks > 1) {
            HSQ_CUDA(ctx, cudaEventRecord(ev_k, ctx->stream));
            HSQ_CUDA(ctx, cudaStreamWaitEvent(sout, ev_k, 0));
        }
        if (o1 > o0) {
            HSQ_CUDA(ctx, cudaMemcpyAsync(out_seq + o0, d_out_seq + o0, (size_t)(o1 - o0), cudaMemcpyDeviceToHost, sout));
            HSQ_CUDA(ctx, cudaMemcpyAsync(out_qual + o0, d_out_qual + o0, (size_t)(o1 - o0), cudaMemcpyDeviceToHost, sout));
        }
    }
    // (with more than one chunk the copy kernels wait for copies in: hsq_last_spans then holds the phase-1 kernels only)
    HSQ_CUDA(ctx, cudaGetLastError());
    if (n_chunks > 1) {  // the call's stream ends after the last copy back
        HSQ_CUDA(ctx, cudaEventRecord(ctx->trim_events[0], ctx->s_out));
        HSQ_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->trim_events[0], 0));
    }
    return timer.finish();
}
