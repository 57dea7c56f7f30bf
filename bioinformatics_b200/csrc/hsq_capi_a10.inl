// hsq_capi_a10.inl -- SURVEY 8(a) row A10, first half: SAMRecordUtil.getVariableLengthUid (HSQ/utils/SAMRecordUtil.java:223-261), dead in
// the CLI (allowVariableLengthUids = false, HSQ/cli/DeduplicationCli.java:120).  One NeedlemanWunschGlobalAlignment(read, primer, scorer)
// per item; the UID is the read prefix before the first column that holds a primer base (getIndexOfFirstMatchInReference reads the QUERY
// alignment, NeedlemanWunschGlobalAlignment.java:412-428); it is dropped when insertions + deletions + substitutions >= primer length / 4.
// The reference never increments insertsSinceLastInsertionOrMismatch (:238-250), so the deletion count includes the read bases around the
// primer: every column that is not a match counts, edit = columns - matches.  Kept as it is.
// (Second half of A10, mergePairs = mergeSequences + an alignment of the merged read to the capture target, is hsq_merge_pairs followed
// by hsq_align_batch: tests/test_a10.py.)
namespace hsq {

struct VarUidSink {  // columns arrive last first
    static constexpr bool HAS_FAST = false;
    int n = 0, matches = 0, last_q = -1;
    __device__ inline void push(uint32_t c) {
        const uint32_t rc = c & 0xFFu, qc = c >> 8;
        if (qc != CODE_GAP) last_q = n;
        if ((rc & qc & MATCH_MASK) != 0) matches++;  // '=' column (a gap never meets a gap in an alignment column)
        n++;
    }
};

__global__ void __launch_bounds__(DP_BLOCK) varuid_kernel(DeviceSeqs reads, DeviceSeqs primers, int64_t n, ScoreParams sp, int pen_start, int pen_end,
                                                         ScratchPool pool, int32_t* __restrict__ uid_len, unsigned long long* cells_total) {
    __shared__ int4 prof[PROF_INT4_PER_THREAD * DP_BLOCK];
    DpScratch sc = thread_scratch(pool, prof);
    int64_t cells = 0;
    for (int64_t k = sc.tid; k < n; k += sc.nthreads) {
        const int64_t r0 = reads.offsets[k], q0 = primers.offsets[k];
        SeqView ref = make_view(reads.codes + r0, (int)(reads.offsets[k + 1] - r0), false, false);
        SeqView qry = make_view(primers.codes + q0, (int)(primers.offsets[k + 1] - q0), false, false);
        dp_fill(ref, qry, sp, pen_start, pen_end, sc);
        cells += (int64_t)ref.len * qry.len;
        VarUidSink s;
        dp_traceback_stream(ref, qry, sc, s);
        int end = s.last_q >= 0 ? s.n - 1 - s.last_q : -1;  // first forward column with a primer base
        if (s.n - s.matches >= qry.len / 4) end = -1;
        uid_len[k] = end;
    }
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
}

}  // namespace hsq

extern "C" int32_t hsq_variable_length_uid(hsq_ctx* ctx, const hsq_seqs* reads, const hsq_seqs* primers, const hsq_scorer* scorer, uint32_t flags,
                                           int32_t* uid_len) {
    if (!ctx || !reads || !primers || !scorer) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (reads->n != primers->n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "reads and primers differ in n");
    const int64_t n = reads->n;
    if (n > 0 && !uid_len) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null output");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = reset_status(ctx);
    if (rc) return rc;
    if ((rc = upload_seqs(ctx, reads, &ctx->sa, d_status(ctx), d_maxlen(ctx, 0)))) return rc;
    if ((rc = upload_seqs(ctx, primers, &ctx->sb, d_status(ctx), d_maxlen(ctx, 1)))) return rc;
    int w[3];
    if ((rc = read_status(ctx, w, 3))) return rc;
    if (w[0] != 0) return fail(ctx, w[0], std::string("hsq_variable_length_uid: ") + status_message(w[0]));
    ScoreParams sp = make_score_params(scorer->match, scorer->mismatch, scorer->gap_extend, scorer->gap_start);
    int grid = 1;
    if ((rc = persistent_grid(ctx, varuid_kernel, DP_BLOCK, n, &grid))) return rc;
    ScratchPool pool;
    if ((rc = alloc_scratch(ctx, (int64_t)grid * DP_BLOCK, 0, w[2], w[1], &pool, nullptr))) return rc;
    HSQ_CUDA(ctx, ctx->out_i32.ensure((size_t)n * 4 + 16));
    if (n > 0) {
        varuid_kernel<<<grid, DP_BLOCK, 0, ctx->stream>>>(ctx->sa.view(), ctx->sb.view(), n, sp, (flags & HSQ_ALIGN_PENALIZE_START) ? 1 : 0,
                                                         (flags & HSQ_ALIGN_PENALIZE_END) ? 1 : 0, pool, ctx->out_i32.as<int32_t>(),
                                                         ctx->cells.as<unsigned long long>());
        ctx->launches++;
        HSQ_CUDA(ctx, cudaGetLastError());
        if ((rc = d2h(ctx, uid_len, ctx->out_i32.as<int32_t>(), (size_t)n))) return rc;
    }
    return timer.finish();
}
