// hsq_capi_align.inl -- hsq_align_batch: the operator seam (one NeedlemanWunschGlobalAlignment per item).
namespace {

// i32 outputs live in one device block: field f of item k at out_i32[f * n + k]
int run_align(hsq_ctx* ctx, const hsq_seqs* refs, const hsq_seqs* qrys, const hsq_scorer* scorer, uint32_t flags, hsq_align_out* o) {
    const int64_t n = refs->n;
    int rc = reset_status(ctx);
    if (rc) return rc;
    if ((rc = upload_seqs(ctx, refs, &ctx->sa, d_status(ctx), d_maxlen(ctx, 0)))) return rc;
    if ((rc = upload_seqs(ctx, qrys, &ctx->sb, d_status(ctx), d_maxlen(ctx, 1)))) return rc;
    int w[3];
    if ((rc = read_status(ctx, w, 3))) return rc;
    if (w[0] != 0) return fail(ctx, w[0], std::string("hsq_align_batch: ") + status_message(w[0]));
    const int r_max = w[1], q_max = w[2];
    ScoreParams sp = make_score_params(scorer->match, scorer->mismatch, scorer->gap_extend, scorer->gap_start);
    int grid = 1;
    if ((rc = persistent_grid(ctx, align_kernel, DP_BLOCK, n, &grid))) return rc;
    ScratchPool pool;
    if ((rc = alloc_scratch(ctx, (int64_t)grid * DP_BLOCK, 0, q_max, r_max, &pool, nullptr))) return rc;
    const size_t nn = (size_t)n;
    HSQ_CUDA(ctx, ctx->out_i32.ensure(17 * nn * sizeof(int32_t) + 16));
    HSQ_CUDA(ctx, ctx->out_u8.ensure(nn + 16));
    HSQ_CUDA(ctx, ctx->out_slots.ensure(nn * (3 * (size_t)o->cols_cap + 2 * (size_t)o->md_cap) + 16));
    int32_t* i32 = ctx->out_i32.as<int32_t>();
    uint8_t* slots = ctx->out_slots.as<uint8_t>();
    AlignDeviceOut d;
    d.cols_cap = o->cols_cap;
    d.md_cap = o->md_cap;
    d.score = i32; d.n_cols = i32 + nn; d.first_q = i32 + 2 * nn; d.last_q = i32 + 3 * nn; d.first_r = i32 + 4 * nn;
    d.last_r = i32 + 5 * nn; d.matches = i32 + 6 * nn; d.mismatches = i32 + 7 * nn; d.non_terminal_gaps = i32 + 8 * nn;
    d.edit_distance = i32 + 9 * nn; d.md_len = i32 + 10 * nn; d.rmd_len = i32 + 11 * nn;
    d.ref_aln = slots; d.qry_aln = slots + nn * o->cols_cap; d.ops = slots + 2 * nn * o->cols_cap;
    d.md = slots + 3 * nn * o->cols_cap; d.rmd = d.md + nn * o->md_cap;
    d.overflow = ctx->out_u8.as<uint8_t>();
    d.find_counts = o->find_counts ? i32 + 12 * nn : nullptr;
    const int ps = (flags & HSQ_ALIGN_PENALIZE_START) ? 1 : 0, pe = (flags & HSQ_ALIGN_PENALIZE_END) ? 1 : 0;
    if (n > 0) {
        align_kernel<<<grid, DP_BLOCK, 0, ctx->stream>>>(ctx->sa.view(), ctx->sb.view(), n, sp, ps, pe, pool, d, d_status(ctx),
                                                        ctx->cells.as<unsigned long long>());
        ctx->launches++;
        HSQ_CUDA(ctx, cudaGetLastError());
    }
    int32_t* hosts[12] = {o->score, o->n_cols, o->first_q, o->last_q, o->first_r, o->last_r, o->matches, o->mismatches,
                          o->non_terminal_gaps, o->edit_distance, o->md_len, o->rmd_len};
    for (int f = 0; f < 12; f++)
        if ((rc = d2h(ctx, hosts[f], i32 + f * nn, nn))) return rc;
    if ((rc = d2h(ctx, o->ref_aln, d.ref_aln, nn * o->cols_cap))) return rc;
    if ((rc = d2h(ctx, o->qry_aln, d.qry_aln, nn * o->cols_cap))) return rc;
    if ((rc = d2h(ctx, o->ops, d.ops, nn * o->cols_cap))) return rc;
    if ((rc = d2h(ctx, o->md, d.md, nn * o->md_cap))) return rc;
    if ((rc = d2h(ctx, o->rmd, d.rmd, nn * o->md_cap))) return rc;
    if ((rc = d2h(ctx, o->overflow, d.overflow, nn))) return rc;
    if ((rc = d2h(ctx, o->find_counts, i32 + 12 * nn, 5 * nn))) return rc;
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->h_cells, ctx->cells.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    if ((rc = read_status(ctx, w, 1))) return rc;
    o->cells = (int64_t)ctx->h_cells[0];
    if (w[0] != 0) return fail(ctx, w[0], std::string("hsq_align_batch: ") + status_message(w[0]));
    return HSQ_OK;
}

}  // namespace

extern "C" int32_t hsq_align_batch(hsq_ctx* ctx, const hsq_seqs* refs, const hsq_seqs* queries, const hsq_scorer* scorer,
                                   uint32_t flags, hsq_align_out* out) {
    if (!ctx || !refs || !queries || !scorer || !out) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (refs->n != queries->n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "refs and queries differ in n");
    if (out->cols_cap < 0 || out->md_cap < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative slot capacity");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = run_align(ctx, refs, queries, scorer, flags, out);
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}
