// hsq_capi_resolve.inl -- hsq_resolve_multi_probe: the tie-break for reads whose coordinates fall inside >1 probe
// (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:704-765).
namespace hsq {

// one thread per (candidate, mate): NW(reference = read minus UID, query = probe sequence | its reverse complement),
// default scorer (:722,:731); metrics of :723-734
__global__ void __launch_bounds__(DP_BLOCK) tie_break_kernel(DeviceSeqs r1, DeviceSeqs r2, DeviceSeqs probe_full, const int64_t* __restrict__ cand_offsets,
                                                        const int32_t* __restrict__ cand_probe, int64_t n_reads, int64_t n_cand, int u, int v,
                                                        ScoreParams sp, ScratchPool pool, int32_t* __restrict__ matches,
                                                        int32_t* __restrict__ edit, unsigned long long* cells_total) {
    __shared__ int4 prof[PROF_INT4_PER_THREAD * DP_BLOCK];
    DpScratch sc = thread_scratch(pool, prof);
    int64_t cells = 0;
    for (int64_t t = sc.tid; t < 2 * n_cand; t += sc.nthreads) {
        const int64_t j = t >> 1;
        const int mate = (int)(t & 1);
        int64_t lo = 0, hi = n_reads;  // read k with cand_offsets[k] <= j < cand_offsets[k+1]
        while (hi - lo > 1) {
            int64_t mid = (lo + hi) >> 1;
            if (cand_offsets[mid] <= j) lo = mid; else hi = mid;
        }
        const int64_t k = lo;
        const int pid = cand_probe[j];
        const int64_t p0 = probe_full.offsets[pid];
        const int pl = (int)(probe_full.offsets[pid + 1] - p0);
        const DeviceSeqs& rs = mate ? r2 : r1;
        const int skip = mate ? v : u;
        const int64_t s0 = rs.offsets[k];
        const int rl = (int)(rs.offsets[k + 1] - s0);
        SeqView read = make_view(rs.codes + s0, rl, false, false);
        SeqView ref = sub_view(read, skip, max(0, rl - skip));           // subSequence(uidLength) (:718,:729)
        SeqView qry = make_view(probe_full.codes + p0, pl, mate != 0, mate != 0);  // getReverseCompliment for read two (:731)
        dp_fill(ref, qry, sp, 0, 0, sc);
        cells += (int64_t)ref.len * qry.len;
        Columns cv{sc.colstore(), dp_traceback(ref, qry, sc)};
        PairIndices ix = pair_indices(cv);
        int m, x, rg, qg;
        count_columns(cv, 0, cv.n - 1, &m, &x, &rg, &qg);
        matches[t] = m;
        edit[t] = ix.first_q + x + non_terminal_gaps(cv, ix);  // getFirstNonInsertQueryMatchInReference + mismatches + non-terminal gaps
    }
    for (int o = 16; o > 0; o >>= 1) cells += __shfl_xor_sync(0xffffffffu, cells, o);
    if ((threadIdx.x & 31) == 0 && cells) atomicAdd(cells_total, (unsigned long long)cells);
}

// per read: most matches, then lowest edit distance; more than one best probe -> -1 (:735-765)
__global__ void tie_break_select_kernel(const int64_t* __restrict__ cand_offsets, const int32_t* __restrict__ cand_probe, int64_t n_reads,
                                        const int32_t* __restrict__ matches2, const int32_t* __restrict__ edit2,
                                        int32_t* __restrict__ best_probe, int32_t* __restrict__ cand_matches, int32_t* __restrict__ cand_edit) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_reads) return;
    int lowest = INT32_MAX, most = 0, n_best = 0, best = -1;
    for (int64_t j = cand_offsets[k]; j < cand_offsets[k + 1]; j++) {
        const int m = matches2[2 * j] + matches2[2 * j + 1];
        const int e = edit2[2 * j] + edit2[2 * j + 1];
        cand_matches[j] = m;
        cand_edit[j] = e;
        if (m > most) { n_best = 1; best = cand_probe[j]; lowest = e; most = m; }
        else if (m == most) {
            if (e == lowest) n_best++;
            else if (e < lowest) { n_best = 1; best = cand_probe[j]; lowest = e; most = m; }
        }
    }
    best_probe[k] = (n_best == 1) ? best : -1;
}

}  // namespace hsq

extern "C" int32_t hsq_resolve_multi_probe(hsq_ctx* ctx, const hsq_candidates* c, int32_t u, int32_t v, int32_t* best_probe,
                                           int32_t* cand_matches, int32_t* cand_edit) {
    if (!ctx || !c || !best_probe || !c->cand_offsets) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (!ctx->have_probes) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_set_probes has not been called");
    if (c->r1.n != c->n || c->r2.n != c->n || u < 0 || v < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "bad candidate batch");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    const int64_t n = c->n;
    const int64_t nc = c->cand_offsets[n];
    if (nc > 0 && !c->cand_probe) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "cand_probe is null");
    for (int64_t j = 0; j < nc; j++)
        if (c->cand_probe[j] < 0 || c->cand_probe[j] >= ctx->n_probes) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "candidate probe out of range");
    int rc = reset_status(ctx);
    if (rc) return rc;
    if ((rc = upload_seqs(ctx, &c->r1, &ctx->r1, d_status(ctx), d_maxlen(ctx, 0)))) return rc;
    if ((rc = upload_seqs(ctx, &c->r2, &ctx->r2, d_status(ctx), d_maxlen(ctx, 1)))) return rc;
    int w[3];
    if ((rc = read_status(ctx, w, 3))) return rc;
    if (w[0] != 0) return fail(ctx, w[0], std::string("hsq_resolve_multi_probe: ") + status_message(w[0]));
    HSQ_CUDA(ctx, ctx->sort_a.ensure((size_t)(n + 1) * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_b.ensure((size_t)nc * 4 + 16));
    HSQ_CUDA(ctx, ctx->out_i32.ensure(((size_t)nc * 6 + (size_t)n) * 4 + 16));
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->sort_a.p, c->cand_offsets, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
    if (nc) HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->sort_b.p, c->cand_probe, (size_t)nc * 4, cudaMemcpyHostToDevice, ctx->stream));
    int32_t* m2 = ctx->out_i32.as<int32_t>();
    int32_t* e2 = m2 + 2 * nc;
    int32_t* cm = e2 + 2 * nc;
    int32_t* ce = cm + nc;
    int32_t* bp = ce + nc;
    ScoreParams sp = make_score_params(1, -4, -1, -6);  // new NeedlemanWunschGlobalAlignment(ref, query): default scorer
    if (nc > 0) {
        int grid = 1;
        if ((rc = persistent_grid(ctx, tie_break_kernel, DP_BLOCK, 2 * nc, &grid))) return rc;
        ScratchPool pool;
        if ((rc = alloc_scratch(ctx, (int64_t)grid * DP_BLOCK, 0, ctx->p_full.max_len, std::max(w[1], w[2]), &pool, nullptr))) return rc;
        tie_break_kernel<<<grid, DP_BLOCK, 0, ctx->stream>>>(ctx->r1.view(), ctx->r2.view(), ctx->p_full.view(), ctx->sort_a.as<int64_t>(),
                                                            ctx->sort_b.as<int32_t>(), n, nc, u, v, sp, pool, m2, e2,
                                                            ctx->cells.as<unsigned long long>());
        ctx->launches++;
    }
    if (n > 0) {
        tie_break_select_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(ctx->sort_a.as<int64_t>(), ctx->sort_b.as<int32_t>(), n, m2, e2,
                                                                                      bp, cm, ce);
        ctx->launches++;
    }
    HSQ_CUDA(ctx, cudaGetLastError());
    if ((rc = d2h(ctx, best_probe, bp, (size_t)n))) return rc;
    if ((rc = d2h(ctx, cand_matches, cm, (size_t)nc))) return rc;
    if ((rc = d2h(ctx, cand_edit, ce, (size_t)nc))) return rc;
    return timer.finish();
}
