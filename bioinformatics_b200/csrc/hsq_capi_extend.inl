// hsq_capi_extend.inl -- hsq_extend_batch: ExtendReadsToPrimer.extendReadsToPrimers for a batch spanning probes.
namespace {

__global__ void check_probe_ids_kernel(const int32_t* __restrict__ probe_id, int64_t n, int64_t n_probes, int* status) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride)
        if (probe_id[k] < 0 || probe_id[k] >= n_probes) atomicMin(status, HSQ_ERR_INVALID_ARGUMENT);
}

// ASCII -> codes, or packed words -> codes (hsq_capi_packed.inl), or nothing when the codes of the staged batch are already valid
int prepare_codes(hsq_ctx* ctx, DevSeqsOwned* d, int* d_maxlen);

int upload_pairs_raw(hsq_ctx* ctx, const hsq_pair_batch* pb) {
    int rc;
    const size_t n = (size_t)pb->n;
    if (pb->r1.n != pb->n || pb->r2.n != pb->n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "pair batch arrays disagree on n");
    if (n && !pb->probe_id) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "probe_id is null");
    HSQ_CUDA(ctx, ctx->probe_id.ensure(n * 4 + 16));
    if (n) HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->probe_id.p, pb->probe_id, n * 4, cudaMemcpyHostToDevice, ctx->io ? ctx->io : ctx->stream));
    if ((rc = copy_seqs(ctx, &pb->r1, &ctx->r1))) return rc;
    if ((rc = copy_seqs(ctx, &pb->r2, &ctx->r2))) return rc;
    return HSQ_OK;
}

// stage: host -> device copies of one pair batch (raw ASCII reads, probe ids)
int extend_stage(hsq_ctx* ctx, const hsq_pair_batch* pb) {
    int rc;
    ctx->ext_n = -1;
    if ((rc = upload_pairs_raw(ctx, pb))) return rc;
    ctx->ext_n = pb->n;
    return HSQ_OK;
}

// What the capture phase of a staged batch needs from its prep phase (raw device pointers of the buffer set that was active
// during prep: the capture phase may run while another set is the active one).
struct ExtendPlan {
    int64_t n = 0;
    bool packed2 = false;
    int grid2 = 1;
    DeviceProbes pv;
    const int32_t* pid = nullptr;
    DeviceSeqs v1, v2;
    const uint32_t* order = nullptr;
    const uint32_t* n_active = nullptr;
    ScoreParams sp_user;
    ScoreParams2 sp_user2;
    ScratchPool pool;
    ScratchPool2 pool2;
    ExtendDeviceOut d;
    uint8_t* pair_ok = nullptr;
    unsigned long long* cells = nullptr;
    unsigned long long* work = nullptr;
    // banded capture stage (hsq_config.band > 0 and the packed kernel runs): what the second, full pass needs to sort again
    int band = 0;
    BucketKeys keys{};
    uint32_t* hist = nullptr;
    uint32_t* cursor = nullptr;
    uint32_t* n_active_w = nullptr;
    uint32_t* order_w = nullptr;
    int n_keys = 0;
    size_t order_bytes = 0;
    int sblocks = 1;
};

// counting sort of the 2n keys in keys.key into `order` (buckets padded to even sizes); *n_active = entries before the skip bucket
int bucket_sort_launch(hsq_ctx* ctx, const BucketKeys& keys, int64_t n2, int nk, uint32_t* hist, uint32_t* cursor, uint32_t* n_active,
                       uint32_t* order, size_t order_bytes, int sblocks) {
    HSQ_CUDA(ctx, cudaMemsetAsync(hist, 0, (size_t)nk * 4, ctx->stream));
    HSQ_CUDA(ctx, cudaMemsetAsync(order, 0xFF, order_bytes, ctx->stream));
    bucket_hist_kernel<<<sblocks, 256, 0, ctx->stream>>>(keys.key, n2, hist);
    bucket_scan_kernel<<<1, 1024, 0, ctx->stream>>>(hist, nk, cursor);
    HSQ_CUDA(ctx, cudaMemcpyAsync(n_active, cursor + keys.skip_key, 4, cudaMemcpyDeviceToDevice, ctx->stream));
    bucket_scatter_kernel<<<sblocks, 256, 0, ctx->stream>>>(keys.key, n2, cursor, order);
    ctx->launches += 3;
    return HSQ_OK;
}

// prep: encode + primer DP + the two bucket sorts of the staged batch, on ctx->stream.
// host_read_max < 0: the longest read and the probe-id check come from the device (one synchronisation);
// host_read_max >= 0: the caller measured the reads and validated the probe ids on the host, nothing here blocks.
int extend_prep(hsq_ctx* ctx, int cigar_cap, int md_cap, int host_read_max, ExtendPlan* plan) {
    const int64_t n = ctx->ext_n;
    if (n < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "no staged batch");
    const size_t n2 = (size_t)(2 * n);
    int rc;
    int read_max = host_read_max;
    if (host_read_max < 0) {
        if ((rc = reset_status(ctx))) return rc;
        if ((rc = prepare_codes(ctx, &ctx->r1, d_maxlen(ctx, 0)))) return rc;
        if ((rc = prepare_codes(ctx, &ctx->r2, d_maxlen(ctx, 1)))) return rc;
        if (n > 0) {
            int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
            check_probe_ids_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), n, ctx->n_probes, d_status(ctx));
            ctx->launches++;
        }
        int w[3];
        if ((rc = read_status(ctx, w, 3))) return rc;
        if (w[0] != 0) return fail(ctx, w[0], std::string("hsq_extend: ") + (w[0] == HSQ_ERR_INVALID_ARGUMENT ? "probe_id out of range" : status_message(w[0])));
        read_max = std::max(w[1], w[2]);
    } else {
        if ((rc = prepare_codes(ctx, &ctx->r1, nullptr))) return rc;
        if ((rc = prepare_codes(ctx, &ctx->r2, nullptr))) return rc;
    }
    const int primer_max = std::max(ctx->p_ext.max_len, ctx->p_lig.max_len);
    if (2 * primer_max + PRIMER_ALIGNMENT_BUFFER + 2 > POW95_N) return fail(ctx, HSQ_ERR_UNSUPPORTED, "primer longer than the 0.95^d table supports");
    // DP#1: rows = primer, cols = read prefix (<= primer+26); DP#2: rows = read remainder, cols = capture target
    const int q_max = std::max(primer_max, read_max);
    const int r_max = std::max(ctx->p_target.max_len, std::min(read_max, primer_max + PRIMER_ALIGNMENT_BUFFER + 1));
    ScoreParams sp_default = make_score_params(1, -4, -1, -6);  // SimpleAlignmentScorer defaults (ExtendReadsToPrimer.java:290-291)
    ScoreParams sp_user = make_score_params(ctx->scorer.match, ctx->scorer.mismatch, ctx->scorer.gap_extend, ctx->scorer.gap_start);
    // packed u16x2 cells when every score fits 14 bits (always, for HSQUtils' scorers at these lengths), else 32-bit cells
    const int prefix_max = std::min(read_max, primer_max + PRIMER_ALIGNMENT_BUFFER + 1);
    const bool packed1 = !ctx->cfg.force_dp32 && dp2_score_range_ok(1, -4, -1, -6, primer_max, prefix_max);
    const bool packed2 = !ctx->cfg.force_dp32 &&
                         dp2_score_range_ok(ctx->scorer.match, ctx->scorer.mismatch, ctx->scorer.gap_extend, ctx->scorer.gap_start, q_max, r_max);
    ScoreParams2 sp_default2 = make_score_params2(sp_default);
    ScoreParams2 sp_user2 = make_score_params2(sp_user);
    int grid1 = 1, grid2 = 1;
    if (packed1) rc = persistent_grid(ctx, primer2_kernel, DP2_BLOCK, n + 1, &grid1);
    else rc = persistent_grid(ctx, primer_kernel, DP_BLOCK, 2 * n, &grid1);
    if (rc) return rc;
    if (packed2) rc = persistent_grid(ctx, capture2_kernel, DP2_BLOCK, n + 1, &grid2);
    else rc = persistent_grid(ctx, capture_kernel, DP_BLOCK, 2 * n, &grid2);
    if (rc) return rc;
    ScratchPool pool;
    ScratchPool2 pool2;
    const int64_t threads32 = (int64_t)std::max(packed1 ? 0 : grid1, packed2 ? 0 : grid2) * DP_BLOCK;
    const int64_t threads16 = (int64_t)std::max(packed1 ? grid1 : 0, packed2 ? grid2 : 0) * DP2_BLOCK;
    if ((rc = alloc_scratch(ctx, threads32, threads16, q_max, r_max, &pool, &pool2))) return rc;
    // Bucket keys = exact DP shape (rows, columns); the last key of the DP#2 sort collects the read ends that stop after
    // DP#1.  Buckets are padded to even sizes (holes = ORDER_HOLE) so a packed thread's two read ends share one shape.
    BucketKeys keys1, keys;
    keys1.key_stride = prefix_max + 1;
    const int n_keys1 = (primer_max + 1) * keys1.key_stride + 1;
    keys1.skip_key = (uint32_t)(n_keys1 - 1);
    keys.key_stride = ctx->p_target.max_len + 1;
    const int n_keys = (read_max + 1) * keys.key_stride + 1;
    keys.skip_key = (uint32_t)(n_keys - 1);
    const int n_keys_max = std::max(n_keys, n_keys1);
    const size_t order_bytes = (n2 + (size_t)n_keys_max + 2) * 4;
    HSQ_CUDA(ctx, ctx->sort_a.ensure(n2 * 4 + 16));
    HSQ_CUDA(ctx, ctx->sort_b.ensure(order_bytes + 16));
    HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)(2 * n_keys_max + 4) * 4));
    keys.key = keys1.key = ctx->sort_a.as<uint32_t>();
    uint32_t* order = ctx->sort_b.as<uint32_t>();
    uint32_t* hist = ctx->sort_c.as<uint32_t>();
    uint32_t* cursor = hist + n_keys_max;
    uint32_t* n_active = cursor + n_keys_max;
    HSQ_CUDA(ctx, ctx->out_i32.ensure(10 * n2 * sizeof(int32_t) + 16));
    HSQ_CUDA(ctx, ctx->out_u8.ensure(2 * n2 + 16));
    HSQ_CUDA(ctx, ctx->out_slots.ensure(n2 * ((size_t)cigar_cap + (size_t)md_cap) + 16));
    HSQ_CUDA(ctx, ctx->pair_ok.ensure((size_t)n + 16));
    int32_t* i32 = ctx->out_i32.as<int32_t>();
    ExtendDeviceOut& d = ctx->ext_out;
    d.cigar_cap = cigar_cap;
    d.md_cap = md_cap;
    d.primer_end = i32; d.score = i32 + n2; d.pos = i32 + 2 * n2; d.read_start = i32 + 3 * n2; d.read_stop = i32 + 4 * n2;
    d.nm = i32 + 5 * n2; d.tlen = i32 + 6 * n2; d.cigar_len = i32 + 7 * n2; d.md_len = i32 + 8 * n2; d.ref_len = i32 + 9 * n2;
    d.ok = ctx->out_u8.as<uint8_t>();
    d.overflow = d.ok + n2;
    d.cigar = ctx->out_slots.as<uint8_t>();
    d.md = d.cigar + n2 * (size_t)cigar_cap;
    if (n > 0) {
        DeviceProbes pv = probe_view(ctx);
        const int32_t* pid = ctx->probe_id.as<int32_t>();
        const DeviceSeqs v1 = ctx->r1.view(), v2 = ctx->r2.view();
        unsigned long long* cells = ctx->cells.as<unsigned long long>();
        unsigned long long* work = cells + 1 + 2 * ctx->work_slot;
        const int sblocks = (int)std::min<int64_t>((2 * n + 255) / 256, (int64_t)ctx->sm_count * 8);
        auto bucket_sort = [&](int nk, uint32_t skip_key) -> int {
            BucketKeys k = keys;
            k.skip_key = skip_key;
            return bucket_sort_launch(ctx, k, 2 * n, nk, hist, cursor, n_active, order, order_bytes, sblocks);
        };
        if (packed1) {
            primer_key_kernel<<<sblocks, 256, 0, ctx->stream>>>(pv, pid, v1, v2, n, keys1);
            ctx->launches++;
            if ((rc = bucket_sort(n_keys1, keys1.skip_key))) return rc;
            HSQ_CUDA(ctx, cudaMemsetAsync(work, 0, 16, ctx->stream));  // work counters of the two packed kernels
            pool2.nthreads = (uint32_t)grid1 * DP2_BLOCK;
            primer2_kernel<<<grid1, DP2_BLOCK, 0, ctx->stream>>>(pv, pid, v1, v2, order, n_active, sp_default2, pool2, d, keys, d_status(ctx),
                                                                cells, work);
        } else {
            primer_kernel<<<grid1, DP_BLOCK, 0, ctx->stream>>>(pv, pid, v1, v2, n, sp_default, pool, d, keys, d_status(ctx), cells);
        }
        ctx->launches++;
        if ((rc = bucket_sort(n_keys, keys.skip_key))) return rc;
        if (packed2 && !packed1) HSQ_CUDA(ctx, cudaMemsetAsync(work, 0, 16, ctx->stream));
        HSQ_CUDA(ctx, cudaGetLastError());
        plan->pv = pv;
        plan->pid = pid;
        plan->v1 = v1;
        plan->v2 = v2;
        plan->order = order;
        plan->n_active = n_active;
        plan->cells = cells;
        plan->work = work;
        plan->keys = keys;
        plan->hist = hist;
        plan->cursor = cursor;
        plan->n_active_w = n_active;
        plan->order_w = order;
        plan->n_keys = n_keys;
        plan->order_bytes = order_bytes;
        plan->sblocks = sblocks;
    }
    // the band applies to the packed capture kernel only, and only when an uncomputed cell has a representable value
    plan->band = 0;
    if (packed2 && ctx->cfg.band > 0) {
        sp_user2.neg2 = dp2_band_neg(ctx->scorer.match, ctx->scorer.mismatch, ctx->scorer.gap_extend, ctx->scorer.gap_start, q_max, r_max) * DP2_REP;
        if (sp_user2.neg2) plan->band = ctx->cfg.band;
    }
    plan->n = n;
    plan->packed2 = packed2;
    plan->grid2 = grid2;
    plan->sp_user = sp_user;
    plan->sp_user2 = sp_user2;
    plan->pool = pool;
    plan->pool2 = pool2;
    plan->pool2.nthreads = (uint32_t)grid2 * DP2_BLOCK;
    plan->d = d;
    plan->pair_ok = ctx->pair_ok.as<uint8_t>();
    return HSQ_OK;
}

// capture: the capture-target DP of the read ends that passed the primer stage + the per-pair finalize, on ctx->stream
int extend_capture(hsq_ctx* ctx, const ExtendPlan& p) {
    if (p.n > 0 && p.packed2 && ctx->cfg.l2_persist_rows && ctx->l2_window_max > 0 && ctx->l2_persist_max > 0) {
        // the boundary rows are rewritten and re-read once per strip while 8 B per cell column of pointer words stream past them
        const size_t bytes = (size_t)p.pool2.nthreads * (size_t)p.pool2.r_stride * 4;
        if (!ctx->l2_limit_set) {
            HSQ_CUDA(ctx, cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, std::min(ctx->l2_persist_max, bytes + (8u << 20))));
            ctx->l2_limit_set = true;
        }
        cudaStreamAttrValue a{};
        a.accessPolicyWindow.base_ptr = p.pool2.brow;
        a.accessPolicyWindow.num_bytes = std::min(bytes, ctx->l2_window_max);
        a.accessPolicyWindow.hitRatio = 1.0f;
        a.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        a.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        HSQ_CUDA(ctx, cudaStreamSetAttribute(ctx->stream, cudaStreamAttributeAccessPolicyWindow, &a));
    }
    if (p.n > 0) {
        if (p.packed2 && p.band > 0) {
            // pass 1: banded; the read ends it accepts move to the skip bucket.  pass 2: the others, full matrices.
            capture2_band_kernel<<<p.grid2, DP2_BLOCK, 0, ctx->stream>>>(p.pv, p.pid, p.v1, p.v2, p.order, p.n_active, p.sp_user2, p.pool2, p.d, p.cells,
                                                                        p.work + 1, p.band, p.keys, p.cells + BAND_STATS_SLOT);
            ctx->launches++;
            HSQ_CUDA(ctx, cudaGetLastError());
            int rc = bucket_sort_launch(ctx, p.keys, 2 * p.n, p.n_keys, p.hist, p.cursor, p.n_active_w, p.order_w, p.order_bytes, p.sblocks);
            if (rc) return rc;
            HSQ_CUDA(ctx, cudaMemsetAsync(p.work + 1, 0, 8, ctx->stream));
            capture2_kernel<<<p.grid2, DP2_BLOCK, 0, ctx->stream>>>(p.pv, p.pid, p.v1, p.v2, p.order, p.n_active, p.sp_user2, p.pool2, p.d, p.cells, p.work + 1);
        } else if (p.packed2)
            capture2_kernel<<<p.grid2, DP2_BLOCK, 0, ctx->stream>>>(p.pv, p.pid, p.v1, p.v2, p.order, p.n_active, p.sp_user2, p.pool2, p.d, p.cells, p.work + 1);
        else
            capture_kernel<<<p.grid2, DP_BLOCK, 0, ctx->stream>>>(p.pv, p.pid, p.v1, p.v2, p.order, p.n_active, p.sp_user, p.pool, p.d, p.cells);
        ctx->launches++;
        HSQ_CUDA(ctx, cudaGetLastError());
        finalize_extend_kernel<<<(unsigned)((p.n + 255) / 256), 256, 0, ctx->stream>>>(p.pv, p.pid, p.n, p.d, p.pair_ok, d_status(ctx));
        ctx->launches++;
        HSQ_CUDA(ctx, cudaGetLastError());
    }
    return HSQ_OK;
}

// run: prep + capture of the staged batch on one stream; results stay on the device.
int extend_run(hsq_ctx* ctx, int cigar_cap, int md_cap, int host_read_max = -1) {
    ExtendPlan plan;
    int rc = extend_prep(ctx, cigar_cap, md_cap, host_read_max, &plan);
    if (!rc) rc = extend_capture(ctx, plan);
    if (!rc) ctx->ext_ran = true;
    return rc;
}

// fetch, part 1: device -> host copies of the results of the last run (asynchronous, on the context's copy stream)
int extend_fetch_copies(hsq_ctx* ctx, hsq_extend_out* o) {
    if (!ctx->ext_ran || ctx->ext_n < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "no results to fetch");
    const int64_t n = ctx->ext_n;
    const size_t n2 = (size_t)(2 * n);
    const ExtendDeviceOut& d = ctx->ext_out;
    if (o->cigar_cap != d.cigar_cap || o->md_cap != d.md_cap) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "slot capacities differ from the run");
    int rc;
    int32_t* i32 = ctx->out_i32.as<int32_t>();
    int32_t* hosts[9] = {o->primer_end, o->score, o->pos, o->read_start, o->read_stop, o->nm, o->tlen, o->cigar_len, o->md_len};
    for (int f = 0; f < 9; f++)
        if ((rc = d2h(ctx, hosts[f], i32 + f * n2, n2))) return rc;
    if ((rc = d2h(ctx, o->ok, d.ok, n2))) return rc;
    if ((rc = d2h(ctx, o->overflow, d.overflow, n2))) return rc;
    if ((rc = d2h(ctx, o->cigar, d.cigar, n2 * (size_t)d.cigar_cap))) return rc;
    if ((rc = d2h(ctx, o->md, d.md, n2 * (size_t)d.md_cap))) return rc;
    if ((rc = d2h(ctx, o->pair_ok, ctx->pair_ok.as<uint8_t>(), (size_t)n))) return rc;
    return HSQ_OK;
}

// fetch, part 2: cell counter and status of everything run since the counters were last reset (synchronises)
int extend_counters(hsq_ctx* ctx, int64_t* cells_out) {
    int rc;
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->h_cells, ctx->cells.p, (BAND_STATS_SLOT + 2) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    int w[4];
    if ((rc = read_status(ctx, w, 4))) return rc;
    if (ctx->io && ctx->io != ctx->stream) HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->io));
    *cells_out = (int64_t)ctx->h_cells[0];
    ctx->band_stats[0] = (int64_t)ctx->h_cells[BAND_STATS_SLOT];
    ctx->band_stats[1] = (int64_t)ctx->h_cells[BAND_STATS_SLOT + 1];
    if (w[0] != 0) return fail(ctx, w[0], std::string("hsq_extend: ") + status_message(w[0]));
    if (w[3] != 0) return fail(ctx, HSQ_ERR_OVERFLOW, std::string("hsq_extend: ") + status_message(HSQ_ERR_OVERFLOW));
    return HSQ_OK;
}

int extend_fetch_finish(hsq_ctx* ctx, hsq_extend_out* o) { return extend_counters(ctx, &o->cells); }

int extend_fetch(hsq_ctx* ctx, hsq_extend_out* o) {
    int rc = extend_fetch_copies(ctx, o);
    return rc ? rc : extend_fetch_finish(ctx, o);
}

// The extend-batch buffers exist three times; the members of hsq_ctx hold the active set, the others are parked.
void swap_extend_set(hsq_ctx* ctx, hsq_ctx::ExtendSet& st) {
    std::swap(ctx->r1, st.r1);
    std::swap(ctx->r2, st.r2);
    std::swap(ctx->probe_id, st.probe_id);
    std::swap(ctx->out_i32, st.out_i32);
    std::swap(ctx->out_u8, st.out_u8);
    std::swap(ctx->out_slots, st.out_slots);
    std::swap(ctx->pair_ok, st.pair_ok);
    std::swap(ctx->ptr_words, st.ptr_words);
    std::swap(ctx->brow, st.brow);
    std::swap(ctx->cols, st.cols);
    std::swap(ctx->colinfo, st.colinfo);
    std::swap(ctx->slow, st.slow);
    std::swap(ctx->sort_a, st.sort_a);
    std::swap(ctx->sort_b, st.sort_b);
    std::swap(ctx->sort_c, st.sort_c);
    std::swap(ctx->ext_out, st.ext_out);
    std::swap(ctx->ext_n, st.ext_n);
}

void activate_extend_set(hsq_ctx* ctx, int k) {
    if (ctx->active_set == k) return;
    swap_extend_set(ctx, ctx->parked[ctx->active_set]);  // park the current members
    swap_extend_set(ctx, ctx->parked[k]);                // load set k
    ctx->active_set = k;
}

hsq_seqs seqs_window(const hsq_seqs& s, int64_t a, int64_t n) {
    hsq_seqs w;
    w.data = s.data;
    w.offsets = s.offsets + a;
    w.n = n;
    return w;
}

// Longest read of a window of a ragged batch and the probe-id range check, on the host (what extend_run otherwise gets
// from the device at the price of a synchronisation).
int host_chunk_dims(hsq_ctx* ctx, const hsq_pair_batch& w, int* read_max) {
    int64_t mx = 0;
    bool bad = false, order = false;
    for (int64_t k = 0; k < w.n; k++) {
        const int64_t l1 = w.r1.offsets[k + 1] - w.r1.offsets[k], l2 = w.r2.offsets[k + 1] - w.r2.offsets[k];
        order |= (l1 < 0) | (l2 < 0);
        mx = std::max(mx, std::max(l1, l2));
        bad |= (w.probe_id[k] < 0) | (w.probe_id[k] >= ctx->n_probes);
    }
    if (order) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_extend: offsets must not decrease");
    if (bad) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_extend: probe_id out of range");
    if (mx > INT32_MAX) return fail(ctx, HSQ_ERR_UNSUPPORTED, "hsq_extend: read too long");
    *read_max = (int)mx;
    return HSQ_OK;
}

// Chunk boundaries of the host pipeline: sizes ramp up and down (weights 1,2,4,8 towards the middle) because the first H2D
// and the last D2H are the only copies nothing overlaps, so the edge chunks are small and the middle ones large.
std::vector<int64_t> chunk_starts(const hsq_ctx* ctx, int64_t n, int n_chunks) {
    std::vector<int64_t> starts((size_t)n_chunks + 1, 0);
    std::vector<double> wgt((size_t)n_chunks);
    double total = 0;
    const bool ramp = !ctx->cfg.uniform_chunks;
    for (int c = 0; c < n_chunks; c++) {
        const int edge = std::min(c, n_chunks - 1 - c);
        wgt[(size_t)c] = ramp ? (double)(1 << std::min(edge, 3)) : 1.0;
        total += wgt[(size_t)c];
    }
    double acc = 0;
    for (int c = 0; c < n_chunks; c++) {
        acc += wgt[(size_t)c];
        starts[(size_t)c + 1] = std::min<int64_t>(n, (int64_t)((double)n * acc / total + 0.5));
    }
    starts[(size_t)n_chunks] = n;
    return starts;
}

// chunks of >= 256 k pairs on average, at most 12 (hsq_config.extend_chunks overrides; 1 = no pipelining)
int pick_chunks(const hsq_ctx* ctx, int64_t n) {
    int n_chunks = (int)std::min<int64_t>(12, n / 262144);
    if (ctx->cfg.extend_chunks > 0) n_chunks = ctx->cfg.extend_chunks;
    if (n_chunks < 1) n_chunks = 1;
    if (n < n_chunks) n_chunks = (int)std::max<int64_t>(1, n);
    return n_chunks;
}

// What differs between the callers of the host pipeline (ASCII batches with slot output, packed batches with pooled output).
struct ExtendChunkIO {
    // host-side size / range checks of chunk [a, a+m): longest read
    std::function<int(int64_t a, int64_t m, int* read_max)> dims;
    // host -> device copies of the chunk into the active buffer set, on ctx->io
    std::function<int(int64_t a, int64_t m)> stage;
    // device -> host delivery of the active set's results for the chunk, on ctx->io (the set's capture kernels have finished)
    std::function<int(int64_t a, int64_t m)> fetch;
    // optional: runs on the prep stream after the chunk's copies have arrived and before its prep; may replace the staged pairs by a
    // subset of them (hsq_dedup_extend_packed_keys: the unique pairs) and says how many pairs the prep is to see
    std::function<int(int64_t a, int64_t m, int64_t* m_out)> filter;
};

// A large batch flows through
//     H2D (s_in) | prep: encode, primer DP, bucket sorts (s_prep) | capture DP (s_cap[chunk & 1]) | D2H (s_out)
// with three complete buffer sets, so that three chunks are in flight: while the capture kernel of chunk c runs, the prep
// of chunk c+1 is already done and chunk c+2 is being copied / prepared.  When a capture kernel drains, the next one is
// ready to take the freed SMs at once (no chain of small kernels in between), and its tail overlaps that start because
// consecutive capture kernels sit on different streams.  The host never blocks inside the loop; events order the streams.
// Every exit goes through the join at the end: no copy is in flight and the home buffer set is active when this returns.
int extend_pipelined(hsq_ctx* ctx, int64_t n, int cigar_cap, int md_cap, int n_chunks, const ExtendChunkIO& io, int64_t* cells_out) {
    constexpr int NS = hsq_ctx::N_EXTEND_SETS;
    const std::vector<int64_t> starts = chunk_starts(ctx, n, n_chunks);
    int rc = HSQ_OK;
    cudaStream_t main_stream = ctx->stream;
    ctx->keep_counters = false;
    if ((rc = reset_status(ctx))) return rc;
    HSQ_CUDA(ctx, cudaStreamSynchronize(main_stream));  // counters are zero before any other stream touches them
    const int home_set = ctx->active_set;
    // the capture phase of chunk c is issued after the prep of chunk c+1, so that the prep stream always runs one chunk ahead
    std::vector<ExtendPlan> plans((size_t)n_chunks);
#define HSQ_STEP(expr)                                                                              \
    do {                                                                                            \
        cudaError_t _e = (expr);                                                                    \
        if (_e != cudaSuccess) return hsq::fail(ctx, HSQ_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
    } while (0)
    auto issue_capture = [&](int c) -> int {
        const int k = c % NS;
        cudaStream_t st = ctx->s_cap[c & 1];
        HSQ_STEP(cudaStreamWaitEvent(st, ctx->ev_prep[k], 0));
        ctx->stream = st;
        int r = extend_capture(ctx, plans[(size_t)c]);
        ctx->stream = main_stream;
        if (r) return r;
        HSQ_STEP(cudaEventRecord(ctx->ev_cap[k], st));
        // delivery of this chunk's results
        HSQ_STEP(cudaStreamWaitEvent(ctx->s_out, ctx->ev_cap[k], 0));
        activate_extend_set(ctx, k);
        ctx->ext_ran = true;
        ctx->io = ctx->s_out;
        r = io.fetch(starts[(size_t)c], starts[(size_t)c + 1] - starts[(size_t)c]);
        ctx->io = nullptr;
        if (r) return r;
        HSQ_STEP(cudaEventRecord(ctx->ev_out[k], ctx->s_out));
        return HSQ_OK;
    };
    auto issue_front = [&](int c) -> int {  // H2D + prep of chunk c
        const int k = c % NS;
        const int64_t a = starts[(size_t)c], m = starts[(size_t)c + 1] - a;
        int read_max = 0, r;
        if ((r = io.dims(a, m, &read_max))) return r;
        activate_extend_set(ctx, k);
        // H2D into this set once the chunk that used it before (c - NS) has delivered its results
        if (c >= NS) HSQ_STEP(cudaStreamWaitEvent(ctx->s_in, ctx->ev_out[k], 0));
        ctx->io = ctx->s_in;
        ctx->ext_n = -1;
        r = io.stage(a, m);
        ctx->io = nullptr;
        if (r) return r;
        ctx->ext_n = m;
        HSQ_STEP(cudaEventRecord(ctx->ev_in[k], ctx->s_in));
        HSQ_STEP(cudaStreamWaitEvent(ctx->s_prep, ctx->ev_in[k], 0));
        ctx->stream = ctx->s_prep;
        ctx->work_slot = k;
        if (io.filter) {
            int64_t m_eff = m;
            r = io.filter(a, m, &m_eff);
            if (r) {
                ctx->stream = main_stream;
                ctx->work_slot = 0;
                return r;
            }
            ctx->ext_n = m_eff;
        }
        r = extend_prep(ctx, cigar_cap, md_cap, read_max, &plans[(size_t)c]);
        ctx->stream = main_stream;
        ctx->work_slot = 0;
        if (r) return r;
        HSQ_STEP(cudaEventRecord(ctx->ev_prep[k], ctx->s_prep));
        return HSQ_OK;
    };
#undef HSQ_STEP
    for (int c = 0; c < n_chunks && !rc; c++) {
        rc = issue_front(c);
        if (!rc && c >= 1) rc = issue_capture(c - 1);
    }
    if (!rc) rc = issue_capture(n_chunks - 1);
    // everything joins the main stream, which then reads the counters
    ctx->stream = main_stream;
    ctx->work_slot = 0;
    cudaStreamSynchronize(ctx->s_prep);
    for (int k = 0; k < 2; k++) cudaStreamSynchronize(ctx->s_cap[k]);
    cudaStreamSynchronize(ctx->s_in);
    cudaStreamSynchronize(ctx->s_out);
    activate_extend_set(ctx, home_set);
    ctx->io = nullptr;
    int rc2 = rc ? HSQ_OK : extend_counters(ctx, cells_out);
    for (auto& ps : ctx->parked) ps.ext_n = -1;
    ctx->ext_n = -1;  // no single staged batch remains
    ctx->ext_ran = false;
    return rc ? rc : rc2;
}

}  // namespace

static int extend_args_ok(hsq_ctx* ctx) {
    if (!ctx->have_probes) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_set_probes has not been called");
    if (!ctx->have_scorer) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_set_scorer has not been called");
    return HSQ_OK;
}

extern "C" int32_t hsq_extend_batch(hsq_ctx* ctx, const hsq_pair_batch* pairs, hsq_extend_out* out) {
    if (!ctx || !pairs || !out) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    int rc = extend_args_ok(ctx);
    if (rc) return rc;
    if (pairs->n < 0 || out->cigar_cap < 0 || out->md_cap < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative size");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    const int n_chunks = pick_chunks(ctx, pairs->n);
    if (n_chunks > 1) {
        ExtendChunkIO io;
        auto window = [&](int64_t a, int64_t m) {
            hsq_pair_batch w = *pairs;
            w.n = m;
            w.probe_id = pairs->probe_id + a;
            w.read_index = pairs->read_index ? pairs->read_index + a : nullptr;
            w.r1 = seqs_window(pairs->r1, a, m);
            w.r2 = seqs_window(pairs->r2, a, m);
            return w;
        };
        io.dims = [&](int64_t a, int64_t m, int* read_max) { return host_chunk_dims(ctx, window(a, m), read_max); };
        io.stage = [&](int64_t a, int64_t m) {
            hsq_pair_batch w = window(a, m);
            return upload_pairs_raw(ctx, &w);
        };
        io.fetch = [&](int64_t a, int64_t) {
            hsq_extend_out w = *out;
            int32_t** f32[9] = {&w.primer_end, &w.score, &w.pos, &w.read_start, &w.read_stop, &w.nm, &w.tlen, &w.cigar_len, &w.md_len};
            for (auto f : f32) if (*f) *f += 2 * a;
            if (w.ok) w.ok += 2 * a;
            if (w.overflow) w.overflow += 2 * a;
            if (w.cigar) w.cigar += 2 * a * (int64_t)out->cigar_cap;
            if (w.md) w.md += 2 * a * (int64_t)out->md_cap;
            if (w.pair_ok) w.pair_ok += a;
            return extend_fetch_copies(ctx, &w);
        };
        rc = extend_pipelined(ctx, pairs->n, out->cigar_cap, out->md_cap, n_chunks, io, &out->cells);
    } else {
        rc = extend_stage(ctx, pairs);
        if (!rc) rc = extend_run(ctx, out->cigar_cap, out->md_cap);
        if (!rc) rc = extend_fetch(ctx, out);
    }
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}

extern "C" int32_t hsq_extend_stage(hsq_ctx* ctx, const hsq_pair_batch* pairs) {
    if (!ctx || !pairs || pairs->n < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = extend_stage(ctx, pairs);
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}

extern "C" int32_t hsq_extend_run(hsq_ctx* ctx, int32_t cigar_cap, int32_t md_cap) {
    if (!ctx || cigar_cap < 0 || md_cap < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "bad argument");
    int rc = extend_args_ok(ctx);
    if (rc) return rc;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    rc = extend_run(ctx, cigar_cap, md_cap);
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}

extern "C" int32_t hsq_extend_fetch(hsq_ctx* ctx, hsq_extend_out* out) {
    if (!ctx || !out) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = extend_fetch(ctx, out);
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}
