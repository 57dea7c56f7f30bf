// hsq_capi_pipeline.inl -- hsq_dedup_extend_batch: the body of one probe task for all probes of a batch
// (HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:1172-1216): FilterByUid.reduceReadsByProbeAndUid, then
// ExtendReadsToPrimer.extendReadsToPrimers on the unique pairs (and on the duplicates when they are kept or marked).
// The reads cross PCIe once.  Without duplicate extension the unique pairs are compacted on the device (flag scan -> list,
// length scan -> offsets, warp-per-read byte copy) into the second buffer set and the DP kernels see only them.
namespace {

__global__ void flags_to_i64_kernel(const uint8_t* __restrict__ f, int64_t n, int64_t* __restrict__ out) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) out[k] = f[k] ? 1 : 0;
}

// survivors list: list[pos[k]] = k for flagged k; their probe ids and read lengths in list order
__global__ void survivor_list_kernel(const uint8_t* __restrict__ f, const int64_t* __restrict__ pos, int64_t n, const int32_t* __restrict__ probe_id,
                                     const int64_t* __restrict__ off1, const int64_t* __restrict__ off2, int32_t* __restrict__ list,
                                     int32_t* __restrict__ probe_out, int64_t* __restrict__ len1, int64_t* __restrict__ len2) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) {
        if (!f[k]) continue;
        const int64_t j = pos[k];
        list[j] = (int32_t)k;
        probe_out[j] = probe_id[k];
        len1[j] = off1[k + 1] - off1[k];
        len2[j] = off2[k + 1] - off2[k];
    }
}

// one warp per surviving read: raw bytes to their compacted place
__global__ void survivor_copy_kernel(const uint8_t* __restrict__ src, const int64_t* __restrict__ src_off, const int32_t* __restrict__ list,
                                     const int64_t* __restrict__ dst_off, int64_t m, uint8_t* __restrict__ dst) {
    const int lane = threadIdx.x & 31;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; w < m; w += stride) {
        const int64_t s = src_off[list[w]], d = dst_off[w];
        const int len = (int)(dst_off[w + 1] - d);
        for (int i = lane; i < len; i += 32) dst[d + i] = src[s + i];
    }
}

// per-probe "unable to extend" counters (:1189,:1208): slot j holds pair list[j] (or j itself when list is null)
__global__ void unable_to_extend_kernel(const uint8_t* __restrict__ pair_ok, const int32_t* __restrict__ list, const uint8_t* __restrict__ best,
                                        const int32_t* __restrict__ probe_of_slot, int64_t m, int32_t* __restrict__ unable_unique,
                                        int32_t* __restrict__ unable_dup) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; j < m; j += stride) {
        if (pair_ok[j]) continue;
        const int64_t k = list ? list[j] : j;
        atomicAdd((best[k] ? unable_unique : unable_dup) + probe_of_slot[j], 1);
    }
}

__global__ void iota_kernel(int32_t* __restrict__ out, int64_t n) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) out[k] = (int32_t)k;
}

void swap_read_buffers(hsq_ctx* ctx) {
    std::swap(ctx->r1, ctx->alt_r1);
    std::swap(ctx->r2, ctx->alt_r2);
    std::swap(ctx->probe_id, ctx->alt_probe_id);
}

// compacts the flagged pairs of the staged batch (ctx->r1/r2/probe_id) into the second buffer set; returns their number
// known_m >= 0: the caller knows how many pairs are flagged (and the batch holds device codes): nothing here waits for the device
int compact_survivors(hsq_ctx* ctx, int64_t n, const uint8_t* d_keep, int32_t* d_list, int64_t* m_out, int64_t known_m = -1) {
    const size_t nn = (size_t)n;
    int rc;
    const int64_t n_tiles = (n + 1 + SCAN_TILE - 1) / SCAN_TILE;
    HSQ_CUDA(ctx, ctx->sort_a.ensure((nn + 1) * 8));                     // flags as int64, then lengths of R1
    HSQ_CUDA(ctx, ctx->sort_b.ensure((nn + 1) * 8));                     // positions, then lengths of R2
    HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)n_tiles * 8 + 16));         // tile sums
    int64_t* a = ctx->sort_a.as<int64_t>();
    int64_t* b = ctx->sort_b.as<int64_t>();
    int64_t* tiles = ctx->sort_c.as<int64_t>();
    const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
    flags_to_i64_kernel<<<gblocks, 256, 0, ctx->stream>>>(d_keep, n, a);
    ctx->launches++;
    if ((rc = scan_i64(ctx, a, n, tiles, b))) return rc;
    int64_t* h_m = reinterpret_cast<int64_t*>(ctx->h_cells) + 1;  // pinned
    if (known_m < 0) {
        HSQ_CUDA(ctx, cudaMemcpyAsync(h_m, b + n, 8, cudaMemcpyDeviceToHost, ctx->stream));
        HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    const int64_t m = known_m < 0 ? *h_m : known_m;
    *m_out = m;
    const size_t mm = (size_t)m;
    HSQ_CUDA(ctx, ctx->alt_probe_id.ensure(mm * 4 + 16));
    HSQ_CUDA(ctx, ctx->alt_r1.offsets.ensure((mm + 1) * 8));
    HSQ_CUDA(ctx, ctx->alt_r2.offsets.ensure((mm + 1) * 8));
    if (m > 0) {
        // the length arrays may overwrite the flag array, not the positions: lengths go to a and a second array
        HSQ_CUDA(ctx, ctx->flags.ensure((mm + 1) * 8 + 16));
        int64_t* len1 = ctx->flags.as<int64_t>();
        HSQ_CUDA(ctx, ctx->slow.ensure((mm + 1) * 8 + 16));
        int64_t* len2 = ctx->slow.as<int64_t>();
        survivor_list_kernel<<<gblocks, 256, 0, ctx->stream>>>(d_keep, b, n, ctx->probe_id.as<int32_t>(), ctx->r1.offsets.as<int64_t>(),
                                                              ctx->r2.offsets.as<int64_t>(), d_list, ctx->alt_probe_id.as<int32_t>(), len1, len2);
        ctx->launches++;
        if ((rc = scan_i64(ctx, len1, m, tiles, ctx->alt_r1.offsets.as<int64_t>()))) return rc;
        if ((rc = scan_i64(ctx, len2, m, tiles, ctx->alt_r2.offsets.as<int64_t>()))) return rc;
    } else {
        HSQ_CUDA(ctx, cudaMemsetAsync(ctx->alt_r1.offsets.p, 0, 8, ctx->stream));
        HSQ_CUDA(ctx, cudaMemsetAsync(ctx->alt_r2.offsets.p, 0, 8, ctx->stream));
    }
    // the compacted reads are never larger than the batch: size the buffers without a second synchronisation
    struct { DevSeqsOwned* src; DevSeqsOwned* dst; } sides[2] = {{&ctx->r1, &ctx->alt_r1}, {&ctx->r2, &ctx->alt_r2}};
    for (auto& s : sides) {
        HSQ_CUDA(ctx, s.dst->ascii.ensure((size_t)s.src->total + 32));
        HSQ_CUDA(ctx, s.dst->codes.ensure((size_t)s.src->total + 32));
        s.dst->n = m;
        // a batch that arrived packed holds device codes already (packed_bits < 0): those are what is compacted, and the extension
        // will not decode again; an ASCII batch is compacted as ASCII and encoded afterwards
        const bool coded = s.src->packed_bits < 0;
        s.dst->packed_bits = coded ? -1 : 0;
        if (m > 0) {
            const int blocks = (int)std::min<int64_t>((m * 32 + 255) / 256, (int64_t)ctx->sm_count * 16);
            survivor_copy_kernel<<<blocks, 256, 0, ctx->stream>>>(coded ? s.src->codes.as<uint8_t>() : s.src->ascii.as<uint8_t>(), s.src->offsets.as<int64_t>(),
                                                                 d_list, s.dst->offsets.as<int64_t>(), m,
                                                                 coded ? s.dst->codes.as<uint8_t>() : s.dst->ascii.as<uint8_t>());
            ctx->launches++;
        }
    }
    HSQ_CUDA(ctx, cudaGetLastError());
    if (known_m >= 0 && ctx->r1.packed_bits < 0 && ctx->r2.packed_bits < 0) {
        // device codes are not encoded again: the totals only size buffers, and the uncompacted totals bound them
        ctx->alt_r1.total = ctx->r1.total;
        ctx->alt_r2.total = ctx->r2.total;
        return HSQ_OK;
    }
    // totals of the compacted sides (needed by the encoder's launch size)
    int64_t* h_tot = reinterpret_cast<int64_t*>(ctx->h_cells) + 2;
    HSQ_CUDA(ctx, cudaMemcpyAsync(h_tot, ctx->alt_r1.offsets.as<int64_t>() + m, 8, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaMemcpyAsync(h_tot + 1, ctx->alt_r2.offsets.as<int64_t>() + m, 8, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->alt_r1.total = h_tot[0];
    ctx->alt_r2.total = h_tot[1];
    return HSQ_OK;
}

}  // namespace

namespace {

// the pairs of a probe-task call: an ASCII batch or a packed one
struct PairSource {
    const hsq_pair_batch* a = nullptr;
    const hsq_packed_pairs* p = nullptr;
    int64_t n() const { return a ? a->n : p->n; }
};

int dedup_extend_args(hsq_ctx* ctx, const PairSource& src, int32_t u, int32_t v, const hsq_dedup_extend_out* out, bool pooled) {
    if (!ctx || (!src.a && !src.p) || !out || !out->is_best || !out->extended_pair) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (pooled ? !out->extend_pooled : !out->extend) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null extension output");
    const int64_t n = src.n();
    if (src.a && (n < 0 || src.a->r1.n != n || src.a->r2.n != n)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "pair batch arrays disagree on n");
    if (src.p) {
        if (n < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative size");
        if (n > 0 && !src.p->probe_id) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "probe_id is null");
        int rc;
        if ((rc = check_packed_reads(ctx, &src.p->r1, n, "hsq_dedup_extend_packed")) || (rc = check_packed_reads(ctx, &src.p->r2, n, "hsq_dedup_extend_packed"))) return rc;
    }
    if (n >= (1ll << 31)) return fail(ctx, HSQ_ERR_UNSUPPORTED, "more than 2^31 pairs in one batch");
    if (u < 0 || v < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative UID length");
    if (u + v > 32) return fail(ctx, HSQ_ERR_UNSUPPORTED, "UID longer than 32 bases in total is outside the accelerated domain");
    if (!out->stats || out->stats->n_probes <= 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "probe stats (with n_probes) are required");
    if (pooled ? (out->extend_pooled->cigar_cap < 0 || out->extend_pooled->md_cap < 0) : (out->extend->cigar_cap < 0 || out->extend->md_cap < 0))
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative size");
    int rc = extend_args_ok(ctx);
    if (rc) return rc;
    if (out->stats->n_probes != ctx->n_probes) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "stats->n_probes differs from the probe table");
    return HSQ_OK;
}

// The probe task body for this context's pairs.  sharded: the reduction runs over the union of all ranks' pairs (collective, see
// hsq_capi_comm.inl) and the per-probe counters are those of the whole batch; the extension runs on this rank's own pairs.
int dedup_extend_run(hsq_ctx* ctx, const PairSource& src, int32_t u, int32_t v, int32_t extend_duplicates, hsq_dedup_extend_out* out, bool sharded,
                     bool pooled) {
    const int64_t n = src.n(), np = out->stats->n_probes;
    const size_t nn = (size_t)n;
    out->n_extended = 0;
    int rc;
    const int cigar_cap = pooled ? (out->extend_pooled->cigar_cap ? out->extend_pooled->cigar_cap : 64) : out->extend->cigar_cap;
    const int md_cap = pooled ? (out->extend_pooled->md_cap ? out->extend_pooled->md_cap : 96) : out->extend->md_cap;
    PoolState ps;
    if (pooled) {
        out->extend_pooled->pool_bytes = 0;
        out->extend_pooled->cells = 0;
        if ((rc = pool_begin(ctx, out->extend_pooled, &ps)) && !sharded) return rc;
    }
    int bad = src.a ? dedup_upload_keygen(ctx, src.a, u, v, np) : dedup_upload_keygen_packed(ctx, src.p, u, v, np);
    if (bad && !sharded) return bad;
    const std::string bad_msg = ctx->error;
    HSQ_CUDA(ctx, ctx->dd_u8.ensure(nn + 32));
    HSQ_CUDA(ctx, ctx->dd_i32.ensure((2 * nn + 5 * (size_t)np) * 4 + 16));
    uint8_t* d_best = ctx->dd_u8.as<uint8_t>();
    int32_t* d_gs = ctx->dd_i32.as<int32_t>();
    int32_t* d_list = d_gs + nn;
    int32_t* d_stats = d_list + nn;
    int32_t* d_unable = d_stats + 3 * np;
    if (sharded) {  // a rank whose own batch is bad still takes part in the collectives (with no pairs), or the others would hang
        if (bad) HSQ_CUDA(ctx, cudaMemsetAsync(d_status(ctx), 0, 4, ctx->stream));
        rc = dedup_exchange_reduce(ctx, bad ? 0 : n, np, u + v, d_best, d_gs, d_stats);
    } else {
        rc = dedup_reduce_device(ctx, n, np, u + v, d_best, d_gs, d_stats);
    }
    if (rc && !sharded) return rc;
    const bool skip = bad || rc;  // sharded: the second collective below must still be entered
    if (!skip) {
        HSQ_CUDA(ctx, cudaMemsetAsync(d_unable, 0, 2 * (size_t)np * 4, ctx->stream));
        if ((rc = d2h(ctx, out->is_best, d_best, nn))) return rc;
        if ((rc = d2h(ctx, out->group_size, d_gs, nn))) return rc;
        if ((rc = fetch_stats(ctx, out->stats, d_stats))) return rc;
    }
    int64_t m = skip ? 0 : n;
    const bool compact = !extend_duplicates && m > 0;
    bool swapped = false;
    int rce = HSQ_OK;
    if (!skip) {
        span_begin(ctx, 2);
        if (compact) {
            if ((rce = compact_survivors(ctx, n, d_best, d_list, &m))) m = 0;
            else {
                swap_read_buffers(ctx);  // the compacted pairs are the staged batch now
                swapped = true;
            }
        } else if (n > 0) {
            const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
            iota_kernel<<<gblocks, 256, 0, ctx->stream>>>(d_list, n);
            ctx->launches++;
        }
        span_end(ctx, 2);
        if (!rce) rce = d2h(ctx, out->extended_pair, d_list, (size_t)m);
        if (!rce) {
            ctx->ext_n = m;
            ctx->keep_counters = false;
            span_begin(ctx, 3);
            rce = extend_run(ctx, cigar_cap, md_cap);
            span_end(ctx, 3);
        }
        if (!rce && m > 0) {
            const int gblocks = (int)std::min<int64_t>((m + 255) / 256, (int64_t)ctx->sm_count * 16);
            unable_to_extend_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->pair_ok.as<uint8_t>(), d_list, d_best, ctx->probe_id.as<int32_t>(), m,
                                                                     d_unable, d_unable + np);
            ctx->launches++;
        }
    } else {
        HSQ_CUDA(ctx, cudaMemsetAsync(d_unable, 0, 2 * (size_t)np * 4, ctx->stream));
    }
    if (sharded) {  // "unable to extend" is counted where the reads are: sum over ranks
        ncclResult_t r = nccl_api()->AllReduce(d_unable, d_unable, (size_t)(2 * np), ncclInt32, ncclSum, static_cast<ncclComm_t>(ctx->comm), ctx->stream);
        if (r != ncclSuccess && !rce) rce = fail(ctx, HSQ_ERR_CUDA, std::string("ncclAllReduce: ") + nccl_api()->GetErrorString(r));
    }
    if (!skip) {
        if (!rce) rce = d2h(ctx, out->unique_unable_to_extend, d_unable, (size_t)np);
        if (!rce) rce = d2h(ctx, out->duplicate_unable_to_extend, d_unable + np, (size_t)np);
        if (!rce && !pooled) rce = extend_fetch(ctx, out->extend);
        if (!rce && pooled) {
            rce = fetch_pooled_chunk(ctx, 0, m, out->extend_pooled, ps);
            if (!rce) rce = extend_counters(ctx, &out->extend_pooled->cells);
            if (!rce || rce == HSQ_ERR_OVERFLOW) {
                const int r3 = pool_end(ctx, out->extend_pooled, ps);
                if (!rce) rce = r3;
            }
        }
    }
    if (swapped) swap_read_buffers(ctx);
    ctx->r1.packed_bits = ctx->r2.packed_bits = ctx->alt_r1.packed_bits = ctx->alt_r2.packed_bits = 0;
    ctx->ext_n = -1;
    ctx->ext_ran = false;
    if (bad) return fail(ctx, bad, bad_msg);
    if (rc) return rc;
    if (rce) return rce;
    out->n_extended = m;
    return HSQ_OK;
}

}  // namespace

extern "C" int32_t hsq_dedup_extend_batch(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t u, int32_t v, int32_t extend_duplicates,
                                          hsq_dedup_extend_out* out) {
    PairSource src;
    src.a = pairs;
    int rc = dedup_extend_args(ctx, src, u, v, out, false);
    if (rc) return rc;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    rc = dedup_extend_run(ctx, src, u, v, extend_duplicates, out, false, false);
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}

// hsq_dedup_extend_batch over the union of all ranks' pairs: collective, every rank of the communicator calls it with its own pairs.
extern "C" int32_t hsq_dedup_extend_sharded(hsq_ctx* ctx, const hsq_pair_batch* pairs, int32_t u, int32_t v, int32_t extend_duplicates,
                                            hsq_dedup_extend_out* out) {
    PairSource src;
    src.a = pairs;
    int rc = dedup_extend_args(ctx, src, u, v, out, false);
    if (!rc) rc = comm_ready(ctx);
    if (rc) return rc;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    rc = dedup_extend_run(ctx, src, u, v, extend_duplicates, out, true, false);
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}

// The probe task for a packed batch, results pooled (out->extend_pooled): what a host should use for large batches -- with host-made
// quality sums about 100 bytes per pair cross PCIe instead of 624.  sharded != 0: collective over the context's communicator.
extern "C" int32_t hsq_dedup_extend_packed(hsq_ctx* ctx, const hsq_packed_pairs* pairs, int32_t u, int32_t v, int32_t extend_duplicates,
                                           int32_t sharded, hsq_dedup_extend_out* out) {
    PairSource src;
    src.p = pairs;
    int rc = dedup_extend_args(ctx, src, u, v, out, true);
    if (!rc && sharded) rc = comm_ready(ctx);
    if (rc) return rc;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    rc = dedup_extend_run(ctx, src, u, v, extend_duplicates, out, sharded != 0, true);
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}

// One process, n GPUs (the reference's shape: one JVM, probe tasks fanned out to a pool, …Identification.java:836-846): pairs[k] /
// outs[k] belong to ctxs[k]; one host thread per context runs hsq_dedup_extend_sharded.  The contexts get a communicator on first use.
extern "C" int32_t hsq_dedup_extend_multi_gpu(hsq_ctx** ctxs, int32_t n, const hsq_pair_batch* pairs, int32_t u, int32_t v,
                                              int32_t extend_duplicates, hsq_dedup_extend_out* outs) {
    if (!ctxs || n < 1 || !pairs || !outs) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    for (int k = 0; k < n; k++) {
        if (!ctxs[k]) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "null context");
        PairSource src;
        src.a = &pairs[k];
        int rc = dedup_extend_args(ctxs[k], src, u, v, &outs[k], false);  // argument errors before any collective starts
        if (rc) return rc;
    }
    bool have = true;
    for (int k = 0; k < n; k++) have = have && ctxs[k]->comm && ctxs[k]->comm_world == n && ctxs[k]->comm_rank == k;
    if (!have) {
        int rc = hsq_comm_init_all(ctxs, n);
        if (rc) return rc;
    }
    std::vector<int> rcs((size_t)n, 0);
    std::vector<std::thread> th;
    for (int k = 0; k < n; k++)
        th.emplace_back([&, k] { rcs[(size_t)k] = hsq_dedup_extend_sharded(ctxs[k], &pairs[k], u, v, extend_duplicates, &outs[k]); });
    for (auto& t : th) t.join();
    for (int k = 0; k < n; k++)
        if (rcs[(size_t)k]) return rcs[(size_t)k];  // hsq_last_error(ctxs[k]) holds the message
    return HSQ_OK;
}

// ExtendReadsToPrimer.extendReadsToPrimers on the pairs the caller flags (keep[k] != 0) -- e.g. the unique pairs of a dedup
// that ran elsewhere (sharded over several GPUs): same upload-once / compact-on-device path as hsq_dedup_extend_batch.
extern "C" int32_t hsq_extend_selected(hsq_ctx* ctx, const hsq_pair_batch* pairs, const uint8_t* keep, int64_t* n_extended,
                                       int32_t* extended_pair, hsq_extend_out* out) {
    if (!ctx || !pairs || !n_extended || !extended_pair || !out) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (pairs->n < 0 || pairs->r1.n != pairs->n || pairs->r2.n != pairs->n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "pair batch arrays disagree on n");
    if (pairs->n > 0 && !keep) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null keep flags");
    if (pairs->n >= (1ll << 31)) return fail(ctx, HSQ_ERR_UNSUPPORTED, "more than 2^31 pairs in one batch");
    if (out->cigar_cap < 0 || out->md_cap < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative size");
    int rc = extend_args_ok(ctx);
    if (rc) return rc;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    const int64_t n = pairs->n;
    const size_t nn = (size_t)n;
    *n_extended = 0;
    auto run = [&]() -> int {
        int rc;
        if ((rc = reset_status(ctx))) return rc;
        if ((rc = upload_pairs_raw(ctx, pairs))) return rc;
        HSQ_CUDA(ctx, ctx->dd_u8.ensure(nn + 32));
        HSQ_CUDA(ctx, ctx->dd_i32.ensure(nn * 4 + 16));
        uint8_t* d_keep = ctx->dd_u8.as<uint8_t>();
        int32_t* d_list = ctx->dd_i32.as<int32_t>();
        int64_t m = 0;
        if (n > 0) {
            HSQ_CUDA(ctx, cudaMemcpyAsync(d_keep, keep, nn, cudaMemcpyHostToDevice, ctx->stream));
            const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
            check_probe_ids_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), n, ctx->n_probes, d_status(ctx));
            ctx->launches++;
            if ((rc = compact_survivors(ctx, n, d_keep, d_list, &m))) return rc;  // synchronises: the status word is final
            int w[1];
            if ((rc = read_status(ctx, w, 1))) return rc;
            if (w[0] != 0) return fail(ctx, w[0], "hsq_extend_selected: probe_id out of range");
            swap_read_buffers(ctx);
        }
        if ((rc = d2h(ctx, extended_pair, d_list, (size_t)m))) return rc;
        ctx->ext_n = m;
        ctx->keep_counters = false;
        rc = extend_run(ctx, out->cigar_cap, out->md_cap);
        if (!rc) rc = extend_fetch(ctx, out);
        if (n > 0) swap_read_buffers(ctx);
        ctx->ext_n = -1;
        ctx->ext_ran = false;
        if (!rc) *n_extended = m;
        return rc;
    };
    rc = run();
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}

namespace {
// one warp per pair: any 'N' / 'n' in either read
__global__ void flag_n_kernel(const uint8_t* __restrict__ r1, const int64_t* __restrict__ off1, const uint8_t* __restrict__ r2,
                              const int64_t* __restrict__ off2, int64_t n, uint8_t* __restrict__ has_n) {
    const int lane = threadIdx.x & 31;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; w < n; w += stride) {
        bool hit = false;
        for (int64_t i = off1[w] + lane; i < off1[w + 1]; i += 32) hit |= ((r1[i] | 0x20) == 'n');
        for (int64_t i = off2[w] + lane; i < off2[w + 1]; i += 32) hit |= ((r2[i] | 0x20) == 'n');
        hit = __any_sync(0xffffffffu, hit);
        if (lane == 0) has_n[w] = hit ? 1 : 0;
    }
}
}  // namespace

extern "C" int32_t hsq_flag_n_pairs(hsq_ctx* ctx, const hsq_pair_batch* pairs, uint8_t* has_n) {
    if (!ctx || !pairs) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    const int64_t n = pairs->n;
    if (n < 0 || pairs->r1.n != n || pairs->r2.n != n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "pair batch arrays disagree on n");
    if (n > 0 && !has_n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null output");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc;
    if ((rc = copy_seqs(ctx, &pairs->r1, &ctx->r1))) return rc;
    if ((rc = copy_seqs(ctx, &pairs->r2, &ctx->r2))) return rc;
    HSQ_CUDA(ctx, ctx->out_u8.ensure((size_t)n + 16));
    if (n > 0) {
        const int blocks = (int)std::min<int64_t>((n * 32 + 255) / 256, (int64_t)ctx->sm_count * 16);
        flag_n_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->r1.ascii.as<uint8_t>(), ctx->r1.offsets.as<int64_t>(), ctx->r2.ascii.as<uint8_t>(),
                                                      ctx->r2.offsets.as<int64_t>(), n, ctx->out_u8.as<uint8_t>());
        ctx->launches++;
        HSQ_CUDA(ctx, cudaGetLastError());
        if ((rc = d2h(ctx, has_n, ctx->out_u8.as<uint8_t>(), (size_t)n))) return rc;
    }
    return timer.finish();
}

// ---- host helper: per-pair reason codes from the outputs of a probe task ------------------------------------------------------------
extern "C" int32_t hsq_pair_reasons(const hsq_dedup_extend_out* out, int64_t n, int32_t extend_duplicates, uint8_t* reason) {
    if (!out || n < 0 || (n > 0 && (!reason || !out->is_best || !out->extended_pair))) return HSQ_ERR_INVALID_ARGUMENT;
    const int32_t* primer_end = out->extend_pooled ? out->extend_pooled->primer_end : (out->extend ? out->extend->primer_end : nullptr);
    const uint8_t* ok = out->extend_pooled ? out->extend_pooled->ok : (out->extend ? out->extend->ok : nullptr);
    if (out->n_extended > 0 && (!primer_end || !ok)) return HSQ_ERR_INVALID_ARGUMENT;  // the two columns this needs were not fetched
    if (out->n_extended < 0 || out->n_extended > n) return HSQ_ERR_INVALID_ARGUMENT;
    for (int64_t k = 0; k < n; k++) reason[k] = HSQ_PAIR_DUPLICATE_DROPPED;  // a pair that was not extended: not the best of its group
    for (int64_t j = 0; j < out->n_extended; j++) {
        const int64_t k = out->extended_pair[j];
        if (k < 0 || k >= n) return HSQ_ERR_INVALID_ARGUMENT;
        uint8_t r;
        const int32_t e1 = primer_end[2 * j], e2 = primer_end[2 * j + 1];
        if (e1 == HSQ_PRIMER_NOT_FOUND || e1 < 0) r = HSQ_PAIR_PRIMER_NOT_FOUND_ONE;
        else if (e2 == HSQ_PRIMER_NOT_FOUND || e2 < 0) r = HSQ_PAIR_PRIMER_NOT_FOUND_TWO;
        else if (!ok[2 * j]) r = HSQ_PAIR_NOT_ALIGNED_ONE;
        else if (!ok[2 * j + 1]) r = HSQ_PAIR_NOT_ALIGNED_TWO;
        else r = out->is_best[k] ? HSQ_PAIR_WRITTEN_UNIQUE : HSQ_PAIR_WRITTEN_DUPLICATE;
        reason[k] = r;
    }
    (void)extend_duplicates;  // (an extended pair that is not the best of its group exists only when duplicates are extended)
    return HSQ_OK;
}
