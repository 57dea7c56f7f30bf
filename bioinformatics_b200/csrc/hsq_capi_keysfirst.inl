// hsq_capi_keysfirst.inl -- hsq_dedup_extend_packed_keys: the probe task of hsq_dedup_extend_packed with the reduction fed by keys.
//
// hsq_dedup_extend_packed is sequential by construction: all reads go to the device (23 ms per 12.5 M pairs at 55 GB/s), then the keys,
// the reduction, the extension of the unique pairs, the copies back -- the reduction needs every key before the first extension can
// start, so nothing overlaps.  But the reduction needs 20 bytes of a pair, not its reads: the UID code (hsq_uid_keys, host helper, 8 B),
// the two quality sums (hsq_quality_sums, 2 x 2 B), probe id and read index.  This entry point takes those first, reduces, and then
// runs the extension through the chunk pipeline of hsq_extend_packed: the reads of chunk c + 1 are copied in while chunk c is aligned,
// each chunk is unpacked, cut down to its unique pairs on the device (compact_survivors) and extended.  Results are those of
// hsq_dedup_extend_packed, slot for slot (tests/test_packed.py).  Reference: the task body ...Identification.java:1172-1216.
namespace hsq {

__global__ void kf_keygen_kernel(const int16_t* __restrict__ qs1, const int16_t* __restrict__ qs2, const int32_t* __restrict__ read_index, int64_t n,
                                 uint64_t* __restrict__ rank_out, uint32_t* __restrict__ idx_out) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) {
        const int16_t qsum = (int16_t)((int)qs1[k] + (int)qs2[k]);  // Java short arithmetic wraps mod 2^16 (as dedup_keygen_kernel)
        rank_out[k] = ((uint64_t)(uint16_t)(32767 - (int)qsum) << 32) | (uint64_t)((uint32_t)read_index[k] ^ 0x80000000u);
        idx_out[k] = (uint32_t)k;
    }
}

__global__ void kf_list_kernel(const uint8_t* __restrict__ keep, const int64_t* __restrict__ pos, int64_t n, int32_t* __restrict__ list) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride)
        if (keep[k]) list[pos[k]] = (int32_t)k;
}

__global__ void kf_gather_i64_kernel(const int64_t* __restrict__ src, const int64_t* __restrict__ at, int n, int64_t* __restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) out[k] = src[at[k]];
}

}  // namespace hsq

// ---- host helper: the UID code of every pair from its packed reads -------------------------------------------------------------------
extern "C" int32_t hsq_uid_keys(const hsq_packed_reads* r1, const hsq_packed_reads* r2, int32_t u, int32_t v, uint64_t* out) {
    if (!r1 || !r2 || u < 0 || v < 0 || r1->n != r2->n || (r1->n > 0 && !out)) return HSQ_ERR_INVALID_ARGUMENT;
    if (u + v > 32) return HSQ_ERR_UNSUPPORTED;
    const int64_t n = r1->n;
    const hsq_packed_reads* sides[2] = {r1, r2};
    const int want[2] = {u, v};
    for (int s = 0; s < 2; s++) {
        const hsq_packed_reads* r = sides[s];
        if (r->bits != 2 && r->bits != 4) return HSQ_ERR_INVALID_ARGUMENT;
        if (n > 0 && (!r->len || (want[s] > 0 && !r->data))) return HSQ_ERR_INVALID_ARGUMENT;
        // 2 bit: a base that is not A/C/G/T was packed as 0 and listed as an exception; one inside a UID is outside the accelerated domain
        for (int64_t e = 0; e < r->n_exceptions; e++)
            if ((int64_t)r->exc_pos[e] < want[s]) return HSQ_ERR_UNSUPPORTED;
    }
    // pieces of the batch on the host's cores (the reads of a side lie back to back: a piece needs the word offset of its first read)
    const int n_threads = (int)std::max<int64_t>(1, std::min<int64_t>(std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 32u), n / 65536));
    std::vector<int64_t> first_word[2];
    for (int s = 0; s < 2; s++) {
        first_word[s].assign((size_t)n_threads + 1, 0);
        int64_t w = 0;
        int t = 0;
        for (int64_t k = 0; k < n; k++) {
            while (t < n_threads && k == n * t / n_threads) first_word[s][(size_t)t++] = w;
            w += hsq::packed_words_of(sides[s]->len[k] < 0 ? 0 : sides[s]->len[k], sides[s]->bits);
        }
        while (t <= n_threads) first_word[s][(size_t)t++] = w;
    }
    std::vector<int> status((size_t)n_threads, HSQ_OK);
    auto piece = [&](int t) {
        const int64_t k0 = n * t / n_threads, k1 = n * (t + 1) / n_threads;
        for (int64_t k = k0; k < k1; k++) out[k] = 0;
        for (int s = 0; s < 2; s++) {
            const hsq_packed_reads* r = sides[s];
            const int L = want[s], bits = r->bits, per = 32 / bits;
            const uint32_t* words = reinterpret_cast<const uint32_t*>(r->data);
            int64_t w0 = first_word[s][(size_t)t];
            for (int64_t k = k0; k < k1; k++) {
                const int len = r->len[k];
                if (len < L) {  // a read shorter than its UID
                    status[(size_t)t] = HSQ_ERR_OUT_OF_RANGE;
                    return;
                }
                uint64_t code = 0;
                for (int i = 0; i < L; i++) {
                    const uint32_t x = (words[w0 + i / per] >> (bits * (i % per))) & (bits == 2 ? 3u : 15u);
                    uint32_t c;
                    if (bits == 2) c = x ^ (x >> 1);  // packed A C T G = 0 1 2 3 -> key order A C G T
                    else {
                        if (x != 1 && x != 2 && x != 4 && x != 8) {  // an IUPAC code inside a UID
                            status[(size_t)t] = HSQ_ERR_UNSUPPORTED;
                            return;
                        }
                        c = x == 1 ? 0u : x == 2 ? 1u : x == 4 ? 2u : 3u;
                    }
                    code = (code << 2) | c;
                }
                out[k] = (out[k] << (2 * L)) | code;
                w0 += hsq::packed_words_of(len, bits);
            }
        }
    };
    if (n_threads == 1) piece(0);
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < n_threads; t++) pool.emplace_back(piece, t);
        for (auto& th : pool) th.join();
    }
    for (int st : status)
        if (st != HSQ_OK) return st;
    return HSQ_OK;
}

extern "C" int32_t hsq_dedup_extend_packed_keys(hsq_ctx* ctx, const hsq_packed_pairs* pairs, const uint64_t* uid_keys, int32_t u, int32_t v,
                                                int32_t extend_duplicates, int32_t sharded, hsq_dedup_extend_out* out) {
    PairSource src;
    src.p = pairs;
    int rc = dedup_extend_args(ctx, src, u, v, out, true);
    if (rc) return rc;
    const int64_t n = pairs->n, np = out->stats->n_probes;
    const size_t nn = (size_t)n;
    if (n > 0 && (!uid_keys || !pairs->qsum1 || !pairs->qsum2 || !pairs->read_index))
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_dedup_extend_packed_keys: uid keys, quality sums and read indexes are required");
    if (sharded && (rc = comm_ready(ctx))) return rc;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    hsq_extend_pooled_out* po = out->extend_pooled;
    const int cigar_cap = po->cigar_cap ? po->cigar_cap : 64, md_cap = po->md_cap ? po->md_cap : 96;
    out->n_extended = 0;
    po->pool_bytes = 0;
    po->cells = 0;
    if ((rc = reset_status(ctx))) return rc;
    // ---- phase A: the reduction, from 20 bytes per pair ------------------------------------------------------------------------------
    HSQ_CUDA(ctx, ctx->probe_id.ensure(nn * 4 + 16));
    HSQ_CUDA(ctx, ctx->read_index.ensure(nn * 4 + 16));
    HSQ_CUDA(ctx, ctx->q1.ensure(nn * 2 + 16));
    HSQ_CUDA(ctx, ctx->q2.ensure(nn * 2 + 16));
    HSQ_CUDA(ctx, ctx->sort_a.ensure(nn * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_b.ensure(nn * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_d.ensure(nn * 8 + 16));
    HSQ_CUDA(ctx, ctx->dd_u8.ensure(2 * nn + 64));
    HSQ_CUDA(ctx, ctx->dd_i32.ensure((2 * nn + 5 * (size_t)np) * 4 + 16));
    uint8_t* d_best = ctx->dd_u8.as<uint8_t>();
    uint8_t* d_keep = d_best + ((nn + 31) & ~(size_t)15);
    int32_t* d_gs = ctx->dd_i32.as<int32_t>();
    int32_t* d_list = d_gs + nn;
    int32_t* d_stats = d_list + nn;
    int32_t* d_unable = d_stats + 3 * np;
    if (n > 0) {
        HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->probe_id.p, pairs->probe_id, nn * 4, cudaMemcpyHostToDevice, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->read_index.p, pairs->read_index, nn * 4, cudaMemcpyHostToDevice, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->q1.p, pairs->qsum1, nn * 2, cudaMemcpyHostToDevice, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->q2.p, pairs->qsum2, nn * 2, cudaMemcpyHostToDevice, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->sort_a.p, uid_keys, nn * 8, cudaMemcpyHostToDevice, ctx->stream));
        const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
        span_begin(ctx, 0);
        check_probe_ids_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), n, np, d_status(ctx));
        kf_keygen_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->q1.as<int16_t>(), ctx->q2.as<int16_t>(), ctx->read_index.as<int32_t>(), n,
                                                           ctx->sort_b.as<uint64_t>(), ctx->sort_d.as<uint32_t>());
        span_end(ctx, 0);
        ctx->launches += 2;
    }
    int bad = HSQ_OK;
    {
        int w[1];
        if ((rc = read_status(ctx, w, 1))) return rc;
        if (w[0] != 0) bad = w[0];
    }
    if (bad && !sharded) return fail(ctx, bad, "hsq_dedup_extend_packed_keys: probe_id out of range");
    if (sharded) {  // a rank whose own batch is bad still takes part in the collectives (with no pairs), or the others would hang
        if (bad) HSQ_CUDA(ctx, cudaMemsetAsync(d_status(ctx), 0, 4, ctx->stream));
        rc = dedup_exchange_reduce(ctx, bad ? 0 : n, np, u + v, d_best, d_gs, d_stats);
    } else {
        rc = dedup_reduce_device(ctx, n, np, u + v, d_best, d_gs, d_stats);
    }
    if (rc && !sharded) return rc;
    const bool skip = bad || rc;
    HSQ_CUDA(ctx, cudaMemsetAsync(d_unable, 0, 2 * (size_t)np * 4, ctx->stream));
    int rce = HSQ_OK;
    int64_t m_total = 0;
    PoolState ps;
    if (!skip) {
        if ((rce = d2h(ctx, out->is_best, d_best, nn))) return rce;
        if ((rce = d2h(ctx, out->group_size, d_gs, nn))) return rce;
        if ((rce = fetch_stats(ctx, out->stats, d_stats))) return rce;
        // ---- phase B: which pairs are extended, and where their results go -------------------------------------------------------------
        span_begin(ctx, 2);
        const int n_chunks = std::min(pick_chunks(ctx, n), 63);
        PackedChunks pc;
        pc.starts = chunk_starts(ctx, n, n_chunks);
        pc.init();
        const int64_t n_tiles = (n + 1 + SCAN_TILE - 1) / SCAN_TILE;
        HSQ_CUDA(ctx, ctx->kf_pos.ensure((2 * (nn + 1) + (size_t)n_tiles + 2 * (size_t)(n_chunks + 2)) * 8 + 64));
        int64_t* d_flags64 = ctx->kf_pos.as<int64_t>();
        int64_t* d_pos = d_flags64 + nn + 1;
        int64_t* d_tiles = d_pos + nn + 1;
        int64_t* d_at = d_tiles + n_tiles;
        int64_t* d_cut = d_at + n_chunks + 2;
        std::vector<int64_t> cut((size_t)n_chunks + 1, 0);
        if (n > 0) {
            const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
            if (extend_duplicates) HSQ_CUDA(ctx, cudaMemsetAsync(d_keep, 1, nn, ctx->stream));  // every pair is extended: nothing to cut down
            else HSQ_CUDA(ctx, cudaMemcpyAsync(d_keep, d_best, nn, cudaMemcpyDeviceToDevice, ctx->stream));
            flags_to_i64_kernel<<<gblocks, 256, 0, ctx->stream>>>(d_keep, n, d_flags64);
            ctx->launches++;
            if ((rce = scan_i64(ctx, d_flags64, n, d_tiles, d_pos))) return rce;
            kf_list_kernel<<<gblocks, 256, 0, ctx->stream>>>(d_keep, d_pos, n, d_list);
            HSQ_CUDA(ctx, cudaMemcpyAsync(d_at, pc.starts.data(), (size_t)(n_chunks + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
            kf_gather_i64_kernel<<<1, 64, 0, ctx->stream>>>(d_pos, d_at, std::min(n_chunks + 1, 64), d_cut);
            ctx->launches += 2;
            HSQ_CUDA(ctx, cudaMemcpyAsync(cut.data(), d_cut, (size_t)(n_chunks + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
            span_end(ctx, 2);
            HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // starts (host vector) has been read, cut is filled
            m_total = cut[(size_t)n_chunks];
            if ((rce = d2h(ctx, out->extended_pair, d_list, (size_t)m_total))) return rce;
        }
        // ---- the extension of the kept pairs, chunk by chunk through the pipeline of hsq_extend_packed -----------------------------------
        rce = pool_begin(ctx, po, &ps);
        if (!rce && n > 0) {
            const int32_t* list_base = d_list;
            ExtendChunkIO io;
            io.dims = [&](int64_t a, int64_t, int* read_max) {
                const int c = pc.chunk_of(a);
                const int r = scan_packed_chunk(ctx, pairs, false, &pc, c);
                *read_max = pc.read_max[(size_t)c];
                return r;
            };
            io.stage = [&](int64_t a, int64_t) { return stage_packed_chunk(ctx, pairs, pc, pc.chunk_of(a)); };
            if (!extend_duplicates) io.filter = [&](int64_t a, int64_t m, int64_t* m_out) -> int {
                // on the prep stream, after the chunk's copies: codes of all m pairs, then the kept ones only (second buffer set, swapped in)
                int r;
                if ((r = prepare_codes(ctx, &ctx->r1, nullptr)) || (r = prepare_codes(ctx, &ctx->r2, nullptr))) return r;
                ctx->r1.packed_bits = ctx->r2.packed_bits = -1;  // the staged chunk holds device codes now: they are what is cut down
                HSQ_CUDA(ctx, ctx->kf_local.ensure((size_t)m * 4 + 16));
                const int c = pc.chunk_of(a);
                int64_t kept = 0;  // known from the positions of the whole batch: the compaction does not wait for the device
                if ((r = compact_survivors(ctx, m, d_keep + a, ctx->kf_local.as<int32_t>(), &kept, cut[(size_t)c + 1] - cut[(size_t)c]))) return r;
                swap_read_buffers(ctx);
                *m_out = kept;
                return HSQ_OK;
            };
            io.fetch = [&](int64_t a, int64_t) -> int {
                const int c = pc.chunk_of(a);
                const int64_t j0 = cut[(size_t)c], mk = cut[(size_t)c + 1] - j0;
                if (mk > 0) {  // per-probe "unable to extend" counters of this chunk's slots (on the delivery stream: the capture has finished)
                    const int gblocks = (int)std::min<int64_t>((mk + 255) / 256, (int64_t)ctx->sm_count * 16);
                    unable_to_extend_kernel<<<gblocks, 256, 0, ctx->io ? ctx->io : ctx->stream>>>(ctx->pair_ok.as<uint8_t>(), list_base + j0, d_best,
                                                                                                 ctx->probe_id.as<int32_t>(), mk, d_unable, d_unable + np);
                    ctx->launches++;
                }
                return fetch_pooled_chunk(ctx, j0, mk, po, ps);
            };
            rce = extend_pipelined(ctx, n, cigar_cap, md_cap, n_chunks, io, &po->cells);
        }
    }
    if (sharded) {  // "unable to extend" is counted where the reads are: sum over ranks
        ncclResult_t r = nccl_api()->AllReduce(d_unable, d_unable, (size_t)(2 * np), ncclInt32, ncclSum, static_cast<ncclComm_t>(ctx->comm), ctx->stream);
        if (r != ncclSuccess && !rce) rce = fail(ctx, HSQ_ERR_CUDA, std::string("ncclAllReduce: ") + nccl_api()->GetErrorString(r));
    }
    if (!skip) {
        if (!rce) rce = d2h(ctx, out->unique_unable_to_extend, d_unable, (size_t)np);
        if (!rce) rce = d2h(ctx, out->duplicate_unable_to_extend, d_unable + np, (size_t)np);
        if (!rce || rce == HSQ_ERR_OVERFLOW) {
            const int r3 = pool_end(ctx, po, ps);
            if (!rce) rce = r3;
        }
    }
    ctx->r1.packed_bits = ctx->r2.packed_bits = ctx->alt_r1.packed_bits = ctx->alt_r2.packed_bits = 0;
    for (auto& st : ctx->parked) st.r1.packed_bits = st.r2.packed_bits = 0;
    ctx->ext_n = -1;
    ctx->ext_ran = false;
    const int rc2 = timer.finish();
    if (bad) return fail(ctx, bad, "hsq_dedup_extend_packed_keys: probe_id out of range");
    if (rc) return rc;
    if (rce) return rce;
    if (rc2) return rc2;
    out->n_extended = m_total;
    return HSQ_OK;
}
