// hsq_capi_assign.inl -- coordinate read -> probe assignment:
//   RangeMap.getObjectsThatContainRangeInclusive                HSQ/process/RangeMap.java:92-153
//   ReadToProbeAssigner.assignProbesToReads + alternative hits   …Identification.java:439-576
// Device plan: one thread per query (record or alternative hit) runs the reference's binary search + leftward walk on the
// (contig, strand) slice of the start-sorted window table; one thread per record unions its own hits with its alternative
// hits' (sorted, unique); records are radix-sorted by read name (stable, so input order survives inside a name) and the
// head of every name pairs the first two records, intersecting their sets.
namespace hsq {

struct RangeTable {
    const int32_t *start, *end, *value;  // windows sorted by (contig, strand, start), stable
    const int64_t* grp_off;              // [2 * n_contigs + 1]: slice of group contig * 2 + strand
    int32_t n_contigs;
};

// The search of RangeMap.java:103-150 on one slice.  Returns -1 when the contig has no range map (null, :462,:565).
__device__ inline int contain_query(const RangeTable& t, int contig, int negative, int a, int b, int32_t* out, int cap) {
    if (contig < 0 || contig >= t.n_contigs) return -1;
    if (t.grp_off[2 * contig + 2] == t.grp_off[2 * contig]) return -1;  // no probe on this contig at all
    const int64_t base = t.grp_off[2 * contig + negative];
    const int size = (int)(t.grp_off[2 * contig + negative + 1] - base);
    const int32_t* st = t.start + base;
    const int32_t* en = t.end + base;
    const int32_t* va = t.value + base;
    const int find_start = min(a, b), find_stop = max(a, b);
    int n = 0;
    int lo = 1, hi = size;
    bool done = false;
    while (!done) {
        const int after = lo + ((hi - lo) / 2);
        if (after <= size) {
            const int after_loc = (after < size) ? st[after] : INT32_MAX;
            if (after_loc >= find_start) {
                if (st[after - 1] <= find_start) {
                    int cur = after - 1;
                    // "this currently assumes that no probe completely overlaps another probe" (:125): stop at the first
                    // window that ends before the query
                    while (cur >= 0 && en[cur] >= find_stop) {
                        if (st[cur] <= find_start && en[cur] >= find_stop) {
                            if (n < cap) out[n] = va[cur];
                            n++;
                        }
                        cur--;
                    }
                    done = true;
                } else {
                    hi = after - 1;
                }
            } else {
                lo = after + 1;
            }
            if (hi < lo) done = true;
        } else {
            done = true;
        }
    }
    return n;
}

struct DeviceRecords {
    const int32_t* contig;
    const uint8_t *negative, *second;
    const int32_t *start, *end, *read_index;
    const int64_t* alt_off;  // may be null
    const int32_t* alt_contig;
    const uint8_t* alt_negative;
    const int32_t *alt_start, *alt_stop;
};

// inserts v into the ascending unique list s[0..n) (capacity cap); returns the new size, sets *ovf when it does not fit
__device__ inline int sorted_insert(int32_t* s, int n, int cap, int32_t v, uint8_t* ovf) {
    int p = 0;
    while (p < n && s[p] < v) p++;
    if (p < n && s[p] == v) return n;
    if (n >= cap) { *ovf = 1; return n; }
    for (int k = n; k > p; k--) s[k] = s[k - 1];
    s[p] = v;
    return n + 1;
}

// one thread per record: own hits + alternative hits -> ascending unique probe list; sort key of the record
__global__ void assign_union_kernel(RangeTable t, DeviceRecords r, int64_t n, int K, int32_t* __restrict__ n_probes,
                                    int32_t* __restrict__ probes, uint8_t* __restrict__ overflow, uint8_t* __restrict__ usable,
                                    uint64_t* __restrict__ key, uint32_t* __restrict__ idx, int32_t* __restrict__ tmp) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) {
        int32_t* mine = probes + k * K;
        int32_t* scratch = tmp + k * K;
        uint8_t ovf = 0;
        int neg = r.negative[k] ? 1 : 0;
        if (r.second[k]) neg ^= 1;  // read two has the opposite strandedness of its probe (:452-455)
        int cnt = 0;
        int c = contain_query(t, r.contig[k], neg, r.start[k], r.end[k], scratch, K);
        const bool ok = c >= 0;
        if (ok) {
            if (c > K) { ovf = 1; c = K; }
            for (int q = 0; q < c; q++) cnt = sorted_insert(mine, cnt, K, scratch[q], &ovf);
            if (r.alt_off) {
                for (int64_t a = r.alt_off[k]; a < r.alt_off[k + 1]; a++) {  // getPotentialAlternativeHits (:547-576)
                    int an = r.alt_negative[a] ? 1 : 0;
                    if (r.second[k]) an ^= 1;
                    int ca = contain_query(t, r.alt_contig[a], an, r.alt_start[a], r.alt_stop[a], scratch, K);
                    if (ca > K) { ovf = 1; ca = K; }
                    for (int q = 0; q < ca; q++) cnt = sorted_insert(mine, cnt, K, scratch[q], &ovf);
                }
            }
        }
        n_probes[k] = cnt;
        overflow[k] = ovf;
        usable[k] = ok ? 1 : 0;
        // usable records group by read name; ignored ones get keys of their own (bit 32 set)
        key[k] = ok ? (uint64_t)((uint32_t)r.read_index[k] ^ 0x80000000u) : ((1ull << 32) | (uint64_t)(uint32_t)k);
        idx[k] = (uint32_t)k;
    }
}

// sorted position i is the head of a name when its key differs from the previous one: the head walks its records in input
// order and applies the first-found / second-found logic of :470-540 (run on one thread): 1 record -> unpaired; 2 or 3 ->
// the first two are the pair (a third stays unpaired); 4 or more -> the name completed a second pair and is removed (:392-400)
__global__ void assign_pair_kernel(const uint64_t* __restrict__ key, const uint32_t* __restrict__ idx, int64_t n, int K,
                                   const int32_t* __restrict__ n_probes, const int32_t* __restrict__ probes,
                                   const uint8_t* __restrict__ usable, int32_t* __restrict__ mate, int32_t* __restrict__ n_common,
                                   int32_t* __restrict__ common, uint8_t* __restrict__ overflow) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        if (i > 0 && key[i] == key[i - 1]) continue;
        int64_t e = i + 1;
        while (e < n && key[e] == key[i]) e++;
        const int64_t cnt = e - i;
        const uint32_t r0 = idx[i];
        if (!usable[r0]) { mate[r0] = HSQ_ASSIGN_IGNORED; n_common[r0] = 0; continue; }
        if (cnt >= 4) {
            for (int64_t k = i; k < e; k++) { mate[idx[k]] = HSQ_ASSIGN_AMBIGUOUS_NAME; n_common[idx[k]] = 0; }
            continue;
        }
        for (int64_t k = i; k < e; k++) { mate[idx[k]] = HSQ_ASSIGN_UNPAIRED; n_common[idx[k]] = 0; }
        if (cnt >= 2) {
            const uint32_t r1 = idx[i + 1];
            const int32_t* a = probes + (int64_t)r0 * K;
            const int32_t* b = probes + (int64_t)r1 * K;
            const int na = min(n_probes[r0], K), nb = min(n_probes[r1], K);
            int32_t* c0 = common + (int64_t)r0 * K;
            int32_t* c1 = common + (int64_t)r1 * K;
            int pa = 0, pb = 0, nc = 0;
            while (pa < na && pb < nb) {  // both ascending: containedProbesInBothReads (:478-483)
                if (a[pa] < b[pb]) pa++;
                else if (a[pa] > b[pb]) pb++;
                else { c0[nc] = a[pa]; c1[nc] = a[pa]; nc++; pa++; pb++; }
            }
            mate[r0] = (int32_t)r1;
            mate[r1] = (int32_t)r0;
            n_common[r0] = nc;
            n_common[r1] = nc;
            if (overflow[r0] | overflow[r1]) { overflow[r0] = 1; overflow[r1] = 1; }  // the intersection is incomplete too
        }
    }
}

}  // namespace hsq

using namespace hsq;

extern "C" int32_t hsq_set_probe_ranges(hsq_ctx* ctx, const hsq_probe_ranges* pr) {
    if (!ctx || !pr) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    const int64_t n = pr->n;
    if (n < 0 || n > INT32_MAX) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_set_probe_ranges: bad n");
    if (n > 0 && (!pr->contig || !pr->reverse_strand || !pr->start || !pr->stop)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null array");
    int32_t n_contigs = 0;
    for (int64_t p = 0; p < n; p++) {
        if (pr->contig[p] < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_set_probe_ranges: negative contig id");
        if (pr->contig[p] >= (1 << 28)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_set_probe_ranges: contig id too large");
        n_contigs = std::max(n_contigs, pr->contig[p] + 1);
    }
    // RangeMap.put normalises (min, max) (RangeMap.java:62-67); Collections.sort by start is stable (:98-101), so probes
    // with equal starts keep probe-file order inside their (contig, strand) map
    std::vector<int32_t> order((size_t)n);
    for (int64_t p = 0; p < n; p++) order[(size_t)p] = (int32_t)p;
    auto group = [&](int32_t p) { return (int64_t)pr->contig[p] * 2 + (pr->reverse_strand[p] ? 1 : 0); };
    auto lo = [&](int32_t p) { return std::min(pr->start[p], pr->stop[p]); };
    std::stable_sort(order.begin(), order.end(), [&](int32_t x, int32_t y) {
        const int64_t gx = group(x), gy = group(y);
        if (gx != gy) return gx < gy;
        return lo(x) < lo(y);
    });
    std::vector<int32_t> st((size_t)n), en((size_t)n), va((size_t)n);
    std::vector<int64_t> off((size_t)n_contigs * 2 + 1, 0);
    for (int64_t k = 0; k < n; k++) {
        const int32_t p = order[(size_t)k];
        st[(size_t)k] = lo(p);
        en[(size_t)k] = std::max(pr->start[p], pr->stop[p]);
        va[(size_t)k] = p;
        off[(size_t)group(p) + 1]++;
    }
    for (size_t g = 0; g + 1 < off.size(); g++) off[g + 1] += off[g];
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    struct { DevBuf* d; const void* h; size_t bytes; } up[] = {{&ctx->rg_start, st.data(), (size_t)n * 4}, {&ctx->rg_end, en.data(), (size_t)n * 4},
                                                             {&ctx->rg_value, va.data(), (size_t)n * 4}, {&ctx->rg_off, off.data(), off.size() * 8}};
    for (auto& c : up) {
        HSQ_CUDA(ctx, c.d->ensure(c.bytes + 16));
        if (c.bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(c.d->p, c.h, c.bytes, cudaMemcpyHostToDevice, ctx->stream));
    }
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the host vectors go away
    ctx->rg_n = n;
    ctx->rg_contigs = n_contigs;
    return HSQ_OK;
}

extern "C" int32_t hsq_assign_probes(hsq_ctx* ctx, const hsq_mapped_records* rec, hsq_assign_out* out) {
    if (!ctx || !rec || !out) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (ctx->rg_n < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_assign_probes: call hsq_set_probe_ranges first");
    const int64_t n = rec->n;
    const int K = out->max_probes;
    if (n < 0 || n >= (1ll << 31)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_assign_probes: bad n");
    if (K < 1 || K > 64) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_assign_probes: max_probes must be in 1..64");
    if (n == 0) return HSQ_OK;
    if (!rec->contig || !rec->negative_strand || !rec->second_of_pair || !rec->start || !rec->end || !rec->read_index)
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_assign_probes: null record array");
    if (!out->n_probes || !out->probes || !out->mate || !out->n_common || !out->common || !out->overflow)
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_assign_probes: null output array");
    const int64_t n_alt = rec->alt_offsets ? rec->alt_offsets[n] - rec->alt_offsets[0] : 0;
    if (n_alt < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_assign_probes: alt_offsets not ascending");
    if (n_alt > 0 && (!rec->alt_contig || !rec->alt_negative || !rec->alt_start || !rec->alt_stop))
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_assign_probes: null alternative-hit array");
    if (rec->alt_offsets && rec->alt_offsets[0] != 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_assign_probes: alt_offsets[0] must be 0");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    const size_t nn = (size_t)n, na = (size_t)n_alt;
    // inputs: one arena in ctx->misc (4-byte arrays first, then the offsets, then bytes)
    auto al = [](size_t b) { return (b + 15) & ~(size_t)15; };
    const size_t o_contig = 0, o_start = o_contig + al(nn * 4), o_end = o_start + al(nn * 4), o_ri = o_end + al(nn * 4);
    const size_t o_aoff = o_ri + al(nn * 4), o_ac = o_aoff + al((nn + 1) * 8), o_as = o_ac + al(na * 4), o_ae = o_as + al(na * 4);
    const size_t o_neg = o_ae + al(na * 4), o_sec = o_neg + al(nn), o_aneg = o_sec + al(nn), in_bytes = o_aneg + al(na);
    HSQ_CUDA(ctx, ctx->misc.ensure(in_bytes + 16));
    uint8_t* in = ctx->misc.as<uint8_t>();
    struct { size_t off; const void* h; size_t bytes; } up[] = {
        {o_contig, rec->contig, nn * 4}, {o_start, rec->start, nn * 4}, {o_end, rec->end, nn * 4}, {o_ri, rec->read_index, nn * 4},
        {o_aoff, rec->alt_offsets, rec->alt_offsets ? (nn + 1) * 8 : 0}, {o_ac, rec->alt_contig, na * 4}, {o_as, rec->alt_start, na * 4},
        {o_ae, rec->alt_stop, na * 4}, {o_neg, rec->negative_strand, nn}, {o_sec, rec->second_of_pair, nn}, {o_aneg, rec->alt_negative, na}};
    for (auto& c : up)
        if (c.bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(in + c.off, c.h, c.bytes, cudaMemcpyHostToDevice, ctx->stream));
    DeviceRecords dr{(const int32_t*)(in + o_contig), in + o_neg, in + o_sec, (const int32_t*)(in + o_start), (const int32_t*)(in + o_end),
                     (const int32_t*)(in + o_ri), rec->alt_offsets ? (const int64_t*)(in + o_aoff) : nullptr, (const int32_t*)(in + o_ac),
                     in + o_aneg, (const int32_t*)(in + o_as), (const int32_t*)(in + o_ae)};
    RangeTable rt{ctx->rg_start.as<int32_t>(), ctx->rg_end.as<int32_t>(), ctx->rg_value.as<int32_t>(), ctx->rg_off.as<int64_t>(), ctx->rg_contigs};
    // outputs: out_i32 = n_probes | mate | n_common | probes[n][K] | common[n][K] | tmp[n][K]; out_u8 = overflow | usable
    const size_t slots = nn * (size_t)K;
    HSQ_CUDA(ctx, ctx->out_i32.ensure((3 * nn + 3 * slots) * 4 + 64));
    HSQ_CUDA(ctx, ctx->out_u8.ensure(2 * nn + 16));
    int32_t* d_np = ctx->out_i32.as<int32_t>();
    int32_t* d_mate = d_np + nn;
    int32_t* d_nc = d_mate + nn;
    int32_t* d_probes = d_nc + nn;
    int32_t* d_common = d_probes + slots;
    int32_t* d_tmp = d_common + slots;
    uint8_t* d_ovf = ctx->out_u8.as<uint8_t>();
    uint8_t* d_usable = d_ovf + nn;
    HSQ_CUDA(ctx, ctx->sort_a.ensure(nn * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_b.ensure(nn * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_c.ensure(nn * 4 + 16));
    HSQ_CUDA(ctx, ctx->sort_d.ensure(nn * 4 + 16));
    uint64_t* key_in = ctx->sort_a.as<uint64_t>();
    uint64_t* key_out = ctx->sort_b.as<uint64_t>();
    uint32_t* idx_in = ctx->sort_c.as<uint32_t>();
    uint32_t* idx_out = ctx->sort_d.as<uint32_t>();
    const int blocks = (int)std::min<int64_t>((n + 127) / 128, (int64_t)ctx->sm_count * 16);
    assign_union_kernel<<<blocks, 128, 0, ctx->stream>>>(rt, dr, n, K, d_np, d_probes, d_ovf, d_usable, key_in, idx_in, d_tmp);
    ctx->launches++;
    HSQ_CUDA(ctx, cudaGetLastError());
    int rc = sort_pass(ctx, key_in, key_out, idx_in, idx_out, n, 33);  // stable: input order survives inside a read name
    if (rc) return rc;
    assign_pair_kernel<<<blocks, 128, 0, ctx->stream>>>(key_out, idx_out, n, K, d_np, d_probes, d_usable, d_mate, d_nc, d_common, d_ovf);
    ctx->launches++;
    HSQ_CUDA(ctx, cudaGetLastError());
    struct { void* h; const void* d; size_t bytes; } down[] = {{out->n_probes, d_np, nn * 4},   {out->mate, d_mate, nn * 4},
                                                               {out->n_common, d_nc, nn * 4},   {out->probes, d_probes, slots * 4},
                                                               {out->common, d_common, slots * 4}, {out->overflow, d_ovf, nn}};
    for (auto& c : down) HSQ_CUDA(ctx, cudaMemcpyAsync(c.h, c.d, c.bytes, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return timer.finish();
}
