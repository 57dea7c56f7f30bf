// hsq_capi_probes.inl -- hsq_set_probes: the parsed probe file becomes a device-resident table.
namespace hsq {

// Probe.getProbeSequence (COMMON/utils/probeinfo/Probe.java:83-88): extension primer ++ capture target ++ ligation primer
__global__ void concat_probe_kernel(DeviceSeqs ext, DeviceSeqs target, DeviceSeqs lig, int64_t n, uint8_t* __restrict__ out,
                                    int64_t* __restrict__ out_offsets) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k > n) return;
    int64_t base = ext.offsets[k] + target.offsets[k] + lig.offsets[k];
    out_offsets[k] = base;
    if (k == n) return;
    const DeviceSeqs* parts[3] = {&ext, &target, &lig};
    for (int p = 0; p < 3; p++) {
        int64_t b = parts[p]->offsets[k], e = parts[p]->offsets[k + 1];
        for (int64_t i = b; i < e; i++) out[base++] = parts[p]->codes[i];
    }
}

}  // namespace hsq

extern "C" int32_t hsq_set_probes(hsq_ctx* ctx, const hsq_probe_table* t) {
    if (!ctx || !t) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (t->n < 0 || t->ext.n != t->n || t->lig.n != t->n || t->target.n != t->n)
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "probe table arrays disagree on n");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    ctx->have_probes = false;
    int rc = reset_status(ctx);
    if (rc) return rc;
    const size_t n = (size_t)t->n;
    struct { hsq::DevBuf* d; const void* h; size_t bytes; } cols[] = {
        {&ctx->p_rev, t->reverse_strand, n},
        {&ctx->p_ext_start, t->ext_start, n * 4}, {&ctx->p_ext_stop, t->ext_stop, n * 4},
        {&ctx->p_lig_start, t->lig_start, n * 4}, {&ctx->p_lig_stop, t->lig_stop, n * 4}};
    for (auto& c : cols) {
        if (n && !c.h) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null probe column");
        HSQ_CUDA(ctx, c.d->ensure(c.bytes + 16));
        if (c.bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(c.d->p, c.h, c.bytes, cudaMemcpyHostToDevice, ctx->stream));
    }
    if ((rc = upload_seqs(ctx, &t->ext, &ctx->p_ext, d_status(ctx), d_maxlen(ctx, 0)))) return rc;
    if ((rc = upload_seqs(ctx, &t->lig, &ctx->p_lig, d_status(ctx), d_maxlen(ctx, 1)))) return rc;
    if ((rc = upload_seqs(ctx, &t->target, &ctx->p_target, d_status(ctx), d_maxlen(ctx, 2)))) return rc;
    ctx->p_full.n = t->n;
    ctx->p_full.total = ctx->p_ext.total + ctx->p_lig.total + ctx->p_target.total;
    HSQ_CUDA(ctx, ctx->p_full.codes.ensure((size_t)ctx->p_full.total + 32));
    HSQ_CUDA(ctx, ctx->p_full.offsets.ensure((n + 1) * sizeof(int64_t)));
    concat_probe_kernel<<<(unsigned)((n + 1 + 127) / 128), 128, 0, ctx->stream>>>(ctx->p_ext.view(), ctx->p_target.view(), ctx->p_lig.view(),
                                                                                  t->n, ctx->p_full.codes.as<uint8_t>(),
                                                                                  ctx->p_full.offsets.as<int64_t>());
    int w[4];
    if ((rc = read_status(ctx, w, 4))) return rc;
    if (w[0] != 0) return fail(ctx, w[0], std::string("hsq_set_probes: ") + status_message(w[0]));
    ctx->p_ext.max_len = w[1];
    ctx->p_lig.max_len = w[2];
    ctx->p_target.max_len = w[3];
    ctx->p_full.max_len = w[1] + w[2] + w[3];
    ctx->n_probes = t->n;
    ctx->have_probes = true;
    return HSQ_OK;
}
