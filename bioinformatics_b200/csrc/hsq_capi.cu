// hsq_capi.cu -- the C ABI of include/hsq_b200.h.  Host buffers in, host buffers out; all compute on the GPU.
// There is no CPU fallback: every entry point needs a context, and a context needs a CUDA device.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include "hsq_host.cuh"
#include "hsq_extend.cuh"

using namespace hsq;

namespace {

struct CallTimer {  // brackets the kernels + copies of one batch call with events on the context's stream
    hsq_ctx* ctx;
    explicit CallTimer(hsq_ctx* c) : ctx(c) {
        ctx->launches = 0;
        for (int k = 0; k < hsq_ctx::N_SPANS; k++) ctx->span_open[k] = ctx->span_done[k] = false;
        cudaEventRecord(ctx->ev0, ctx->stream);
    }
    int finish() {
        cudaEventRecord(ctx->ev1, ctx->stream);
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) return fail(ctx, HSQ_ERR_CUDA, std::string("stream sync: ") + cudaGetErrorString(e));
        cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->ev0, ctx->ev1);
        for (int k = 0; k < hsq_ctx::N_SPANS; k++) {
            ctx->span_ms[k] = -1.f;
            if (ctx->span_done[k]) cudaEventElapsedTime(&ctx->span_ms[k], ctx->ev_span[k][0], ctx->ev_span[k][1]);
        }
        ctx->last_launches = ctx->launches;
        return HSQ_OK;
    }
};

// Kernel-only spans: the events sit on the context's stream between the copies and the kernels of a call, so a span is the device
// time of the kernels between them (the H2D copies before span_begin have completed when its event fires).
void span_begin(hsq_ctx* ctx, int k) {
    cudaEventRecord(ctx->ev_span[k][0], ctx->stream);
    ctx->span_open[k] = true;
}
void span_end(hsq_ctx* ctx, int k) {
    if (!ctx->span_open[k]) return;
    cudaEventRecord(ctx->ev_span[k][1], ctx->stream);
    ctx->span_done[k] = true;
}

int reset_status(hsq_ctx* ctx) {
    HSQ_CUDA(ctx, ctx->status.ensure(64));
    HSQ_CUDA(ctx, ctx->cells.ensure(128));
    // words 0..2 (error, max lengths) always; word 3 (slot overflow) and the cell counter only at the start of a call
    HSQ_CUDA(ctx, cudaMemsetAsync(ctx->status.p, 0, ctx->keep_counters ? 12 : 64, ctx->stream));
    if (!ctx->keep_counters) HSQ_CUDA(ctx, cudaMemsetAsync(ctx->cells.p, 0, 128, ctx->stream));
    return HSQ_OK;
}

// status word 0 = error code, words 1..2 = max lengths, word 3 = "an output slot was too small"
int* d_status(hsq_ctx* ctx) { return ctx->status.as<int>(); }
int* d_maxlen(hsq_ctx* ctx, int slot) { return ctx->status.as<int>() + 1 + slot; }

int read_status(hsq_ctx* ctx, int* host_words, int n_words) {
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->h_status, ctx->status.p, sizeof(int) * (size_t)n_words, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < n_words; k++) host_words[k] = ctx->h_status[k];
    return HSQ_OK;
}

// resident threads of a persistent launch
template <class K>
int persistent_grid(hsq_ctx* ctx, K kernel, int block, int64_t n_tasks, int* grid_out) {
    int per_sm = 0;
    HSQ_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block, 0));
    if (per_sm < 1) per_sm = 1;
    int64_t grid = (int64_t)per_sm * ctx->sm_count;
    int64_t need = (n_tasks + block - 1) / block;
    if (need < 1) need = 1;
    *grid_out = (int)std::min(grid, need);
    return HSQ_OK;
}

// Scratch of the DP kernels: `nthreads` resident threads of the 32-bit kernels and `nthreads2` of the packed kernels
// (which run one after the other on the context's stream and share the same allocations).
int alloc_scratch(hsq_ctx* ctx, int64_t nthreads, int64_t nthreads2, int q_max, int r_max, ScratchPool* pool, ScratchPool2* pool2) {
    size_t n_strips = (size_t)(q_max + KROWS - 1) / KROWS;
    if (n_strips < 1) n_strips = 1;
    const size_t rs = (size_t)r_max + 1, t1 = (size_t)nthreads, t2 = (size_t)nthreads2;
    HSQ_CUDA(ctx, ctx->ptr_words.ensure(std::max(t1 * sizeof(uint32_t), t2 * sizeof(uint2)) * n_strips * rs));
    HSQ_CUDA(ctx, ctx->brow.ensure(std::max(t1, t2) * (rs + 2) * sizeof(int32_t)));  // two spare columns: the packed fill prefetches the row above two columns ahead without a bounds test
    HSQ_CUDA(ctx, ctx->cols.ensure(t1 * (size_t)(q_max + r_max + 1) * sizeof(uint16_t) + 16));
    HSQ_CUDA(ctx, ctx->colinfo.ensure(t2 * ((rs + 3) / 4 + 2) * sizeof(uint32_t) + 16));  // one byte per column, packed four to a word
    HSQ_CUDA(ctx, ctx->slow.ensure(t2 * KROWS * sizeof(uint32_t) + 16));
    pool->r_stride = r_max + 1;
    pool->ptr_words = ctx->ptr_words.as<uint32_t>();
    pool->brow = ctx->brow.as<int32_t>();
    pool->cols = ctx->cols.as<uint16_t>();
    if (pool2) {
        pool2->r_stride = r_max + 1;
        pool2->ptr_words = ctx->ptr_words.as<uint2>();
        pool2->brow = ctx->brow.as<uint32_t>();
        pool2->colinfo = ctx->colinfo.as<uint32_t>();
        pool2->nthreads = 0;  // set at every launch site
        pool2->slow = ctx->slow.as<uint32_t>();
    }
    return HSQ_OK;
}

DeviceProbes probe_view(const hsq_ctx* ctx) {
    DeviceProbes p;
    p.reverse_strand = ctx->p_rev.as<uint8_t>();
    p.ext_start = ctx->p_ext_start.as<int32_t>();
    p.ext_stop = ctx->p_ext_stop.as<int32_t>();
    p.lig_start = ctx->p_lig_start.as<int32_t>();
    p.lig_stop = ctx->p_lig_stop.as<int32_t>();
    p.ext = ctx->p_ext.view();
    p.lig = ctx->p_lig.view();
    p.target = ctx->p_target.view();
    return p;
}

void comm_release(hsq_ctx* ctx);  // hsq_capi_comm.inl

template <class T>
int d2h(hsq_ctx* ctx, T* host, const T* dev, size_t count) {
    if (!host || count == 0) return HSQ_OK;
    HSQ_CUDA(ctx, cudaMemcpyAsync(host, dev, count * sizeof(T), cudaMemcpyDeviceToHost, ctx->io ? ctx->io : ctx->stream));
    return HSQ_OK;
}

}  // namespace

extern "C" {

const char* hsq_version(void) { return "hsq-b200 0.1 (sm_100a)"; }

const char* hsq_last_error(const hsq_ctx* ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

static int ctx_init(hsq_ctx* ctx, const cudaDeviceProp& prop) {
    ctx->sm_count = prop.multiProcessorCount;
    ctx->l2_window_max = (size_t)prop.accessPolicyMaxWindowSize;
    ctx->l2_persist_max = (size_t)prop.persistingL2CacheMaxSize;
    HSQ_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    HSQ_CUDA(nullptr, cudaEventCreate(&ctx->ev0));
    HSQ_CUDA(nullptr, cudaEventCreate(&ctx->ev1));
    HSQ_CUDA(nullptr, cudaEventCreate(&ctx->ev_user0));
    HSQ_CUDA(nullptr, cudaEventCreate(&ctx->ev_user1));
    for (int k = 0; k < hsq_ctx::N_SPANS; k++)
        for (int e = 0; e < 2; e++) HSQ_CUDA(nullptr, cudaEventCreate(&ctx->ev_span[k][e]));
    HSQ_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking));
    HSQ_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking));
    ctx->trim_events.assign(17, nullptr);
    for (auto& ev : ctx->trim_events) HSQ_CUDA(nullptr, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    {   // the prep stream's short kernels should get SM slots ahead of the capture kernels' CTAs
        int lo = 0, hi = 0;
        HSQ_CUDA(nullptr, cudaDeviceGetStreamPriorityRange(&lo, &hi));
        HSQ_CUDA(nullptr, cudaStreamCreateWithPriority(&ctx->s_prep, cudaStreamNonBlocking, hi));
    }
    for (int k = 0; k < 2; k++) HSQ_CUDA(nullptr, cudaStreamCreateWithFlags(&ctx->s_cap[k], cudaStreamNonBlocking));
    for (int k = 0; k < hsq_ctx::N_EXTEND_SETS; k++) {
        HSQ_CUDA(nullptr, cudaEventCreateWithFlags(&ctx->ev_in[k], cudaEventDisableTiming));
        HSQ_CUDA(nullptr, cudaEventCreateWithFlags(&ctx->ev_prep[k], cudaEventDisableTiming));
        HSQ_CUDA(nullptr, cudaEventCreateWithFlags(&ctx->ev_cap[k], cudaEventDisableTiming));
        HSQ_CUDA(nullptr, cudaEventCreateWithFlags(&ctx->ev_out[k], cudaEventDisableTiming));
    }
    HSQ_CUDA(nullptr, cudaMallocHost((void**)&ctx->h_status, 64));
    HSQ_CUDA(nullptr, cudaMallocHost((void**)&ctx->h_cells, 128));
    std::vector<double> tbl(POW95_N);
    for (int d = 0; d < POW95_N; d++) tbl[(size_t)d] = std::pow(0.95, (double)d);
    HSQ_CUDA(nullptr, cudaMemcpyToSymbol(c_pow95, tbl.data(), sizeof(double) * POW95_N));
    return HSQ_OK;
}

static int apply_config(hsq_ctx* ctx, const hsq_config* config) {
    hsq_config c{};
    if (config) {
        if (config->struct_bytes <= 0 || config->struct_bytes > (int32_t)sizeof(hsq_config))
            return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_config.struct_bytes must be sizeof(hsq_config) of the caller's header");
        std::memcpy(&c, config, (size_t)config->struct_bytes);  // an older, shorter struct leaves the newer fields at their defaults
    }
    if (c.extend_chunks < 0 || c.band < 0 || c.trim_chunks < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative value in hsq_config");
    ctx->cfg = c;
    return HSQ_OK;
}

int32_t hsq_ctx_create_ex(int32_t device, const hsq_config* config, hsq_ctx** out) {
    if (!out) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "out is null");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, HSQ_ERR_NO_DEVICE, std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "count is 0") +
                                                    " (this library has no CPU fallback)");
    if (device < 0 || device >= count) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "device index out of range");
    cudaDeviceProp prop;
    HSQ_CUDA(nullptr, cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail(nullptr, HSQ_ERR_NO_DEVICE, std::string("device ") + prop.name + " is not sm_100 class; kernels are built for sm_100a only");
    HSQ_CUDA(nullptr, cudaSetDevice(device));
    hsq_ctx* ctx = new hsq_ctx();
    ctx->device = device;
    int rc = ctx_init(ctx, prop);
    if (!rc && (rc = apply_config(ctx, config))) g_create_error = ctx->error;
    if (rc) {  // nothing created so far may leak
        hsq_ctx_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return HSQ_OK;
}

int32_t hsq_ctx_create(int32_t device, hsq_ctx** out) { return hsq_ctx_create_ex(device, nullptr, out); }

int32_t hsq_ctx_configure(hsq_ctx* ctx, const hsq_config* config) {
    if (!ctx) return HSQ_ERR_INVALID_ARGUMENT;
    return apply_config(ctx, config);
}

void hsq_ctx_destroy(hsq_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    hsq::DevBuf* bufs[] = {&ctx->p_rev, &ctx->p_ext_start, &ctx->p_ext_stop, &ctx->p_lig_start, &ctx->p_lig_stop, &ctx->q1, &ctx->q2,
                           &ctx->probe_id, &ctx->read_index, &ctx->ptr_words, &ctx->brow, &ctx->cols, &ctx->status, &ctx->cells,
                           &ctx->misc, &ctx->out_i32, &ctx->out_u8, &ctx->out_slots, &ctx->pair_ok, &ctx->sort_a, &ctx->sort_b,
                           &ctx->sort_c, &ctx->sort_d, &ctx->sort_tmp, &ctx->flags, &ctx->slow, &ctx->colinfo, &ctx->kf_pos, &ctx->kf_local, &ctx->alt_probe_id, &ctx->rg_start, &ctx->rg_end,
                           &ctx->rg_value, &ctx->rg_off, &ctx->dd_u8, &ctx->dd_i32, &ctx->fi_start, &ctx->fi_len, &ctx->fi_refs, &ctx->fi_size, &ctx->pool_tmp, &ctx->pool_cursor, &ctx->pool_dev,
                           &ctx->cm_send, &ctx->cm_recv, &ctx->cm_i32, &ctx->cm_cnt, &ctx->cm_tie, &ctx->cm_probe};
    for (auto* b : bufs) b->release();
    for (auto& b : ctx->ing) b.release();
    ctx->dd_seg.release();
    ctx->dd_vseg.release();
    hsq::DevSeqsOwned* seqs[] = {&ctx->p_ext, &ctx->p_lig, &ctx->p_target, &ctx->p_full, &ctx->r1, &ctx->r2, &ctx->sa, &ctx->sb, &ctx->alt_r1, &ctx->alt_r2, &ctx->fs};
    for (auto* s : seqs) s->release();
    for (auto& ps : ctx->parked) {
        hsq::DevBuf* pb[] = {&ps.probe_id, &ps.out_i32, &ps.out_u8, &ps.out_slots, &ps.pair_ok, &ps.ptr_words, &ps.brow, &ps.cols, &ps.colinfo,
                             &ps.slow, &ps.sort_a, &ps.sort_b, &ps.sort_c};
        for (auto* b : pb) b->release();
        ps.r1.release();
        ps.r2.release();
    }
    if (ctx->h_status) cudaFreeHost(ctx->h_status);
    if (ctx->h_cells) cudaFreeHost(ctx->h_cells);
    for (cudaEvent_t ev : {ctx->ev0, ctx->ev1, ctx->ev_user0, ctx->ev_user1})
        if (ev) cudaEventDestroy(ev);
    for (int k = 0; k < hsq_ctx::N_SPANS; k++)
        for (int e = 0; e < 2; e++)
            if (ctx->ev_span[k][e]) cudaEventDestroy(ctx->ev_span[k][e]);
    for (int k = 0; k < hsq_ctx::N_EXTEND_SETS; k++)
        for (cudaEvent_t ev : {ctx->ev_in[k], ctx->ev_prep[k], ctx->ev_cap[k], ctx->ev_out[k]})
            if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : ctx->trim_events)
        if (ev) cudaEventDestroy(ev);
    for (cudaStream_t st : {ctx->s_cap[0], ctx->s_cap[1], ctx->s_prep, ctx->s_in, ctx->s_out, ctx->stream})
        if (st) cudaStreamDestroy(st);
    comm_release(ctx);
    delete ctx;
}

int32_t hsq_pinned_alloc(hsq_ctx* ctx, size_t bytes, void** out) {
    if (!ctx || !out) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    HSQ_CUDA(ctx, cudaMallocHost(out, bytes ? bytes : 1));
    return HSQ_OK;
}

int32_t hsq_pinned_free(hsq_ctx* ctx, void* p) {
    if (!ctx) return HSQ_ERR_INVALID_ARGUMENT;
    HSQ_CUDA(ctx, cudaFreeHost(p));
    return HSQ_OK;
}

int32_t hsq_set_scorer(hsq_ctx* ctx, const hsq_scorer* scorer) {
    if (!ctx || !scorer) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    ctx->scorer = *scorer;
    ctx->have_scorer = true;
    return HSQ_OK;
}

int32_t hsq_last_spans(const hsq_ctx* ctx, float* ms4) {
    if (!ctx || !ms4) return HSQ_ERR_INVALID_ARGUMENT;
    for (int k = 0; k < hsq_ctx::N_SPANS; k++) ms4[k] = ctx->span_ms[k];
    return HSQ_OK;
}

int32_t hsq_band_stats(const hsq_ctx* ctx, int64_t* banded, int64_t* redone) {
    if (!ctx) return HSQ_ERR_INVALID_ARGUMENT;
    if (banded) *banded = ctx->band_stats[0];
    if (redone) *redone = ctx->band_stats[1];
    return HSQ_OK;
}

int32_t hsq_last_timing(const hsq_ctx* ctx, float* kernel_ms, int32_t* launches) {
    if (!ctx) return HSQ_ERR_INVALID_ARGUMENT;
    if (kernel_ms) *kernel_ms = ctx->last_kernel_ms;
    if (launches) *launches = ctx->last_launches;
    return HSQ_OK;
}

}  // extern "C"

#include "hsq_capi_probes.inl"
#include "hsq_capi_align.inl"
#include "hsq_capi_a10.inl"
#include "hsq_capi_extend.inl"
#include "hsq_capi_trim.inl"
#include "hsq_capi_dedup.inl"
#include "hsq_capi_merge.inl"
#include "hsq_capi_resolve.inl"
#include "hsq_capi_assign.inl"
#include "hsq_capi_packed.inl"
#include "hsq_capi_comm.inl"
#include "hsq_capi_pipeline.inl"
#include "hsq_capi_keysfirst.inl"
#include "hsq_capi_records.inl"
#include "hsq_capi_find.inl"
#include "hsq_capi_ingest.inl"
#include "hsq_capi_bam.inl"
#include "hsq_capi_peak.inl"
#include "hsq_capi_noalloc.inl"
