// hsq_host.cuh -- host-side plumbing shared by the C ABI entry points: context, grow-only device buffers,
// error capture.  No compute.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <algorithm>
#include <string>
#include <vector>
#include <functional>
#include <cstdio>
#include "hsq_extend.cuh"

namespace hsq {

struct DevBuf {  // grow-only device allocation reused across batch calls
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct DevSeqsOwned {  // encoded ragged batch resident on the device
    DevBuf ascii, codes, offsets;
    int64_t n = 0, total = 0;
    int max_len = 0;
    // packed staging (hsq_packed_reads): `ascii` then holds the packed words, `lens` the read lengths, `word_off` the first word of
    // every read (device scan), `exc` the exception records of a 2-bit batch
    DevBuf lens, word_off, scan_tmp, exc;
    int packed_bits = 0;       // 0: `ascii` holds ASCII
    int64_t packed_words = 0;  // 32-bit words of packed data
    int64_t n_exc = 0;
    int64_t exc_read_base = 0; // exception read numbers are relative to the caller's batch: first read of this window
    DeviceSeqs view() const { return DeviceSeqs{codes.as<uint8_t>(), offsets.as<int64_t>()}; }
    void release() { ascii.release(); codes.release(); offsets.release(); lens.release(); word_off.release(); scan_tmp.release(); exc.release(); }
};

}  // namespace hsq

static constexpr int BAND_STATS_SLOT = 8;  // words 8, 9 of hsq_ctx::cells: read ends aligned in the band, of them redone in full

struct hsq_ctx {
    int device = 0;
    int sm_count = 0;
    size_t l2_window_max = 0, l2_persist_max = 0;  // cudaDeviceProp: accessPolicyMaxWindowSize, persistingL2CacheMaxSize
    bool l2_limit_set = false;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_user0 = nullptr, ev_user1 = nullptr;
    std::string error;
    // kernel-only spans of the last batch call (events around groups of kernels, copies excluded): hsq_last_spans
    static constexpr int N_SPANS = 4;
    cudaEvent_t ev_span[N_SPANS][2] = {};
    bool span_open[N_SPANS] = {}, span_done[N_SPANS] = {};
    float span_ms[N_SPANS] = {};
    float last_kernel_ms = 0.f;
    int last_launches = 0;
    int launches = 0;  // running count inside a call
    bool have_scorer = false, have_probes = false;
    hsq_scorer scorer{1, -4, -1, -6};
    // probe table
    int64_t n_probes = 0;
    hsq::DevBuf p_rev, p_ext_start, p_ext_stop, p_lig_start, p_lig_stop;
    hsq::DevSeqsOwned p_ext, p_lig, p_target, p_full;  // p_full = ext ++ target ++ lig (Probe.getProbeSequence)
    // batch inputs
    hsq::DevSeqsOwned r1, r2, sa, sb;
    hsq::DevBuf q1, q2, probe_id, read_index;
    // scratch + outputs
    hsq::DevBuf ptr_words, brow, cols, status, cells, misc, out_i32, out_u8, out_slots, pair_ok;
    hsq::DevBuf sort_a, sort_b, sort_c, sort_d, sort_tmp, flags, slow, colinfo;
    hsq_config cfg{};  // hsq_ctx_create_ex / hsq_ctx_configure
    // probe search windows sorted by (contig, strand, start) for hsq_assign_probes
    hsq::DevBuf rg_start, rg_end, rg_value, rg_off;
    hsq::DevBuf dd_seg;         // dedup segment path: per-probe counters, offsets
    hsq::DevBuf dd_vseg;        // the same per UID-hash bucket when oversized probes are split
    // in-library multi-GPU exchange (hsq_capi_comm.inl): NCCL communicator of this context, exchange buffers
    void* comm = nullptr;
    int comm_rank = 0, comm_world = 1;
    hsq::DevBuf cm_send, cm_recv, cm_i32, cm_cnt, cm_tie, cm_probe;
    int64_t* h_counts = nullptr;  // pinned: the world x (world + 1) count matrix
    hsq::DevBuf pool_tmp, pool_cursor, pool_dev;  // pooled CIGAR / MD delivery (hsq_extend_packed)
    hsq::DevBuf kf_pos, kf_local;  // hsq_dedup_extend_packed_keys: positions of the kept pairs, a chunk's local survivor list
    hsq::DevBuf dd_u8, dd_i32;  // hsq_dedup_extend_batch: dedup answers kept while the extend stage reuses the output buffers
    int64_t rg_n = -1;
    int32_t rg_contigs = 0;
    // k-mer index of the sequences to find (hsq_set_find_sequences): refs of k-mer c are fi_refs[fi_start[c] .. + fi_len[c])
    hsq::DevSeqsOwned fs;
    hsq::DevBuf fi_start, fi_len, fi_refs, fi_size;
    int64_t fi_n = -1;
    hsq::DevBuf ing[9];  // hsq_join_names: inputs, counters and one working arena
    int32_t fi_k = 0, fi_max_len = 0;
    // hsq_extend_batch pipelines chunks through  H2D | prep (encode, primer DP, bucket sorts) | capture DP | D2H  with three
    // complete buffer sets; the two sets not in the members above are parked here (see activate_extend_set)
    struct ExtendSet {
        hsq::DevSeqsOwned r1, r2;
        hsq::DevBuf probe_id, out_i32, out_u8, out_slots, pair_ok, ptr_words, brow, cols, colinfo, slow, sort_a, sort_b, sort_c;
        hsq::ExtendDeviceOut ext_out{};
        int64_t ext_n = -1;
    };
    static constexpr int N_EXTEND_SETS = 3;
    ExtendSet parked[N_EXTEND_SETS];
    int active_set = 0;
    hsq::DevSeqsOwned alt_r1, alt_r2;  // destination of the device-side compaction (hsq_dedup_extend_batch)
    hsq::DevBuf alt_probe_id;
    cudaStream_t s_cap[2] = {nullptr, nullptr};  // capture kernels of even / odd chunks (a tail overlaps the next start)
    cudaStream_t s_prep = nullptr;               // everything before the capture kernel, one chunk ahead
    int work_slot = 0;                           // which pair of work counters the packed kernels of this run use
    cudaStream_t io = nullptr;      // stream the host<->device copies of the current call go to (default: stream)
    cudaStream_t s_in = nullptr, s_out = nullptr;
    std::vector<cudaEvent_t> trim_events;       // hsq_trim_batch's chunk pipeline: [0] phase boundary, then (copied in, kernel done) per chunk
    cudaEvent_t ev_in[N_EXTEND_SETS] = {}, ev_prep[N_EXTEND_SETS] = {}, ev_cap[N_EXTEND_SETS] = {}, ev_out[N_EXTEND_SETS] = {};
    bool keep_counters = false;     // chunks after the first: keep the cell counter and the overflow flag
    int64_t ext_n = -1;  // staged extend batch
    bool ext_ran = false;
    hsq::ExtendDeviceOut ext_out{};
    int* h_status = nullptr;  // pinned
    unsigned long long* h_cells = nullptr;
    int64_t band_stats[2] = {0, 0};              // last extend call: read ends aligned in the band, of them redone with full matrices
};

namespace hsq {

static thread_local std::string g_create_error;

inline int fail(hsq_ctx* ctx, int code, const std::string& msg) {
    if (ctx) ctx->error = msg;
    else g_create_error = msg;
    return code;
}

#define HSQ_CUDA(ctx, expr)                                                                               \
    do {                                                                                                  \
        cudaError_t _e = (expr);                                                                          \
        if (_e != cudaSuccess)                                                                            \
            return hsq::fail(ctx, HSQ_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));    \
    } while (0)

inline const char* status_message(int st) {
    switch (st) {
        case HSQ_ERR_BAD_CHARACTER: return "a sequence holds a character that is not an IUPAC nucleotide code";
        case HSQ_ERR_OUT_OF_RANGE: return "index out of range (the reference would throw here)";
        case HSQ_ERR_OVERFLOW: return "an output slot was too small; see the overflow flags";
        case HSQ_ERR_UNSUPPORTED: return "input outside the accelerated domain";
        default: return "error";
    }
}

__global__ void rebase_offsets_kernel(int64_t* __restrict__ offsets, int64_t n, int64_t base) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) offsets[i] -= base;
}

// host -> device copy of a ragged ASCII batch on the context's copy stream.  offsets need not start at 0 (a window of a
// larger batch): the device copy is rebased.
inline int copy_seqs(hsq_ctx* ctx, const hsq_seqs* s, DevSeqsOwned* d) {
    if (!s || s->n < 0 || !s->offsets) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null sequence batch");
    const int64_t base = s->offsets[0];
    if (base < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "offsets must not be negative");
    if (s->n > 0 && (s->offsets[1] < base || s->offsets[s->n] < s->offsets[s->n - 1]))  // the interior is checked on the device (max_len_kernel / host_chunk_dims)
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "offsets must not decrease");
    d->n = s->n;
    d->total = s->offsets[s->n] - base;
    if (d->total < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative batch size");
    if (d->total > 0 && !s->data) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null sequence data");
    d->packed_bits = 0;
    // 32 bytes of slack: the strip / walker loads of the DP kernels read whole aligned words around a sequence's ends
    HSQ_CUDA(ctx, d->ascii.ensure((size_t)d->total + 32));
    HSQ_CUDA(ctx, d->codes.ensure((size_t)d->total + 32));
    HSQ_CUDA(ctx, d->offsets.ensure((size_t)(s->n + 1) * sizeof(int64_t)));
    cudaStream_t st = ctx->io ? ctx->io : ctx->stream;
    if (d->total > 0) HSQ_CUDA(ctx, cudaMemcpyAsync(d->ascii.p, s->data + base, (size_t)d->total, cudaMemcpyHostToDevice, st));
    HSQ_CUDA(ctx, cudaMemcpyAsync(d->offsets.p, s->offsets, (size_t)(s->n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st));
    if (base != 0) {
        const int blocks = (int)std::min<int64_t>((s->n + 256) / 256, (int64_t)ctx->sm_count * 4);
        rebase_offsets_kernel<<<blocks, 256, 0, st>>>(d->offsets.as<int64_t>(), s->n + 1, base);
        ctx->launches++;
    }
    return HSQ_OK;
}

// ASCII -> device codes of a staged batch (+ optional max length into a device int)
inline int encode_seqs(hsq_ctx* ctx, DevSeqsOwned* d, int* d_status, int* d_maxlen) {
    if (d->total > 0) {
        int64_t threads = (d->total + 15) / 16;
        int blocks = (int)std::min<int64_t>((threads + 255) / 256, (int64_t)ctx->sm_count * 16);
        encode_kernel<<<blocks, 256, 0, ctx->stream>>>(d->ascii.as<uint8_t>(), d->codes.as<uint8_t>(), d->total, d_status);
        ctx->launches++;
    }
    if (d_maxlen && d->n > 0) {
        int blocks = (int)std::min<int64_t>((d->n + 255) / 256, (int64_t)ctx->sm_count * 8);
        max_len_kernel<<<blocks, 256, 0, ctx->stream>>>(d->offsets.as<int64_t>(), d->n, d_maxlen);
        ctx->launches++;
    }
    return HSQ_OK;
}

inline int upload_seqs(hsq_ctx* ctx, const hsq_seqs* s, DevSeqsOwned* d, int* d_status, int* d_maxlen) {
    int rc = copy_seqs(ctx, s, d);
    if (rc) return rc;
    return encode_seqs(ctx, d, d_status, d_maxlen);
}

}  // namespace hsq
