// hsq_capi_noalloc.inl -- hsq_align_scores_noalloc: the operator seam's DP (align_kernel's fill + traceback) for callers that must
// not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has no
// counterpart in the reference.  Every other entry point of this library sizes
// grow-only device buffers with cudaMalloc on first use; this one calls no cudaMalloc / cudaFree / allocating routine at all: the
// DP scratch (pointer words, boundary rows) and the encoded copies of the inputs live in ONE file-scope __device__ array of fixed
// size, which every GPU that loads the library holds until the process exits.
//   - the array is shared by every call on a device and keeps the previous call's contents; nothing in it needs to start at zero (the
//     fill writes every pointer word and boundary cell before the traceback reads it, the encoder writes every code byte before the DP
//     reads it); the caller's status word is cleared at the start of each call;
//   - two calls that overlap on the same GPU (two contexts / two host threads) RACE on the array: serialise them.
// In host code the array's NAME is a host-side stand-in, not a device address; the device address comes from cudaGetSymbolAddress,
// once per device, with that device current (noalloc_scratch_base below).
#include <mutex>

namespace hsq {

static constexpr int NOALLOC_GRID = 64;                                     // CTAs of DP_BLOCK threads: 8192 resident alignments
static constexpr int64_t NOALLOC_THREADS = (int64_t)NOALLOC_GRID * DP_BLOCK;
static constexpr int64_t NOALLOC_CODE_BYTES = 4ll << 20;                    // per side: encoded references / encoded queries
static constexpr int64_t NOALLOC_WORDS_PER_THREAD = 1792;                   // (r_max + 1) * (strips + 1) 32-bit words: 160 x 160 bases fit
static constexpr int64_t NOALLOC_SCRATCH_BYTES = 2 * NOALLOC_CODE_BYTES + NOALLOC_THREADS * NOALLOC_WORDS_PER_THREAD * 4;  // 64 MiB

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

// ASCII -> device codes, one byte per thread step (the inputs are the caller's device buffers: no alignment is assumed); also checks
// that the offsets end where the caller said the data ends
__global__ void noalloc_encode_kernel(const uint8_t* __restrict__ in, int64_t n_bytes, const int64_t* __restrict__ offsets, int64_t n,
                                      uint8_t* __restrict__ out, int* status) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    if (i == 0 && (offsets[0] != 0 || offsets[n] != n_bytes)) atomicMin(status, HSQ_ERR_INVALID_ARGUMENT);
    bool bad = false;
    for (; i < n_bytes; i += stride) {
        const uint8_t c = encode_ascii(in[i]);
        bad |= (c == CODE_BAD);
        out[i] = c;
    }
    if (bad) atomicMin(status, HSQ_ERR_BAD_CHARACTER);
}

struct CountSink {  // matches / mismatches as AlignmentPair.getNumberOfMatches / getNumberOfMismatches count them (AlignmentPair.java:246-290)
    static constexpr bool HAS_FAST = false;
    int n = 0, matches = 0, mismatches = 0;
    __device__ inline void push(uint32_t c) {
        const uint32_t rc = c & 0xFFu, qc = c >> 8;
        if (rc != CODE_GAP && qc != CODE_GAP) {
            if ((rc & qc & MATCH_MASK) != 0) matches++;
            else mismatches++;
        }
        n++;
    }
};

// align_kernel's DP with the scratch pool handed in by address and only scalar results (streamed traceback: no column store)
__global__ void __launch_bounds__(DP_BLOCK) noalloc_align_kernel(DeviceSeqs refs, DeviceSeqs qrys, int64_t n, ScoreParams sp, int pen_start, int pen_end,
                                                                ScratchPool pool, int max_ref, int max_qry, int32_t* __restrict__ score,
                                                                int32_t* __restrict__ n_cols, int32_t* __restrict__ matches,
                                                                int32_t* __restrict__ mismatches, int* status) {
    __shared__ int4 prof[PROF_INT4_PER_THREAD * DP_BLOCK];
    DpScratch sc = thread_scratch(pool, prof);
    for (int64_t k = sc.tid; k < n; k += sc.nthreads) {
        const int64_t r0 = refs.offsets[k], q0 = qrys.offsets[k];
        const int64_t rl = refs.offsets[k + 1] - r0, ql = qrys.offsets[k + 1] - q0;
        if (rl < 0 || ql < 0 || rl > max_ref || ql > max_qry) {  // longer than the caller declared: the scratch slice would not hold it
            atomicMin(status, HSQ_ERR_INVALID_ARGUMENT);
            score[k] = 0; n_cols[k] = 0; matches[k] = 0; mismatches[k] = 0;
            continue;
        }
        SeqView ref = make_view(refs.codes + r0, (int)rl, false, false);
        SeqView qry = make_view(qrys.codes + q0, (int)ql, false, false);
        const int s = dp_fill(ref, qry, sp, pen_start, pen_end, sc);
        CountSink sink;
        dp_traceback_stream(ref, qry, sc, sink);
        score[k] = s;
        n_cols[k] = sink.n;
        matches[k] = sink.matches;
        mismatches[k] = sink.mismatches;
    }
}

}  // namespace hsq

namespace {

// device address of g_noalloc_scratch on `device` (which must be current): looked up once per device
int noalloc_scratch_base(hsq_ctx* ctx, int device, unsigned char** out) {
    static std::mutex mu;
    static unsigned char* base[64] = {};
    if (device < 0 || device >= 64) return fail(ctx, HSQ_ERR_UNSUPPORTED, "hsq_align_scores_noalloc: device index above 63");
    std::lock_guard<std::mutex> lock(mu);
    if (!base[device]) {
        unsigned char* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail(ctx, HSQ_ERR_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
        base[device] = p;
    }
    *out = base[device];
    return HSQ_OK;
}

}  // namespace

extern "C" int32_t hsq_align_scores_noalloc(hsq_ctx* ctx, const uint8_t* d_ref_data, const int64_t* d_ref_offsets, int64_t ref_bytes,
                                            const uint8_t* d_qry_data, const int64_t* d_qry_offsets, int64_t qry_bytes, int64_t n, int32_t max_ref_len,
                                            int32_t max_qry_len, const hsq_scorer* scorer, uint32_t flags, int32_t* d_score, int32_t* d_n_cols,
                                            int32_t* d_matches, int32_t* d_mismatches, int32_t* d_status) {
    if (!ctx || !scorer || !d_status) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    if (n < 0 || ref_bytes < 0 || qry_bytes < 0 || max_ref_len < 0 || max_qry_len < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative size");
    if (!d_ref_offsets || !d_qry_offsets || (n > 0 && (!d_score || !d_n_cols || !d_matches || !d_mismatches)) || (ref_bytes > 0 && !d_ref_data) ||
        (qry_bytes > 0 && !d_qry_data))
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null device buffer");
    // the fixed scratch: reject what it cannot hold
    if (ref_bytes > NOALLOC_CODE_BYTES || qry_bytes > NOALLOC_CODE_BYTES)
        return fail(ctx, HSQ_ERR_UNSUPPORTED, "hsq_align_scores_noalloc: more than 4 MiB of reference or of query bases in one call (HSQ_NOALLOC_MAX_BYTES_PER_SIDE)");
    const int64_t strips = std::max<int64_t>(1, ((int64_t)max_qry_len + KROWS - 1) / KROWS);
    const int64_t r_stride = (int64_t)max_ref_len + 1;
    if (r_stride * (strips + 1) > NOALLOC_WORDS_PER_THREAD)
        return fail(ctx, HSQ_ERR_UNSUPPORTED,
                    "hsq_align_scores_noalloc: (max_ref_len + 1) * (ceil(max_qry_len / 16) + 1) exceeds 1792 scratch words per alignment "
                    "(e.g. 160 x 160 bases fit); use hsq_align_batch for longer sequences");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    int rc = noalloc_scratch_base(ctx, ctx->device, &base);
    if (rc) return rc;
    CallTimer timer(ctx);
    // layout: [encoded references | encoded queries | pointer words of all threads | boundary rows of all threads]
    uint8_t* ref_codes = base;
    uint8_t* qry_codes = base + NOALLOC_CODE_BYTES;
    uint32_t* ptr_words = reinterpret_cast<uint32_t*>(base + 2 * NOALLOC_CODE_BYTES);
    int32_t* brow = reinterpret_cast<int32_t*>(ptr_words + NOALLOC_THREADS * strips * r_stride);
    HSQ_CUDA(ctx, cudaMemsetAsync(d_status, 0, sizeof(int32_t), ctx->stream));  // the one word that must start at zero
    const int eblocks = ctx->sm_count * 4;
    noalloc_encode_kernel<<<eblocks, 256, 0, ctx->stream>>>(d_ref_data, ref_bytes, d_ref_offsets, n, ref_codes, d_status);
    noalloc_encode_kernel<<<eblocks, 256, 0, ctx->stream>>>(d_qry_data, qry_bytes, d_qry_offsets, n, qry_codes, d_status);
    ctx->launches += 2;
    if (n > 0) {
        ScratchPool pool;
        pool.ptr_words = ptr_words;
        pool.brow = brow;
        pool.cols = nullptr;  // streamed traceback: no column store
        pool.r_stride = (int)r_stride;
        const ScoreParams sp = make_score_params(scorer->match, scorer->mismatch, scorer->gap_extend, scorer->gap_start);
        noalloc_align_kernel<<<NOALLOC_GRID, DP_BLOCK, 0, ctx->stream>>>(DeviceSeqs{ref_codes, d_ref_offsets}, DeviceSeqs{qry_codes, d_qry_offsets}, n, sp,
                                                                        (flags & HSQ_ALIGN_PENALIZE_START) ? 1 : 0, (flags & HSQ_ALIGN_PENALIZE_END) ? 1 : 0,
                                                                        pool, max_ref_len, max_qry_len, d_score, d_n_cols, d_matches, d_mismatches, d_status);
        ctx->launches++;
    }
    HSQ_CUDA(ctx, cudaGetLastError());
    return timer.finish();  // synchronises the context's stream: the outputs and *d_status are final when this returns
}
