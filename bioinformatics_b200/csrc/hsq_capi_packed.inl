// hsq_capi_packed.inl -- north_star subsystem (1): read batches packed to 2 / 4 bits per base cross PCIe, results come back as
// per-read-end integer columns plus one CIGAR/MD byte pool.
//   host side   hsq_pack_reads / hsq_quality_sums (plain CPU code: they prepare buffers, they compute nothing of the path)
//   device side packed words -> 1-byte device codes (what the DP kernels read), exception patch, pool gather
//   entry points hsq_extend_packed (the chunked H2D | prep | capture | D2H pipeline of hsq_extend_batch), hsq_unpack_reads
// The reference keeps its reads packed as well: COMMON/sequence/NucleotideCodeSequence.java:32-48 (2 bit),
// COMMON/sequence/IupacNucleotideCodeSequence.java:34-58 (IUPAC).
#include <thread>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace hsq {

// words of a read of `len` bases
__host__ __device__ inline int64_t packed_words_of(int64_t len, int bits) { return (len * bits + 31) / 32; }

__global__ void packed_lens_kernel(const int32_t* __restrict__ len, int64_t n, int bits, int64_t* __restrict__ bases, int64_t* __restrict__ words,
                                   int* status) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) {
        int l = len[k];
        if (l < 0) {
            atomicMin(status, HSQ_ERR_INVALID_ARGUMENT);
            l = 0;
        }
        bases[k] = l;
        words[k] = packed_words_of(l, bits);
    }
}

// caller-supplied totals of a packed batch against the scans of its lengths
__global__ void packed_totals_kernel(const int64_t* __restrict__ offsets, const int64_t* __restrict__ word_off, int64_t n, int64_t bases, int64_t words,
                                     int* status) {
    if (offsets[n] != bases || word_off[n] != words) atomicMin(status, HSQ_ERR_INVALID_ARGUMENT);
}

// 2 bit: A=0 C=1 T=2 G=3 -> device codes 1, 2, 8, 4;  4 bit: the nucleotide set itself, 0 = gap
__device__ __forceinline__ uint8_t packed_to_code(uint32_t v, int bits) {
    if (bits == 2) return (uint8_t)((0x4821u >> (4 * v)) & 0xFu);
    return v ? (uint8_t)v : CODE_GAP;
}

// One warp per read: lane i writes bases i, i + 32, ...  (a warp store covers 32 consecutive code bytes; the packed words of a
// read are read by all lanes and stay in L1).  HBM: bits/8 B read + 1 B written per base.
template <int bits>
__global__ void unpack_kernel(const uint32_t* __restrict__ words, const int64_t* __restrict__ word_off, const int64_t* __restrict__ offsets,
                              int64_t n, uint8_t* __restrict__ codes) {
    const int lane = threadIdx.x & 31;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    constexpr int per = 32 / bits;
    constexpr uint32_t mask = (1u << bits) - 1u;
    for (; w < n; w += stride) {
        const uint32_t* src = words + word_off[w];
        const int64_t dst = offsets[w];
        const int len = (int)(offsets[w + 1] - dst);
        for (int i = lane; i < len; i += 32) {
            const uint32_t wd = src[i / per];
            codes[dst + i] = packed_to_code((wd >> (bits * (i % per))) & mask, bits);
        }
    }
}

// four 2-bit bases (the low 8 bits of v) -> four device codes, one per byte: spread the pairs of bits over the bytes, e = v ^ (v >> 1)
// turns A C T G = 0 1 2 3 into the exponents 0 1 3 2 of the codes 1 2 8 4, and 1 << e is built without a per-byte shift
__device__ __forceinline__ uint32_t codes4_of_2bit(uint32_t v) {
    uint32_t t = (v | (v << 12)) & 0x000F000Fu;
    t = (t | (t << 6)) & 0x03030303u;
    const uint32_t e = t ^ ((t >> 1) & 0x01010101u);
    const uint32_t m1 = 0x01010101u + (e & 0x01010101u);
    const uint32_t hi = ((e >> 1) & 0x01010101u) * 0xFFu;
    return m1 + 3u * (m1 & hi);
}

// 2 bit, one warp per read: the 16-byte aligned chunks of the destination that lie inside the read are one uint4 store each (16 bases =
// 32 bits taken from two packed words with a funnel shift, expanded four bases at a time); the up to 15 bases before the first and
// after the last such chunk are one per lane.  (The byte-per-lane kernel above moved 4 M reads in 650 us.)
__global__ void unpack2_kernel(const uint32_t* __restrict__ words, const int64_t* __restrict__ word_off, const int64_t* __restrict__ offsets,
                               int64_t n, uint8_t* __restrict__ codes) {
    const int lane = threadIdx.x & 31;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; w < n; w += stride) {
        const uint32_t* src = words + word_off[w];
        const int64_t dst = offsets[w], end = offsets[w + 1];
        const int64_t body_lo = (dst + 15) & ~(int64_t)15, body_hi = end & ~(int64_t)15;
        if (body_lo <= body_hi) {
            for (int64_t c = body_lo + 16 * lane; c < body_hi; c += 512) {
                const int b0 = (int)(c - dst);  // first base of the chunk; the chunk's 16 bases lie inside the read
                const uint32_t lo = src[b0 >> 4], sh = 2u * (uint32_t)(b0 & 15);
                const uint32_t hi = sh ? src[(b0 >> 4) + 1] : 0u;
                const uint32_t v = __funnelshift_r(lo, hi, sh);
                *reinterpret_cast<uint4*>(codes + c) = make_uint4(codes4_of_2bit(v & 0xFFu), codes4_of_2bit((v >> 8) & 0xFFu),
                                                                   codes4_of_2bit((v >> 16) & 0xFFu), codes4_of_2bit(v >> 24));
            }
            if (lane < 16) {
                const int64_t h = dst + lane, t = body_hi + lane;
                if (h < body_lo) {
                    const int i = lane;
                    codes[h] = packed_to_code((src[i >> 4] >> (2 * (i & 15))) & 3u, 2);
                }
                if (t < end) {
                    const int i = (int)(t - dst);
                    codes[t] = packed_to_code((src[i >> 4] >> (2 * (i & 15))) & 3u, 2);
                }
            }
        } else if (lane < 16) {  // fewer than 16 bases, none of them on a chunk boundary
            const int64_t h = dst + lane;
            if (h < end) codes[h] = packed_to_code((src[lane >> 4] >> (2 * (lane & 15))) & 3u, 2);
        }
    }
}

// exceptions of a 2-bit batch: the listed positions get the code of their ASCII character
__global__ void patch_exceptions_kernel(const uint32_t* __restrict__ exc_read, const uint32_t* __restrict__ exc_pos, const uint8_t* __restrict__ exc_char,
                                        int64_t n_exc, int64_t read_base, int64_t n, const int64_t* __restrict__ offsets, uint8_t* __restrict__ codes,
                                        int* status) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n_exc; k += stride) {
        const int64_t r = (int64_t)exc_read[k] - read_base;
        if (r < 0 || r >= n || (int64_t)exc_pos[k] >= offsets[r + 1] - offsets[r]) {
            atomicMin(status, HSQ_ERR_INVALID_ARGUMENT);
            continue;
        }
        const uint8_t c = encode_ascii(exc_char[k]);
        if (c == CODE_BAD) atomicMin(status, HSQ_ERR_BAD_CHARACTER);
        codes[offsets[r] + exc_pos[k]] = c;
    }
}

__global__ void decode_kernel(const uint8_t* __restrict__ codes, int64_t n, uint8_t* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = decode_code(codes[i]);
}

// ---- pooled CIGAR / MD output ----------------------------------------------------------------------------------------------
__global__ void pool_lens_kernel(ExtendDeviceOut d, int64_t n_ends, int64_t* __restrict__ tot) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; e < n_ends; e += stride) tot[e] = (int64_t)min(d.cigar_len[e], d.cigar_cap) + (int64_t)min(d.md_len[e], d.md_cap);
}

static constexpr int POOL_BLOCK = 256;
static constexpr int POOL_SMEM = 24 * 1024;

// A block takes 256 consecutive read ends: their strings are first laid out in shared memory exactly as they will sit in the pool
// (same 16-byte phase), then written with aligned 16-byte stores -- the pool may be pinned host memory, where byte-sized stores of
// single threads would each become their own PCIe write.  str_off gets the absolute pool offset of every read end.
__global__ void __launch_bounds__(POOL_BLOCK) pool_gather_kernel(ExtendDeviceOut d, int64_t n_ends, const int64_t* __restrict__ rel,
                                                                 const int64_t* __restrict__ cursor, int64_t pool_cap, uint8_t* __restrict__ pool,
                                                                 int64_t* __restrict__ str_off) {
    __shared__ __align__(16) uint8_t buf[POOL_SMEM];
    const int tid = threadIdx.x;
    const int64_t base = *cursor;
    for (int64_t e0 = (int64_t)blockIdx.x * POOL_BLOCK; e0 < n_ends; e0 += (int64_t)gridDim.x * POOL_BLOCK) {
        const int64_t e1 = min(n_ends, e0 + POOL_BLOCK);
        const int64_t e = e0 + tid;
        const int64_t g0 = base + rel[e0], g1 = base + rel[e1];
        const int64_t bytes = g1 - g0;
        int64_t mine = 0;
        int clen = 0, mlen = 0;
        if (e < e1) {
            mine = base + rel[e];
            str_off[e] = mine;
            clen = min(d.cigar_len[e], d.cigar_cap);
            mlen = min(d.md_len[e], d.md_cap);
        }
        if (g1 > pool_cap) continue;  // block-uniform; the host sees cursor > pool_cap at the end of the call
        const int shift = (int)((reinterpret_cast<uintptr_t>(pool) + (uintptr_t)g0) & 15u);
        const uint8_t* cs = d.cigar + e * d.cigar_cap;
        const uint8_t* ms = d.md + e * d.md_cap;
        if (bytes + shift <= POOL_SMEM) {
            uint8_t* at = buf + shift + (mine - g0);
            for (int i = 0; i < clen; i++) at[i] = cs[i];
            at += clen;
            for (int i = 0; i < mlen; i++) at[i] = ms[i];
            __syncthreads();
            const int total = shift + (int)bytes;
            uint8_t* gp = pool + g0 - shift;  // 16-byte aligned
            for (int lo = tid * 16; lo < total; lo += POOL_BLOCK * 16) {
                const int hi = lo + 16;
                if (lo >= shift && hi <= total) {
                    *reinterpret_cast<uint4*>(gp + lo) = *reinterpret_cast<const uint4*>(buf + lo);
                } else {
                    for (int b = max(lo, shift); b < min(hi, total); b++) gp[b] = buf[b];
                }
            }
            __syncthreads();
        } else if (e < e1) {  // strings longer than the staging buffer takes: straight copies
            uint8_t* at = pool + mine;
            for (int i = 0; i < clen; i++) at[i] = cs[i];
            at += clen;
            for (int i = 0; i < mlen; i++) at[i] = ms[i];
        }
    }
}

__global__ void pool_advance_kernel(int64_t* cursor, const int64_t* __restrict__ rel, int64_t n_ends) { *cursor += rel[n_ends]; }

}  // namespace hsq

namespace {

// ---- host side: packing ------------------------------------------------------------------------------------------------------
inline int ascii_2bit(uint8_t ch) {  // A C G T in either case -> 0 1 3 2, anything else -> -1
    switch (ch) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'T': case 't': return 2;
        case 'G': case 'g': return 3;
        default: return -1;
    }
}

struct PackExc { uint32_t read, pos; uint8_t ch; };

#if defined(__x86_64__)
// 16 upper-case A/C/G/T bytes -> 32 bits (bits 1-2 of every letter), or false when any byte is something else
__attribute__((target("bmi2,sse4.1"))) inline bool pack16_fast(const uint8_t* p, uint32_t* out) {
    const __m128i v = _mm_loadu_si128(reinterpret_cast<const __m128i*>(p));
    const __m128i ok = _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(v, _mm_set1_epi8('A')), _mm_cmpeq_epi8(v, _mm_set1_epi8('C'))),
                                    _mm_or_si128(_mm_cmpeq_epi8(v, _mm_set1_epi8('G')), _mm_cmpeq_epi8(v, _mm_set1_epi8('T'))));
    if (_mm_movemask_epi8(ok) != 0xFFFF) return false;
    const uint64_t lo = (uint64_t)_mm_cvtsi128_si64(v), hi = (uint64_t)_mm_extract_epi64(v, 1);
    *out = (uint32_t)_pext_u64(lo, 0x0606060606060606ull) | ((uint32_t)_pext_u64(hi, 0x0606060606060606ull) << 16);
    return true;
}
#endif

// reads [k0, k1) of a ragged ASCII batch -> packed words at their final place; 0 or an HSQ_ERR code
int pack_range(const hsq_seqs* s, int bits, int64_t k0, int64_t k1, const int64_t* word_off, uint32_t* words, int32_t* len_out,
               std::vector<PackExc>* exc, bool fast) {
    const int per = 32 / bits;
    for (int64_t k = k0; k < k1; k++) {
        const uint8_t* p = s->data + s->offsets[k];
        const int64_t len = s->offsets[k + 1] - s->offsets[k];
        if (len < 0 || len > INT32_MAX) return HSQ_ERR_INVALID_ARGUMENT;
        len_out[k] = (int32_t)len;
        uint32_t* w = words + word_off[k];
        const int64_t nw = hsq::packed_words_of(len, bits);
        for (int64_t j = 0; j < nw; j++) {
            const int64_t i0 = j * per;
            const int cnt = (int)std::min<int64_t>(per, len - i0);
            uint32_t acc = 0;
            if (bits == 2) {
#if defined(__x86_64__)
                if (fast && cnt == 16 && pack16_fast(p + i0, &acc)) {
                    w[j] = acc;
                    continue;
                }
#endif
                for (int b = 0; b < cnt; b++) {
                    const uint8_t ch = p[i0 + b];
                    int v = ascii_2bit(ch);
                    if (v < 0) {
                        if (hsq::encode_ascii(ch) == hsq::CODE_BAD) return HSQ_ERR_BAD_CHARACTER;
                        exc->push_back(PackExc{(uint32_t)k, (uint32_t)(i0 + b), ch});
                        v = 0;
                    }
                    acc |= (uint32_t)v << (2 * b);
                }
            } else {
                for (int b = 0; b < cnt; b++) {
                    const uint8_t c = hsq::encode_ascii(p[i0 + b]);
                    if (c == hsq::CODE_BAD) return HSQ_ERR_BAD_CHARACTER;
                    if (c == 0x18) return HSQ_ERR_UNSUPPORTED;  // 'U'
                    acc |= (uint32_t)(c & 15u) << (4 * b);
                }
            }
            w[j] = acc;
        }
    }
    return HSQ_OK;
}

int host_threads(int32_t threads) {
    int t = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    return std::max(1, std::min(t, 64));
}

}  // namespace

extern "C" int64_t hsq_packed_bytes(const int64_t* offsets, int64_t n, int32_t bits) {
    if (!offsets || n < 0 || (bits != 2 && bits != 4)) return -1;
    int64_t words = 0;
    for (int64_t k = 0; k < n; k++) words += hsq::packed_words_of(offsets[k + 1] - offsets[k], bits);
    return 4 * words;
}

extern "C" int32_t hsq_pack_reads(const hsq_seqs* reads, int32_t bits, uint8_t* data_out, int32_t* len_out, int64_t exc_cap,
                                  int64_t* n_exceptions, uint32_t* exc_read, uint32_t* exc_pos, uint8_t* exc_char, int32_t threads) {
    if (!reads || reads->n < 0 || !reads->offsets || (reads->n && (!len_out)))
        return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "hsq_pack_reads: null argument");
    if (bits != 2 && bits != 4) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "hsq_pack_reads: bits must be 2 or 4");
    if (reinterpret_cast<uintptr_t>(data_out) & 3u) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "hsq_pack_reads: data_out must be 4-byte aligned");
    const int64_t n = reads->n;
    if (n_exceptions) *n_exceptions = 0;
    if (n >= (1ll << 32)) return fail(nullptr, HSQ_ERR_UNSUPPORTED, "hsq_pack_reads: more than 2^32 reads");
    std::vector<int64_t> word_off((size_t)n + 1, 0);
    for (int64_t k = 0; k < n; k++) {
        const int64_t len = reads->offsets[k + 1] - reads->offsets[k];
        if (len < 0) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "hsq_pack_reads: offsets must not decrease");
        word_off[(size_t)k + 1] = word_off[(size_t)k] + hsq::packed_words_of(len, bits);
    }
    if (word_off[(size_t)n] > 0 && (!data_out || !reads->data)) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "hsq_pack_reads: null buffer");
    const int nt = (int)std::min<int64_t>(host_threads(threads), std::max<int64_t>(1, n / 4096));
    std::vector<std::vector<PackExc>> exc((size_t)nt);
    std::vector<int> rcs((size_t)nt, 0);
    bool fast = false;
#if defined(__x86_64__)
    fast = __builtin_cpu_supports("bmi2") && __builtin_cpu_supports("sse4.1");
#endif
    auto work = [&](int t) {
        const int64_t k0 = n * t / nt, k1 = n * (t + 1) / nt;
        rcs[(size_t)t] = pack_range(reads, bits, k0, k1, word_off.data(), reinterpret_cast<uint32_t*>(data_out), len_out, &exc[(size_t)t], fast);
    };
    if (nt == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < nt; t++) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    for (int rc : rcs)
        if (rc) return fail(nullptr, rc, rc == HSQ_ERR_BAD_CHARACTER ? "hsq_pack_reads: a read holds a character that is not an IUPAC nucleotide code"
                                         : rc == HSQ_ERR_UNSUPPORTED ? "hsq_pack_reads: 'U' has no 4-bit code"
                                                                     : "hsq_pack_reads: bad offsets");
    int64_t total = 0;
    for (auto& v : exc) total += (int64_t)v.size();
    if (n_exceptions) *n_exceptions = total;
    if (total > 0) {
        if (total > exc_cap || !exc_read || !exc_pos || !exc_char)
            return fail(nullptr, HSQ_ERR_OVERFLOW, "hsq_pack_reads: more exceptions than exc_cap (see *n_exceptions)");
        int64_t at = 0;
        for (auto& v : exc)  // thread ranges ascend, so the concatenation is ordered by read, then position
            for (auto& x : v) {
                exc_read[at] = x.read;
                exc_pos[at] = x.pos;
                exc_char[at] = x.ch;
                at++;
            }
    }
    return HSQ_OK;
}

// (short) sum(q - 33) per read, exactly BamFileUtil.getQualityScore (HSQ/utils/BamFileUtil.java:74-85; Java short arithmetic wraps
// mod 2^16).  Optional host helper for callers that would rather ship 2 bytes per read than the quality string to hsq_dedup_packed.
extern "C" int32_t hsq_quality_sums(const uint8_t* quals, const int64_t* offsets, int64_t n, int16_t* out, int32_t threads) {
    if (n < 0 || !offsets || (n && !out)) return fail(nullptr, HSQ_ERR_INVALID_ARGUMENT, "hsq_quality_sums: null argument");
    const int nt = (int)std::min<int64_t>(host_threads(threads), std::max<int64_t>(1, n / 4096));
    auto work = [&](int t) {
        for (int64_t k = n * t / nt; k < n * (t + 1) / nt; k++) {
            int s = 0;
            for (int64_t i = offsets[k]; i < offsets[k + 1]; i++) s += (int)quals[i] - 33;
            out[k] = (int16_t)s;
        }
    };
    if (nt == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < nt; t++) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    return HSQ_OK;
}

namespace {

// host -> device copies of (a window of) a packed batch into d: packed words, lengths, exceptions.  `a` = first read of the
// window, word_a = its first word; total_bases / total_words of the window come from the caller's host pass.
int copy_packed(hsq_ctx* ctx, const hsq_packed_reads* s, int64_t a, int64_t m, int64_t word_a, int64_t total_words, int64_t total_bases,
                int64_t exc_a, int64_t exc_b, DevSeqsOwned* d) {
    const size_t mm = (size_t)m;
    d->n = m;
    d->total = total_bases;
    d->packed_bits = s->bits;
    d->packed_words = total_words;
    d->n_exc = exc_b - exc_a;
    d->exc_read_base = a;
    const int64_t n_tiles = (m + 1 + SCAN_TILE - 1) / SCAN_TILE;
    HSQ_CUDA(ctx, d->ascii.ensure((size_t)total_words * 4 + 32));
    HSQ_CUDA(ctx, d->codes.ensure((size_t)total_bases + 32));
    HSQ_CUDA(ctx, d->offsets.ensure((mm + 1) * 8));
    HSQ_CUDA(ctx, d->word_off.ensure((mm + 1) * 8));
    HSQ_CUDA(ctx, d->lens.ensure(mm * 4 + 16));
    HSQ_CUDA(ctx, d->scan_tmp.ensure((2 * (mm + 1) + (size_t)n_tiles) * 8 + 16));
    cudaStream_t st = ctx->io ? ctx->io : ctx->stream;
    if (total_words) HSQ_CUDA(ctx, cudaMemcpyAsync(d->ascii.p, s->data + 4 * word_a, (size_t)total_words * 4, cudaMemcpyHostToDevice, st));
    if (m) HSQ_CUDA(ctx, cudaMemcpyAsync(d->lens.p, s->len + a, mm * 4, cudaMemcpyHostToDevice, st));
    if (d->n_exc > 0) {
        const size_t ne = (size_t)d->n_exc;
        HSQ_CUDA(ctx, d->exc.ensure(ne * 9 + 16));
        uint8_t* base = d->exc.as<uint8_t>();
        HSQ_CUDA(ctx, cudaMemcpyAsync(base, s->exc_read + exc_a, ne * 4, cudaMemcpyHostToDevice, st));
        HSQ_CUDA(ctx, cudaMemcpyAsync(base + ne * 4, s->exc_pos + exc_a, ne * 4, cudaMemcpyHostToDevice, st));
        HSQ_CUDA(ctx, cudaMemcpyAsync(base + ne * 8, s->exc_char + exc_a, ne, cudaMemcpyHostToDevice, st));
    }
    return HSQ_OK;
}

// packed words -> codes of a staged packed batch, on ctx->stream: lengths -> two scans (base offsets, word offsets) -> unpack
// -> exception patch
int unpack_seqs(hsq_ctx* ctx, DevSeqsOwned* d) {
    const int64_t n = d->n;
    int rc;
    int64_t* bases = d->scan_tmp.as<int64_t>();
    int64_t* words = bases + n + 1;
    int64_t* tiles = words + n + 1;
    if (n > 0) {
        const int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 8);
        packed_lens_kernel<<<blocks, 256, 0, ctx->stream>>>(d->lens.as<int32_t>(), n, d->packed_bits, bases, words, d_status(ctx));
        ctx->launches++;
    }
    if ((rc = scan_i64(ctx, bases, n, tiles, d->offsets.as<int64_t>()))) return rc;
    if ((rc = scan_i64(ctx, words, n, tiles, d->word_off.as<int64_t>()))) return rc;
    packed_totals_kernel<<<1, 1, 0, ctx->stream>>>(d->offsets.as<int64_t>(), d->word_off.as<int64_t>(), n, d->total, d->packed_words, d_status(ctx));
    ctx->launches++;
    if (n > 0) {
        const int blocks = (int)std::min<int64_t>((n * 32 + 255) / 256, (int64_t)ctx->sm_count * 16);
        if (d->packed_bits == 2)
            unpack2_kernel<<<blocks, 256, 0, ctx->stream>>>(d->ascii.as<uint32_t>(), d->word_off.as<int64_t>(), d->offsets.as<int64_t>(), n, d->codes.as<uint8_t>());
        else
            unpack_kernel<4><<<blocks, 256, 0, ctx->stream>>>(d->ascii.as<uint32_t>(), d->word_off.as<int64_t>(), d->offsets.as<int64_t>(), n, d->codes.as<uint8_t>());
        ctx->launches++;
    }
    if (d->n_exc > 0) {
        const size_t ne = (size_t)d->n_exc;
        const uint8_t* base = d->exc.as<uint8_t>();
        const int blocks = (int)std::min<int64_t>((d->n_exc + 255) / 256, (int64_t)ctx->sm_count * 4);
        patch_exceptions_kernel<<<blocks, 256, 0, ctx->stream>>>(reinterpret_cast<const uint32_t*>(base), reinterpret_cast<const uint32_t*>(base + ne * 4),
                                                                base + ne * 8, d->n_exc, d->exc_read_base, n, d->offsets.as<int64_t>(),
                                                                d->codes.as<uint8_t>(), d_status(ctx));
        ctx->launches++;
    }
    HSQ_CUDA(ctx, cudaGetLastError());
    return HSQ_OK;
}

int prepare_codes(hsq_ctx* ctx, DevSeqsOwned* d, int* d_maxlen) {
    if (d->packed_bits != 0) {  // < 0: the codes are already in place (device-side compaction of an unpacked batch)
        if (d->packed_bits > 0) {
            int rc = unpack_seqs(ctx, d);
            if (rc) return rc;
        }
        if (d_maxlen && d->n > 0) {
            const int blocks = (int)std::min<int64_t>((d->n + 255) / 256, (int64_t)ctx->sm_count * 8);
            max_len_kernel<<<blocks, 256, 0, ctx->stream>>>(d->offsets.as<int64_t>(), d->n, d_maxlen);
            ctx->launches++;
        }
        return HSQ_OK;
    }
    return encode_seqs(ctx, d, d_status(ctx), d_maxlen);
}

int check_packed_reads(hsq_ctx* ctx, const hsq_packed_reads* r, int64_t n, const char* what) {
    if (r->n != n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, std::string(what) + ": packed batch arrays disagree on n");
    if (r->bits != 2 && r->bits != 4) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, std::string(what) + ": bits must be 2 or 4");
    if (n > 0 && !r->len) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, std::string(what) + ": null length array");
    if (r->n_exceptions < 0 || (r->n_exceptions > 0 && (r->bits != 2 || !r->exc_read || !r->exc_pos || !r->exc_char)))
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, std::string(what) + ": bad exception list");
    if (reinterpret_cast<uintptr_t>(r->data) & 3u) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, std::string(what) + ": packed data must be 4-byte aligned");
    if (r->data_bytes < 0 || (r->data_bytes & 3) || r->n_bases < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, std::string(what) + ": bad totals");
    return HSQ_OK;
}

// Host pass over a packed pair batch, one chunk at a time and in chunk order (so that it overlaps the device work of the chunks
// before): first word / exception of both sides, totals, longest read; validates lengths and probe ids.  (Sizes only -- what the
// copies need to be issued.)
struct PackedChunks {
    std::vector<int64_t> starts;           // n_chunks + 1 pair indexes
    std::vector<int64_t> word[2], base[2]; // cumulative words / bases at every chunk start (n_chunks + 1)
    std::vector<int64_t> exc[2];           // first exception at every chunk start
    std::vector<int> read_max;             // per chunk
    int done = 0;                          // chunks scanned so far
    int chunk_of(int64_t a) const { return (int)(std::upper_bound(starts.begin(), starts.end(), a) - starts.begin()) - 1; }
    void init() {
        const size_t nc = starts.size() - 1;
        read_max.assign(nc, 0);
        for (int s = 0; s < 2; s++) {
            word[s].assign(nc + 1, 0);
            base[s].assign(nc + 1, 0);
            exc[s].assign(nc + 1, 0);
        }
        done = 0;
    }
};

int scan_packed_chunk(hsq_ctx* ctx, const hsq_packed_pairs* pb, bool check_probes, PackedChunks* pc, int c, bool need_read_max = true) {
    for (; pc->done <= c; pc->done++) {
        const size_t cc = (size_t)pc->done;
        const int64_t k0 = pc->starts[cc], k1 = pc->starts[cc + 1];
        const hsq_packed_reads* sides[2] = {&pb->r1, &pb->r2};
        if (!need_read_max && !check_probes && k0 == 0 && k1 == pb->n && pb->r1.data_bytes > 0 && pb->r2.data_bytes > 0 && pb->r1.n_bases > 0 &&
            pb->r2.n_bases > 0) {
            // the whole batch in one piece with the caller's totals: no pass over the lengths (the device checks the totals)
            for (int s = 0; s < 2; s++) {
                pc->word[s][cc + 1] = sides[s]->data_bytes / 4;
                pc->base[s][cc + 1] = sides[s]->n_bases;
                pc->exc[s][cc + 1] = sides[s]->n_exceptions;
            }
            continue;
        }
        int mx = 0;
        for (int s = 0; s < 2; s++) {
            const hsq_packed_reads* r = sides[s];
            const int bits = r->bits;
            int64_t w = 0, b = 0;
            int lo = 0;
            for (int64_t k = k0; k < k1; k++) {
                const int32_t l = r->len[k];
                lo = std::min(lo, l);
                mx = std::max(mx, l);
                w += ((int64_t)l * bits + 31) >> 5;
                b += l;
            }
            if (lo < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "packed batch: negative read length");
            pc->word[s][cc + 1] = pc->word[s][cc] + w;
            pc->base[s][cc + 1] = pc->base[s][cc] + b;
            pc->exc[s][cc + 1] = r->n_exceptions == 0 ? 0
                : (int64_t)(std::lower_bound(r->exc_read + pc->exc[s][cc], r->exc_read + r->n_exceptions, (uint32_t)std::min<int64_t>(k1, UINT32_MAX)) - r->exc_read);
        }
        pc->read_max[cc] = mx;
        if (check_probes) {
            bool bad = false;
            for (int64_t k = k0; k < k1; k++) bad |= (pb->probe_id[k] < 0) | (pb->probe_id[k] >= ctx->n_probes);
            if (bad) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "packed batch: probe_id out of range");
        }
    }
    return HSQ_OK;
}

// stage chunk c of a packed pair batch into the active buffer set (probe ids, both sides)
int stage_packed_chunk(hsq_ctx* ctx, const hsq_packed_pairs* pb, const PackedChunks& pc, int c) {
    const int64_t a = pc.starts[(size_t)c], m = pc.starts[(size_t)c + 1] - a;
    int rc;
    HSQ_CUDA(ctx, ctx->probe_id.ensure((size_t)m * 4 + 16));
    if (m) HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->probe_id.p, pb->probe_id + a, (size_t)m * 4, cudaMemcpyHostToDevice, ctx->io ? ctx->io : ctx->stream));
    const hsq_packed_reads* sides[2] = {&pb->r1, &pb->r2};
    DevSeqsOwned* dst[2] = {&ctx->r1, &ctx->r2};
    for (int s = 0; s < 2; s++)
        if ((rc = copy_packed(ctx, sides[s], a, m, pc.word[s][(size_t)c], pc.word[s][(size_t)c + 1] - pc.word[s][(size_t)c],
                              pc.base[s][(size_t)c + 1] - pc.base[s][(size_t)c], pc.exc[s][(size_t)c], pc.exc[s][(size_t)c + 1], dst[s])))
            return rc;
    return HSQ_OK;
}

// Delivery of the active set's results for read ends [2a, 2a + 2m) in pooled form, on ctx->io: string lengths -> scan -> gather into
// the pool at the running cursor -> copies of the integer columns the caller asked for.
struct PoolState {
    uint8_t* pool_dst = nullptr;  // device-visible pool (mapped host memory, or ctx->pool_dev)
    bool direct = false;          // pool_dst is the caller's pinned buffer
    int64_t* cursor = nullptr;    // device
};

int fetch_pooled_chunk(hsq_ctx* ctx, int64_t a, int64_t m, hsq_extend_pooled_out* o, const PoolState& ps) {
    if (!ctx->ext_ran || ctx->ext_n != m) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "no results to fetch");
    const size_t n2 = (size_t)(2 * m);
    const ExtendDeviceOut& d = ctx->ext_out;
    cudaStream_t st = ctx->io ? ctx->io : ctx->stream;
    cudaStream_t saved = ctx->stream;
    int rc = HSQ_OK;
    const int64_t n_tiles = ((int64_t)n2 + 1 + SCAN_TILE - 1) / SCAN_TILE;
    HSQ_CUDA(ctx, ctx->pool_tmp.ensure((3 * (n2 + 1) + (size_t)n_tiles) * 8 + 16));
    int64_t* tot = ctx->pool_tmp.as<int64_t>();
    int64_t* rel = tot + n2 + 1;
    int64_t* soff = rel + n2 + 1;
    int64_t* tiles = soff + n2 + 1;
    if (m > 0) {
        const int blocks = (int)std::min<int64_t>(((int64_t)n2 + 255) / 256, (int64_t)ctx->sm_count * 8);
        pool_lens_kernel<<<blocks, 256, 0, st>>>(d, (int64_t)n2, tot);
        ctx->launches++;
        ctx->stream = st;
        rc = scan_i64(ctx, tot, (int64_t)n2, tiles, rel);
        ctx->stream = saved;
        if (rc) return rc;
        const int gblocks = (int)std::min<int64_t>(((int64_t)n2 + POOL_BLOCK - 1) / POOL_BLOCK, (int64_t)ctx->sm_count * 4);
        pool_gather_kernel<<<gblocks, POOL_BLOCK, 0, st>>>(d, (int64_t)n2, rel, ps.cursor, o->pool_cap, ps.pool_dst, soff);
        pool_advance_kernel<<<1, 1, 0, st>>>(ps.cursor, rel, (int64_t)n2);
        ctx->launches += 2;
        HSQ_CUDA(ctx, cudaGetLastError());
    }
    int32_t* i32 = ctx->out_i32.as<int32_t>();
    int32_t* hosts[9] = {o->primer_end, o->score, o->pos, o->read_start, o->read_stop, o->nm, o->tlen, o->cigar_len, o->md_len};
    for (int f = 0; f < 9; f++)
        if (hosts[f] && (rc = d2h(ctx, hosts[f] + 2 * a, i32 + f * n2, n2))) return rc;
    if (o->ok && (rc = d2h(ctx, o->ok + 2 * a, d.ok, n2))) return rc;
    if (o->str_off && (rc = d2h(ctx, o->str_off + 2 * a, soff, n2))) return rc;
    if (o->pair_ok && (rc = d2h(ctx, o->pair_ok + a, ctx->pair_ok.as<uint8_t>(), (size_t)m))) return rc;
    return HSQ_OK;
}

// cursor set-up before, pool delivery after a pooled run
int pool_begin(hsq_ctx* ctx, hsq_extend_pooled_out* o, PoolState* ps) {
    if (o->pool_cap < 0 || (o->pool_cap > 0 && !o->pool)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_extend_packed: bad pool");
    HSQ_CUDA(ctx, ctx->pool_cursor.ensure(64));
    ps->cursor = ctx->pool_cursor.as<int64_t>();
    HSQ_CUDA(ctx, cudaMemsetAsync(ps->cursor, 0, 8, ctx->stream));
    cudaPointerAttributes at{};
    ps->direct = false;
    if (o->pool && cudaPointerGetAttributes(&at, o->pool) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) {
        ps->direct = true;
        ps->pool_dst = static_cast<uint8_t*>(at.devicePointer);
    } else {
        cudaGetLastError();  // a pageable pointer leaves an error state on older drivers
        HSQ_CUDA(ctx, ctx->pool_dev.ensure((size_t)o->pool_cap + 32));
        ps->pool_dst = ctx->pool_dev.as<uint8_t>();
    }
    return HSQ_OK;
}

int pool_end(hsq_ctx* ctx, hsq_extend_pooled_out* o, const PoolState& ps) {
    int64_t* h = reinterpret_cast<int64_t*>(ctx->h_cells) + 4;
    HSQ_CUDA(ctx, cudaMemcpyAsync(h, ps.cursor, 8, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    o->pool_bytes = *h;
    if (o->pool_bytes > o->pool_cap) return fail(ctx, HSQ_ERR_OVERFLOW, "hsq_extend_packed: the CIGAR / MD pool is too small (pool_bytes holds the size needed)");
    if (!ps.direct && o->pool_bytes > 0) {
        HSQ_CUDA(ctx, cudaMemcpyAsync(o->pool, ps.pool_dst, (size_t)o->pool_bytes, cudaMemcpyDeviceToHost, ctx->stream));
        HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return HSQ_OK;
}

}  // namespace

extern "C" int32_t hsq_extend_packed(hsq_ctx* ctx, const hsq_packed_pairs* pairs, hsq_extend_pooled_out* out) {
    if (!ctx || !pairs || !out) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    int rc = extend_args_ok(ctx);
    if (rc) return rc;
    const int64_t n = pairs->n;
    if (n < 0 || out->cigar_cap < 0 || out->md_cap < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative size");
    if (n > 0 && !pairs->probe_id) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "probe_id is null");
    if ((rc = check_packed_reads(ctx, &pairs->r1, n, "hsq_extend_packed")) || (rc = check_packed_reads(ctx, &pairs->r2, n, "hsq_extend_packed"))) return rc;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    const int cigar_cap = out->cigar_cap ? out->cigar_cap : 64, md_cap = out->md_cap ? out->md_cap : 96;
    out->pool_bytes = 0;
    out->cells = 0;
    PackedChunks pc;
    const int n_chunks = pick_chunks(ctx, n);
    pc.starts = chunk_starts(ctx, n, n_chunks);
    pc.init();
    PoolState ps;
    rc = pool_begin(ctx, out, &ps);
    if (!rc && n > 0) {
        ExtendChunkIO io;
        io.dims = [&](int64_t a, int64_t, int* read_max) {
            const int c = pc.chunk_of(a);
            const int r = scan_packed_chunk(ctx, pairs, true, &pc, c);
            *read_max = pc.read_max[(size_t)c];
            return r;
        };
        io.stage = [&](int64_t a, int64_t) { return stage_packed_chunk(ctx, pairs, pc, pc.chunk_of(a)); };
        io.fetch = [&](int64_t a, int64_t m) { return fetch_pooled_chunk(ctx, a, m, out, ps); };
        rc = extend_pipelined(ctx, n, cigar_cap, md_cap, n_chunks, io, &out->cells);
    }
    if (!rc || rc == HSQ_ERR_OVERFLOW) {
        const int rc3 = pool_end(ctx, out, ps);
        if (!rc) rc = rc3;
    }
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}

extern "C" int32_t hsq_unpack_reads(hsq_ctx* ctx, const hsq_packed_reads* reads, uint8_t* out_data, int64_t* out_offsets) {
    if (!ctx || !reads || !out_offsets) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    int rc = check_packed_reads(ctx, reads, reads->n, "hsq_unpack_reads");
    if (rc) return rc;
    const int64_t n = reads->n;
    int64_t words = 0, bases = 0;
    for (int64_t k = 0; k < n; k++) {
        if (reads->len[k] < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_unpack_reads: negative read length");
        words += hsq::packed_words_of(reads->len[k], reads->bits);
        bases += reads->len[k];
    }
    if (bases > 0 && !out_data) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null output");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    if ((rc = reset_status(ctx))) return rc;
    DevSeqsOwned* d = &ctx->sa;
    if ((rc = copy_packed(ctx, reads, 0, n, 0, words, bases, 0, reads->n_exceptions, d))) return rc;
    rc = unpack_seqs(ctx, d);
    d->packed_bits = 0;
    if (rc) return rc;
    HSQ_CUDA(ctx, ctx->sb.ascii.ensure((size_t)bases + 32));
    if (bases > 0) {
        const int blocks = (int)std::min<int64_t>((bases + 255) / 256, (int64_t)ctx->sm_count * 16);
        decode_kernel<<<blocks, 256, 0, ctx->stream>>>(d->codes.as<uint8_t>(), bases, ctx->sb.ascii.as<uint8_t>());
        ctx->launches++;
        if ((rc = d2h(ctx, out_data, ctx->sb.ascii.as<uint8_t>(), (size_t)bases))) return rc;
    }
    if ((rc = d2h(ctx, out_offsets, d->offsets.as<int64_t>(), (size_t)n + 1))) return rc;
    int w[1];
    if ((rc = read_status(ctx, w, 1))) return rc;
    if (w[0] != 0) return fail(ctx, w[0], std::string("hsq_unpack_reads: ") + (w[0] == HSQ_ERR_INVALID_ARGUMENT ? "an exception lies outside its read" : status_message(w[0])));
    return timer.finish();
}

// ---- hsq_dedup_packed: FilterByUid key generation from a packed batch ----------------------------------------------------------
namespace {

// host -> device copies of a packed pair batch, unpack, key generation: uid -> sort_a, rank -> sort_b, pair index -> sort_d[0..n)
int dedup_upload_keygen_packed(hsq_ctx* ctx, const hsq_packed_pairs* pb, int u, int v, int64_t np) {
    const int64_t n = pb->n;
    const size_t nn = (size_t)n;
    int rc = reset_status(ctx);
    if (rc) return rc;
    PackedChunks pc;
    pc.starts = {0, n};
    pc.init();
    if ((rc = scan_packed_chunk(ctx, pb, false, &pc, 0, false))) return rc;
    if ((rc = stage_packed_chunk(ctx, pb, pc, 0))) return rc;
    const bool sums = pb->qsum1 && pb->qsum2;
    struct { hsq::DevBuf* d; const void* h; size_t bytes; } up[] = {
        {&ctx->q1, sums ? (const void*)pb->qsum1 : (const void*)pb->q1, sums ? nn * 2 : (size_t)pc.base[0][1]},
        {&ctx->q2, sums ? (const void*)pb->qsum2 : (const void*)pb->q2, sums ? nn * 2 : (size_t)pc.base[1][1]},
        {&ctx->read_index, pb->read_index, nn * 4}};
    for (auto& c : up) {
        if (c.bytes && !c.h) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_dedup_packed: null input array");
        HSQ_CUDA(ctx, c.d->ensure(c.bytes + 16));
        if (c.bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(c.d->p, c.h, c.bytes, cudaMemcpyHostToDevice, ctx->stream));
    }
    if ((rc = unpack_seqs(ctx, &ctx->r1)) || (rc = unpack_seqs(ctx, &ctx->r2))) return rc;
    HSQ_CUDA(ctx, ctx->sort_a.ensure(nn * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_b.ensure(nn * 8 + 16));
    HSQ_CUDA(ctx, ctx->sort_d.ensure(nn * 8 + 16));
    if (n > 0) {
        const int gblocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)ctx->sm_count * 16);
        const int kblocks = (int)std::min<int64_t>((n + 127) / 128, (int64_t)ctx->sm_count * 16);
        span_begin(ctx, 0);
        check_probe_ids_kernel<<<gblocks, 256, 0, ctx->stream>>>(ctx->probe_id.as<int32_t>(), n, np > 0 ? np : INT32_MAX, d_status(ctx));
        dedup_keygen_kernel<true><<<kblocks, 128, 0, ctx->stream>>>(
            ctx->r1.codes.as<uint8_t>(), ctx->r1.offsets.as<int64_t>(), ctx->q1.as<uint8_t>(), ctx->r2.codes.as<uint8_t>(), ctx->r2.offsets.as<int64_t>(),
            ctx->q2.as<uint8_t>(), sums ? ctx->q1.as<int16_t>() : nullptr, sums ? ctx->q2.as<int16_t>() : nullptr, ctx->read_index.as<int32_t>(), n, u, v,
            ctx->sort_a.as<uint64_t>(), ctx->sort_b.as<uint64_t>(), ctx->sort_d.as<uint32_t>(), d_status(ctx));
        span_end(ctx, 0);
        ctx->launches += 2;
        int w[1];
        if ((rc = read_status(ctx, w, 1))) return rc;
        if (w[0] != 0)
            return fail(ctx, w[0], w[0] == HSQ_ERR_UNSUPPORTED ? "hsq_dedup_packed: a UID holds a base other than A,C,G,T (filter N reads first, "
                                                                 "PrimerReadExtensionAndPcrDuplicateIdentification.java:815-821)"
                                   : w[0] == HSQ_ERR_INVALID_ARGUMENT ? "hsq_dedup_packed: probe_id out of range, negative length or a misplaced exception"
                                   : w[0] == HSQ_ERR_BAD_CHARACTER ? "hsq_dedup_packed: an exception character is not an IUPAC code"
                                                                   : "hsq_dedup_packed: a read is shorter than its UID");
    }
    // the staged batch now holds valid codes: a following extend stage must not decode again
    ctx->r1.packed_bits = ctx->r2.packed_bits = -1;
    return HSQ_OK;
}

}  // namespace

extern "C" int32_t hsq_dedup_packed(hsq_ctx* ctx, const hsq_packed_pairs* pairs, int32_t u, int32_t v, uint8_t* is_best, int32_t* group_size,
                                    hsq_probe_stats* stats) {
    if (!ctx || !pairs || (pairs->n > 0 && !is_best)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "null argument");
    const int64_t n = pairs->n;
    if (n < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative size");
    if (n > 0 && !pairs->probe_id) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "probe_id is null");
    if (u < 0 || v < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "negative UID length");
    if (u + v > 32) return fail(ctx, HSQ_ERR_UNSUPPORTED, "UID longer than 32 bases in total is outside the accelerated domain");
    if (!stats || stats->n_probes <= 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "probe stats (with n_probes) are required");
    int rc;
    if ((rc = check_packed_reads(ctx, &pairs->r1, n, "hsq_dedup_packed")) || (rc = check_packed_reads(ctx, &pairs->r2, n, "hsq_dedup_packed"))) return rc;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    const size_t nn = (size_t)n;
    const int64_t np = stats->n_probes;
    auto run = [&]() -> int {
        int rc = dedup_upload_keygen_packed(ctx, pairs, u, v, np);
        if (rc) return rc;
        HSQ_CUDA(ctx, ctx->out_u8.ensure(nn + 32));
        HSQ_CUDA(ctx, ctx->out_i32.ensure((nn + 3 * (size_t)np) * 4 + 16));
        uint8_t* best = ctx->out_u8.as<uint8_t>();
        int32_t* d_gs = ctx->out_i32.as<int32_t>();
        int32_t* d_stats = d_gs + nn;
        if ((rc = dedup_reduce_device(ctx, n, np, u + v, best, d_gs, d_stats))) return rc;
        if ((rc = d2h(ctx, is_best, best, nn))) return rc;
        if ((rc = d2h(ctx, group_size, d_gs, nn))) return rc;
        return fetch_stats(ctx, stats, d_stats);
    };
    rc = run();
    ctx->r1.packed_bits = ctx->r2.packed_bits = 0;
    int rc2 = timer.finish();
    return rc ? rc : rc2;
}
