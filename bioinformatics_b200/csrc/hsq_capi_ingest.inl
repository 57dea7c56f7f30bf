// hsq_capi_ingest.inl -- SURVEY §8(f) row f2: FASTQ ingest and the FASTQ <-> SAM name join in front of the dedup path.
//   hsq_index_fastq : record index of FASTQ text (htsjdk FastqReader grammar; call sites HSQ/utils/FastqSorter.java:77,98)
//   hsq_join_names  : FastqAndBamFileMerger.createBamFileWithDataFromRawFastqFiles without the codecs
//                     (HSQ/process/FastqAndBamFileMerger.java:107-158, MergedSamIterator :213-345, trim check :65-94,366-373)
//   hsq_gather_rows : the raw read / quality strings the merged records take (:386-396)
//   hsq_index_sam / hsq_sam_mapped_records : the SAMRecord getters of the merge and of ReadToProbeAssigner.assignProbesToReads over SAM
//                     text (...Identification.java:439-474; XA entries as HSQ/utils/SAMRecordUtil.java:440-468 keeps them)
//   hsq_trim_fastq  : FastqReadTrimmer.trimReads, the file loop of `HSQutils trim`, text to text (HSQ/process/FastqReadTrimmer.java:262-307)
//   hsq_coordinate_order : order of the final coordinate sort (...Identification.java:897, htsjdk SAMRecordCoordinateComparator)
// The reference's merge is a sequential two-cursor loop over two sorted iterators.  Here both sides are sorted on the device
// (LSD radix over big-endian 8-byte chunks of the zero-padded keys = String.compareTo for ASCII), and the loop becomes, for
// the j-th primary SAM record in sorted order with unique id s:  c = number of earlier not-first-of-pair records with the
// same id (each of them, once matched, moves the FASTQ cursor on, :304-309);  p = first sorted FASTQ position whose id is >= s;
// the record is returned iff FASTQ position p + c exists and carries id s.  That equals the loop whenever ids ascend along both
// sorted sides; if they do not (a sort key orders differently from the id it contains) the call fails with HSQ_ERR_UNSUPPORTED.
#include <cub/device/device_scan.cuh>
#include <climits>
namespace hsq {

// ---------------------------------------------------------------------------------------------------------------------------
// FASTQ text -> line starts -> records
static constexpr int NL_BLOCK = 256, NL_BYTES = 16, NL_TILE = NL_BLOCK * NL_BYTES;

__device__ inline uint32_t newline_mask16(const uint8_t* __restrict__ text, int64_t base, int64_t n, int* lone_cr) {
    __align__(16) uint8_t b[2 * NL_BYTES];
    if (base + NL_BYTES <= n) {
        const uint4 v = *reinterpret_cast<const uint4*>(text + base);  // base is a multiple of 16, text is cudaMalloc aligned
        *reinterpret_cast<uint4*>(b) = v;
        b[NL_BYTES] = base + NL_BYTES < n ? text[base + NL_BYTES] : (uint8_t)'\n';
    } else {
        for (int k = 0; k <= NL_BYTES; k++) b[k] = base + k < n ? text[base + k] : (uint8_t)'\n';
    }
    uint32_t m = 0;
    for (int k = 0; k < NL_BYTES; k++) {
        if (base + k >= n) break;
        if (b[k] == '\n') m |= 1u << k;
        if (b[k] == '\r' && b[k + 1] != '\n') *lone_cr = 1;  // end of text after '\r' counts as a terminator too
    }
    return m;
}

__global__ void nl_count_kernel(const uint8_t* __restrict__ text, int64_t n, int64_t* __restrict__ tile_counts, int* status) {
    const int64_t base = (int64_t)blockIdx.x * NL_TILE + (int64_t)threadIdx.x * NL_BYTES;
    int lone = 0;
    const uint32_t m = base < n ? newline_mask16(text, base, n, &lone) : 0u;
    if (lone) atomicMin(status, HSQ_ERR_UNSUPPORTED);
    int64_t total;
    block_exclusive_scan((int64_t)__popc(m), &total);
    if (threadIdx.x == 0) tile_counts[blockIdx.x] = total;
}

// line_start[0] = 0; line_start[k] = byte after the k-th newline
__global__ void nl_write_kernel(const uint8_t* __restrict__ text, int64_t n, const int64_t* __restrict__ tile_excl, int64_t* __restrict__ line_start) {
    const int64_t base = (int64_t)blockIdx.x * NL_TILE + (int64_t)threadIdx.x * NL_BYTES;
    int lone = 0;
    uint32_t m = base < n ? newline_mask16(text, base, n, &lone) : 0u;
    int64_t total;
    int64_t rank = tile_excl[blockIdx.x] + block_exclusive_scan((int64_t)__popc(m), &total) + 1;
    if (blockIdx.x == 0 && threadIdx.x == 0) line_start[0] = 0;
    while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        line_start[rank++] = base + k + 1;
    }
}

__global__ void set_i64_kernel(int64_t* p, int64_t v) { *p = v; }

__global__ void fastq_record_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ line_start, int64_t n_rec,
                                    int64_t* __restrict__ h_off, int32_t* __restrict__ h_len, int64_t* __restrict__ s_off, int32_t* __restrict__ s_len,
                                    int64_t* __restrict__ q_off, int32_t* __restrict__ q_len, int* status, unsigned long long* first_bad,
                                    int64_t* __restrict__ p_off = nullptr, int32_t* __restrict__ p_len = nullptr) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n_rec; r += stride) {
        int64_t off[4];
        int32_t len[4];
        bool bad = false;
        for (int k = 0; k < 4; k++) {
            off[k] = line_start[4 * r + k];
            int64_t l = line_start[4 * r + k + 1] - 1 - off[k];
            if (l > 0 && text[off[k] + l - 1] == '\r') l--;
            len[k] = (int32_t)l;
            bool blank = true;  // StringUtil.isBlank: FastqReader rejects such a line
            for (int64_t i = 0; i < l && blank; i++) blank = text[off[k] + i] <= ' ';
            bad |= blank;
        }
        if (!bad) bad = text[off[0]] != '@' || text[off[2]] != '+' || len[1] != len[3];
        if (bad) {
            atomicMin(status, HSQ_ERR_MALFORMED);
            atomicMin(first_bad, (unsigned long long)r);
        }
        h_off[r] = off[0] + 1; h_len[r] = len[0] - 1;
        s_off[r] = off[1]; s_len[r] = len[1];
        q_off[r] = off[3]; q_len[r] = len[3];
        if (p_off) { p_off[r] = off[2] + 1; p_len[r] = len[2] - 1; }  // quality header without the '+'
    }
}

// FastqReadTrimmer.trimReads: output bytes per record, flags the reads Java throws on (FastqReadTrimmer.java:310-312)
__global__ void trimfq_len_kernel(const int32_t* __restrict__ h_len, const int32_t* __restrict__ s_len, const int32_t* __restrict__ p_len, int64_t n,
                                  int first, int last, int three_prime, int64_t* __restrict__ out_len, int32_t* __restrict__ keep, int* status) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n; r += stride) {
        const int len = s_len[r];
        int lb = len - 1;
        if (three_prime) lb = min(last, len - 1);
        int kp = lb + 1 - first;
        if (first >= len || kp < 0) { atomicMin(status, HSQ_ERR_OUT_OF_RANGE); kp = 0; }
        keep[r] = kp;
        out_len[r] = (int64_t)h_len[r] + p_len[r] + 2 * (int64_t)kp + 6;
    }
}

// one warp per record: "@" header "\n" bases "\n+" quality header "\n" qualities "\n"
__global__ void trimfq_write_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ h_off, const int32_t* __restrict__ h_len,
                                    const int64_t* __restrict__ s_off, const int64_t* __restrict__ q_off, const int64_t* __restrict__ p_off,
                                    const int32_t* __restrict__ p_len, const int32_t* __restrict__ keep, const int64_t* __restrict__ out_off, int64_t n,
                                    int first, int64_t cap, uint8_t* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; r < n; r += stride) {
        const int64_t d = out_off[r];
        const int total = (int)(out_off[r + 1] - d);
        if (d + total > cap) continue;
        const int hl = h_len[r], pl = p_len[r], kp = keep[r];
        const int64_t ho = h_off[r] - 1, so = s_off[r] + first - (hl + 2), po = p_off[r] - (hl + kp + 4), qo = q_off[r] + first - (hl + kp + pl + 5);
        const int e_head = hl + 1, e_seq = e_head + 1 + kp, e_plus = e_seq + 2 + pl, e_qual = e_plus + 1 + kp;
        for (int k = lane; k < total; k += 32) {
            uint8_t c;
            if (k == 0) c = '@';
            else if (k < e_head) c = text[ho + k];
            else if (k == e_head) c = '\n';
            else if (k < e_seq) c = text[so + k];
            else if (k == e_seq) c = '\n';
            else if (k == e_seq + 1) c = '+';
            else if (k < e_plus) c = text[po + k];
            else if (k == e_plus) c = '\n';
            else if (k < e_qual) c = text[qo + k];
            else c = '\n';
            out[d + k] = c;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------------
// SAM text: header lines are the leading '@' lines; one warp per alignment line finds the tabs with ballots
__global__ void sam_first_line_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ line_start, int64_t n_lines,
                                      unsigned long long* first_alignment_line) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n_lines; i += stride) {
        const int64_t s = line_start[i];
        if (!(s < line_start[i + 1] - 1 && text[s] == '@')) atomicMin(first_alignment_line, (unsigned long long)i);
    }
}

__device__ inline bool parse_i32_field(const uint8_t* __restrict__ p, int len, int32_t* out) {  // Integer.parseInt
    int i = 0;
    bool neg = false;
    if (len > 0 && (p[0] == '-' || p[0] == '+')) { neg = p[0] == '-'; i = 1; }
    if (i >= len) return false;
    int64_t v = 0;
    for (; i < len; i++) {
        if (p[i] < '0' || p[i] > '9') return false;
        v = v * 10 + (p[i] - '0');
        if (v > (int64_t)INT_MAX + 1) return false;
    }
    if (neg) v = -v;
    if (v > INT_MAX || v < INT_MIN) return false;
    *out = (int32_t)v;
    return true;
}

__device__ inline bool cigar_reference_length(const uint8_t* __restrict__ p, int len, int32_t* out) {  // Cigar.getReferenceLength
    if (len == 1 && p[0] == '*') { *out = 0; return true; }
    if (len == 0) return false;
    int64_t rl = 0, count = 0;
    bool digits = false;
    for (int k = 0; k < len; k++) {
        const uint8_t c = p[k];
        if (c >= '0' && c <= '9') {
            count = count * 10 + (c - '0');
            digits = true;
            if (count > INT_MAX) return false;
            continue;
        }
        if (!digits) return false;
        if (c == 'M' || c == 'D' || c == 'N' || c == '=' || c == 'X') rl += count;
        else if (!(c == 'I' || c == 'S' || c == 'H' || c == 'P')) return false;
        count = 0;
        digits = false;
    }
    if (digits || rl > INT_MAX) return false;
    *out = (int32_t)rl;
    return true;
}

__global__ void sam_line_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ line_start, int64_t first_line, int64_t n_rec,
                                int64_t* __restrict__ off, int32_t* __restrict__ val, int* status, unsigned long long* first_bad) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    // column this lane extracts: lanes 0-4 the string columns QNAME, RNAME, CIGAR, SEQ, QUAL; lanes 8-12 FLAG, POS, MAPQ, PNEXT, TLEN
    const int col = lane == 1 ? 2 : lane == 2 ? 5 : lane == 3 ? 9 : lane == 4 ? 10 : lane == 8 ? 1 : lane == 9 ? 3 : lane == 10 ? 4 : lane == 11 ? 7 : lane == 12 ? 8 : 0;
    int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; r < n_rec; r += stride) {
        const int64_t s = line_start[first_line + r];
        int64_t e = line_start[first_line + r + 1] - 1;
        if (e > s && text[e - 1] == '\r') e--;
        int64_t my_tab = e;  // lane k keeps the position of tab k (k < 11): no per-thread array
        int ntab = 0, total_tabs = 0;
        int64_t xa_s = -1, xa_e = -1;
        for (int64_t base = s; base < e; base += 32) {
            const int64_t pos = base + lane;
            const uint8_t c = pos < e ? text[pos] : (uint8_t)0;
            const unsigned mask = __ballot_sync(full, c == '\t');
            const int ordinal = total_tabs + __popc(mask & ((1u << lane) - 1u));  // tabs before this byte
            const bool is_xa = c == '\t' && ordinal >= 10 && pos + 6 <= e && text[pos + 1] == 'X' && text[pos + 2] == 'A' && text[pos + 3] == ':' &&
                               text[pos + 4] == 'Z' && text[pos + 5] == ':';
            const unsigned xmask = __ballot_sync(full, is_xa);
            if (xa_s < 0 && xmask) xa_s = base + (__ffs(xmask) - 1) + 6;
            if (xa_s >= 0 && xa_e < 0) {
                unsigned m2 = mask;
                if (xa_s > base) m2 = xa_s - base >= 32 ? 0u : (mask & ~((1u << (int)(xa_s - base)) - 1u));
                if (m2) xa_e = base + (__ffs(m2) - 1);
            }
            unsigned m = mask;
            while (m && ntab < 11) {
                if (lane == ntab) my_tab = base + (__ffs(m) - 1);
                ntab++;
                m &= m - 1;
            }
            total_tabs += __popc(mask);
            if (ntab >= 11 && xa_e >= 0) break;  // nothing left to find on this line
        }
        const int64_t t_prev = __shfl_sync(full, my_tab, col > 0 ? col - 1 : 0), t_cur = __shfl_sync(full, my_tab, col), t10 = __shfl_sync(full, my_tab, 10);
        bool bad = e == s || text[s] == '@' || ntab < 10;
        int64_t f_off = col == 0 ? s : t_prev + 1;
        int32_t f_len = (int32_t)((col < ntab ? t_cur : e) - f_off);
        if (lane == HSQ_SAM_TAGS) {
            f_off = ntab >= 11 ? t10 + 1 : e;
            f_len = (int32_t)(e - f_off);
        } else if (lane == HSQ_SAM_XA) {
            if (xa_s >= 0 && xa_e < 0) xa_e = e;
            const bool have = xa_s >= 0 && xa_e > xa_s;
            f_off = have ? xa_s : 0;
            f_len = have ? (int32_t)(xa_e - xa_s) : 0;
        }
        int32_t num = 0, ref_len = 0;
        if (!bad) {
            bool ok = true;
            if (lane >= 8 && lane <= 12) ok = parse_i32_field(text + f_off, f_len, &num);
            else if (lane == HSQ_SAM_CIGAR) ok = cigar_reference_length(text + f_off, f_len, &ref_len);
            bad = __any_sync(full, !ok);
        }
        if (bad) {
            if (lane == 0) {
                atomicMin(status, HSQ_ERR_MALFORMED);
                atomicMin(first_bad, (unsigned long long)r);
            }
            continue;
        }
        if (lane < HSQ_SAM_N_FIELDS) {
            off[(int64_t)lane * n_rec + r] = f_off;
            val[(int64_t)lane * n_rec + r] = f_len;
        }
        if (lane >= 8 && lane <= 12) val[(int64_t)(HSQ_SAM_FLAG + lane - 8) * n_rec + r] = num;
        if (lane == HSQ_SAM_CIGAR) val[(int64_t)HSQ_SAM_REF_LEN * n_rec + r] = ref_len;
    }
}

// ---------------------------------------------------------------------------------------------------------------------------
// SAM columns -> mapped records (hsq_sam_mapped_records)
struct NameDict {  // contig names in the caller's numbering
    const uint8_t* data;
    const int64_t* off;
    int n;
};
__device__ inline int dict_lookup(const NameDict& d, const uint8_t* __restrict__ p, int len) {
    for (int c = 0; c < d.n; c++) {
        const int64_t o = d.off[c];
        if ((int)(d.off[c + 1] - o) != len) continue;
        int i = 0;
        while (i < len && d.data[o + i] == p[i]) i++;
        if (i == len) return c;
    }
    return -1;
}

struct AltSink {
    int32_t* contig;
    uint8_t* negative;
    int32_t* start;
    int32_t* stop;
};

// AlternativeHit.fromAttributeText (HSQ/utils/SAMRecordUtil.java:440-468) over the XA value; returns the number of hits kept and
// writes them from slot `at` on when sink.contig is not null
__device__ inline int xa_hits(const uint8_t* __restrict__ p, int len, const NameDict& dict, const AltSink& sink, int64_t at, bool* bad_cigar) {
    int count = 0;
    int es = 0;
    while (es < len) {
        int ee = es;
        while (ee < len && p[ee] != ';') ee++;
        int f_start[4] = {0, 0, 0, 0}, f_len[4] = {0, 0, 0, 0};
        int nf = 0, last_nonempty = -1, fs = es;
        for (int k = es; k <= ee; k++) {
            if (k == ee || p[k] == ',') {
                if (nf < 4) { f_start[nf] = fs; f_len[nf] = k - fs; }
                if (k > fs) last_nonempty = nf;
                nf++;
                fs = k + 1;
            }
        }
        if (last_nonempty == 3) {  // String.split drops trailing empty strings: exactly four fields remain
            int32_t start_and_strand = 0, rl = 0;
            if (parse_i32_field(p + f_start[1], f_len[1], &start_and_strand)) {
                if (!cigar_reference_length(p + f_start[2], f_len[2], &rl)) *bad_cigar = true;
                else {
                    if (sink.contig) {
                        const int32_t st = start_and_strand < 0 ? -start_and_strand : start_and_strand;
                        sink.contig[at + count] = dict_lookup(dict, p + f_start[0], f_len[0]);
                        sink.negative[at + count] = start_and_strand < 0 ? 1 : 0;
                        sink.start[at + count] = st;
                        sink.stop[at + count] = st + rl - 1;
                    }
                    count++;
                }
            }
        }
        es = ee + 1;
    }
    return count;
}

// per alignment line: kept (both mates mapped, ...Identification.java:446) and its number of alternative hits
__global__ void samrec_count_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ off, const int32_t* __restrict__ val, int64_t n_rec,
                                    NameDict dict, int64_t* __restrict__ keep, int64_t* __restrict__ n_alt, int* status) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n_rec; r += stride) {
        const int f = val[(int64_t)HSQ_SAM_FLAG * n_rec + r];
        if (!(f & 0x1)) atomicMin(status, HSQ_ERR_UNSUPPORTED);  // getMateUnmappedFlag on an unpaired record throws
        const bool k = !(f & 0x4) && !(f & 0x8);
        int c = 0;
        const int xl = val[(int64_t)HSQ_SAM_XA * n_rec + r];
        if (k && xl > 0) {
            bool bad = false;
            c = xa_hits(text + off[(int64_t)HSQ_SAM_XA * n_rec + r], xl, dict, AltSink{nullptr, nullptr, nullptr, nullptr}, 0, &bad);
            if (bad) atomicMin(status, HSQ_ERR_MALFORMED);
        }
        keep[r] = k ? 1 : 0;
        n_alt[r] = c;
    }
}

__global__ void samrec_fill_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ off, const int32_t* __restrict__ val, int64_t n_rec,
                                   NameDict dict, const int64_t* __restrict__ keep_pos, const int64_t* __restrict__ alt_pos, int64_t cap, int64_t alt_cap,
                                   int32_t* __restrict__ sam_line, int32_t* __restrict__ contig, uint8_t* __restrict__ negative,
                                   uint8_t* __restrict__ second, int32_t* __restrict__ start, int32_t* __restrict__ end, int32_t* __restrict__ read_index,
                                   int64_t* __restrict__ alt_offsets, AltSink sink, int* status) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r <= n_rec; r += stride) {
        const int64_t j = keep_pos[r];
        if (r == n_rec) {
            if (j <= cap) alt_offsets[j] = alt_pos[r];
            continue;
        }
        if (keep_pos[r + 1] == j || j >= cap) continue;
        const int f = val[(int64_t)HSQ_SAM_FLAG * n_rec + r];
        const int c = dict_lookup(dict, text + off[(int64_t)HSQ_SAM_RNAME * n_rec + r], val[(int64_t)HSQ_SAM_RNAME * n_rec + r]);
        int32_t ri = 0;
        if (c >= 0 && !parse_i32_field(text + off[(int64_t)HSQ_SAM_QNAME * n_rec + r], val[(int64_t)HSQ_SAM_QNAME * n_rec + r], &ri))
            atomicMin(status, HSQ_ERR_BAD_CHARACTER);  // "Expecting all read names to be numbers" (:471-474)
        const int32_t pos = val[(int64_t)HSQ_SAM_POS * n_rec + r];
        sam_line[j] = (int32_t)r;
        contig[j] = c;
        negative[j] = (f & 0x10) ? 1 : 0;
        second[j] = (f & 0x80) ? 1 : 0;
        start[j] = pos;
        end[j] = pos + val[(int64_t)HSQ_SAM_REF_LEN * n_rec + r] - 1;
        read_index[j] = ri;
        alt_offsets[j] = alt_pos[r];
        const int xl = val[(int64_t)HSQ_SAM_XA * n_rec + r];
        if (xl > 0 && alt_pos[r + 1] <= alt_cap) {
            bool bad = false;
            xa_hits(text + off[(int64_t)HSQ_SAM_XA * n_rec + r], xl, dict, sink, alt_pos[r], &bad);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------------
// keys of the final coordinate sort (htsjdk SAMRecordCoordinateComparator): signed integers biased to unsigned order
__global__ void coord_keys_kernel(const int32_t* __restrict__ ref, const int32_t* __restrict__ pos, const int32_t* __restrict__ flags,
                                  const int32_t* __restrict__ mapq, const int32_t* __restrict__ mate_ref, const int32_t* __restrict__ mate_pos,
                                  const int32_t* __restrict__ tlen, int64_t n, uint64_t* __restrict__ k_place, uint64_t* __restrict__ k_strand,
                                  uint64_t* __restrict__ k_flags, uint64_t* __restrict__ k_mapq_mref, uint64_t* __restrict__ k_mpos_tlen) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    auto b = [](int32_t v) { return (uint64_t)((uint32_t)v ^ 0x80000000u); };
    for (; i < n; i += stride) {
        const uint64_t rk = ref[i] == -1 ? 0xFFFFFFFFull : b(ref[i]);           // records without a reference go last
        k_place[i] = (rk << 32) | (ref[i] == -1 ? 0ull : b(pos[i]));              // ... and are not ordered by position
        k_strand[i] = (flags[i] & 0x10) ? 1 : 0;
        k_flags[i] = b(flags[i]);
        k_mapq_mref[i] = (b(mapq[i]) << 32) | b(mate_ref[i]);
        k_mpos_tlen[i] = (b(mate_pos[i]) << 32) | b(tlen[i]);
    }
}

// ---------------------------------------------------------------------------------------------------------------------------
// ragged row gather
__global__ void rows_len_kernel(const int64_t* __restrict__ off, const int32_t* __restrict__ len, int64_t n, int64_t n_bytes,
                                int64_t* __restrict__ len64, int* status) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) {
        int64_t l = len[k];
        if (l < 0 || off[k] < 0 || off[k] + l > n_bytes) { atomicMin(status, HSQ_ERR_OUT_OF_RANGE); l = 0; }
        len64[k] = l;
    }
}

__global__ void rows_copy_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ off, const int64_t* __restrict__ dst_off, int64_t n,
                                 int64_t cap, uint8_t* __restrict__ dst) {
    const int lane = threadIdx.x & 31;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; w < n; w += stride) {
        const int64_t s = off[w], d = dst_off[w];
        const int len = (int)(dst_off[w + 1] - d);
        if (d + len > cap) continue;
        for (int i = lane; i < len; i += 32) dst[d + i] = text[s + i];
    }
}

// ---------------------------------------------------------------------------------------------------------------------------
// name join
enum { JC_ORDER = 0, JC_MISMATCH = 1, JC_KMAX = 2, JC_TRIM = 3, JC_EMITTED = 4, JC_PAIRS = 5, JC_PARSE = 6, JC_NONPRIMARY = 7,
       JC_MAX_SAM = 8, JC_MAX_FQ1 = 9, JC_MAX_FQ2 = 10, JC_WORDS = 16 };

__device__ inline int unique_id_len(const uint8_t* __restrict__ p, int len) {  // IlluminaFastQReadNameUtil.java:40-58
    for (int i = 0; i < len; i++)
        if (p[i] == ' ' || p[i] == '/') return i;
    return len;
}
// String.compareTo for ASCII: first differing byte, else the shorter string first
__device__ inline int str_cmp(const uint8_t* __restrict__ a, int la, const uint8_t* __restrict__ b, int lb) {
    const int m = la < lb ? la : lb;
    for (int i = 0; i < m; i++) {
        const int d = (int)a[i] - (int)b[i];
        if (d) return d;
    }
    return la - lb;
}

// FASTQ sort keys: header [+ " _ " + ordinal] (FastqSorter.java:298-302), zero padded, as big-endian 8-byte chunks, chunk-major
__global__ void fq_keys_kernel(const uint8_t* __restrict__ data, const int64_t* __restrict__ off, int64_t n, int n_chunks, int append_index,
                               uint64_t* __restrict__ keys, int32_t* __restrict__ uid_len) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n; r += stride) {
        const uint8_t* p = data + off[r];
        const int len = (int)(off[r + 1] - off[r]);
        uid_len[r] = unique_id_len(p, len);
        if (!keys) continue;
        uint8_t suf[16];
        int sl = 0;
        if (append_index) {
            suf[0] = ' '; suf[1] = '_'; suf[2] = ' ';
            uint8_t dig[12];
            int nd = 0;
            int64_t v = r;
            do { dig[nd++] = (uint8_t)('0' + v % 10); v /= 10; } while (v);
            sl = 3;
            while (nd) suf[sl++] = dig[--nd];
        }
        const int total = len + sl;
        for (int c = 0; c < n_chunks; c++) {
            uint64_t w = 0;
            for (int b = 0; b < 8; b++) {
                const int pos = c * 8 + b;
                const uint8_t byte = pos < len ? p[pos] : (pos < total ? suf[pos - len] : (uint8_t)0);
                w = (w << 8) | byte;
            }
            keys[(int64_t)c * n + r] = w;
        }
    }
}

// SAM sort keys by name (BamSorter.java:348-361): name chunks, a first-of-pair-first key, a primary-records-first key
__global__ void sam_keys_kernel(const uint8_t* __restrict__ data, const int64_t* __restrict__ off, const int32_t* __restrict__ flags, int64_t n,
                                int n_chunks, uint64_t* __restrict__ keys, uint64_t* __restrict__ k_nonfirst, uint64_t* __restrict__ k_nonprimary,
                                int32_t* __restrict__ uid_len, int* ctr) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n; r += stride) {
        const uint8_t* p = data + off[r];
        const int len = (int)(off[r + 1] - off[r]);
        uid_len[r] = unique_id_len(p, len);
        const int f = flags[r];
        k_nonfirst[r] = (f & 0x40) ? 0 : 1;
        k_nonprimary[r] = (f & 0x100) ? 1 : 0;
        if (f & 0x100) atomicAdd(ctr + JC_NONPRIMARY, 1);
        for (int c = 0; c < n_chunks; c++) {
            uint64_t w = 0;
            for (int b = 0; b < 8; b++) {
                const int pos = c * 8 + b;
                w = (w << 8) | (pos < len ? p[pos] : (uint8_t)0);
            }
            keys[(int64_t)c * n + r] = w;
        }
    }
}

// SAM sort key by index (BamSorter.java:322-346): Integer.parseInt(name), first of pair first, primary records first
__global__ void sam_intkeys_kernel(const uint8_t* __restrict__ data, const int64_t* __restrict__ off, const int32_t* __restrict__ flags, int64_t n,
                                   uint64_t* __restrict__ key, int32_t* __restrict__ val, int32_t* __restrict__ uid_len, int* ctr) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; r < n; r += stride) {
        const uint8_t* p = data + off[r];
        const int len = (int)(off[r + 1] - off[r]);
        int i = 0;
        bool neg = false, ok = true;
        if (len > 0 && (p[0] == '-' || p[0] == '+')) { neg = p[0] == '-'; i = 1; }
        if (i >= len) ok = false;
        int64_t v = 0;
        for (; i < len && ok; i++) {
            if (p[i] < '0' || p[i] > '9') ok = false;
            else {
                v = v * 10 + (p[i] - '0');
                if (v > (int64_t)INT_MAX + 1) ok = false;
            }
        }
        if (neg) v = -v;
        if (v > INT_MAX || v < INT_MIN) ok = false;
        if (!ok) { atomicExch(ctr + JC_PARSE, 1); v = 0; }
        const int f = flags[r];
        val[r] = (int32_t)v;
        uid_len[r] = len;
        if (f & 0x100) atomicAdd(ctr + JC_NONPRIMARY, 1);
        key[r] = ((uint64_t)((f & 0x100) ? 1 : 0) << 33) | ((uint64_t)(uint32_t)((int32_t)v ^ INT_MIN) << 1) | ((f & 0x40) ? 0u : 1u);
    }
}

// live[c] = 1 when chunk c differs between records (a chunk all records share needs no sort pass)
__global__ void chunk_live_kernel(const uint64_t* __restrict__ keys, int64_t n, int n_chunks, int* __restrict__ live) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t total = n * n_chunks;
    for (; i < total; i += stride) {
        const int64_t c = i / n;
        if (keys[i] != keys[c * n] && !live[c]) live[c] = 1;
    }
}

__global__ void iota_u32_kernel(uint32_t* __restrict__ out, int64_t n) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; k < n; k += stride) out[k] = (uint32_t)k;
}

// sorted view of one FASTQ file: where the unique id of sorted position q sits
__global__ void fq_sorted_view_kernel(const uint32_t* __restrict__ perm, const int64_t* __restrict__ off, const int32_t* __restrict__ uid_len, int64_t n,
                                      int64_t* __restrict__ s_off, int32_t* __restrict__ s_len) {
    int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; q < n; q += stride) {
        const int64_t r = perm ? perm[q] : q;
        s_off[q] = off[r];
        s_len[q] = uid_len[r];
    }
}

struct SortedIds {
    const uint8_t* data;
    const int64_t* off;  // by sorted position
    const int32_t* len;
};

// ids must ascend along the sorted FASTQ-one side; FASTQ one / two ids must agree position by position (:266-273)
__global__ void fq_check_kernel(SortedIds a, SortedIds b, int64_t n1, int64_t nf, int by_index, int* ctr) {
    int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; q < n1; q += stride) {
        if (!by_index && q > 0 && str_cmp(a.data + a.off[q - 1], a.len[q - 1], a.data + a.off[q], a.len[q]) > 0) atomicExch(ctr + JC_ORDER, 1);
        if (q < nf && str_cmp(a.data + a.off[q], a.len[q], b.data + b.off[q], b.len[q]) != 0) atomicMin(ctr + JC_MISMATCH, (int)q);
    }
}

// per primary SAM record in sorted order: start of its id group and whether it is a not-first-of-pair record
__global__ void sam_heads_kernel(const uint8_t* __restrict__ data, const int64_t* __restrict__ off, const int32_t* __restrict__ uid_len,
                                 const int32_t* __restrict__ flags, const uint32_t* __restrict__ perm, int64_t np, const int32_t* __restrict__ val,
                                 int32_t* __restrict__ headpos, int32_t* __restrict__ nonfirst, int* ctr) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; j < np; j += stride) {
        const uint32_t s = perm[j];
        int cmp = 1;
        if (j > 0) {
            const uint32_t t = perm[j - 1];
            if (val) cmp = val[s] > val[t] ? 1 : (val[s] < val[t] ? -1 : 0);
            else cmp = str_cmp(data + off[s], uid_len[s], data + off[t], uid_len[t]);
            if (cmp < 0) atomicExch(ctr + JC_ORDER, 1);
        }
        headpos[j] = cmp != 0 ? (int32_t)j : 0;
        nonfirst[j] = (flags[s] & 0x40) ? 0 : 1;
    }
}

struct TrimParams {
    int check, three_prime, r1_first, r1_stop, r2_first, r2_stop;
};

__device__ inline int64_t fq_lower_bound(const SortedIds& f, int64_t nf, const uint8_t* s, int ls, bool upper) {
    int64_t lo = 0, hi = nf;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        const int c = str_cmp(f.data + f.off[mid], f.len[mid], s, ls);
        if (c < 0 || (upper && c == 0)) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

__global__ void join_match_kernel(const uint8_t* __restrict__ data, const int64_t* __restrict__ off, const int32_t* __restrict__ uid_len,
                                  const int32_t* __restrict__ flags, const int32_t* __restrict__ seq_len, const int32_t* __restrict__ qual_len,
                                  const uint32_t* __restrict__ perm, int64_t np, const int32_t* __restrict__ val, const int32_t* __restrict__ headpos,
                                  const int32_t* __restrict__ excl_nonfirst, SortedIds f1, int64_t nf, const uint32_t* __restrict__ perm1,
                                  const uint32_t* __restrict__ perm2, const int32_t* __restrict__ fq1_len, const int32_t* __restrict__ fq2_len,
                                  TrimParams tp, int32_t* __restrict__ emit, int32_t* __restrict__ used, int32_t* __restrict__ nlen,
                                  int32_t* __restrict__ target_of, int* ctr) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; j < np; j += stride) {
        const uint32_t s = perm[j];
        const uint8_t* id = data + off[s];
        const int ls = uid_len[s];
        const int64_t cnt = (int64_t)excl_nonfirst[j] - excl_nonfirst[headpos[j]];
        const bool first = (flags[s] & 0x40) != 0;
        int64_t p0, target;
        bool matched;
        if (val) {  // FASTQ key of position k is k itself (:274-277)
            p0 = val[s] < 0 ? 0 : val[s];
            target = p0 + cnt;
            matched = val[s] >= 0 && cnt == 0 && target < nf;
        } else {
            p0 = fq_lower_bound(f1, nf, id, ls, false);
            target = p0 + cnt;
            matched = target < nf && str_cmp(f1.data + f1.off[target], f1.len[target], id, ls) == 0;
        }
        int e = 0, u = -1, nl = -1;
        if (matched) {
            e = 1;
            u = (int32_t)(first ? (perm1 ? perm1[target] : target) : (perm2 ? perm2[target] : target));
            nl = f1.len[target];
            if (!first) atomicAdd(ctr + JC_PAIRS, 1);
            if (tp.check) {  // isTrimAmountCorrect (:65-94) over FastqReadTrimmer.trim (FastqReadTrimmer.java:309-328)
                const int len = first ? fq1_len[u] : fq2_len[u];
                const int t_first = first ? tp.r1_first : tp.r2_first, t_stop = first ? tp.r1_stop : tp.r2_stop;
                int last_base = len - 1;
                if (tp.three_prime) last_base = min(t_stop, len - 1);
                const int expected = last_base + 1 - t_first;
                if (t_first >= len || expected < 0) atomicMin((unsigned*)(ctr + JC_TRIM), (unsigned)j * 2u);
                else if (seq_len[s] > expected || qual_len[s] > expected) atomicMin((unsigned*)(ctr + JC_TRIM), (unsigned)j * 2u + 1u);
            }
        }
        emit[j] = e;
        used[j] = u;
        nlen[j] = nl;
        target_of[j] = (int32_t)(target < nf ? target : nf);
        if (j == np - 1) {  // last FASTQ position the loop fetches: where the cursor stands at the last primary record
            int64_t cur = target;
            if (!matched) {
                if (val) cur = val[s] < 0 ? 0 : (cnt == 0 ? target : p0 + 1);
                else {
                    const int64_t ub = fq_lower_bound(f1, nf, id, ls, true);
                    cur = target < ub ? target : ub;
                }
            }
            ctr[JC_KMAX] = (int)(cur < nf ? cur : nf - 1);
        }
    }
}

__global__ void join_scatter_kernel(const uint32_t* __restrict__ perm, int64_t n, int64_t np, const int32_t* __restrict__ emit,
                                    const int32_t* __restrict__ epos, const int32_t* __restrict__ used, const int32_t* __restrict__ nlen,
                                    int32_t* __restrict__ sam_record, int32_t* __restrict__ fastq_record, int32_t* __restrict__ name_len,
                                    uint8_t* __restrict__ status, int* ctr) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; j < n; j += stride) {
        const uint32_t s = perm[j];
        if (j >= np) { status[s] = 0; continue; }
        if (emit[j]) {
            const int32_t e = epos[j];
            sam_record[e] = (int32_t)s;
            fastq_record[e] = used[j];
            name_len[e] = nlen[j];
            status[s] = 1;
        } else {
            status[s] = 2;
        }
        if (j == np - 1) ctr[JC_EMITTED] = epos[j] + emit[j];
    }
}

}  // namespace hsq

using namespace hsq;

namespace {

inline int grid_for(hsq_ctx* ctx, int64_t n, int block, int per_sm) {
    return (int)std::max<int64_t>(1, std::min<int64_t>((n + block - 1) / block, (int64_t)ctx->sm_count * per_sm));
}

// LSD radix sort of chunk-major keys: `perm` (sorted position -> record) ends up in *result (one of pa / pb)
int lsd_chunk_sort(hsq_ctx* ctx, const uint64_t* keys, int n_chunks, const int* live, int64_t n, uint32_t** cur, uint32_t** other, uint64_t* ka,
                   uint64_t* kb) {
    const int g = grid_for(ctx, n, 256, 16);
    for (int c = n_chunks - 1; c >= 0; c--) {
        if (!live[c]) continue;
        gather_u64_kernel<uint64_t><<<g, 256, 0, ctx->stream>>>(keys + (int64_t)c * n, *cur, n, ka);
        ctx->launches++;
        int rc = sort_pass(ctx, ka, kb, *cur, *other, n, 64);
        if (rc) return rc;
        std::swap(*cur, *other);
    }
    return HSQ_OK;
}

int bit_sort(hsq_ctx* ctx, const uint64_t* key, int64_t n, uint32_t** cur, uint32_t** other, uint64_t* ka, uint64_t* kb, int end_bit = 1) {
    gather_u64_kernel<uint64_t><<<grid_for(ctx, n, 256, 16), 256, 0, ctx->stream>>>(key, *cur, n, ka);
    ctx->launches++;
    int rc = sort_pass(ctx, ka, kb, *cur, *other, n, end_bit);
    if (rc) return rc;
    std::swap(*cur, *other);
    return HSQ_OK;
}

struct Arena {  // bump allocator over one DevBuf, sized up front
    size_t need = 0;
    uint8_t* base = nullptr;
    size_t reserve(size_t bytes) {
        const size_t o = need;
        need += (bytes + 255) & ~(size_t)255;
        return o;
    }
    template <class T> T* at(size_t o) const { return reinterpret_cast<T*>(base + o); }
};

int upload_strings(hsq_ctx* ctx, const hsq_seqs* s, DevBuf* data, DevBuf* off, const char* what) {
    if (s->n < 0 || s->n >= (1ll << 31) || !s->offsets) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, std::string("hsq_join_names: bad batch: ") + what);
    if (s->offsets[0] != 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, std::string("hsq_join_names: offsets must start at 0: ") + what);
    const int64_t total = s->offsets[s->n];
    if (total < 0 || (total > 0 && !s->data)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, std::string("hsq_join_names: null data: ") + what);
    HSQ_CUDA(ctx, data->ensure((size_t)total + 16));
    HSQ_CUDA(ctx, off->ensure((size_t)(s->n + 1) * 8));
    if (total) HSQ_CUDA(ctx, cudaMemcpyAsync(data->p, s->data, (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
    HSQ_CUDA(ctx, cudaMemcpyAsync(off->p, s->offsets, (size_t)(s->n + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
    return HSQ_OK;
}

// text -> device copy + line starts (line i = [lines[i], lines[i+1] - 1), '\n' excluded; a last line without terminator counts)
int build_line_index(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, uint8_t** d_text_out, int64_t** d_lines_out, int64_t* n_lines_out) {
    const int64_t n_tiles = (n_bytes + NL_TILE - 1) / NL_TILE;
    HSQ_CUDA(ctx, ctx->r1.ascii.ensure((size_t)n_bytes + 32));
    HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)(n_tiles + 1) * 8 + 16));
    uint8_t* d_text = ctx->r1.ascii.as<uint8_t>();
    int64_t* d_tiles = ctx->sort_c.as<int64_t>();
    HSQ_CUDA(ctx, cudaMemcpyAsync(d_text, text, (size_t)n_bytes, cudaMemcpyHostToDevice, ctx->stream));
    HSQ_CUDA(ctx, cudaMemsetAsync(d_tiles + n_tiles, 0, 8, ctx->stream));
    HSQ_CUDA(ctx, cudaMemsetAsync(ctx->cells.p, 0xff, 16, ctx->stream));  // first malformed record | first alignment line
    nl_count_kernel<<<(unsigned)n_tiles, NL_BLOCK, 0, ctx->stream>>>(d_text, n_bytes, d_tiles, d_status(ctx));
    scan_sums_kernel<<<1, SCAN_BLOCK, 0, ctx->stream>>>(d_tiles, n_tiles + 1);  // in place: exclusive; entry n_tiles = total
    ctx->launches += 2;
    HSQ_CUDA(ctx, cudaGetLastError());
    int64_t n_newlines = 0;
    HSQ_CUDA(ctx, cudaMemcpyAsync(&n_newlines, d_tiles + n_tiles, 8, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const bool open_tail = text[n_bytes - 1] != '\n';  // a last line without terminator is still a line (BufferedReader.readLine)
    const int64_t n_lines = n_newlines + (open_tail ? 1 : 0);
    HSQ_CUDA(ctx, ctx->sort_a.ensure((size_t)(n_lines + 2) * 8));
    int64_t* d_lines = ctx->sort_a.as<int64_t>();
    nl_write_kernel<<<(unsigned)n_tiles, NL_BLOCK, 0, ctx->stream>>>(d_text, n_bytes, d_tiles, d_lines);
    ctx->launches++;
    if (open_tail) {
        set_i64_kernel<<<1, 1, 0, ctx->stream>>>(d_lines + n_lines, n_bytes + 1);
        ctx->launches++;
    }
    *d_text_out = d_text;
    *d_lines_out = d_lines;
    *n_lines_out = n_lines;
    return HSQ_OK;
}

}  // namespace

extern "C" int32_t hsq_index_fastq(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, hsq_fastq_index* out) {
    if (!ctx || !out || n_bytes < 0 || (n_bytes > 0 && !text)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_index_fastq: null argument");
    out->n = 0;
    if (n_bytes == 0) return HSQ_OK;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = reset_status(ctx);
    if (rc) return rc;
    uint8_t* d_text;
    int64_t* d_lines;
    int64_t n_lines;
    if ((rc = build_line_index(ctx, text, n_bytes, &d_text, &d_lines, &n_lines))) return rc;
    const int64_t n_rec = n_lines / 4;
    out->n = n_rec;
    int st[4];
    if (n_lines % 4 != 0) {
        if ((rc = read_status(ctx, st, 1))) return rc;
        if (st[0] == HSQ_ERR_UNSUPPORTED) return fail(ctx, st[0], "hsq_index_fastq: a carriage return that is not followed by a line feed");
        return fail(ctx, HSQ_ERR_MALFORMED, "hsq_index_fastq: File is too short: the last record has fewer than four lines");
    }
    if (n_rec > out->cap) return fail(ctx, HSQ_ERR_OVERFLOW, "hsq_index_fastq: more records than out->cap");
    if (n_rec > 0 && (!out->header_off || !out->header_len || !out->seq_off || !out->seq_len || !out->qual_off || !out->qual_len))
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_index_fastq: null output array");
    const size_t nr = (size_t)n_rec;
    HSQ_CUDA(ctx, ctx->out_i32.ensure(nr * 12 + 64));
    HSQ_CUDA(ctx, ctx->sort_b.ensure(nr * 24 + 64));
    int64_t* d_o64 = ctx->sort_b.as<int64_t>();
    int32_t* d_o32 = ctx->out_i32.as<int32_t>();
    if (n_rec > 0) {
        fastq_record_kernel<<<grid_for(ctx, n_rec, 128, 16), 128, 0, ctx->stream>>>(d_text, d_lines, n_rec, d_o64, d_o32, d_o64 + nr, d_o32 + nr,
                                                                                 d_o64 + 2 * nr, d_o32 + 2 * nr, d_status(ctx),
                                                                                 ctx->cells.as<unsigned long long>());
        ctx->launches++;
    }
    HSQ_CUDA(ctx, cudaGetLastError());
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->h_cells, ctx->cells.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    if ((rc = read_status(ctx, st, 1))) return rc;
    if (st[0] == HSQ_ERR_UNSUPPORTED) return fail(ctx, st[0], "hsq_index_fastq: a carriage return that is not followed by a line feed");
    if (st[0]) return fail(ctx, st[0], "hsq_index_fastq: record " + std::to_string(ctx->h_cells[0]) +
                                           " is not '@' header / bases / '+' header / qualities of the same length (or holds a blank line)");
    struct { void* h; const void* d; size_t bytes; } down[] = {{out->header_off, d_o64, nr * 8}, {out->seq_off, d_o64 + nr, nr * 8},
                                                               {out->qual_off, d_o64 + 2 * nr, nr * 8}, {out->header_len, d_o32, nr * 4},
                                                               {out->seq_len, d_o32 + nr, nr * 4}, {out->qual_len, d_o32 + 2 * nr, nr * 4}};
    for (auto& c : down)
        if (c.bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(c.h, c.d, c.bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return timer.finish();
}

extern "C" int32_t hsq_trim_fastq(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, int32_t first, int32_t last, int32_t three_prime,
                                  uint8_t* out_text, int64_t out_cap, int64_t* out_bytes, int64_t* n_records) {
    if (!ctx || !out_bytes || n_bytes < 0 || out_cap < 0 || (n_bytes > 0 && !text)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_trim_fastq: null argument");
    if (first < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_trim_fastq: First base to keep must be greater than zero");  // FastqReadTrimmer.java:264-266
    if (last <= first) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_trim_fastq: Last base to keep must be greater than the first base to keep");  // :268-270
    *out_bytes = 0;
    if (n_records) *n_records = 0;
    if (n_bytes == 0) return HSQ_OK;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = reset_status(ctx);
    if (rc) return rc;
    uint8_t* d_text;
    int64_t* d_lines;
    int64_t n_lines;
    if ((rc = build_line_index(ctx, text, n_bytes, &d_text, &d_lines, &n_lines))) return rc;
    int st[1];
    if (n_lines % 4 != 0) {
        if ((rc = read_status(ctx, st, 1))) return rc;
        if (st[0] == HSQ_ERR_UNSUPPORTED) return fail(ctx, st[0], "hsq_trim_fastq: a carriage return that is not followed by a line feed");
        return fail(ctx, HSQ_ERR_MALFORMED, "hsq_trim_fastq: File is too short: the last record has fewer than four lines");
    }
    const int64_t n_rec = n_lines / 4;
    const size_t nr = (size_t)n_rec;
    if (n_records) *n_records = n_rec;
    const int64_t n_tiles = (n_rec + 1 + SCAN_TILE - 1) / SCAN_TILE;
    DevBuf* ib = ctx->ing;
    HSQ_CUDA(ctx, ib[4].ensure(nr * 8 * 4 + (nr + 1) * 16 + 64));  // four offset columns, output lengths, output offsets
    HSQ_CUDA(ctx, ib[5].ensure(nr * 4 * 5 + 64));                  // four length columns, kept bases
    HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)n_tiles * 8 + 16));
    HSQ_CUDA(ctx, ctx->out_u8.ensure((size_t)out_cap + 16));
    int64_t* o64 = ib[4].as<int64_t>();
    int32_t* o32 = ib[5].as<int32_t>();
    int64_t *h_off = o64, *s_off = o64 + nr, *q_off = o64 + 2 * nr, *p_off = o64 + 3 * nr, *out_len = o64 + 4 * nr, *out_off = out_len + nr + 1;
    int32_t *h_len = o32, *s_len = o32 + nr, *q_len = o32 + 2 * nr, *p_len = o32 + 3 * nr, *keep = o32 + 4 * nr;
    fastq_record_kernel<<<grid_for(ctx, n_rec, 128, 16), 128, 0, ctx->stream>>>(d_text, d_lines, n_rec, h_off, h_len, s_off, s_len, q_off, q_len, d_status(ctx),
                                                                             ctx->cells.as<unsigned long long>(), p_off, p_len);
    trimfq_len_kernel<<<grid_for(ctx, n_rec, 256, 16), 256, 0, ctx->stream>>>(h_len, s_len, p_len, n_rec, first, last, three_prime, out_len, keep, d_status(ctx));
    ctx->launches += 2;
    if ((rc = scan_i64(ctx, out_len, n_rec, ctx->sort_c.as<int64_t>(), out_off))) return rc;
    trimfq_write_kernel<<<grid_for(ctx, n_rec * 32, 256, 16), 256, 0, ctx->stream>>>(d_text, h_off, h_len, s_off, q_off, p_off, p_len, keep, out_off, n_rec, first,
                                                                                     out_cap, ctx->out_u8.as<uint8_t>());
    ctx->launches++;
    HSQ_CUDA(ctx, cudaGetLastError());
    HSQ_CUDA(ctx, cudaMemcpyAsync(out_bytes, out_off + nr, 8, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->h_cells, ctx->cells.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    if ((rc = read_status(ctx, st, 1))) return rc;
    if (st[0] == HSQ_ERR_UNSUPPORTED) return fail(ctx, st[0], "hsq_trim_fastq: a carriage return that is not followed by a line feed");
    if (st[0] == HSQ_ERR_MALFORMED) return fail(ctx, st[0], "hsq_trim_fastq: record " + std::to_string(ctx->h_cells[0]) + " is not a FASTQ record");
    if (st[0]) return fail(ctx, st[0], "hsq_trim_fastq: Unable to trim: a read is not longer than the first base to keep");
    if (*out_bytes > out_cap) return fail(ctx, HSQ_ERR_OVERFLOW, "hsq_trim_fastq: out_cap too small; *out_bytes holds the size");
    if (*out_bytes > 0) {
        if (!out_text) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_trim_fastq: null out_text");
        HSQ_CUDA(ctx, cudaMemcpyAsync(out_text, ctx->out_u8.p, (size_t)*out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    }
    return timer.finish();
}

namespace {

// text -> device-resident column planes (stride n_rec): d_off [HSQ_SAM_N_FIELDS][n_rec] in ctx->sort_b, d_val [HSQ_SAM_N_VALUES][n_rec] in
// ctx->out_i32; the text stays in ctx->r1.ascii
int sam_index_device(hsq_ctx* ctx, const char* who, const uint8_t* text, int64_t n_bytes, uint8_t** d_text_out, int64_t** d_off_out, int32_t** d_val_out,
                     int64_t* n_rec_out, int64_t* n_header_out) {
    int rc = reset_status(ctx);
    if (rc) return rc;
    uint8_t* d_text;
    int64_t* d_lines;
    int64_t n_lines;
    if ((rc = build_line_index(ctx, text, n_bytes, &d_text, &d_lines, &n_lines))) return rc;
    unsigned long long* d_first = ctx->cells.as<unsigned long long>() + 1;
    sam_first_line_kernel<<<grid_for(ctx, n_lines, 256, 16), 256, 0, ctx->stream>>>(d_text, d_lines, n_lines, d_first);
    ctx->launches++;
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->h_cells, ctx->cells.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int64_t n_header = ctx->h_cells[1] == ~0ull ? n_lines : (int64_t)ctx->h_cells[1];
    const int64_t n_rec = n_lines - n_header;
    if (n_rec >= (1ll << 31)) return fail(ctx, HSQ_ERR_UNSUPPORTED, std::string(who) + ": more than 2^31 alignment lines in one call");
    const size_t nr = (size_t)n_rec;
    HSQ_CUDA(ctx, ctx->sort_b.ensure(nr * 8 * HSQ_SAM_N_FIELDS + 64));
    HSQ_CUDA(ctx, ctx->out_i32.ensure(nr * 4 * HSQ_SAM_N_VALUES + 64));
    int64_t* d_off = ctx->sort_b.as<int64_t>();
    int32_t* d_val = ctx->out_i32.as<int32_t>();
    if (n_rec > 0) {
        sam_line_kernel<<<grid_for(ctx, n_rec * 32, 256, 16), 256, 0, ctx->stream>>>(d_text, d_lines, n_header, n_rec, d_off, d_val, d_status(ctx),
                                                                                     ctx->cells.as<unsigned long long>());
        ctx->launches++;
    }
    HSQ_CUDA(ctx, cudaGetLastError());
    HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->h_cells, ctx->cells.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
    int st[1];
    if ((rc = read_status(ctx, st, 1))) return rc;
    if (st[0] == HSQ_ERR_UNSUPPORTED) return fail(ctx, st[0], std::string(who) + ": a carriage return that is not followed by a line feed");
    if (st[0]) return fail(ctx, st[0], std::string(who) + ": alignment line " + std::to_string(ctx->h_cells[0]) +
                                           " is empty, a header line, has fewer than 11 columns, a non-numeric FLAG/POS/MAPQ/PNEXT/TLEN or a bad CIGAR");
    *d_text_out = d_text;
    *d_off_out = d_off;
    *d_val_out = d_val;
    *n_rec_out = n_rec;
    *n_header_out = n_header;
    return HSQ_OK;
}

}  // namespace

extern "C" int32_t hsq_index_sam(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, hsq_sam_index* out) {
    if (!ctx || !out || n_bytes < 0 || (n_bytes > 0 && !text)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_index_sam: null argument");
    out->n = 0;
    out->n_header_lines = 0;
    if (n_bytes == 0) return HSQ_OK;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    uint8_t* d_text;
    int64_t* d_off;
    int32_t* d_val;
    int64_t n_rec, n_header;
    int rc = sam_index_device(ctx, "hsq_index_sam", text, n_bytes, &d_text, &d_off, &d_val, &n_rec, &n_header);
    if (rc) return rc;
    out->n = n_rec;
    out->n_header_lines = n_header;
    if (n_rec > out->cap) return fail(ctx, HSQ_ERR_OVERFLOW, "hsq_index_sam: more alignment lines than out->cap");
    if (n_rec > 0 && (!out->off || !out->val)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_index_sam: null output plane");
    const size_t nr = (size_t)n_rec;
    for (int f = 0; f < HSQ_SAM_N_FIELDS && nr; f++)
        HSQ_CUDA(ctx, cudaMemcpyAsync(out->off + (int64_t)f * out->cap, d_off + (size_t)f * nr, nr * 8, cudaMemcpyDeviceToHost, ctx->stream));
    for (int v = 0; v < HSQ_SAM_N_VALUES && nr; v++)
        HSQ_CUDA(ctx, cudaMemcpyAsync(out->val + (int64_t)v * out->cap, d_val + (size_t)v * nr, nr * 4, cudaMemcpyDeviceToHost, ctx->stream));
    return timer.finish();
}

extern "C" int32_t hsq_sam_mapped_records(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, const hsq_seqs* contig_names, hsq_sam_records_out* out) {
    if (!ctx || !out || !contig_names || n_bytes < 0 || (n_bytes > 0 && !text)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_sam_mapped_records: null argument");
    out->n = 0;
    out->n_alt = 0;
    if (contig_names->n < 0 || contig_names->n > 100000 || !contig_names->offsets || contig_names->offsets[0] != 0)
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_sam_mapped_records: bad contig_names");
    if (out->cap >= 0 && out->alt_offsets) out->alt_offsets[0] = 0;
    if (n_bytes == 0) return HSQ_OK;
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    uint8_t* d_text;
    int64_t* d_off;
    int32_t* d_val;
    int64_t n_rec, n_header;
    int rc = sam_index_device(ctx, "hsq_sam_mapped_records", text, n_bytes, &d_text, &d_off, &d_val, &n_rec, &n_header);
    if (rc) return rc;
    if (n_rec == 0) return timer.finish();
    DevBuf* ib = ctx->ing;
    const int64_t dict_bytes = contig_names->offsets[contig_names->n];
    HSQ_CUDA(ctx, ib[0].ensure((size_t)dict_bytes + 16));
    HSQ_CUDA(ctx, ib[1].ensure((size_t)(contig_names->n + 1) * 8));
    if (dict_bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(ib[0].p, contig_names->data, (size_t)dict_bytes, cudaMemcpyHostToDevice, ctx->stream));
    HSQ_CUDA(ctx, cudaMemcpyAsync(ib[1].p, contig_names->offsets, (size_t)(contig_names->n + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
    NameDict dict{ib[0].as<uint8_t>(), ib[1].as<int64_t>(), (int)contig_names->n};
    const size_t nr = (size_t)n_rec;
    const int64_t n_tiles = (n_rec + 1 + SCAN_TILE - 1) / SCAN_TILE;
    HSQ_CUDA(ctx, ib[2].ensure((nr + 1) * 8 * 4 + 64));
    HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)n_tiles * 8 + 16));
    int64_t* d_keep = ib[2].as<int64_t>();
    int64_t* d_nalt = d_keep + nr + 1;
    int64_t* d_keep_pos = d_nalt + nr + 1;
    int64_t* d_alt_pos = d_keep_pos + nr + 1;
    const int g = grid_for(ctx, n_rec + 1, 128, 16);
    samrec_count_kernel<<<g, 128, 0, ctx->stream>>>(d_text, d_off, d_val, n_rec, dict, d_keep, d_nalt, d_status(ctx));
    ctx->launches++;
    if ((rc = scan_i64(ctx, d_keep, n_rec, ctx->sort_c.as<int64_t>(), d_keep_pos))) return rc;
    if ((rc = scan_i64(ctx, d_nalt, n_rec, ctx->sort_c.as<int64_t>(), d_alt_pos))) return rc;
    int64_t totals[2] = {0, 0};
    HSQ_CUDA(ctx, cudaMemcpyAsync(&totals[0], d_keep_pos + nr, 8, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaMemcpyAsync(&totals[1], d_alt_pos + nr, 8, cudaMemcpyDeviceToHost, ctx->stream));
    int st[1];
    if ((rc = read_status(ctx, st, 1))) return rc;
    if (st[0] == HSQ_ERR_UNSUPPORTED) return fail(ctx, st[0], "hsq_sam_mapped_records: a record without the paired flag (htsjdk: Inappropriate call if not paired read)");
    if (st[0]) return fail(ctx, st[0], "hsq_sam_mapped_records: malformed CIGAR in an XA entry");
    out->n = totals[0];
    out->n_alt = totals[1];
    if (out->n > out->cap || out->n_alt > out->alt_cap) return fail(ctx, HSQ_ERR_OVERFLOW, "hsq_sam_mapped_records: more records or alternative hits than cap / alt_cap");
    const size_t nk = (size_t)out->n, na = (size_t)out->n_alt;
    if (nk && (!out->sam_line || !out->contig || !out->negative_strand || !out->second_of_pair || !out->start || !out->end || !out->read_index || !out->alt_offsets))
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_sam_mapped_records: null output array");
    if (na && (!out->alt_contig || !out->alt_negative || !out->alt_start || !out->alt_stop))
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_sam_mapped_records: null alternative-hit array");
    // outputs: 5 int32 columns + 2 byte columns + offsets; 3 int32 + 1 byte per hit
    HSQ_CUDA(ctx, ib[3].ensure(nk * 4 * 5 + (nk + 1) * 8 + nk * 2 + na * 13 + 256));
    uint8_t* ob = ib[3].as<uint8_t>();
    int64_t* o_altoff = reinterpret_cast<int64_t*>(ob);
    int32_t* o_line = reinterpret_cast<int32_t*>(o_altoff + nk + 1);
    int32_t *o_contig = o_line + nk, *o_start = o_contig + nk, *o_end = o_start + nk, *o_ri = o_end + nk;
    int32_t *o_ac = o_ri + nk, *o_as = o_ac + na, *o_ae = o_as + na;
    uint8_t *o_neg = reinterpret_cast<uint8_t*>(o_ae + na), *o_sec = o_neg + nk, *o_an = o_sec + nk;
    samrec_fill_kernel<<<g, 128, 0, ctx->stream>>>(d_text, d_off, d_val, n_rec, dict, d_keep_pos, d_alt_pos, (int64_t)nk, (int64_t)na, o_line, o_contig, o_neg,
                                                   o_sec, o_start, o_end, o_ri, o_altoff, AltSink{o_ac, o_an, o_as, o_ae}, d_status(ctx));
    ctx->launches++;
    HSQ_CUDA(ctx, cudaGetLastError());
    if ((rc = read_status(ctx, st, 1))) return rc;
    if (st[0]) return fail(ctx, st[0], "hsq_sam_mapped_records: Expecting all read names to be numbers (the merged BAM names its records by FASTQ ordinal)");
    struct { void* h; const void* d; size_t bytes; } down[] = {
        {out->sam_line, o_line, nk * 4}, {out->contig, o_contig, nk * 4}, {out->start, o_start, nk * 4}, {out->end, o_end, nk * 4},
        {out->read_index, o_ri, nk * 4}, {out->negative_strand, o_neg, nk}, {out->second_of_pair, o_sec, nk}, {out->alt_offsets, o_altoff, (nk + 1) * 8},
        {out->alt_contig, o_ac, na * 4}, {out->alt_start, o_as, na * 4}, {out->alt_stop, o_ae, na * 4}, {out->alt_negative, o_an, na}};
    for (auto& c : down)
        if (c.bytes && c.h) HSQ_CUDA(ctx, cudaMemcpyAsync(c.h, c.d, c.bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return timer.finish();
}

extern "C" int32_t hsq_gather_rows(hsq_ctx* ctx, const uint8_t* text, int64_t n_bytes, const int64_t* off, const int32_t* len, int64_t n,
                                   uint8_t* out_data, int64_t* out_offsets, int64_t out_cap) {
    if (!ctx || n < 0 || n_bytes < 0 || !out_offsets || out_cap < 0) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_gather_rows: bad argument");
    if (n > 0 && (!off || !len)) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_gather_rows: null row array");
    if (n_bytes > 0 && !text) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_gather_rows: null text");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc = reset_status(ctx);
    if (rc) return rc;
    const size_t nn = (size_t)n;
    const int64_t n_tiles = (n + 1 + SCAN_TILE - 1) / SCAN_TILE;
    HSQ_CUDA(ctx, ctx->r1.ascii.ensure((size_t)n_bytes + 32));
    HSQ_CUDA(ctx, ctx->sort_a.ensure((nn + 1) * 8));  // row offsets in
    HSQ_CUDA(ctx, ctx->sort_b.ensure((nn + 1) * 8));  // lengths (int64)
    HSQ_CUDA(ctx, ctx->sort_d.ensure((nn + 1) * 8));  // offsets out
    HSQ_CUDA(ctx, ctx->sort_c.ensure((size_t)n_tiles * 8 + 16));
    HSQ_CUDA(ctx, ctx->out_i32.ensure(nn * 4 + 16));
    HSQ_CUDA(ctx, ctx->out_u8.ensure((size_t)out_cap + 16));
    if (n_bytes) HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->r1.ascii.p, text, (size_t)n_bytes, cudaMemcpyHostToDevice, ctx->stream));
    if (n) {
        HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->sort_a.p, off, nn * 8, cudaMemcpyHostToDevice, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(ctx->out_i32.p, len, nn * 4, cudaMemcpyHostToDevice, ctx->stream));
        rows_len_kernel<<<grid_for(ctx, n, 256, 16), 256, 0, ctx->stream>>>(ctx->sort_a.as<int64_t>(), ctx->out_i32.as<int32_t>(), n, n_bytes,
                                                                             ctx->sort_b.as<int64_t>(), d_status(ctx));
        ctx->launches++;
    }
    if ((rc = scan_i64(ctx, ctx->sort_b.as<int64_t>(), n, ctx->sort_c.as<int64_t>(), ctx->sort_d.as<int64_t>()))) return rc;
    if (n) {
        rows_copy_kernel<<<grid_for(ctx, n * 32, 256, 16), 256, 0, ctx->stream>>>(ctx->r1.ascii.as<uint8_t>(), ctx->sort_a.as<int64_t>(),
                                                                                  ctx->sort_d.as<int64_t>(), n, out_cap, ctx->out_u8.as<uint8_t>());
        ctx->launches++;
    }
    HSQ_CUDA(ctx, cudaGetLastError());
    HSQ_CUDA(ctx, cudaMemcpyAsync(out_offsets, ctx->sort_d.p, (nn + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    int st[1];
    if ((rc = read_status(ctx, st, 1))) return rc;
    if (st[0]) return fail(ctx, st[0], "hsq_gather_rows: a row reaches outside the text");
    const int64_t total = out_offsets[n];
    if (total > out_cap) return fail(ctx, HSQ_ERR_OVERFLOW, "hsq_gather_rows: out_cap too small; out_offsets holds the sizes");
    if (total > 0) {
        if (!out_data) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_gather_rows: null out_data");
        HSQ_CUDA(ctx, cudaMemcpyAsync(out_data, ctx->out_u8.p, (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
    }
    return timer.finish();
}

extern "C" int32_t hsq_join_names(hsq_ctx* ctx, const hsq_name_join* in, hsq_name_join_out* out) {
    if (!ctx || !in || !out) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_join_names: null argument");
    const int64_t ns = in->sam_names.n, n1 = in->fq1_headers.n, n2 = in->fq2_headers.n;
    const int64_t nf = std::min(n1, n2);  // the loop ends when either FASTQ iterator runs out (:214-215)
    out->n_emitted = 0;
    out->total_matching_pairs = 0;
    const bool by_index = in->by_index != 0;
    const bool check = !in->trimming_skipped && in->check_trim;
    if (ns > 0 && (!in->sam_flags || !in->sam_seq_len || !out->sam_record || !out->fastq_record || !out->name_len || !out->status))
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_join_names: null array");
    if (check && ((n1 > 0 && !in->fq1_seq_len) || (n2 > 0 && !in->fq2_seq_len)))
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_join_names: the trim check needs the FASTQ read lengths");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc;
    DevBuf* ib = ctx->ing;
    // inputs
    if ((rc = upload_strings(ctx, &in->sam_names, &ib[0], &ib[1], "sam_names"))) return rc;
    if ((rc = upload_strings(ctx, &in->fq1_headers, &ib[2], &ib[3], "fq1_headers"))) return rc;
    if ((rc = upload_strings(ctx, &in->fq2_headers, &ib[4], &ib[5], "fq2_headers"))) return rc;
    const size_t nss = (size_t)ns, n1s = (size_t)n1, n2s = (size_t)n2;
    HSQ_CUDA(ctx, ib[6].ensure((nss * 3 + n1s + n2s) * 4 + 64));
    int32_t* d_flags = ib[6].as<int32_t>();
    int32_t* d_seq_len = d_flags + nss;
    int32_t* d_qual_len = d_seq_len + nss;
    int32_t* d_fq1_len = d_qual_len + nss;
    int32_t* d_fq2_len = d_fq1_len + n1s;
    if (ns) {
        HSQ_CUDA(ctx, cudaMemcpyAsync(d_flags, in->sam_flags, nss * 4, cudaMemcpyHostToDevice, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(d_seq_len, in->sam_seq_len, nss * 4, cudaMemcpyHostToDevice, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(d_qual_len, in->sam_qual_len ? in->sam_qual_len : in->sam_seq_len, nss * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    if (check && n1) HSQ_CUDA(ctx, cudaMemcpyAsync(d_fq1_len, in->fq1_seq_len, n1s * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (check && n2) HSQ_CUDA(ctx, cudaMemcpyAsync(d_fq2_len, in->fq2_seq_len, n2s * 4, cudaMemcpyHostToDevice, ctx->stream));
    // counters
    HSQ_CUDA(ctx, ib[7].ensure(JC_WORDS * 4 + 64 * 4 * 3));
    int* d_ctr = ib[7].as<int>();
    int* d_live = d_ctr + JC_WORDS;  // up to 64 chunks per key set, three key sets
    HSQ_CUDA(ctx, cudaMemsetAsync(d_ctr, 0, JC_WORDS * 4 + 64 * 4 * 3, ctx->stream));
    HSQ_CUDA(ctx, cudaMemsetAsync(d_ctr + JC_MISMATCH, 0x7f, 4, ctx->stream));  // 0x7f7f7f7f: larger than any position
    HSQ_CUDA(ctx, cudaMemsetAsync(d_ctr + JC_TRIM, 0xff, 4, ctx->stream));
    const int64_t* off_s = ib[1].as<int64_t>();
    const int64_t* off_1 = ib[3].as<int64_t>();
    const int64_t* off_2 = ib[5].as<int64_t>();
    if (ns) max_len_kernel<<<grid_for(ctx, ns, 256, 8), 256, 0, ctx->stream>>>(off_s, ns, d_ctr + JC_MAX_SAM);
    if (n1) max_len_kernel<<<grid_for(ctx, n1, 256, 8), 256, 0, ctx->stream>>>(off_1, n1, d_ctr + JC_MAX_FQ1);
    if (n2) max_len_kernel<<<grid_for(ctx, n2, 256, 8), 256, 0, ctx->stream>>>(off_2, n2, d_ctr + JC_MAX_FQ2);
    ctx->launches += 3;
    int h_ctr[JC_WORDS + 64 * 3];
    HSQ_CUDA(ctx, cudaMemcpyAsync(h_ctr, d_ctr, JC_WORDS * 4, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    auto digits = [](int64_t v) { int d = 1; while (v >= 10) { v /= 10; d++; } return d; };
    const int ch_s = by_index ? 0 : std::max(1, (h_ctr[JC_MAX_SAM] + 7) / 8);
    const int ch_1 = by_index ? 0 : (h_ctr[JC_MAX_FQ1] + 3 + digits(std::max<int64_t>(n1 - 1, 0)) + 7) / 8;
    const int ch_2 = by_index ? 0 : (h_ctr[JC_MAX_FQ2] + 3 + digits(std::max<int64_t>(n2 - 1, 0)) + 7) / 8;
    if (ch_s > 64 || ch_1 > 64 || ch_2 > 64) return fail(ctx, HSQ_ERR_UNSUPPORTED, "hsq_join_names: read names longer than 490 bytes");
    // working set
    const size_t nmax = std::max(nss, std::max(n1s, n2s));
    Arena ar;
    const size_t o_keys_s = ar.reserve(nss * 8 * (size_t)std::max(ch_s, 1)), o_keys_1 = ar.reserve(n1s * 8 * (size_t)ch_1), o_keys_2 = ar.reserve(n2s * 8 * (size_t)ch_2);
    const size_t o_kf = ar.reserve(nss * 8), o_kp = ar.reserve(nss * 8), o_ka = ar.reserve(nmax * 8), o_kb = ar.reserve(nmax * 8);
    const size_t o_ps_a = ar.reserve(nss * 4), o_ps_b = ar.reserve(nss * 4), o_p1_a = ar.reserve(n1s * 4), o_p1_b = ar.reserve(n1s * 4);
    const size_t o_p2_a = ar.reserve(n2s * 4), o_p2_b = ar.reserve(n2s * 4);
    const size_t o_uid_s = ar.reserve(nss * 4), o_uid_1 = ar.reserve(n1s * 4), o_uid_2 = ar.reserve(n2s * 4), o_val = ar.reserve(nss * 4);
    const size_t o_so1 = ar.reserve(n1s * 8), o_sl1 = ar.reserve(n1s * 4), o_so2 = ar.reserve(n2s * 8), o_sl2 = ar.reserve(n2s * 4);
    const size_t o_head = ar.reserve(nss * 4), o_headmax = ar.reserve(nss * 4), o_nf = ar.reserve(nss * 4), o_excl = ar.reserve(nss * 4);
    const size_t o_emit = ar.reserve(nss * 4), o_epos = ar.reserve(nss * 4), o_used = ar.reserve(nss * 4), o_nlen = ar.reserve(nss * 4), o_tgt = ar.reserve(nss * 4);
    const size_t o_out = ar.reserve(nss * 13 + 64);
    HSQ_CUDA(ctx, ib[8].ensure(ar.need + 256));
    ar.base = ib[8].as<uint8_t>();
    uint64_t *ka = ar.at<uint64_t>(o_ka), *kb = ar.at<uint64_t>(o_kb);
    uint32_t *ps = ar.at<uint32_t>(o_ps_a), *ps_o = ar.at<uint32_t>(o_ps_b);
    uint32_t *p1 = ar.at<uint32_t>(o_p1_a), *p1_o = ar.at<uint32_t>(o_p1_b), *p2 = ar.at<uint32_t>(o_p2_a), *p2_o = ar.at<uint32_t>(o_p2_b);
    int32_t *uid_s = ar.at<int32_t>(o_uid_s), *uid_1 = ar.at<int32_t>(o_uid_1), *uid_2 = ar.at<int32_t>(o_uid_2), *val = ar.at<int32_t>(o_val);
    // keys
    if (n1) fq_keys_kernel<<<grid_for(ctx, n1, 128, 16), 128, 0, ctx->stream>>>(ib[2].as<uint8_t>(), off_1, n1, ch_1, 1, by_index ? nullptr : ar.at<uint64_t>(o_keys_1), uid_1);
    if (n2) fq_keys_kernel<<<grid_for(ctx, n2, 128, 16), 128, 0, ctx->stream>>>(ib[4].as<uint8_t>(), off_2, n2, ch_2, 1, by_index ? nullptr : ar.at<uint64_t>(o_keys_2), uid_2);
    ctx->launches += 2;
    if (ns) {
        if (by_index)
            sam_intkeys_kernel<<<grid_for(ctx, ns, 128, 16), 128, 0, ctx->stream>>>(ib[0].as<uint8_t>(), off_s, d_flags, ns, ar.at<uint64_t>(o_keys_s), val, uid_s, d_ctr);
        else
            sam_keys_kernel<<<grid_for(ctx, ns, 128, 16), 128, 0, ctx->stream>>>(ib[0].as<uint8_t>(), off_s, d_flags, ns, ch_s, ar.at<uint64_t>(o_keys_s),
                                                                                 ar.at<uint64_t>(o_kf), ar.at<uint64_t>(o_kp), uid_s, d_ctr);
        ctx->launches++;
    }
    if (!by_index) {
        if (ns) chunk_live_kernel<<<grid_for(ctx, ns * ch_s, 256, 16), 256, 0, ctx->stream>>>(ar.at<uint64_t>(o_keys_s), ns, ch_s, d_live);
        if (n1) chunk_live_kernel<<<grid_for(ctx, n1 * ch_1, 256, 16), 256, 0, ctx->stream>>>(ar.at<uint64_t>(o_keys_1), n1, ch_1, d_live + 64);
        if (n2) chunk_live_kernel<<<grid_for(ctx, n2 * ch_2, 256, 16), 256, 0, ctx->stream>>>(ar.at<uint64_t>(o_keys_2), n2, ch_2, d_live + 128);
        ctx->launches += 3;
    }
    HSQ_CUDA(ctx, cudaGetLastError());
    HSQ_CUDA(ctx, cudaMemcpyAsync(h_ctr, d_ctr, sizeof(h_ctr), cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (h_ctr[JC_PARSE])
        return fail(ctx, HSQ_ERR_UNSUPPORTED, "hsq_join_names: by_index needs integer read names (Integer.parseInt would throw NumberFormatException)");
    const int64_t np = ns - h_ctr[JC_NONPRIMARY];
    const int* live = h_ctr + JC_WORDS;
    // sorts
    if (ns) iota_u32_kernel<<<grid_for(ctx, ns, 256, 16), 256, 0, ctx->stream>>>(ps, ns);
    ctx->launches++;
    if (by_index) {
        if (ns) {
            if ((rc = sort_pass(ctx, ar.at<uint64_t>(o_keys_s), kb, ps, ps_o, ns, 34))) return rc;
            std::swap(ps, ps_o);
        }
    } else {
        if (ns) {
            if ((rc = bit_sort(ctx, ar.at<uint64_t>(o_kf), ns, &ps, &ps_o, ka, kb))) return rc;
            if ((rc = lsd_chunk_sort(ctx, ar.at<uint64_t>(o_keys_s), ch_s, live, ns, &ps, &ps_o, ka, kb))) return rc;
            if (np < ns && (rc = bit_sort(ctx, ar.at<uint64_t>(o_kp), ns, &ps, &ps_o, ka, kb))) return rc;
        }
        if (n1) {
            iota_u32_kernel<<<grid_for(ctx, n1, 256, 16), 256, 0, ctx->stream>>>(p1, n1);
            ctx->launches++;
            if ((rc = lsd_chunk_sort(ctx, ar.at<uint64_t>(o_keys_1), ch_1, live + 64, n1, &p1, &p1_o, ka, kb))) return rc;
        }
        if (n2) {
            iota_u32_kernel<<<grid_for(ctx, n2, 256, 16), 256, 0, ctx->stream>>>(p2, n2);
            ctx->launches++;
            if ((rc = lsd_chunk_sort(ctx, ar.at<uint64_t>(o_keys_2), ch_2, live + 128, n2, &p2, &p2_o, ka, kb))) return rc;
        }
    }
    // sorted FASTQ views and their checks
    SortedIds f1{ib[2].as<uint8_t>(), ar.at<int64_t>(o_so1), ar.at<int32_t>(o_sl1)};
    SortedIds f2{ib[4].as<uint8_t>(), ar.at<int64_t>(o_so2), ar.at<int32_t>(o_sl2)};
    if (n1) fq_sorted_view_kernel<<<grid_for(ctx, n1, 256, 16), 256, 0, ctx->stream>>>(by_index ? nullptr : p1, off_1, uid_1, n1, ar.at<int64_t>(o_so1), ar.at<int32_t>(o_sl1));
    if (n2) fq_sorted_view_kernel<<<grid_for(ctx, n2, 256, 16), 256, 0, ctx->stream>>>(by_index ? nullptr : p2, off_2, uid_2, n2, ar.at<int64_t>(o_so2), ar.at<int32_t>(o_sl2));
    if (n1) fq_check_kernel<<<grid_for(ctx, n1, 256, 16), 256, 0, ctx->stream>>>(f1, f2, n1, nf, by_index ? 1 : 0, d_ctr);
    ctx->launches += 3;
    // the SAM side
    int32_t* d_sam_record = ar.at<int32_t>(o_out);
    int32_t* d_fastq_record = d_sam_record + nss;
    int32_t* d_name_len = d_fastq_record + nss;
    uint8_t* d_rec_status = reinterpret_cast<uint8_t*>(d_name_len + nss);
    if (np > 0) {
        const int g = grid_for(ctx, np, 128, 16);
        sam_heads_kernel<<<g, 128, 0, ctx->stream>>>(ib[0].as<uint8_t>(), off_s, uid_s, d_flags, ps, np, by_index ? val : nullptr, ar.at<int32_t>(o_head),
                                                     ar.at<int32_t>(o_nf), d_ctr);
        ctx->launches++;
        size_t tb1 = 0, tb2 = 0;
        HSQ_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tb1, ar.at<int32_t>(o_nf), ar.at<int32_t>(o_excl), (int)np, ctx->stream));
        HSQ_CUDA(ctx, cub::DeviceScan::InclusiveScan(nullptr, tb2, ar.at<int32_t>(o_head), ar.at<int32_t>(o_headmax), cub::Max(), (int)np, ctx->stream));
        HSQ_CUDA(ctx, ctx->sort_tmp.ensure(std::max(tb1, tb2) + 16));
        HSQ_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->sort_tmp.p, tb1, ar.at<int32_t>(o_nf), ar.at<int32_t>(o_excl), (int)np, ctx->stream));
        HSQ_CUDA(ctx, cub::DeviceScan::InclusiveScan(ctx->sort_tmp.p, tb2, ar.at<int32_t>(o_head), ar.at<int32_t>(o_headmax), cub::Max(), (int)np, ctx->stream));
        TrimParams tp{check ? 1 : 0, in->three_prime, in->r1_first, in->r1_stop, in->r2_first, in->r2_stop};
        join_match_kernel<<<g, 128, 0, ctx->stream>>>(ib[0].as<uint8_t>(), off_s, uid_s, d_flags, d_seq_len, d_qual_len, ps, np, by_index ? val : nullptr,
                                                      ar.at<int32_t>(o_headmax), ar.at<int32_t>(o_excl), f1, nf, by_index ? nullptr : p1,
                                                      by_index ? nullptr : p2, d_fq1_len, d_fq2_len, tp, ar.at<int32_t>(o_emit), ar.at<int32_t>(o_used),
                                                      ar.at<int32_t>(o_nlen), ar.at<int32_t>(o_tgt), d_ctr);
        ctx->launches++;
        HSQ_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->sort_tmp.p, tb1, ar.at<int32_t>(o_emit), ar.at<int32_t>(o_epos), (int)np, ctx->stream));
        ctx->launches += 3;  // library scans (approximate)
    }
    if (ns) {
        join_scatter_kernel<<<grid_for(ctx, ns, 256, 16), 256, 0, ctx->stream>>>(ps, ns, np, ar.at<int32_t>(o_emit), ar.at<int32_t>(o_epos), ar.at<int32_t>(o_used),
                                                                                 ar.at<int32_t>(o_nlen), d_sam_record, d_fastq_record, d_name_len, d_rec_status, d_ctr);
        ctx->launches++;
    }
    HSQ_CUDA(ctx, cudaGetLastError());
    HSQ_CUDA(ctx, cudaMemcpyAsync(h_ctr, d_ctr, JC_WORDS * 4, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (h_ctr[JC_ORDER])
        return fail(ctx, HSQ_ERR_UNSUPPORTED, "hsq_join_names: unique ids do not ascend along a sorted side; the reference's result depends on its merge order here");
    // which exception the sequential loop meets first
    const bool have_mismatch = np > 0 && nf > 0 && h_ctr[JC_MISMATCH] <= h_ctr[JC_KMAX];
    const unsigned trim_key = (unsigned)h_ctr[JC_TRIM];
    if (trim_key != 0xffffffffu) {
        int32_t tgt = 0;
        HSQ_CUDA(ctx, cudaMemcpyAsync(&tgt, ar.at<int32_t>(o_tgt) + (trim_key >> 1), 4, cudaMemcpyDeviceToHost, ctx->stream));
        HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (!(have_mismatch && h_ctr[JC_MISMATCH] <= tgt)) {
            if (trim_key & 1u) return fail(ctx, HSQ_ERR_NOT_TRIMMED, "hsq_join_names: a SAM read is longer than its FASTQ read after trimming (sorted primary record " + std::to_string(trim_key >> 1) + ")");
            return fail(ctx, HSQ_ERR_OUT_OF_RANGE, "hsq_join_names: Unable to trim: first base to keep is not inside a FASTQ read");
        }
    }
    if (have_mismatch)
        return fail(ctx, HSQ_ERR_NAME_MISMATCH, "hsq_join_names: The read names within the Fastq files do not match at sorted position " + std::to_string(h_ctr[JC_MISMATCH]));
    if (h_ctr[JC_PAIRS] == 0)
        return fail(ctx, HSQ_ERR_NO_MATCHES, "hsq_join_names: The read names in the input Fastq files do not match the reads names in the provided bam/sam file");
    out->n_emitted = h_ctr[JC_EMITTED];
    out->total_matching_pairs = h_ctr[JC_PAIRS];
    const size_t ne = (size_t)out->n_emitted;
    if (ne) {
        HSQ_CUDA(ctx, cudaMemcpyAsync(out->sam_record, d_sam_record, ne * 4, cudaMemcpyDeviceToHost, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(out->fastq_record, d_fastq_record, ne * 4, cudaMemcpyDeviceToHost, ctx->stream));
        HSQ_CUDA(ctx, cudaMemcpyAsync(out->name_len, d_name_len, ne * 4, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (ns) HSQ_CUDA(ctx, cudaMemcpyAsync(out->status, d_rec_status, nss, cudaMemcpyDeviceToHost, ctx->stream));
    return timer.finish();
}

extern "C" int32_t hsq_coordinate_order(hsq_ctx* ctx, const hsq_coord_sort_in* in, int32_t* order) {
    if (!ctx || !in) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_coordinate_order: null argument");
    const int64_t n = in->n;
    if (n < 0 || n >= (1ll << 31) || in->names.n != n) return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_coordinate_order: bad n");
    if (n == 0) return HSQ_OK;
    if (!order || !in->ref_index || !in->pos || !in->flags || !in->mapq || !in->mate_ref_index || !in->mate_pos || !in->tlen)
        return fail(ctx, HSQ_ERR_INVALID_ARGUMENT, "hsq_coordinate_order: null array");
    HSQ_CUDA(ctx, cudaSetDevice(ctx->device));
    CallTimer timer(ctx);
    int rc;
    DevBuf* ib = ctx->ing;
    if ((rc = upload_strings(ctx, &in->names, &ib[0], &ib[1], "names"))) return rc;
    const size_t nn = (size_t)n;
    HSQ_CUDA(ctx, ib[6].ensure(nn * 4 * 7 + 64));
    int32_t* d_cols = ib[6].as<int32_t>();
    const int32_t* cols[7] = {in->ref_index, in->pos, in->flags, in->mapq, in->mate_ref_index, in->mate_pos, in->tlen};
    for (int c = 0; c < 7; c++) HSQ_CUDA(ctx, cudaMemcpyAsync(d_cols + (size_t)c * nn, cols[c], nn * 4, cudaMemcpyHostToDevice, ctx->stream));
    HSQ_CUDA(ctx, ib[7].ensure(JC_WORDS * 4 + 64 * 4 * 3));
    int* d_ctr = ib[7].as<int>();
    int* d_live = d_ctr + JC_WORDS;
    HSQ_CUDA(ctx, cudaMemsetAsync(d_ctr, 0, JC_WORDS * 4 + 64 * 4 * 3, ctx->stream));
    max_len_kernel<<<grid_for(ctx, n, 256, 8), 256, 0, ctx->stream>>>(ib[1].as<int64_t>(), n, d_ctr + JC_MAX_SAM);
    ctx->launches++;
    int h_ctr[JC_WORDS + 64];
    HSQ_CUDA(ctx, cudaMemcpyAsync(h_ctr, d_ctr, JC_WORDS * 4, cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int ch = std::max(1, (h_ctr[JC_MAX_SAM] + 7) / 8);
    if (ch > 64) return fail(ctx, HSQ_ERR_UNSUPPORTED, "hsq_coordinate_order: read names longer than 512 bytes");
    Arena ar;
    const size_t o_keys = ar.reserve(nn * 8 * (size_t)ch), o_k = ar.reserve(nn * 8 * 5), o_ka = ar.reserve(nn * 8), o_kb = ar.reserve(nn * 8);
    const size_t o_pa = ar.reserve(nn * 4), o_pb = ar.reserve(nn * 4), o_uid = ar.reserve(nn * 4);
    HSQ_CUDA(ctx, ib[8].ensure(ar.need + 256));
    ar.base = ib[8].as<uint8_t>();
    uint64_t *keys = ar.at<uint64_t>(o_keys), *k5 = ar.at<uint64_t>(o_k), *ka = ar.at<uint64_t>(o_ka), *kb = ar.at<uint64_t>(o_kb);
    uint32_t *pa = ar.at<uint32_t>(o_pa), *pb = ar.at<uint32_t>(o_pb);
    const int g = grid_for(ctx, n, 256, 16);
    coord_keys_kernel<<<g, 256, 0, ctx->stream>>>(d_cols, d_cols + nn, d_cols + 2 * nn, d_cols + 3 * nn, d_cols + 4 * nn, d_cols + 5 * nn, d_cols + 6 * nn, n, k5,
                                                  k5 + nn, k5 + 2 * nn, k5 + 3 * nn, k5 + 4 * nn);
    fq_keys_kernel<<<grid_for(ctx, n, 128, 16), 128, 0, ctx->stream>>>(ib[0].as<uint8_t>(), ib[1].as<int64_t>(), n, ch, 0, keys, ar.at<int32_t>(o_uid));
    chunk_live_kernel<<<grid_for(ctx, n * ch, 256, 16), 256, 0, ctx->stream>>>(keys, n, ch, d_live);
    iota_u32_kernel<<<g, 256, 0, ctx->stream>>>(pa, n);
    ctx->launches += 4;
    HSQ_CUDA(ctx, cudaGetLastError());
    HSQ_CUDA(ctx, cudaMemcpyAsync(h_ctr, d_ctr, sizeof(h_ctr), cudaMemcpyDeviceToHost, ctx->stream));
    HSQ_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    // least significant key first, every pass stable
    if ((rc = bit_sort(ctx, k5 + 4 * nn, n, &pa, &pb, ka, kb, 64))) return rc;  // mate start, insert size
    if ((rc = bit_sort(ctx, k5 + 3 * nn, n, &pa, &pb, ka, kb, 64))) return rc;  // mapping quality, mate reference
    if ((rc = bit_sort(ctx, k5 + 2 * nn, n, &pa, &pb, ka, kb, 32))) return rc;  // flags
    if ((rc = lsd_chunk_sort(ctx, keys, ch, h_ctr + JC_WORDS, n, &pa, &pb, ka, kb))) return rc;  // read name
    if ((rc = bit_sort(ctx, k5 + nn, n, &pa, &pb, ka, kb, 1))) return rc;       // strand
    if ((rc = bit_sort(ctx, k5, n, &pa, &pb, ka, kb, 64))) return rc;           // reference, start
    HSQ_CUDA(ctx, cudaMemcpyAsync(order, pa, nn * 4, cudaMemcpyDeviceToHost, ctx->stream));
    return timer.finish();
}
