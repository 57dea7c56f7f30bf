// hsq_oracle.cpp -- TEST INFRASTRUCTURE ONLY (parity oracle / timed CPU baseline); see hsq_oracle_core.hpp.
// C entry points with the same batch structs as include/hsq_b200.h so tests compare array to array.
#include "hsq_oracle_heatseq.hpp"
#include "hsq_oracle_assign.hpp"
#include "hsq_oracle_find.hpp"
#include "hsq_oracle_ingest.hpp"
#include "../include/hsq_b200.h"
#include <cstring>
#include <thread>
#include <atomic>
#include <unordered_map>

using namespace hsqo;

static thread_local std::string g_err;

static seq_t get_seq(const hsq_seqs& s, int64_t k) {
    return encode_string((const char*)s.data + s.offsets[k], s.offsets[k + 1] - s.offsets[k]);
}
static std::string get_str(const uint8_t* data, const int64_t* off, int64_t k) {
    return std::string((const char*)data + off[k], (size_t)(off[k + 1] - off[k]));
}
static void put_slot(const std::string& s, uint8_t* base, int32_t cap, int64_t k, int32_t* len, uint8_t* ovf, uint8_t bit) {
    if (len) len[k] = (int32_t)s.size();
    if ((int)s.size() > cap) { if (ovf) ovf[k] |= bit; }
    if (base) memcpy(base + (size_t)k * cap, s.data(), std::min((size_t)cap, s.size()));
}
static Probe get_probe(const hsq_probe_table& t, int64_t k) {
    Probe p;
    p.reverse_strand = t.reverse_strand[k] != 0;
    p.ext_start = t.ext_start[k]; p.ext_stop = t.ext_stop[k];
    p.lig_start = t.lig_start[k]; p.lig_stop = t.lig_stop[k];
    p.target_start = t.target_start[k]; p.target_stop = t.target_stop[k];
    p.ext = get_seq(t.ext, k); p.lig = get_seq(t.lig, k); p.target = get_seq(t.target, k);
    return p;
}

template <class F>
static int run_parallel(int64_t n, int nthreads, F f) {
    if (nthreads < 1) nthreads = 1;
    std::atomic<int64_t> next(0);
    std::atomic<int> status(0);
    std::vector<std::string> errs((size_t)nthreads);
    auto worker = [&](int t) {
        NwWorkspace ws;
        const int64_t chunk = 64;
        while (true) {
            int64_t b = next.fetch_add(chunk);
            if (b >= n) break;
            int64_t e = std::min(n, b + chunk);
            try {
                for (int64_t k = b; k < e; k++) f(k, ws, t);
            } catch (const std::exception& ex) {
                errs[(size_t)t] = ex.what();
                status = HSQ_ERR_OUT_OF_RANGE;
                if (dynamic_cast<const std::invalid_argument*>(&ex)) status = HSQ_ERR_BAD_CHARACTER;
                break;
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nthreads; t++) th.emplace_back(worker, t);
    worker(0);
    for (auto& x : th) x.join();
    for (auto& e : errs) if (!e.empty()) g_err = e;
    return status;
}

extern "C" {

const char* hsqo_last_error(void) { return g_err.c_str(); }

int32_t hsqo_align_batch(const hsq_seqs* refs, const hsq_seqs* qrys, const hsq_scorer* scorer, uint32_t flags,
                         hsq_align_out* o, int32_t nthreads) {
    Scorer sc;
    sc.match = scorer->match; sc.mismatch = scorer->mismatch; sc.gap_extend = scorer->gap_extend; sc.gap_start = scorer->gap_start;
    sc.penalize_start = (flags & HSQ_ALIGN_PENALIZE_START) != 0;
    sc.penalize_end = (flags & HSQ_ALIGN_PENALIZE_END) != 0;
    int64_t n = refs->n;
    std::vector<int64_t> cells((size_t)std::max(1, nthreads), 0);
    int st = run_parallel(n, nthreads, [&](int64_t k, NwWorkspace& ws, int t) {
        NwResult r = nw_align(get_seq(*refs, k), get_seq(*qrys, k), sc, &ws);
        cells[(size_t)t] += r.cells;
        const AlignmentPair& p = r.pair;
        if (o->overflow) o->overflow[k] = 0;
        if (o->score) o->score[k] = r.score;
        if (o->n_cols) o->n_cols[k] = p.size();
        put_slot(decode_seq(p.ref_aln), o->ref_aln, o->cols_cap, k, nullptr, o->overflow, 1);
        put_slot(decode_seq(p.qry_aln), o->qry_aln, o->cols_cap, k, nullptr, o->overflow, 1);
        put_slot(expanded_cigar(p), o->ops, o->cols_cap, k, nullptr, o->overflow, 1);
        if (o->first_q) o->first_q[k] = p.first_q;
        if (o->last_q) o->last_q[k] = p.last_q;
        if (o->first_r) o->first_r[k] = p.first_r;
        if (o->last_r) o->last_r[k] = p.last_r;
        if (o->matches) o->matches[k] = p.number_of_matches();
        if (o->mismatches) o->mismatches[k] = p.number_of_mismatches();
        if (o->non_terminal_gaps) o->non_terminal_gaps[k] = p.number_of_non_terminal_gaps();
        if (o->edit_distance) o->edit_distance[k] = p.edit_distance();
        put_slot(md_string(p), o->md, o->md_cap, k, o->md_len, o->overflow, 2);
        put_slot(reverse_md_string(p), o->rmd, o->md_cap, k, o->rmd_len, o->overflow, 4);
        if (o->find_counts) {  // FastqFindAndTrimTool.java:930-938 on getAlignmentWithoutEndingAndBeginningQueryInserts (AlignmentPair.java:95-99)
            const seq_t ra = seq_sub(p.ref_aln, p.first_q, p.last_q), qa = seq_sub(p.qry_aln, p.first_q, p.last_q);
            int fx = 0, ig = 0, ib = 0, dg = 0, db = 0;
            bool in_i = false, in_d = false;
            for (size_t c = 0; c < ra.size(); c++) {
                if (ra[c] == GAP) { ib++; if (!in_i) { in_i = true; ig++; } } else in_i = false;      // :190-216
                if (qa[c] == GAP) { db++; if (!in_d) { in_d = true; dg++; } } else in_d = false;      // :218-244
                if (ra[c] != GAP && qa[c] != GAP && !code_matches(ra[c], qa[c])) fx++;                  // :246-256
            }
            int32_t* fc = o->find_counts + 5 * k;
            fc[0] = fx; fc[1] = ig; fc[2] = ib; fc[3] = dg; fc[4] = db;
        }
    });
    o->cells = 0;
    for (auto c : cells) o->cells += c;
    return st;
}

int32_t hsqo_trim_batch(const hsq_seqs* reads, const uint8_t* quals, int32_t first, int32_t last, int32_t three_prime,
                        uint8_t* out_seq, uint8_t* out_qual, int64_t* out_offsets) {
    try {
        int64_t pos = 0;
        out_offsets[0] = 0;
        for (int64_t k = 0; k < reads->n; k++) {
            int64_t b = reads->offsets[k];
            int len = (int)(reads->offsets[k + 1] - b);
            int f, l;
            trim_range(len, first, last, three_prime != 0, &f, &l);
            int keep = l - f + 1;
            memcpy(out_seq + pos, reads->data + b + f, (size_t)keep);
            memcpy(out_qual + pos, quals + b + f, (size_t)keep);
            pos += keep;
            out_offsets[k + 1] = pos;
        }
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return HSQ_ERR_OUT_OF_RANGE;
    }
    return HSQ_OK;
}

// mergeSequences per item; out_offsets always filled; HSQ_ERR_OVERFLOW when the total exceeds out_cap
int32_t hsqo_merge_pairs(const hsq_merge_batch* mb, uint8_t* out_seq, uint8_t* out_qual, int64_t* out_offsets, int64_t out_cap) {
    try {
        int64_t pos = 0;
        out_offsets[0] = 0;
        bool overflow = false;
        for (int64_t k = 0; k < mb->n; k++) {
            std::string ms, mq;
            merge_sequences(mb->negative_strand[k] != 0, get_str(mb->up.data, mb->up.offsets, k), get_str(mb->down.data, mb->down.offsets, k),
                            get_str(mb->up_qual, mb->up.offsets, k), get_str(mb->down_qual, mb->down.offsets, k), mb->up_start[k],
                            mb->down_start[k], &ms, &mq);
            if (pos + (int64_t)ms.size() <= out_cap) {
                memcpy(out_seq + pos, ms.data(), ms.size());
                memcpy(out_qual + pos, mq.data(), mq.size());
            } else {
                overflow = true;
            }
            pos += (int64_t)ms.size();
            out_offsets[k + 1] = pos;
        }
        if (overflow) return HSQ_ERR_OVERFLOW;
    } catch (const std::invalid_argument& ex) {
        g_err = ex.what();
        return HSQ_ERR_BAD_CHARACTER;
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return HSQ_ERR_OUT_OF_RANGE;
    }
    return HSQ_OK;
}

// The SAM record fields of extended pairs: runs extendReadPair (ExtendReadsToPrimer.java:96-198) itself and reports what
// createRecord (:247-284) and SAMRecordUtil.setSAMRecordsAsPair (SAMRecordUtil.java:270-289) put on the two records.
int32_t hsqo_extend_records(const hsq_probe_table* pt, const hsq_scorer* scorer, const hsq_pair_batch* pb, const uint8_t* duplicate,
                            hsq_record_out* o, int64_t out_cap) {
    Scorer sc;
    sc.match = scorer->match; sc.mismatch = scorer->mismatch; sc.gap_extend = scorer->gap_extend; sc.gap_start = scorer->gap_start;
    try {
        std::vector<Probe> probes;
        for (int64_t k = 0; k < pt->n; k++) probes.push_back(get_probe(*pt, k));
        NwWorkspace ws;
        int64_t cells = 0, at = 0;
        bool overflow = false;
        o->offsets[0] = 0;
        for (int64_t k = 0; k < pb->n; k++) {
            const Probe& p = probes[(size_t)pb->probe_id[k]];
            seq_t s1 = get_seq(pb->r1, k), s2 = get_seq(pb->r2, k);
            PairExtension e = extend_read_pair(p, s1, get_str(pb->q1, pb->r1.offsets, k), s2, get_str(pb->q2, pb->r2.offsets, k), sc, &ws, &cells);
            const std::string* seqs[2] = {&e.seq1, &e.seq2};
            const std::string* quals[2] = {&e.qual1, &e.qual2};
            const bool neg[2] = {e.neg1, e.neg2};
            const int pos[2] = {e.one.pos, e.two.pos};
            for (int m = 0; m < 2; m++) {
                const int64_t i = 2 * k + m;
                o->flag[i] = 0;
                o->mate_pos[i] = 0;
                if (e.pair_ok()) {  // only pairs whose two reads extended are written (:364)
                    // setReadPairedFlag, setProperPairFlag, first/second of pair (SAMRecordUtil.java:271-276); read strand (:257);
                    // mate strand = the mate's read strand (:283); duplicate (:266)
                    o->flag[i] = (uint16_t)(0x1 | 0x2 | (m ? 0x80 : 0x40) | (neg[m] ? 0x10 : 0) | (neg[1 - m] ? 0x20 : 0) |
                                            ((duplicate && duplicate[k]) ? 0x400 : 0));
                    o->mate_pos[i] = pos[1 - m];
                    if (at + (int64_t)seqs[m]->size() <= out_cap) {
                        memcpy(o->seq + at, seqs[m]->data(), seqs[m]->size());
                        memcpy(o->qual + at, quals[m]->data(), quals[m]->size());
                    } else {
                        overflow = true;
                    }
                    at += (int64_t)seqs[m]->size();
                }
                o->offsets[i + 1] = at;
            }
        }
        if (overflow) return HSQ_ERR_OVERFLOW;
    } catch (const std::invalid_argument& ex) {
        g_err = ex.what();
        return HSQ_ERR_BAD_CHARACTER;
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return HSQ_ERR_OUT_OF_RANGE;
    }
    return HSQ_OK;
}

// ReadToProbeAssigner.assignProbesToReads + RangeMap (hsq_oracle_assign.hpp)
int32_t hsqo_assign_probes(const hsq_probe_ranges* pr, const hsq_mapped_records* rec, hsq_assign_out* o) {
    try {
        assign_probes(pr, rec, o);
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return HSQ_ERR_INVALID_ARGUMENT;
    }
    return HSQ_OK;
}

// trimReadBasedOnQuality + SimpleMapper.getBestCandidateReferences per read (hsq_oracle_find.hpp)
int32_t hsqo_find_candidates(const hsq_seqs* seqs, int32_t kmer, int32_t max_refs_per_kmer, const hsq_seqs* reads, const uint8_t* quals,
                             int32_t leading_quality, int32_t trailing_quality, int32_t candidate_limit, double min_hit_ratio,
                             hsq_find_out* out) {
    try {
        return find_candidates(seqs, kmer, max_refs_per_kmer, reads, quals, leading_quality, trailing_quality, candidate_limit,
                               min_hit_ratio, out);
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return HSQ_ERR_INVALID_ARGUMENT;
    }
}

// FilterByUid.reduceReadsByProbeAndUid over all probes (HSQ/process/FilterByUid.java:134-203) + findBestData (:281-297)
int32_t hsqo_dedup_batch(const hsq_pair_batch* pb, int32_t u, int32_t v, uint8_t* is_best, int32_t* group_size,
                         hsq_probe_stats* stats) {
    try {
        std::map<std::pair<int32_t, std::string>, std::vector<int64_t>> groups;
        for (int64_t k = 0; k < pb->n; k++) {
            std::string r1 = get_str(pb->r1.data, pb->r1.offsets, k), r2 = get_str(pb->r2.data, pb->r2.offsets, k);
            if ((int)r1.size() < u || (int)r2.size() < v) throw std::out_of_range("UID longer than read");
            std::string uid = r1.substr(0, (size_t)u) + r2.substr(0, (size_t)v);  // :136-138, SAMRecordUtil.java:73-86
            groups[{pb->probe_id[k], uid}].push_back(k);
        }
        if (stats) for (int64_t p = 0; p < stats->n_probes; p++) {
            stats->unique_pairs[p] = 0; stats->duplicate_pairs[p] = 0; stats->max_pairs_per_uid[p] = 0;
        }
        for (auto& g : groups) {
            int64_t best = -1;
            int best_score = INT32_MIN;
            int32_t best_id = 0;
            for (int64_t k : g.second) {
                int score = total_quality_score((const char*)pb->q1 + pb->r1.offsets[k], pb->r1.offsets[k + 1] - pb->r1.offsets[k],
                                                (const char*)pb->q2 + pb->r2.offsets[k], pb->r2.offsets[k + 1] - pb->r2.offsets[k]);
                int32_t id = pb->read_index[k];
                if (score > best_score || (score == best_score && id < best_id)) { best_score = score; best = k; best_id = id; }
            }
            for (int64_t k : g.second) {
                is_best[k] = (k == best) ? 1 : 0;
                if (group_size) group_size[k] = (int32_t)g.second.size();
            }
            if (stats) {
                int32_t p = g.first.first;
                stats->unique_pairs[p] += 1;
                stats->duplicate_pairs[p] += (int32_t)g.second.size() - 1;
                stats->max_pairs_per_uid[p] = std::max(stats->max_pairs_per_uid[p], (int32_t)g.second.size());
            }
        }
    } catch (const std::exception& ex) {
        g_err = ex.what();
        return HSQ_ERR_OUT_OF_RANGE;
    }
    return HSQ_OK;
}

int32_t hsqo_extend_batch(const hsq_probe_table* pt, const hsq_scorer* scorer, const hsq_pair_batch* pb, hsq_extend_out* o,
                          int32_t nthreads) {
    Scorer sc;
    sc.match = scorer->match; sc.mismatch = scorer->mismatch; sc.gap_extend = scorer->gap_extend; sc.gap_start = scorer->gap_start;
    std::vector<Probe> probes;
    try {
        for (int64_t k = 0; k < pt->n; k++) probes.push_back(get_probe(*pt, k));
    } catch (const std::exception& ex) { g_err = ex.what(); return HSQ_ERR_BAD_CHARACTER; }
    std::vector<int64_t> cells((size_t)std::max(1, nthreads), 0);
    int st = run_parallel(pb->n, nthreads, [&](int64_t k, NwWorkspace& ws, int t) {
        const Probe& p = probes[(size_t)pb->probe_id[k]];
        seq_t s1 = get_seq(pb->r1, k), s2 = get_seq(pb->r2, k);
        std::string q1(s1.size(), 'I'), q2(s2.size(), 'I');
        if (pb->q1) q1 = get_str(pb->q1, pb->r1.offsets, k);
        if (pb->q2) q2 = get_str(pb->q2, pb->r2.offsets, k);
        PairExtension e = extend_read_pair(p, s1, q1, s2, q2, sc, &ws, &cells[(size_t)t]);
        const ReadEndDetails* d[2] = {&e.one, &e.two};
        int tl[2] = {e.tlen1, e.tlen2};
        for (int m = 0; m < 2; m++) {
            int64_t i = 2 * k + m;
            o->ok[i] = d[m]->ok;
            o->primer_end[i] = d[m]->primer_end;
            o->score[i] = d[m]->score;
            o->pos[i] = d[m]->ok ? d[m]->pos : 0;
            o->read_start[i] = d[m]->ok ? d[m]->read_start : 0;
            o->read_stop[i] = d[m]->ok ? d[m]->read_stop : 0;
            o->nm[i] = d[m]->ok ? d[m]->nm : 0;
            if (o->tlen) o->tlen[i] = d[m]->ok ? tl[m] : 0;
            o->overflow[i] = 0;
            put_slot(d[m]->cigar, o->cigar, o->cigar_cap, i, o->cigar_len, o->overflow, 1);
            put_slot(d[m]->md, o->md, o->md_cap, i, o->md_len, o->overflow, 2);
        }
        o->pair_ok[k] = e.pair_ok();
    });
    o->cells = 0;
    for (auto c : cells) o->cells += c;
    return st;
}

int32_t hsqo_resolve_multi_probe(const hsq_probe_table* pt, const hsq_candidates* c, int32_t u, int32_t v, int32_t* best_probe,
                                 int32_t* cand_matches, int32_t* cand_edit) {
    try {
        NwWorkspace ws;
        int64_t cells = 0;
        for (int64_t k = 0; k < c->n; k++) {
            seq_t r1 = get_seq(c->r1, k), r2 = get_seq(c->r2, k);
            int lowest = INT32_MAX, most = 0;
            std::vector<int32_t> best;  // …Identification.java:704-751
            for (int64_t j = c->cand_offsets[k]; j < c->cand_offsets[k + 1]; j++) {
                int m, e;
                tie_break_metrics(get_probe(*pt, c->cand_probe[j]), r1, r2, u, v, &ws, &cells, &m, &e);
                if (cand_matches) cand_matches[j] = m;
                if (cand_edit) cand_edit[j] = e;
                if (m > most) { best.clear(); best.push_back(c->cand_probe[j]); lowest = e; most = m; }
                else if (m == most) {
                    if (e == lowest) best.push_back(c->cand_probe[j]);
                    else if (e < lowest) { best.clear(); best.push_back(c->cand_probe[j]); lowest = e; most = m; }
                }
            }
            best_probe[k] = (best.size() == 1) ? best[0] : -1;  // :757-765
        }
    } catch (const std::exception& ex) { g_err = ex.what(); return HSQ_ERR_OUT_OF_RANGE; }
    return HSQ_OK;
}

// single-call helpers for the KAT tests
// config 1 file to file: FastqReadTrimmer.trimReads on FASTQ text (hsq_oracle_ingest.hpp)
int32_t hsqo_trim_fastq(const uint8_t* text, int64_t n_bytes, int32_t first, int32_t last, int32_t three_prime, uint8_t* out_text, int64_t out_cap,
                        int64_t* out_bytes, int64_t* n_records) {
    try {
        const std::string t = trim_fastq_text(text, n_bytes, first, last, three_prime != 0);
        *out_bytes = (int64_t)t.size();
        if (n_records) { int64_t c = 0; for (char ch : t) c += ch == '\n'; *n_records = c / 4; }
        if ((int64_t)t.size() > out_cap) { g_err = "out_cap too small"; return HSQ_ERR_OVERFLOW; }
        memcpy(out_text, t.data(), t.size());
        return HSQ_OK;
    } catch (const FastqMalformed& e) { g_err = e.what(); return HSQ_ERR_MALFORMED;
    } catch (const std::invalid_argument& e) { g_err = e.what(); return HSQ_ERR_INVALID_ARGUMENT;
    } catch (const std::out_of_range& e) { g_err = e.what(); return HSQ_ERR_OUT_OF_RANGE; }
}

// f4: order of the final coordinate sort (hsq_oracle_ingest.hpp)
int32_t hsqo_coordinate_order(const hsq_coord_sort_in* in, int32_t* order) {
    std::vector<CoordRecord> r((size_t)in->n);
    for (int64_t k = 0; k < in->n; k++) {
        CoordRecord& c = r[(size_t)k];
        c.ref = in->ref_index[k]; c.pos = in->pos[k]; c.flags = in->flags[k]; c.mapq = in->mapq[k];
        c.mate_ref = in->mate_ref_index[k]; c.mate_pos = in->mate_pos[k]; c.tlen = in->tlen[k];
        c.name = get_str(in->names.data, in->names.offsets, k);
    }
    const std::vector<int32_t> o = coordinate_order(r);
    for (int64_t k = 0; k < in->n; k++) order[k] = o[(size_t)k];
    return HSQ_OK;
}

// f2: FastqReader record index of FASTQ text (hsq_oracle_ingest.hpp)
int32_t hsqo_index_fastq(const uint8_t* text, int64_t n_bytes, hsq_fastq_index* o) {
    try {
        FastqIndex ix = index_fastq(text, n_bytes);
        o->n = (int64_t)ix.header_off.size();
        if (o->n > o->cap) { g_err = "more records than cap"; return HSQ_ERR_OVERFLOW; }
        for (int64_t k = 0; k < o->n; k++) {
            o->header_off[k] = ix.header_off[(size_t)k]; o->header_len[k] = ix.header_len[(size_t)k];
            o->seq_off[k] = ix.seq_off[(size_t)k]; o->seq_len[k] = ix.seq_len[(size_t)k];
            o->qual_off[k] = ix.qual_off[(size_t)k]; o->qual_len[k] = ix.qual_len[(size_t)k];
        }
        return HSQ_OK;
    } catch (const FastqMalformed& e) { g_err = e.what(); return HSQ_ERR_MALFORMED; }
}

// f2: SAM text field index (hsq_oracle_ingest.hpp)
int32_t hsqo_index_sam(const uint8_t* text, int64_t n_bytes, hsq_sam_index* o) {
    try {
        SamIndex ix = index_sam(text, n_bytes);
        o->n = (int64_t)ix.flag.size();
        o->n_header_lines = ix.n_header_lines;
        if (o->n > o->cap) { g_err = "more alignment lines than cap"; return HSQ_ERR_OVERFLOW; }
        const std::vector<int64_t>* offs[HSQ_SAM_N_FIELDS] = {&ix.qname_off, &ix.rname_off, &ix.cigar_off, &ix.seq_off, &ix.qual_off, &ix.tags_off, &ix.xa_off};
        const std::vector<int32_t>* vals[HSQ_SAM_N_VALUES] = {&ix.qname_len, &ix.rname_len, &ix.cigar_len, &ix.seq_len, &ix.qual_len, &ix.tags_len, &ix.xa_len,
                                                              &ix.flag, &ix.pos, &ix.mapq, &ix.pnext, &ix.tlen, &ix.ref_len};
        for (int f = 0; f < HSQ_SAM_N_FIELDS; f++)
            for (int64_t k = 0; k < o->n; k++) o->off[(int64_t)f * o->cap + k] = (*offs[f])[(size_t)k];
        for (int v = 0; v < HSQ_SAM_N_VALUES; v++)
            for (int64_t k = 0; k < o->n; k++) o->val[(int64_t)v * o->cap + k] = (*vals[v])[(size_t)k];
        return HSQ_OK;
    } catch (const FastqMalformed& e) { g_err = e.what(); return HSQ_ERR_MALFORMED; }
}

// f2 -> f1: SAM text -> mapped records (hsq_oracle_ingest.hpp)
int32_t hsqo_sam_mapped_records(const uint8_t* text, int64_t n_bytes, const hsq_seqs* contig_names, hsq_sam_records_out* o) {
    try {
        std::vector<std::string> names;
        for (int64_t k = 0; k < contig_names->n; k++) names.push_back(get_str(contig_names->data, contig_names->offsets, k));
        SamRecords r = sam_mapped_records(text, n_bytes, names);
        o->n = (int64_t)r.contig.size();
        o->n_alt = (int64_t)r.alt_contig.size();
        if (o->n > o->cap || o->n_alt > o->alt_cap) { g_err = "more records or hits than cap"; return HSQ_ERR_OVERFLOW; }
        for (int64_t k = 0; k < o->n; k++) {
            const size_t i = (size_t)k;
            o->sam_line[k] = r.sam_line[i]; o->contig[k] = r.contig[i]; o->negative_strand[k] = r.negative_strand[i];
            o->second_of_pair[k] = r.second_of_pair[i]; o->start[k] = r.start[i]; o->end[k] = r.end[i]; o->read_index[k] = r.read_index[i];
        }
        for (int64_t k = 0; k <= o->n; k++) o->alt_offsets[k] = r.alt_offsets[(size_t)k];
        for (int64_t k = 0; k < o->n_alt; k++) {
            const size_t i = (size_t)k;
            o->alt_contig[k] = r.alt_contig[i]; o->alt_negative[k] = r.alt_negative[i]; o->alt_start[k] = r.alt_start[i]; o->alt_stop[k] = r.alt_stop[i];
        }
        return HSQ_OK;
    } catch (const SamUnpaired& e) { g_err = e.what(); return HSQ_ERR_UNSUPPORTED;
    } catch (const SamNameNotANumber& e) { g_err = e.what(); return HSQ_ERR_BAD_CHARACTER;
    } catch (const FastqMalformed& e) { g_err = e.what(); return HSQ_ERR_MALFORMED; }
}

// f2: the two sorts and the merge loop of FastqAndBamFileMerger (hsq_oracle_ingest.hpp)
int32_t hsqo_join_names(const hsq_name_join* in, hsq_name_join_out* o) {
    JoinInput ji;
    JoinParams jp;
    const int64_t ns = in->sam_names.n;
    for (int64_t k = 0; k < ns; k++) ji.sam_names.push_back(get_str(in->sam_names.data, in->sam_names.offsets, k));
    ji.sam_flags.assign(in->sam_flags, in->sam_flags + ns);
    ji.sam_seq_len.assign(in->sam_seq_len, in->sam_seq_len + ns);
    const int32_t* ql = in->sam_qual_len ? in->sam_qual_len : in->sam_seq_len;
    ji.sam_qual_len.assign(ql, ql + ns);
    for (int64_t k = 0; k < in->fq1_headers.n; k++) ji.fq1_headers.push_back(get_str(in->fq1_headers.data, in->fq1_headers.offsets, k));
    for (int64_t k = 0; k < in->fq2_headers.n; k++) ji.fq2_headers.push_back(get_str(in->fq2_headers.data, in->fq2_headers.offsets, k));
    if (in->fq1_seq_len) ji.fq1_seq_len.assign(in->fq1_seq_len, in->fq1_seq_len + in->fq1_headers.n);
    else ji.fq1_seq_len.assign((size_t)in->fq1_headers.n, 0);
    if (in->fq2_seq_len) ji.fq2_seq_len.assign(in->fq2_seq_len, in->fq2_seq_len + in->fq2_headers.n);
    else ji.fq2_seq_len.assign((size_t)in->fq2_headers.n, 0);
    jp.by_index = in->by_index != 0;
    jp.trimming_skipped = in->trimming_skipped != 0; jp.check_trim = in->check_trim != 0; jp.three_prime = in->three_prime != 0;
    jp.r1_first = in->r1_first; jp.r1_stop = in->r1_stop; jp.r2_first = in->r2_first; jp.r2_stop = in->r2_stop;
    o->n_emitted = 0;
    o->total_matching_pairs = 0;
    try {
        JoinOutput jo = join_names(ji, jp);
        o->n_emitted = (int64_t)jo.sam_record.size();
        o->total_matching_pairs = jo.total_matching_pairs;
        for (size_t e = 0; e < jo.sam_record.size(); e++) {
            o->sam_record[e] = jo.sam_record[e];
            o->fastq_record[e] = jo.fastq_record[e];
            o->name_len[e] = jo.name_len[e];
        }
        for (int64_t k = 0; k < ns; k++) o->status[k] = jo.status[(size_t)k];
        return HSQ_OK;
    } catch (const JoinNameMismatch& e) { g_err = e.what(); return HSQ_ERR_NAME_MISMATCH;
    } catch (const JoinNotTrimmed& e) { g_err = e.what(); return HSQ_ERR_NOT_TRIMMED;
    } catch (const JoinNoMatches& e) { g_err = e.what(); return HSQ_ERR_NO_MATCHES;
    } catch (const JoinBadNumber& e) { g_err = e.what(); return HSQ_ERR_UNSUPPORTED;
    } catch (const std::out_of_range& e) { g_err = e.what(); return HSQ_ERR_OUT_OF_RANGE; }
}

int32_t hsqo_primer_end(const char* primer, const char* read, double* weighted) {
    int64_t cells = 0;
    NwWorkspace ws;
    return primer_end_index_in_read(encode_string(primer), encode_string(read), &ws, &cells, weighted);
}
int32_t hsqo_index_of_first_match_in_reference(const char* ref, const char* qry) {
    Scorer sc;
    return index_of_first_match_in_reference(nw_align(encode_string(ref), encode_string(qry), sc).pair);
}
// SAMRecordUtil.getVariableLengthUid (HSQ/utils/SAMRecordUtil.java:223-261): length of the UID prefix, -1 for null.
// insertsSinceLastInsertionOrMismatch is never incremented in the reference (:238-250), so every deletion column counts.
int32_t hsqo_variable_length_uid(const char* read, const char* primer, const hsq_scorer* scorer, uint32_t flags) {
    Scorer sc;
    sc.match = scorer->match; sc.mismatch = scorer->mismatch; sc.gap_extend = scorer->gap_extend; sc.gap_start = scorer->gap_start;
    sc.penalize_start = (flags & 1u) != 0; sc.penalize_end = (flags & 2u) != 0;
    const seq_t p = encode_string(primer);
    NwResult r = nw_align(encode_string(read), p, sc);
    int uid_end = index_of_first_match_in_reference(r.pair);
    int ins = 0, del = 0, sub = 0;
    for (char ch : expanded_cigar(r.pair)) {
        if (ch == 'X') sub++;
        else if (ch == 'I') ins++;
        else if (ch == 'D') del++;
    }
    const int edit = ins + del + sub, cutoff = (int)p.size() / 4;
    if (edit >= cutoff) uid_end = -1;
    return uid_end;
}
int32_t hsqo_cigar_edit_distance(const char* ref, const char* qry) {
    Scorer sc;
    return cigar_edit_distance(expanded_cigar(nw_align(encode_string(ref), encode_string(qry), sc).pair));
}
// MD over the begin/end-trimmed alignment with the UNTRIMMED cigar (CigarStringUtilTest.java:75-109); reverse!=0 reverses both.
int32_t hsqo_md_trimmed_with_full_cigar(const char* ref, const char* qry, int32_t reverse, char* out, int32_t cap) {
    Scorer sc;
    NwResult r = nw_align(encode_string(ref), encode_string(qry), sc);
    AlignmentPair t = r.pair.without_ending_and_beginning_query_inserts();
    std::string cig = expanded_cigar(r.pair);
    seq_t ra = t.ref_aln;
    if (reverse) { std::reverse(cig.begin(), cig.end()); ra = seq_reverse(ra); }
    std::string md = md_string(ra, cig);
    snprintf(out, (size_t)cap, "%s", md.c_str());
    return (int32_t)md.size();
}

}  // extern "C"
