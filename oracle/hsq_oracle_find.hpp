// hsq_oracle_find.hpp -- TEST INFRASTRUCTURE ONLY (parity oracle).
// CPU restatement of the candidate stage of FastqFindAndTrimTool:
//   COMMON/fastqtool/FastqFindAndTrimTool.java:1188-1219 (trimReadBasedOnQuality), :904-926 (per entry front half)
//   COMMON/mapping/SimpleMapper.java:90-143 (addReferenceSequence / addSliceToReferenceMap), :154-217 (candidates, tallies)
//   COMMON/mapping/TallyMap.java:61-140 (counts; sort from most tallies to least)
// Pinned on the two SimpleMapperTest cases (COMMON_T/mapping/SimpleMapperTest.java:35-63, tests/test_find.py).
// Where the reference's answer depends on HashMap iteration order (entries with equal tallies) the oracle walks them in
// ascending row order and raises order_dependent when another order could give another list.
#pragma once
#include <algorithm>
#include <cstdint>
#include <map>
#include <set>
#include <stdexcept>
#include <string>
#include <vector>
#include "../include/hsq_b200.h"
#include "hsq_oracle_core.hpp"

namespace hsqo {

struct SimpleMapperO {
    int k = 5, max_refs = 0;
    std::map<std::string, std::set<int>> slices;  // sequenceSliceToReferenceAddressMap
    std::set<std::string> excluded;               // sequencesToExclude
    std::vector<int> ref_size;                    // referenceAddressesToSizeMap

    static std::string nucleotides(const std::string& s) {  // NucleotideCode.getNucleotidesFromString: upper case, A/C/G/T only
        std::string o = s;
        for (char& c : o) {
            if (c >= 'a' && c <= 'z') c = (char)(c - 32);
            if (c != 'A' && c != 'C' && c != 'G' && c != 'T') throw std::invalid_argument("Unable to translate character");
        }
        return o;
    }
    void add_reference(const std::string& raw, int address) {  // SimpleMapper.java:90-101
        const std::string seq = nucleotides(raw);
        ref_size.push_back((int)seq.size());
        if ((int)seq.size() < k) throw std::length_error("comparison sequence size must be less than the size of all sequences");
        for (int start = 0; start < (int)seq.size() - k; start += 1) add_slice(seq.substr((size_t)start, (size_t)k), address);
    }
    void add_slice(const std::string& slice, int address) {  // :125-143
        if (excluded.count(slice)) return;
        std::set<int>& a = slices[slice];
        a.insert(address);
        if (max_refs > 0 && (int)a.size() > max_refs) {
            slices.erase(slice);
            excluded.insert(slice);
        }
    }
    // getReferenceTallyMap :205-217
    std::map<int, int> tallies(const std::string& q) const {
        std::map<int, int> t;
        for (int i = 0; i < (int)q.size() - k; i += 1) {
            auto it = slices.find(q.substr((size_t)i, (size_t)k));
            if (it == slices.end()) continue;
            for (int a : it->second) t[a]++;
        }
        return t;
    }
    // getBestCandidateReferences :154-187; entries of equal tallies in ascending row order
    std::vector<std::pair<int, int>> candidates(const std::string& q, int limit, double min_ratio, bool* order_dependent) const {
        const double available_q = (double)((int)q.size() - k) / 1.0;
        const double limit_q = min_ratio * available_q;
        std::map<int, int> t = tallies(q);
        std::vector<std::pair<int, int>> entries(t.begin(), t.end());  // ascending row
        std::stable_sort(entries.begin(), entries.end(), [](const std::pair<int, int>& a, const std::pair<int, int>& b) { return a.second > b.second; });
        std::vector<std::pair<int, int>> best;
        int last_hits = 0;
        *order_dependent = false;
        for (size_t e = 0; e < entries.size(); e++) {
            const int hits = entries[e].second;
            const double limit_r = min_ratio * ((double)(ref_size[(size_t)entries[e].first] - k) / 1.0);
            const bool below_ratio = (hits < limit_q) && (hits < limit_r);
            const bool past_limit = ((int)best.size() > limit) && (hits < last_hits);
            if (below_ratio || past_limit) {
                if (!past_limit) {  // stopped by the ratio test: would an entry of the same tally have been accepted?
                    for (size_t f = 0; f < entries.size(); f++)
                        if (entries[f].second == hits && !(hits < min_ratio * (double)(ref_size[(size_t)entries[f].first] - k)))
                            *order_dependent = true;
                }
                break;
            }
            best.push_back(entries[e]);
            last_hits = hits;
        }
        return best;
    }
};

inline int32_t find_candidates(const hsq_seqs* seqs, int k, int max_refs, const hsq_seqs* reads, const uint8_t* quals, int lead_q, int trail_q,
                               int limit, double min_ratio, hsq_find_out* o) {
    SimpleMapperO m;
    m.k = k;
    m.max_refs = max_refs;
    for (int64_t r = 0; r < seqs->n; r++) {
        const std::string s((const char*)seqs->data + seqs->offsets[r], (size_t)(seqs->offsets[r + 1] - seqs->offsets[r]));
        try {
            m.add_reference(s, (int)r);
        } catch (const std::invalid_argument&) {
            return HSQ_ERR_BAD_CHARACTER;
        } catch (const std::length_error&) {
            return HSQ_ERR_OUT_OF_RANGE;
        }
    }
    const int K = o->max_candidates;
    int32_t status = HSQ_OK;
    for (int64_t r = 0; r < reads->n; r++) {
        const int64_t a = reads->offsets[r], len = reads->offsets[r + 1] - a;
        const std::string read((const char*)reads->data + a, (size_t)len);
        const uint8_t* q = quals + a;
        for (char c : read)  // new IupacNucleotideCodeSequence(record.getReadString()) :906
            if (encode_char(c) < 0) status = HSQ_ERR_BAD_CHARACTER;
        // trimReadBasedOnQuality (FastqFindAndTrimTool.java:1194-1214)
        int lead = 0;
        while (lead < len && ((int)q[lead] - 33) < lead_q) lead++;
        int trail = 0;
        while (trail < (len - lead) && ((int)q[len - trail - 1] - 33) < trail_q) trail++;
        const std::string trimmed = read.substr((size_t)lead, (size_t)(len - trail - lead));
        o->lead_trimmed[r] = lead;
        o->trail_trimmed[r] = trail;
        o->n_candidates[r] = 0;
        o->overflow[r] = 0;
        o->order_dependent[r] = 0;
        bool has_n = false;
        for (char c : trimmed) has_n |= (c == 'n' || c == 'N');  // toLowerCase().contains("n") :920
        o->has_n[r] = has_n ? 1 : 0;
        if (has_n) continue;
        std::string query;
        try {
            query = SimpleMapperO::nucleotides(trimmed);
        } catch (const std::invalid_argument&) {
            status = HSQ_ERR_BAD_CHARACTER;
            continue;
        }
        bool od = false;
        auto best = m.candidates(query, limit, min_ratio, &od);
        o->order_dependent[r] = od ? 1 : 0;
        o->n_candidates[r] = (int32_t)std::min<size_t>(best.size(), (size_t)K);
        for (size_t c = 0; c < best.size(); c++) {
            if ((int)c < K) {
                o->candidates[r * K + (int64_t)c] = best[c].first;
                o->hits[r * K + (int64_t)c] = best[c].second;
            } else {
                o->overflow[r] = 1;
            }
        }
    }
    return status;
}

}  // namespace hsqo
