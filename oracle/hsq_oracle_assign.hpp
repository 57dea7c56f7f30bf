// hsq_oracle_assign.hpp -- TEST INFRASTRUCTURE ONLY (parity oracle).
// CPU restatement of the coordinate read -> probe assignment of HSQUtils dedup:
//   HSQ/process/RangeMap.java:62-153 (put, stable sort by start, getObjectsThatContainRangeInclusive)
//   HSQ/process/PrimerReadExtensionAndPcrDuplicateIdentification.java:313-331 (one RangeMap per contig and strand),
//   :439-545 (ReadToProbeAssigner.assignProbesToReads, executed here in input order on one thread), :547-576 (alternative hits),
//   :392-400 (read names seen on more than one completed pair are removed).
// Parity unpinned by the reference: it holds no test for this path (RangeMap.main's go3/go4 print without asserting).
#pragma once
#include <algorithm>
#include <climits>
#include <cstdint>
#include <map>
#include <set>
#include <vector>
#include "../include/hsq_b200.h"

namespace hsqo {

struct RangeMapO {
    struct Entry { int start, end, value; };
    std::vector<Entry> v;
    bool sorted = true;
    void put(int a, int b, int value) {  // RangeMap.java:62-67
        v.push_back(Entry{std::min(a, b), std::max(a, b), value});
        sorted = false;
    }
    // RangeMap.java:92-153, statement by statement
    std::vector<int> contain(int a, int b) {
        const int find_start = std::min(a, b), find_stop = std::max(a, b);
        std::vector<int> out;
        if (!sorted) {
            std::stable_sort(v.begin(), v.end(), [](const Entry& x, const Entry& y) { return x.start < y.start; });
            sorted = true;
        }
        const int size = (int)v.size();
        int lo = 1, hi = size;
        bool done = false;
        while (!done) {
            int after = lo + ((hi - lo) / 2);  // Java int division truncates toward zero, like C++
            if (after <= size) {
                int after_loc = INT_MAX;
                if (after < size) after_loc = v[(size_t)after].start;
                if (after_loc >= find_start) {
                    const int before_loc = v[(size_t)(after - 1)].start;  // after >= 1 always holds on this path for size >= 1
                    if (before_loc <= find_start) {
                        int cur = after - 1;  // index of currentIndex
                        while (cur >= 0 && v[(size_t)cur].end >= find_stop) {
                            if (v[(size_t)cur].start <= find_start && v[(size_t)cur].end >= find_stop) out.push_back(v[(size_t)cur].value);
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
        return out;
    }
};

inline void write_set(const std::set<int>& s, int32_t* count, int32_t* slots, int cap, uint8_t* overflow) {
    *count = (int32_t)std::min<size_t>(s.size(), (size_t)cap);  // an overflowing record reports a full slot array
    int k = 0;
    for (int p : s) {
        if (k < cap) slots[k] = p;
        k++;
    }
    if (k > cap) *overflow = 1;
}

inline void assign_probes(const hsq_probe_ranges* pr, const hsq_mapped_records* rec, hsq_assign_out* o) {
    std::map<std::pair<int, int>, RangeMapO> maps;  // (contig, strand)
    std::set<int> contigs;
    for (int64_t p = 0; p < pr->n; p++) {
        contigs.insert(pr->contig[p]);
        maps[{pr->contig[p], 0}];
        maps[{pr->contig[p], 1}];  // both maps exist for every sequence name of the probe file (:315-330)
        maps[{pr->contig[p], pr->reverse_strand[p] ? 1 : 0}].put(pr->start[p], pr->stop[p], (int)p);
    }
    const int K = o->max_probes;
    auto query = [&](int contig, bool negative, int a, int b, std::set<int>* into) {
        if (contig < 0 || !contigs.count(contig)) return false;
        for (int p : maps[{contig, negative ? 1 : 0}].contain(a, b)) into->insert(p);
        return true;
    };
    const int64_t n = rec->n;
    std::vector<std::set<int>> own((size_t)n), uni((size_t)n);
    std::vector<char> usable((size_t)n, 0);
    for (int64_t k = 0; k < n; k++) {
        o->n_probes[k] = 0; o->n_common[k] = 0; o->overflow[k] = 0; o->mate[k] = HSQ_ASSIGN_UNPAIRED;
        bool neg = rec->negative_strand[k] != 0;
        if (rec->second_of_pair[k]) neg = !neg;  // :452-455
        usable[(size_t)k] = query(rec->contig[k], neg, rec->start[k], rec->end[k], &own[(size_t)k]) ? 1 : 0;
        if (!usable[(size_t)k]) { o->mate[k] = HSQ_ASSIGN_IGNORED; continue; }
        uni[(size_t)k] = own[(size_t)k];
        if (rec->alt_offsets)
            for (int64_t a = rec->alt_offsets[k]; a < rec->alt_offsets[k + 1]; a++) {  // :547-576
                bool an = rec->alt_negative[a] != 0;
                if (rec->second_of_pair[k]) an = !an;
                query(rec->alt_contig[a], an, rec->alt_start[a], rec->alt_stop[a], &uni[(size_t)k]);
            }
        write_set(uni[(size_t)k], &o->n_probes[k], o->probes + k * K, K, &o->overflow[k]);
    }
    // the sequential first-found / second-found logic (:470-540) in input order
    std::map<int, int64_t> unpaired;           // read index -> first found record
    std::map<int, std::pair<int64_t, int64_t>> results;
    std::set<int> repeated;
    for (int64_t k = 0; k < n; k++) {
        if (!usable[(size_t)k]) continue;
        const int ri = rec->read_index[k];
        auto it = unpaired.find(ri);
        if (it == unpaired.end()) {
            unpaired[ri] = k;
        } else if (results.count(ri)) {
            repeated.insert(ri);  // :484-486 (the unpaired entry stays)
        } else {
            results[ri] = {it->second, k};
            unpaired.erase(it);
        }
    }
    for (int ri : repeated) {  // :392-400
        for (int64_t k = 0; k < n; k++)
            if (usable[(size_t)k] && rec->read_index[k] == ri) o->mate[k] = HSQ_ASSIGN_AMBIGUOUS_NAME;
        results.erase(ri);
    }
    for (auto& r : results) {
        const int64_t a = r.second.first, b = r.second.second;
        std::set<int> common;
        for (int p : uni[(size_t)a])
            if (uni[(size_t)b].count(p)) common.insert(p);  // :478-483
        o->mate[a] = (int32_t)b;
        o->mate[b] = (int32_t)a;
        write_set(common, &o->n_common[a], o->common + a * K, K, &o->overflow[a]);
        write_set(common, &o->n_common[b], o->common + b * K, K, &o->overflow[b]);
        if (o->overflow[a] || o->overflow[b]) o->overflow[a] = o->overflow[b] = 1;  // an incomplete mate set taints the pair
    }
}

}  // namespace hsqo
