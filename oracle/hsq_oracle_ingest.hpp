// hsq_oracle_ingest.hpp -- TEST INFRASTRUCTURE ONLY (parity oracle / timed CPU baseline).
// CPU restatement of the FASTQ <-> SAM name join in front of the dedup path:
//   HSQ/process/FastqAndBamFileMerger.java:65-94 (isTrimAmountCorrect), :107-158 (driver), :213-345 (the merge loop of
//   MergedSamIterator), :386-396 (storeFastqInfoInRecord); HSQ/utils/FastqSorter.java:90-124,291-318 (index appended to the
//   header with " _ "), :334-339 (FastqRecordNameComparator); HSQ/utils/BamSorter.java:322-361 (the two SAM comparators);
//   COMMON/utils/IlluminaFastQReadNameUtil.java:40-58 (getUniqueIdForReadHeader); HSQ/process/FastqReadTrimmer.java:309-328.
// The FASTQ text grammar is htsjdk's FastqReader (htsjdk 2.9.0, pinned in git_superpom/pom.xml:72-75, source absent from
// /root/reference): four lines per record, '@' header, '+' quality header, equal sequence / quality length, no blank lines.
// Parity: unpinned by the reference (its tests hold nothing for FastqAndBamFileMerger); pinned here on hand traces
// (tests/test_ingest.py) and on a second, dictionary-based statement of the same join.
// Sorting: the chunked external sorts (FastqSorter.java:291-318 + the TreeSet merge :136-141, BamSorter) are restated as one
// stable in-memory sort by the same comparator.  With the index appended every FASTQ key is distinct, so the TreeSet merge
// loses nothing; SAM records that compare equal (same name, same first-of-pair flag) keep input order.
#pragma once
#include <algorithm>
#include <climits>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

namespace hsqo {

static const char* const FASTQ_NAME_AND_INDEX_DELIMITER = " _ ";  // FastqSorter.java:50

struct JoinNameMismatch : std::runtime_error { using std::runtime_error::runtime_error; };   // IllegalStateException :270-273
struct JoinNotTrimmed : std::runtime_error { using std::runtime_error::runtime_error; };     // UnableToMergeFastqAndBamFilesException :370
struct JoinNoMatches : std::runtime_error { using std::runtime_error::runtime_error; };      // IllegalStateException :143-148
struct JoinBadNumber : std::runtime_error { using std::runtime_error::runtime_error; };      // NumberFormatException :276

// IlluminaFastQReadNameUtil.getUniqueIdForReadHeader (:40-58)
inline std::string unique_id_for_read_header(const std::string& h) {
    size_t sp = h.find(' '), sl = h.find('/');
    size_t end = std::min(std::min(sp, sl), h.size());  // npos is the largest value, as Integer.MAX_VALUE is there
    return h.substr(0, end);
}

// the name both FASTQ headers go through before they are compared (:252-263)
inline std::string fastq_join_name(std::string h) {
    size_t us = h.rfind(FASTQ_NAME_AND_INDEX_DELIMITER);
    if (us != std::string::npos) h = h.substr(0, us);
    return unique_id_for_read_header(h);
}

// Integer.parseInt: optional sign, decimal digits, int32 range
inline int32_t java_parse_int(const std::string& s) {
    size_t i = 0;
    bool neg = false;
    if (!s.empty() && (s[0] == '-' || s[0] == '+')) { neg = s[0] == '-'; i = 1; }
    if (i >= s.size()) throw JoinBadNumber("For input string: \"" + s + "\"");
    int64_t v = 0;
    for (; i < s.size(); i++) {
        if (s[i] < '0' || s[i] > '9') throw JoinBadNumber("For input string: \"" + s + "\"");
        v = v * 10 + (s[i] - '0');
        if (v > (int64_t)INT32_MAX + 1) throw JoinBadNumber("For input string: \"" + s + "\"");
    }
    if (neg) v = -v;
    if (v > INT32_MAX || v < INT32_MIN) throw JoinBadNumber("For input string: \"" + s + "\"");
    return (int32_t)v;
}

// length FastqReadTrimmer.trim would keep (FastqReadTrimmer.java:309-328); throws as it does
inline int trimmed_length(int len, int first, int last, bool three_prime) {
    if (first >= len) throw std::out_of_range("Unable to trim bases from the beginning of a sequence");  // :310-312
    int last_base = len - 1;
    if (three_prime) last_base = std::min(last, len - 1);
    if (last_base + 1 < first) throw std::out_of_range("substring end before begin");  // String.substring
    return last_base + 1 - first;
}

struct JoinParams {
    bool by_index = false;          // useFastqIndexesAsFastqReadNamesWhenMerging; also: FASTQ not sorted, no index appended
    bool trimming_skipped = false, check_trim = true, three_prime = false;
    int r1_first = 0, r1_stop = 0, r2_first = 0, r2_stop = 0;
};

struct JoinInput {
    std::vector<std::string> sam_names;
    std::vector<int32_t> sam_flags, sam_seq_len, sam_qual_len;
    std::vector<std::string> fq1_headers, fq2_headers;  // without '@'
    std::vector<int32_t> fq1_seq_len, fq2_seq_len;
};

struct JoinOutput {
    std::vector<int32_t> sam_record, fastq_record, name_len;
    std::vector<uint8_t> status;  // by input record: 0 not primary, 1 emitted, 2 dropped
    int64_t total_matching_pairs = 0;
};

inline JoinOutput join_names(const JoinInput& in, const JoinParams& p) {
    const size_t ns = in.sam_names.size();
    JoinOutput out;
    out.status.assign(ns, 2);
    // FASTQ iterators: (header as the iterator returns it, ordinal in the file)
    struct Fq { std::string header; int32_t ordinal; };
    auto make_fastq = [&](const std::vector<std::string>& hs) {
        std::vector<Fq> v(hs.size());
        for (size_t k = 0; k < hs.size(); k++) {
            v[k].ordinal = (int32_t)k;
            v[k].header = p.by_index ? hs[k] : hs[k] + FASTQ_NAME_AND_INDEX_DELIMITER + std::to_string(k);  // FastqSorter.java:298-302
        }
        if (!p.by_index)  // FastqRecordNameComparator :334-339 (String.compareTo == unsigned byte order for ASCII)
            std::stable_sort(v.begin(), v.end(), [](const Fq& a, const Fq& b) { return a.header < b.header; });
        return v;
    };
    const std::vector<Fq> f1 = make_fastq(in.fq1_headers), f2 = make_fastq(in.fq2_headers);
    // SAM iterator
    std::vector<int32_t> order(ns);
    for (size_t k = 0; k < ns; k++) order[k] = (int32_t)k;
    auto first_of_pair = [&](int32_t r) { return (in.sam_flags[(size_t)r] & 0x40) != 0; };
    if (p.by_index) {  // SamRecordNameAsIndexComparator, BamSorter.java:322-346 (names that are not integers: outside this statement)
        std::vector<int32_t> key(ns);
        for (size_t k = 0; k < ns; k++) key[k] = java_parse_int(in.sam_names[k]);
        std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) {
            if (key[(size_t)a] != key[(size_t)b]) return key[(size_t)a] < key[(size_t)b];
            return first_of_pair(a) && !first_of_pair(b);
        });
    } else {  // SamRecordNameComparator, BamSorter.java:348-361
        std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) {
            int c = in.sam_names[(size_t)a].compare(in.sam_names[(size_t)b]);
            if (c != 0) return c < 0;
            return first_of_pair(a) && !first_of_pair(b);
        });
    }
    // MergedSamIterator.getNextRecordToReturn (:213-345), run to exhaustion
    size_t si = 0, fi1 = 0, fi2 = 0;
    int32_t sam = -1;
    const Fq *one = nullptr, *two = nullptr;
    int32_t fastq_index = -1;
    while (((sam < 0 && si < ns) || sam >= 0) && ((!one && fi1 < f1.size() && !two && fi2 < f2.size()) || (one && two))) {
        if (sam < 0) {
            sam = order[si++];
            if (in.sam_flags[(size_t)sam] & 0x100) {  // :222-225
                out.status[(size_t)sam] = 0;
                sam = -1;
                continue;
            }
        }
        if (!one && !two) {
            fastq_index++;
            one = &f1[fi1++];
            two = &f2[fi2++];
        }
        const std::string sam_name = unique_id_for_read_header(in.sam_names[(size_t)sam]);
        const std::string one_name = fastq_join_name(one->header), two_name = fastq_join_name(two->header);
        if (one_name != two_name)
            throw JoinNameMismatch("The read names within the Fastq files do not match, Fastq1 has read name [" + one_name + "] whereas Fastq2 has read name [" + two_name + "]");
        int cmp;
        if (p.by_index) {
            int32_t v = java_parse_int(sam_name);
            cmp = v < fastq_index ? -1 : (v > fastq_index ? 1 : 0);
        } else {
            cmp = sam_name.compare(one_name);
        }
        if (cmp < 0) {
            sam = -1;  // skipped with a warning (:282-297)
        } else if (cmp > 0) {
            one = two = nullptr;
        } else {
            const bool first = first_of_pair(sam);
            const Fq* used = first ? one : two;
            const int32_t fq_len = first ? in.fq1_seq_len[(size_t)used->ordinal] : in.fq2_seq_len[(size_t)used->ordinal];
            if (!first) {
                one = two = nullptr;
                out.total_matching_pairs++;
            }
            // checkAndStoreFastqInfoInRecord :366-373, isTrimAmountCorrect :65-94 (the !trimmingSkipped branch is the only one reached)
            if (!p.trimming_skipped && p.check_trim) {
                const int expected = first ? trimmed_length(fq_len, p.r1_first, p.r1_stop, p.three_prime)
                                           : trimmed_length(fq_len, p.r2_first, p.r2_stop, p.three_prime);
                if (in.sam_seq_len[(size_t)sam] > expected || in.sam_qual_len[(size_t)sam] > expected)
                    throw JoinNotTrimmed("Readname:" + in.sam_names[(size_t)sam] + " is not trimmed appropriately");
            }
            out.status[(size_t)sam] = 1;
            out.sam_record.push_back(sam);
            out.fastq_record.push_back(used->ordinal);
            out.name_len.push_back((int32_t)one_name.size());
            sam = -1;
        }
    }
    // records the loop never reached stay "dropped"; non-primary ones among them are reported as such
    for (size_t k = 0; k < ns; k++)
        if (in.sam_flags[k] & 0x100) out.status[k] = 0;
    if (out.total_matching_pairs == 0)
        throw JoinNoMatches("The read names in the input Fastq files do not match the reads names in the provided bam/sam file");
    return out;
}

// ---- FASTQ text -> record index (htsjdk FastqReader grammar, see the header) -------------------------------------------
struct FastqIndex {
    std::vector<int64_t> header_off, seq_off, qual_off;
    std::vector<int32_t> header_len, seq_len, qual_len;
};
struct FastqMalformed : std::runtime_error { using std::runtime_error::runtime_error; };

inline FastqIndex index_fastq(const uint8_t* text, int64_t n) {
    FastqIndex ix;
    int64_t pos = 0;
    auto read_line = [&](int64_t* off, int32_t* len) -> bool {  // BufferedReader.readLine for \n and \r\n
        if (pos >= n) return false;
        int64_t e = pos;
        while (e < n && text[e] != '\n') e++;
        int64_t stop = e;
        if (stop > pos && text[stop - 1] == '\r') stop--;
        for (int64_t k = pos; k < stop; k++)
            if (text[k] == '\r') throw FastqMalformed("lone carriage return");  // a third terminator readLine knows: not restated
        *off = pos;
        *len = (int32_t)(stop - pos);
        pos = e < n ? e + 1 : n;
        return true;
    };
    auto blank = [&](int64_t off, int32_t len) {  // StringUtil.isBlank
        for (int32_t k = 0; k < len; k++)
            if (text[off + k] > ' ') return false;
        return true;
    };
    while (true) {
        int64_t o[4];
        int32_t l[4];
        if (!read_line(&o[0], &l[0])) break;
        for (int k = 1; k < 4; k++)
            if (!read_line(&o[k], &l[k])) throw FastqMalformed("File is too short");
        for (int k = 0; k < 4; k++)
            if (blank(o[k], l[k])) throw FastqMalformed("Missing line");
        if (text[o[0]] != '@') throw FastqMalformed("Sequence header must start with @");
        if (text[o[2]] != '+') throw FastqMalformed("Quality header must start with +");
        if (l[1] != l[3]) throw FastqMalformed("Sequence and quality line must be the same length");
        ix.header_off.push_back(o[0] + 1);
        ix.header_len.push_back(l[0] - 1);
        ix.seq_off.push_back(o[1]);
        ix.seq_len.push_back(l[1]);
        ix.qual_off.push_back(o[3]);
        ix.qual_len.push_back(l[3]);
    }
    return ix;
}

// ---- FASTQ text -> trimmed FASTQ text: FastqReadTrimmer.trimReads (HSQ/process/FastqReadTrimmer.java:263-307) ---------------------
// every record keeps its header and quality header; bases and qualities are cut by trim(String, String, int, int, boolean) (:309-328);
// records are written as htsjdk's BasicFastqWriter does: "@" header "\n" bases "\n+" quality-header "\n" qualities "\n".
inline std::string trim_fastq_text(const uint8_t* text, int64_t n, int first, int last, bool three_prime) {
    if (first < 0) throw std::invalid_argument("First base to keep must be greater than zero");                       // :264-266
    if (last <= first) throw std::invalid_argument("Last base to keep must be greater than the first base to keep");  // :268-270
    const FastqIndex ix = index_fastq(text, n);
    std::string out;
    for (size_t k = 0; k < ix.header_off.size(); k++) {
        const int keep = trimmed_length(ix.seq_len[k], first, last, three_prime);
        const int64_t plus = ix.seq_off[k] + ix.seq_len[k];  // line terminator of the bases, then the '+' line up to the qualities
        int64_t p0 = plus;
        while (text[p0] != '+') p0++;
        int64_t p1 = p0;
        while (text[p1] != '\n' && text[p1] != '\r') p1++;
        out += '@';
        out.append((const char*)text + ix.header_off[k], (size_t)ix.header_len[k]);
        out += '\n';
        out.append((const char*)text + ix.seq_off[k] + first, (size_t)keep);
        out += "\n+";
        out.append((const char*)text + p0 + 1, (size_t)(p1 - p0 - 1));
        out += '\n';
        out.append((const char*)text + ix.qual_off[k] + first, (size_t)keep);
        out += '\n';
    }
    return out;
}

// ---- SAM text -> field index ------------------------------------------------------------------------------------------------
// The columns the dedup path reads off a SAMRecord: QNAME / FLAG / SEQ / QUAL lengths (FastqAndBamFileMerger.java:222-330),
// RNAME, POS, strand and alignment end = POS + reference length of the CIGAR - 1, and the XA tag
// (...Identification.java:439-576).  Line grammar as htsjdk's SAMLineParser (2.9.0, absent): '@' header lines first, then lines of
// at least 11 tab separated columns; FLAG, POS, MAPQ, PNEXT, TLEN decimal integers; CIGAR '*' or <count><op> runs.
struct SamIndex {
    int64_t n_header_lines = 0;
    std::vector<int64_t> qname_off, rname_off, cigar_off, seq_off, qual_off, tags_off, xa_off;
    std::vector<int32_t> qname_len, rname_len, cigar_len, seq_len, qual_len, tags_len, xa_len, flag, pos, mapq, pnext, tlen, ref_len;
};

inline SamIndex index_sam(const uint8_t* text, int64_t n) {
    SamIndex ix;
    int64_t p = 0;
    bool in_header = true;
    auto parse_int = [&](int64_t off, int32_t len) -> int32_t {
        try {
            return java_parse_int(std::string((const char*)text + off, (size_t)len));
        } catch (const JoinBadNumber&) { throw FastqMalformed("a numeric SAM column is not an integer"); }
    };
    while (p < n) {
        int64_t e = p;
        while (e < n && text[e] != '\n') e++;
        int64_t stop = e;
        if (stop > p && text[stop - 1] == '\r') stop--;
        const int64_t start = p;
        p = e < n ? e + 1 : n;
        if (stop == start) throw FastqMalformed("empty line");
        if (text[start] == '@') {
            if (!in_header) throw FastqMalformed("header line after the first alignment line");
            ix.n_header_lines++;
            continue;
        }
        in_header = false;
        int64_t off[12];
        int32_t len[12];
        int nf = 0;
        int64_t fs = start;
        for (int64_t k = start; k <= stop && nf < 11; k++)
            if (k == stop || text[k] == '\t') { off[nf] = fs; len[nf] = (int32_t)(k - fs); nf++; fs = k + 1; }
        if (nf < 11) throw FastqMalformed("fewer than 11 columns");
        off[11] = std::min(fs, stop);
        len[11] = (int32_t)(stop - off[11]);
        ix.qname_off.push_back(off[0]); ix.qname_len.push_back(len[0]);
        ix.flag.push_back(parse_int(off[1], len[1]));
        ix.rname_off.push_back(off[2]); ix.rname_len.push_back(len[2]);
        ix.pos.push_back(parse_int(off[3], len[3]));
        ix.mapq.push_back(parse_int(off[4], len[4]));
        ix.cigar_off.push_back(off[5]); ix.cigar_len.push_back(len[5]);
        ix.pnext.push_back(parse_int(off[7], len[7]));
        ix.tlen.push_back(parse_int(off[8], len[8]));
        ix.seq_off.push_back(off[9]); ix.seq_len.push_back(len[9]);
        ix.qual_off.push_back(off[10]); ix.qual_len.push_back(len[10]);
        ix.tags_off.push_back(off[11]); ix.tags_len.push_back(len[11]);
        int64_t rl = 0;  // Cigar.getReferenceLength: M, D, N, =, X consume the reference
        if (!(len[5] == 1 && text[off[5]] == '*')) {
            int64_t count = 0;
            bool digits = false;
            for (int32_t k = 0; k < len[5]; k++) {
                const uint8_t c = text[off[5] + k];
                if (c >= '0' && c <= '9') { count = count * 10 + (c - '0'); digits = true; if (count > INT32_MAX) throw FastqMalformed("CIGAR count"); continue; }
                if (!digits) throw FastqMalformed("CIGAR operator without a count");
                if (c == 'M' || c == 'D' || c == 'N' || c == '=' || c == 'X') rl += count;
                else if (!(c == 'I' || c == 'S' || c == 'H' || c == 'P')) throw FastqMalformed("unknown CIGAR operator");
                count = 0;
                digits = false;
            }
            if (digits || len[5] == 0) throw FastqMalformed("CIGAR ends inside a count");
            if (rl > INT32_MAX) throw FastqMalformed("CIGAR reference length");
        }
        ix.ref_len.push_back((int32_t)rl);
        int64_t xo = off[11];
        int32_t xl = 0;  // first XA:Z: tag
        for (int64_t k = off[11]; k < stop;) {
            int64_t t = k;
            while (t < stop && text[t] != '\t') t++;
            if (t - k >= 5 && text[k] == 'X' && text[k + 1] == 'A' && text[k + 2] == ':' && text[k + 3] == 'Z' && text[k + 4] == ':') {
                xo = k + 5;
                xl = (int32_t)(t - xo);
                break;
            }
            k = t + 1;
        }
        ix.xa_off.push_back(xl ? xo : 0); ix.xa_len.push_back(xl);
    }
    return ix;
}

// ---- SAM text -> the records ReadToProbeAssigner.assignProbesToReads works on ------------------------------------------------
// ...Identification.java:439-470: records with the unmapped or mate-unmapped flag are dropped (:446; htsjdk's getMateUnmappedFlag
// throws IllegalStateException on a record that is not paired), strand / second-of-pair flags, getAlignmentStart / getAlignmentEnd,
// Integer.parseInt(readName) for records on a contig that has probes (:469-474).  Alternative hits:
// SAMRecordUtil.getAlternativeHitsFromAttribute -> AlternativeHit.fromAttributeText (HSQ/utils/SAMRecordUtil.java:362-375,440-468):
// entries split on ';', fields on ',' (String.split: trailing empty strings dropped), exactly four fields, Integer.parseInt of the
// second (sign = strand, abs = start; NumberFormatException: entry skipped), stop = getAlignmentEnd of the copy = start +
// reference length of the entry's CIGAR - 1 (the `stop` computed at :457 is handed to a constructor that ignores it, :399-406).
struct SamUnpaired : std::runtime_error { using std::runtime_error::runtime_error; };
struct SamNameNotANumber : std::runtime_error { using std::runtime_error::runtime_error; };

inline std::vector<std::string> java_split(const std::string& s, char sep) {
    std::vector<std::string> out;
    size_t b = 0;
    for (size_t k = 0; k <= s.size(); k++)
        if (k == s.size() || s[k] == sep) { out.push_back(s.substr(b, k - b)); b = k + 1; }
    while (!out.empty() && out.back().empty()) out.pop_back();
    if (out.empty() && s.empty()) out.push_back("");  // "".split(x) is [""]
    return out;
}

inline bool cigar_ref_len(const std::string& c, int64_t* out) {
    if (c == "*") { *out = 0; return true; }
    if (c.empty()) return false;
    int64_t rl = 0, count = 0;
    bool digits = false;
    for (char ch : c) {
        if (ch >= '0' && ch <= '9') { count = count * 10 + (ch - '0'); digits = true; if (count > INT32_MAX) return false; continue; }
        if (!digits) return false;
        if (ch == 'M' || ch == 'D' || ch == 'N' || ch == '=' || ch == 'X') rl += count;
        else if (!(ch == 'I' || ch == 'S' || ch == 'H' || ch == 'P')) return false;
        count = 0;
        digits = false;
    }
    if (digits || rl > INT32_MAX) return false;
    *out = rl;
    return true;
}

struct SamRecords {
    std::vector<int32_t> sam_line, contig, start, end, read_index, alt_contig, alt_start, alt_stop;
    std::vector<uint8_t> negative_strand, second_of_pair, alt_negative;
    std::vector<int64_t> alt_offsets{0};
};

inline SamRecords sam_mapped_records(const uint8_t* text, int64_t n, const std::vector<std::string>& contig_names) {
    const SamIndex ix = index_sam(text, n);
    SamRecords out;
    auto contig_of = [&](const std::string& name) {
        for (size_t c = 0; c < contig_names.size(); c++)
            if (contig_names[c] == name) return (int32_t)c;
        return (int32_t)-1;
    };
    auto str = [&](int64_t off, int32_t len) { return std::string((const char*)text + off, (size_t)len); };
    for (size_t k = 0; k < ix.flag.size(); k++) {
        const int32_t f = ix.flag[k];
        if (!(f & 0x1)) throw SamUnpaired("Inappropriate call if not paired read");
        if ((f & 0x8) || (f & 0x4)) continue;
        const int32_t contig = contig_of(str(ix.rname_off[k], ix.rname_len[k]));
        int32_t read_index = 0;
        if (contig >= 0) {
            try {
                read_index = java_parse_int(str(ix.qname_off[k], ix.qname_len[k]));
            } catch (const JoinBadNumber&) { throw SamNameNotANumber("Expecting all read names to be numbers because the trim step also replaces names with the read index"); }
        }
        out.sam_line.push_back((int32_t)k);
        out.contig.push_back(contig);
        out.negative_strand.push_back((f & 0x10) ? 1 : 0);
        out.second_of_pair.push_back((f & 0x80) ? 1 : 0);
        out.start.push_back(ix.pos[k]);
        out.end.push_back(ix.pos[k] + ix.ref_len[k] - 1);
        out.read_index.push_back(read_index);
        if (ix.xa_len[k] > 0) {
            for (const std::string& entry : java_split(str(ix.xa_off[k], ix.xa_len[k]), ';')) {
                const std::vector<std::string> fields = java_split(entry, ',');
                if (fields.size() != 4) continue;
                int32_t start_and_strand;
                try {
                    start_and_strand = java_parse_int(fields[1]);
                } catch (const JoinBadNumber&) { continue; }  // warning, hit not used (:462-464)
                int64_t rl;
                if (!cigar_ref_len(fields[2], &rl)) throw FastqMalformed("CIGAR of an alternative hit");
                const int32_t start = start_and_strand < 0 ? -start_and_strand : start_and_strand;
                out.alt_contig.push_back(contig_of(fields[0]));
                out.alt_negative.push_back(start_and_strand < 0 ? 1 : 0);
                out.alt_start.push_back(start);
                out.alt_stop.push_back((int32_t)(start + rl - 1));
            }
        }
        out.alt_offsets.push_back((int64_t)out.alt_contig.size());
    }
    return out;
}

// ---- the final coordinate sort (...Identification.java:897): htsjdk 2.9.0 SAMRecordCoordinateComparator, restated from its published
// source (absent from /root/reference): fileOrderCompare (reference index, -1 last; alignment start), then, for equal strands, read name,
// flags, mapping quality, mate reference index, mate alignment start, inferred insert size; for unequal strands the reverse strand last.
struct CoordRecord {
    int32_t ref, pos, flags, mapq, mate_ref, mate_pos, tlen;
    std::string name;
};
inline std::vector<int32_t> coordinate_order(const std::vector<CoordRecord>& r) {
    std::vector<int32_t> order(r.size());
    for (size_t k = 0; k < r.size(); k++) order[k] = (int32_t)k;
    auto cmp = [](int64_t x, int64_t y) { return x < y ? -1 : (x > y ? 1 : 0); };
    std::stable_sort(order.begin(), order.end(), [&](int32_t x, int32_t y) {
        const CoordRecord &a = r[(size_t)x], &b = r[(size_t)y];
        int c;
        if (a.ref == -1) c = b.ref == -1 ? 0 : 1;
        else if (b.ref == -1) c = -1;
        else if ((c = cmp(a.ref, b.ref)) == 0) c = cmp(a.pos, b.pos);
        if (c != 0) return c < 0;
        const bool na = (a.flags & 0x10) != 0, nb = (b.flags & 0x10) != 0;
        if (na != nb) return !na;  // the record on the reverse strand goes last
        if ((c = a.name.compare(b.name)) != 0) return c < 0;
        if ((c = cmp(a.flags, b.flags)) != 0) return c < 0;
        if ((c = cmp(a.mapq, b.mapq)) != 0) return c < 0;
        if ((c = cmp(a.mate_ref, b.mate_ref)) != 0) return c < 0;
        if ((c = cmp(a.mate_pos, b.mate_pos)) != 0) return c < 0;
        return cmp(a.tlen, b.tlen) < 0;
    });
    return order;
}

}  // namespace hsqo
