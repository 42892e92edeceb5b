// dh_index.cu — BAI/CSI index and region planning for indexed read_bam scans.
//
// Host-only work (KBs–MBs, once per query).  Replaces, for BAM:
//   hts_idx_load3 -> idx_read / idx_read_core, update_loff          htslib/hts.c:2920-3055, 2426-2455
//   hts_parse_region / hts_reglist_create / reg_compact             htslib/hts.c (hts_parse_region), htslib/region.c:84-122,177-260
//   hts_itr_query, hts_itr_multi_bam, reg2bins / reg2intervals      htslib/hts.c:3421-3595, 3597-3745, 3142-3345
// What changes is what is done with the result.  htslib seeks chunk by chunk and tests every record it reads for
// overlap on the CPU (hts_itr_next :4266-4318, hts_itr_multi_next :4320-4604).  Here the chunk list of a query is
// reduced to a few virtual-offset ranges (chunks closer than GAP_MERGE compressed bytes are scanned as one range) and
// the overlap test itself runs inside the GPU boundary walk (bam_decode.cu region_pass), so the rows returned are the
// records in file order that overlap any requested interval — what the iterator returns on a coordinate-sorted file,
// which is the only kind an index can be built for.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "dh_index.h"

namespace dh {

namespace {

struct Reader {
    const uint8_t *p;
    uint64_t n, at = 0;
    bool ok = true;
    bool take(void *dst, uint64_t k) {
        if (!ok || n - at < k) { ok = false; return false; }
        memcpy(dst, p + at, k);
        at += k;
        return true;
    }
    uint32_t u32() { uint32_t v = 0; take(&v, 4); return v; }
    int32_t i32() { int32_t v = 0; take(&v, 4); return v; }
    uint64_t u64() { uint64_t v = 0; take(&v, 8); return v; }
};

inline int64_t bin_first(int l) { return ((1ll << (3 * l)) - 1) / 7; }
inline int64_t bin_parent(int64_t b) { return (b - 1) >> 3; }
inline int bin_level(int64_t b) { int l = 0; for (; b; ++l, b = bin_parent(b)); return l; }
inline int64_t bin_bot(int64_t bin, int n_lvls) { int l = bin_level(bin); return (bin - bin_first(l)) << ((n_lvls - l) * 3); }

}  // namespace

const IdxBin *IdxRef::find(uint32_t bin) const {
    auto it = std::lower_bound(bins.begin(), bins.end(), bin, [](const IdxBin &a, uint32_t b) { return a.bin < b; });
    return (it != bins.end() && it->bin == bin) ? &*it : nullptr;
}

// idx_read_core (hts.c:2920-2983) over an in-memory image; `csi` selects the per-bin loff field and the absence of
// the linear index.  Returns false on a truncated or inconsistent file.
static bool read_core(Reader &rd, HtsIndex &ix, bool csi) {
    for (auto &ref : ix.refs) {
        const int32_t n_bin = rd.i32();
        if (!rd.ok || n_bin < 0) return false;
        ref.bins.resize((size_t)n_bin);
        for (auto &b : ref.bins) {
            b.bin = rd.u32();
            b.loff = csi ? rd.u64() : 0;
            const int32_t n_chunk = rd.i32();
            if (!rd.ok || n_chunk < 0 || (uint64_t)n_chunk * 16 > rd.n - rd.at) return false;
            b.chunks.resize((size_t)n_chunk);
            for (auto &c : b.chunks) { c.u = rd.u64(); c.v = rd.u64(); }
        }
        std::sort(ref.bins.begin(), ref.bins.end(), [](const IdxBin &a, const IdxBin &b) { return a.bin < b.bin; });
        for (size_t i = 1; i < ref.bins.size(); ++i)
            if (ref.bins[i].bin == ref.bins[i - 1].bin) return false;   // duplicate bin number (:2938)
        if (!csi) {
            const int32_t n_intv = rd.i32();
            if (!rd.ok || n_intv < 0 || (uint64_t)n_intv * 8 > rd.n - rd.at) return false;
            ref.lin.resize((size_t)n_intv);
            for (auto &o : ref.lin) o = rd.u64();
            // fill missing values (:2971-2973), then update_loff (:2426-2455)
            size_t k = 0;
            while (k < ref.lin.size() && ref.lin[k] == 0) ++k;
            for (size_t j = ref.lin.size(); j-- > k + 1;)
                if (ref.lin[j - 1] == 0) ref.lin[j - 1] = ref.lin[j];
            for (auto &b : ref.bins) {
                if (b.bin < ix.n_bins) {
                    const int64_t bot = bin_bot(b.bin, ix.n_lvls);
                    b.loff = bot < (int64_t)ref.lin.size() ? ref.lin[(size_t)bot] : 0;
                } else b.loff = 0;
            }
        }
    }
    if (!rd.ok) return false;
    uint64_t nn = 0;
    ix.n_no_coor = rd.take(&nn, 8) ? nn : 0;    // optional trailer (:2976)
    return true;
}

bool index_parse(const uint8_t *data, uint64_t size, HtsIndex &ix, std::string &err) {
    Reader rd{data, size};
    char magic[4] = {0, 0, 0, 0};
    rd.take(magic, 4);
    if (rd.ok && !memcmp(magic, "BAI\1", 4)) {
        ix.csi = false;
        ix.min_shift = 14;
        ix.n_lvls = 5;
        const int32_t n = rd.i32();
        if (!rd.ok || n < 0) { err = "truncated BAI index"; return false; }
        ix.n_bins = (uint32_t)(((1u << (3 * ix.n_lvls + 3)) - 1) / 7);
        ix.refs.assign((size_t)n, IdxRef());
        if (!read_core(rd, ix, false)) { err = "malformed BAI index"; return false; }
        return true;
    }
    if (rd.ok && !memcmp(magic, "CSI\1", 4)) {
        ix.csi = true;
        ix.min_shift = rd.i32();
        ix.n_lvls = rd.i32();
        const int32_t l_aux = rd.i32();
        if (!rd.ok || l_aux < 0 || ix.min_shift < 0 || ix.min_shift > 40 || ix.n_lvls < 0 || ix.n_lvls > 9 || (uint64_t)l_aux > rd.n - rd.at) {
            err = "malformed CSI index header";
            return false;
        }
        rd.at += (uint64_t)l_aux;
        const int32_t n = rd.i32();
        if (!rd.ok || n < 0) { err = "truncated CSI index"; return false; }
        ix.n_bins = (uint32_t)(((1ull << (3 * ix.n_lvls + 3)) - 1) / 7);
        ix.refs.assign((size_t)n, IdxRef());
        if (!read_core(rd, ix, true)) { err = "malformed CSI index"; return false; }
        return true;
    }
    err = "not a BAI or CSI index";
    return false;
}

// ---- region strings ------------------------------------------------------------------------------------
namespace {

// hts_parse_decimal with HTS_PARSE_THOUSANDS_SEP (hts.c: digits, ',' separators, fraction, e/E exponent, k/M/G)
long long parse_decimal(const char *str, const char **strend) {
    unsigned long long n = 0;
    int digits = 0, decimals = 0, e = 0;
    char sign = '+', esign = '+';
    const char *s = str;
    while (*s == ' ' || (*s >= '\t' && *s <= '\r')) ++s;
    if (*s == '+' || *s == '-') sign = *s++;
    while (*s) {
        if (*s >= '0' && *s <= '9') { ++digits; n = n * 10 + (unsigned)(*s++ - '0'); }
        else if (*s == ',') ++s;
        else break;
    }
    if (*s == '.') {
        ++s;
        while (*s >= '0' && *s <= '9') { ++decimals; ++digits; n = n * 10 + (unsigned)(*s++ - '0'); }
    }
    switch (*s) {
    case 'e': case 'E':
        ++s;
        if (*s == '+' || *s == '-') esign = *s++;
        while (*s >= '0' && *s <= '9') e = e * 10 + (*s++ - '0');
        if (esign == '-') e = -e;
        break;
    case 'k': case 'K': e += 3; ++s; break;
    case 'm': case 'M': e += 6; ++s; break;
    case 'g': case 'G': e += 9; ++s; break;
    default: break;
    }
    e -= decimals;
    while (e > 0) { n *= 10; --e; }
    while (e < 0) { n /= 10; ++e; }
    *strend = digits > 0 ? s : str;
    return sign == '+' ? (long long)n : -(long long)n;
}

int name2tid(const std::vector<std::string> &names, const std::string &s) {
    for (size_t i = 0; i < names.size(); ++i)
        if (names[i] == s) return (int)i;
    return -1;
}

// hts_parse_region (no HTS_PARSE_LIST): true = parsed; false = unknown name / malformed (the caller skips it)
bool parse_region(const std::string &str, const std::vector<std::string> &names, int &tid, int64_t &beg, int64_t &end) {
    const char *s = str.c_str();
    size_t s_len = str.size();
    const char *colon = nullptr;
    int quoted = 0;
    if (*s == '{') {
        const char *close = (const char *)memchr(s, '}', s_len);
        if (!close) return false;
        ++s;
        --s_len;
        if (close[1] == ':') colon = close + 1;
        quoted = 1;
    } else {
        for (size_t i = s_len; i-- > 0;)
            if (s[i] == ':') { colon = s + i; break; }
    }
    if (!colon) {
        beg = 0;
        end = INT64_MAX;
        tid = name2tid(names, std::string(s, s_len - (size_t)quoted));
        return tid >= 0;
    }
    if (!quoted) {
        beg = 0;
        end = INT64_MAX;
        if ((tid = name2tid(names, std::string(s, s_len))) >= 0) {
            if (name2tid(names, std::string(s, (size_t)(colon - s))) >= 0) { tid = -1; return false; }   // ambiguous
            return true;
        }
    }
    tid = name2tid(names, std::string(s, (size_t)(colon - s - quoted)));
    if (tid < 0) return false;
    const char *hyphen;
    beg = parse_decimal(colon + 1, &hyphen) - 1;
    if (beg < 0) {
        if (beg != -1 && *hyphen == '-' && colon[1] != '\0') return false;      // coordinate 0
        if ((*hyphen >= '0' && *hyphen <= '9') || *hyphen == '\0' || *hyphen == ',') {
            end = beg == -1 ? INT64_MAX : -(beg + 1);                             // chr:-100 = chr:1-100
            beg = 0;
            return true;
        } else if (beg < -1) return false;
    }
    if (*hyphen == '\0') end = INT64_MAX;
    else if (*hyphen == '-') {
        end = parse_decimal(hyphen + 1, &hyphen);
        if (*hyphen != '\0' && *hyphen != ',') return false;
    } else return false;
    if (end == 0) end = INT64_MAX;
    return beg < end;
}

}  // namespace

int regions_to_intervals(const char *const *regions, int n_regions, const std::vector<std::string> &ref_names,
                         std::vector<RegIntv> &out, std::string &err) {
    out.clear();
    for (int i = 0; i < n_regions; ++i) {
        const std::string r = regions[i] ? regions[i] : "";
        if (r == "." || r == "*") { err = "the special regions '.' and '*' are not available in this build"; return -1; }
        int tid;
        int64_t beg, end;
        if (!parse_region(r, ref_names, tid, beg, end)) continue;   // unknown reference name: htslib warns and continues
        RegIntv v;
        v.tid = tid; v.beg = beg; v.end = end; v.pad = 0;
        out.push_back(v);
    }
    // reg_compact (region.c:84-122): sort by (beg, end) per tid, merge overlapping or adjacent intervals
    std::sort(out.begin(), out.end(), [](const RegIntv &a, const RegIntv &b) {
        if (a.tid != b.tid) return a.tid < b.tid;
        if (a.beg != b.beg) return a.beg < b.beg;
        return a.end < b.end;
    });
    size_t m = 0;
    for (size_t j = 0; j < out.size(); ++j) {
        if (m && out[m - 1].tid == out[j].tid && !(out[m - 1].end < out[j].beg)) {
            if (out[m - 1].end < out[j].end) out[m - 1].end = out[j].end;
        } else out[m++] = out[j];
    }
    out.resize(m);
    return (int)m;
}

// The chunks htslib would visit for [beg, end) on tid (hts_itr_query :3455-3590 / hts_itr_multi_bam :3639-3728 +
// add_to_interval :3216-3245): appended to `chunks` as (u, v) pairs, untrimmed order.
void query_chunks(const HtsIndex &ix, int tid, int64_t beg, int64_t end, std::vector<IdxChunk> &chunks) {
    if (tid < 0 || (size_t)tid >= ix.refs.size()) return;
    const IdxRef &ref = ix.refs[(size_t)tid];
    if (ref.bins.empty()) return;
    if (beg < 0) beg = 0;
    if (end < beg) return;
    const IdxBin *meta = ref.find(ix.n_bins + 1);
    const uint64_t unmapped = (meta && meta->chunks.size() >= 2) ? meta->chunks[1].v : 1;
    const int64_t maxpos = 1ll << (ix.min_shift + 3 * ix.n_lvls);
    if (beg >= maxpos) return;
    const uint64_t rel_off = (uint64_t)(beg >> ix.min_shift);
    // min_off: lowest-level bin containing beg, else the nearest extant bin to the left / above
    int64_t bin = bin_first(ix.n_lvls) + (int64_t)rel_off;
    const IdxBin *k = nullptr;
    do {
        if ((k = ref.find((uint32_t)bin))) break;
        const int64_t first = (bin_parent(bin) << 3) + 1;
        if (bin > first) --bin; else bin = bin_parent(bin);
    } while (bin);
    if (bin == 0) k = ref.find(0);
    uint64_t min_off = k ? k->loff : 0;
    if (!ref.lin.empty() && rel_off < ref.lin.size()) {
        if (min_off < ref.lin[rel_off]) min_off = ref.lin[rel_off];
        if (unmapped) {
            int64_t t = (int64_t)rel_off - 1;
            for (; t >= 0; --t)
                if (ref.lin[(size_t)t] < min_off) { min_off = ref.lin[(size_t)t]; break; }
            if (k && (min_off < k->loff || t < 0)) min_off = k->loff;
        }
    } else if (unmapped) {
        if (k) min_off = k->loff;
    }
    // max_off: first chunk of an extant bin to the right of end
    uint64_t max_off = UINT64_MAX;
    if (end <= maxpos) {
        bin = bin_first(ix.n_lvls) + ((end - 1) >> ix.min_shift) + 1;
        if (bin >= (int64_t)ix.n_bins) bin = 0;
        for (;;) {
            while (bin % 8 == 1) bin = bin_parent(bin);
            if (bin == 0) { max_off = UINT64_MAX; break; }
            const IdxBin *b = ref.find((uint32_t)bin);
            if (b && !b->chunks.empty()) { max_off = b->chunks[0].u; break; }
            ++bin;
        }
    }
    // bins overlapping the region, all levels (reg2bins :3142-3214); the meta bin lies above n_lvls and is skipped
    int64_t e = end;
    if (e >= maxpos) e = maxpos;
    if (beg >= e) return;
    const int64_t e1 = e - 1;
    for (const IdxBin &b : ref.bins) {
        const int level = bin_level(b.bin);
        if (level > ix.n_lvls) continue;
        const int sh = ix.min_shift + 3 * (ix.n_lvls - level);
        const int64_t first = bin_first(level);
        if ((int64_t)b.bin < first + (beg >> sh) || (int64_t)b.bin > first + (e1 >> sh)) continue;
        for (const IdxChunk &c : b.chunks)
            if (c.v > min_off && c.u < max_off) {
                IdxChunk o;
                o.u = std::max(min_off, c.u);
                o.v = std::min(max_off, c.v);
                chunks.push_back(o);
            }
    }
}

void plan_ranges(const HtsIndex &ix, const std::vector<RegIntv> &intervals, std::vector<IdxChunk> &ranges) {
    std::vector<IdxChunk> ch;
    for (const RegIntv &v : intervals) query_chunks(ix, v.tid, v.beg, v.end, ch);
    std::sort(ch.begin(), ch.end(), [](const IdxChunk &a, const IdxChunk &b) { return a.u != b.u ? a.u < b.u : a.v < b.v; });
    ranges.clear();
    constexpr uint64_t GAP_MERGE = 4ull << 20;   // compressed bytes: closer chunks are cheaper to scan through than to restart for
    for (const IdxChunk &c : ch) {
        if (c.v <= c.u) continue;
        if (!ranges.empty() && (c.u >> 16) <= (ranges.back().v >> 16) + GAP_MERGE) {
            if (ranges.back().v < c.v) ranges.back().v = c.v;
        } else ranges.push_back(c);
    }
}

bool contig_span(const HtsIndex &ix, int tid, uint64_t &u, uint64_t &v) {
    std::vector<IdxChunk> ch;
    query_chunks(ix, tid, 0, INT64_MAX, ch);
    if (ch.empty()) return false;
    u = UINT64_MAX;
    v = 0;
    for (auto &c : ch) { u = std::min(u, c.u); v = std::max(v, c.v); }
    return true;
}

}  // namespace dh
