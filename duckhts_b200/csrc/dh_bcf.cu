// dh_bcf.cu — host side of the read_bcf scan behind the C-ABI (include/duckhts_cuda.h, dh_bcf_*): BCF header ->
// dictionaries and column schema (rows B1, B8-B10 of SURVEY.md §8), and the per-batch decode issue.  The segment
// pipeline (H2D -> inflate -> boundary walk -> batches -> D2H) is the one of dh_bam.cu (dh_scan.h); only the two kernels
// and the arena layout differ.
//
// Reference: bcf_hdr_read / bcf_hdr_parse / bcf_hdr_register_hrec (htslib/vcf.c:1710+, 1410+, 831-1010) for the
// dictionaries; bcf_read_bind (src/bcf_reader.c:451-830) and vcf_validate_*_field / vcf_is_list_type
// (src/include/vcf_types.h:25-205) for the schema.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "bcf_decode.cuh"
#include "dh_scan.h"

using namespace dh;

namespace {

enum { VL_FIXED = 0, VL_VAR, VL_A, VL_G, VL_R, VL_OTHER };   // BCF_VL_* (htslib/vcf.h); P/LA/LR/LG/M fold into OTHER
enum { CL_FLT = 0, CL_INFO = 1, CL_FMT = 2 };

struct DictId {
    std::string name;
    bool present = false;          // registered (hdr->id[BCF_DT_ID][i].key != NULL)
    bool has[3] = {false, false, false};
    int type[3] = {HT_STR, HT_STR, HT_STR};
    int vl[3] = {VL_VAR, VL_VAR, VL_VAR};
};
struct Field {
    std::string name;
    int id, type, is_list;
};

}  // namespace

struct dh_bcf_state {
    std::vector<DictId> ids;
    std::vector<std::string> ctg;         // "" + !ctg_present = hole
    std::vector<uint8_t> ctg_present;
    std::vector<std::string> samples;
    std::vector<Field> info, fmt;
    int ver_major = 4, ver_minor = 2;
    bool has_vep = false;
    // device tables
    std::deque<std::string> dict;         // long names: cells point at these
    DevBuf d_ctg, d_ids, d_info_col, d_fmt_col, d_info_meta, d_fmt_meta, d_info_want, d_fmt_want;
    BcfTables tabs;
    uint32_t core_mask = (1u << BCF_CORE_COLS) - 1;
    std::vector<uint8_t> info_want, fmt_want;
    uint64_t bump_cap = 64ull << 20;
    // per batch slot: pointer tables (pinned, read by the kernel) and the host views handed to the caller
    PinBuf h_ptrs[MAX_SLOTS];
    std::vector<const void *> out_info[MAX_SLOTS], out_fmt[MAX_SLOTS];
    std::vector<const uint64_t *> out_info_v[MAX_SLOTS], out_fmt_v[MAX_SLOTS];
    std::vector<const void *> out_fmt_rec[MAX_SLOTS];   // compact layout: per-record {first child, values per sample} of the numeric LIST FORMAT columns
    dh_bcf_batch desc[MAX_SLOTS];
    uint64_t bump_off[MAX_SLOTS] = {};    // where the bump counter sits in the arena
};

struct dh_bcf {
    dh_bam *scan;   // the shared pipeline state (kind = 1)
};

namespace {

StrCell make_cell(dh_bcf_state *st, const std::string &s, bool present) {
    StrCell c;
    memset(&c, 0, sizeof c);
    if (!present) { c.len = 0xffffffffu; return c; }
    c.len = (uint32_t)s.size();
    if (s.size() <= 12) memcpy(c.w, s.data(), s.size());
    else {
        st->dict.push_back(s);
        const char *p = st->dict.back().data();
        memcpy(&c.w[0], p, 4);
        const uint64_t a = (uint64_t)(uintptr_t)p;
        c.w[1] = (uint32_t)a;
        c.w[2] = (uint32_t)(a >> 32);
    }
    return c;
}

// one "##key=<k=v,...>" line -> pairs (bcf_hdr_parse_line, vcf.c:654-829: quoted values may hold commas and \" escapes)
bool parse_struct_line(const std::string &v, std::vector<std::pair<std::string, std::string>> &kv) {
    size_t p = 1;   // after '<'
    while (p < v.size() && v[p] != '>') {
        while (p < v.size() && (v[p] == ' ' || v[p] == ',')) ++p;
        size_t ks = p;
        while (p < v.size() && v[p] != '=' && v[p] != ',' && v[p] != '>') ++p;
        if (p >= v.size() || v[p] != '=') { if (p < v.size() && v[p] == '>') break; if (p >= v.size()) return false; ++p; continue; }
        std::string key = v.substr(ks, p - ks);
        while (!key.empty() && key.back() == ' ') key.pop_back();
        ++p;
        std::string val;
        if (p < v.size() && v[p] == '"') {
            ++p;
            while (p < v.size() && v[p] != '"') {
                if (v[p] == '\\' && p + 1 < v.size()) { val.push_back(v[p]); val.push_back(v[p + 1]); p += 2; continue; }
                val.push_back(v[p++]);
            }
            if (p < v.size()) ++p;
        } else {
            int depth = 0;
            while (p < v.size() && !(depth == 0 && (v[p] == ',' || v[p] == '>'))) {
                if (v[p] == '<') ++depth;
                if (v[p] == '>') --depth;
                val.push_back(v[p++]);
            }
        }
        kv.emplace_back(key, val);
    }
    return true;
}

const char *const INFO_LIST[] = {"AD", "ADF", "ADR", "AC", "AF", "CIGAR", nullptr};
const char *const INFO_SCALAR[] = {"AA", "AN", "BQ", "DB", "DP", "END", "H2", "H3", "MQ", "MQ0", "NS", "SB", "SOMATIC", "VALIDATED", "1000G", nullptr};
const char *const FMT_LIST[] = {"AD", "ADF", "ADR", "EC", "GL", "GP", "PL", "PP", nullptr};
const char *const FMT_SCALAR[] = {"DP", "LEN", "FT", "GQ", "GT", "HQ", "MQ", "PQ", "PS", nullptr};
bool in_list(const char *const *l, const std::string &s) {
    for (; *l; ++l) if (s == *l) return true;
    return false;
}
// vcf_validate_*_field + vcf_is_list_type (vcf_types.h:143-205): the spec table overrides the header's Number
int field_is_list(bool is_fmt, const std::string &name, int vl) {
    if (in_list(is_fmt ? FMT_SCALAR : INFO_SCALAR, name)) return 0;
    if (in_list(is_fmt ? FMT_LIST : INFO_LIST, name)) return 1;
    return vl != VL_FIXED;
}

void register_id(dh_bcf_state *st, int cls, const std::vector<std::pair<std::string, std::string>> &kv) {
    std::string id;
    bool has_id = false;
    long idx = -1;
    int type = -1, vl = -1, num = -1;
    for (auto &e : kv) {
        if (e.first == "ID") { id = e.second; has_id = true; }
        else if (e.first == "IDX") {
            char *end = nullptr;
            idx = strtol(e.second.c_str(), &end, 10);
            if (*end || idx < 0) return;                       // "Error parsing the IDX tag, skipping"
        } else if (e.first == "Type") {
            type = e.second == "Integer" ? HT_INT : e.second == "Float" ? HT_REAL : e.second == "Flag" ? HT_FLAG : HT_STR;
        } else if (e.first == "Number") {
            const std::string &n = e.second;
            if (n == "A") vl = VL_A; else if (n == "R") vl = VL_R; else if (n == "G") vl = VL_G; else if (n == ".") vl = VL_VAR;
            else if (cls == CL_FMT && (n == "P" || n == "LA" || n == "LR" || n == "LG" || n == "M")) vl = VL_OTHER;
            else if (sscanf(n.c_str(), "%d", &num) == 1) vl = VL_FIXED;
        }
    }
    if (cls != CL_FLT) {
        if (type < 0) type = HT_STR;
        if (vl < 0) vl = VL_VAR;
        if (type == HT_FLAG && (vl != VL_FIXED || num != 0)) { vl = VL_FIXED; num = 0; }
    }
    if (!has_id) return;
    int at = -1;
    for (size_t i = 0; i < st->ids.size(); ++i)
        if (st->ids[i].present && st->ids[i].name == id) { at = (int)i; break; }
    if (at >= 0) {
        if (st->ids[at].has[cls]) return;                      // first definition wins
    } else {
        at = idx >= 0 ? (int)idx : (int)st->ids.size();
        if ((size_t)at >= st->ids.size()) st->ids.resize((size_t)at + 1);
        if (st->ids[at].present) return;                       // IDX clash
        st->ids[at].present = true;
        st->ids[at].name = id;
    }
    st->ids[at].has[cls] = true;
    st->ids[at].type[cls] = type < 0 ? HT_STR : type;
    st->ids[at].vl[cls] = vl < 0 ? VL_VAR : vl;
}

void register_contig(dh_bcf_state *st, const std::vector<std::pair<std::string, std::string>> &kv) {
    std::string id;
    bool has_id = false;
    long idx = -1;
    for (auto &e : kv) {
        if (e.first == "ID") { id = e.second; has_id = true; }
        else if (e.first == "IDX") { char *end = nullptr; idx = strtol(e.second.c_str(), &end, 10); if (*end || idx < 0) return; }
        else if (e.first == "length") { char *end = nullptr; long long l = strtoll(e.second.c_str(), &end, 10); if (end == e.second.c_str() || l < 0) return; }
    }
    if (!has_id) return;
    for (size_t i = 0; i < st->ctg.size(); ++i)
        if (st->ctg_present[i] && st->ctg[i] == id) return;
    size_t at = idx >= 0 ? (size_t)idx : st->ctg.size();
    if (at >= st->ctg.size()) { st->ctg.resize(at + 1); st->ctg_present.resize(at + 1, 0); }
    if (st->ctg_present[at]) return;
    st->ctg[at] = id;
    st->ctg_present[at] = 1;
}

// bcf_hdr_parse (vcf.c:1410-1490): PASS first, then every ## line, then the sample line
bool parse_header_text(dh_bcf_state *st, const std::string &text) {
    std::vector<std::pair<std::string, std::string>> kv;
    parse_struct_line("<ID=PASS,Description=\"All filters passed\">", kv);
    register_id(st, CL_FLT, kv);
    size_t p = 0;
    bool have_samples_line = false;
    while (p < text.size()) {
        size_t eol = text.find('\n', p);
        if (eol == std::string::npos) eol = text.size();
        std::string line = text.substr(p, eol - p);
        if (!line.empty() && line.back() == '\r') line.pop_back();
        p = eol + 1;
        if (line.size() >= 2 && line[0] == '#' && line[1] == '#') {
            const size_t eq = line.find('=');
            if (eq == std::string::npos) continue;
            const std::string key = line.substr(2, eq - 2), val = line.substr(eq + 1);
            if (key == "fileformat") { int a = 0, b = 0; if (sscanf(val.c_str(), "VCFv%d.%d", &a, &b) == 2) { st->ver_major = a; st->ver_minor = b; } continue; }
            if (val.empty() || val[0] != '<') continue;
            kv.clear();
            if (!parse_struct_line(val, kv)) continue;
            if (key == "contig") register_contig(st, kv);
            else if (key == "FILTER") register_id(st, CL_FLT, kv);
            else if (key == "INFO") register_id(st, CL_INFO, kv);
            else if (key == "FORMAT") register_id(st, CL_FMT, kv);
        } else if (!line.empty() && line[0] == '#') {
            // #CHROM POS ID REF ALT QUAL FILTER INFO [FORMAT sample...]   (bcf_hdr_parse_sample_line)
            have_samples_line = true;
            size_t f = 0;
            int col = 0;
            while (f <= line.size()) {
                size_t fe = line.find('\t', f);
                if (fe == std::string::npos) fe = line.size();
                if (col >= 9 && fe > f) st->samples.push_back(line.substr(f, fe - f));
                ++col;
                f = fe + 1;
            }
            break;
        }
    }
    return have_samples_line;
}

}  // namespace

namespace {

template <typename T>
int upload(DevBuf &d, const std::vector<T> &v) {
    int rc = d.reserve(std::max<size_t>(1, v.size()) * sizeof(T));
    if (rc != DH_OK) return rc;
    if (!v.empty()) DH_CUDA(cudaMemcpy(d.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return DH_OK;
}

int upload_tables(dh_bam *b) {
    dh_bcf_state *st = b->bcf;
    std::vector<StrCell> ctg, ids;
    for (size_t i = 0; i < st->ctg.size(); ++i) ctg.push_back(make_cell(st, st->ctg[i], st->ctg_present[i] != 0));
    for (auto &d : st->ids) ids.push_back(make_cell(st, d.name, d.present));
    std::vector<int16_t> info_col(st->ids.size(), -1), fmt_col(st->ids.size(), -1);
    std::vector<uint8_t> info_meta, fmt_meta;
    for (size_t c = 0; c < st->info.size(); ++c) {
        info_col[(size_t)st->info[c].id] = (int16_t)c;
        info_meta.push_back((uint8_t)(st->info[c].type | (st->info[c].is_list ? META_LIST : 0)));
    }
    for (size_t c = 0; c < st->fmt.size(); ++c) {
        fmt_col[(size_t)st->fmt[c].id] = (int16_t)c;
        const bool gt = st->fmt[c].name == "GT" && st->fmt[c].type == HT_STR;
        fmt_meta.push_back((uint8_t)(st->fmt[c].type | (st->fmt[c].is_list ? META_LIST : 0) | (gt ? META_GT : 0)));
    }
    int rc;
    if ((rc = upload(st->d_ctg, ctg)) != DH_OK || (rc = upload(st->d_ids, ids)) != DH_OK || (rc = upload(st->d_info_col, info_col)) != DH_OK ||
        (rc = upload(st->d_fmt_col, fmt_col)) != DH_OK || (rc = upload(st->d_info_meta, info_meta)) != DH_OK ||
        (rc = upload(st->d_fmt_meta, fmt_meta)) != DH_OK || (rc = upload(st->d_info_want, st->info_want)) != DH_OK ||
        (rc = upload(st->d_fmt_want, st->fmt_want)) != DH_OK)
        return rc;
    BcfTables &t = st->tabs;
    t.ctg_cells = (const StrCell *)st->d_ctg.p;
    t.id_cells = (const StrCell *)st->d_ids.p;
    t.info_col = (const int16_t *)st->d_info_col.p;
    t.fmt_col = (const int16_t *)st->d_fmt_col.p;
    t.info_meta = (const uint8_t *)st->d_info_meta.p;
    t.fmt_meta = (const uint8_t *)st->d_fmt_meta.p;
    t.info_want = (const uint8_t *)st->d_info_want.p;
    t.fmt_want = (const uint8_t *)st->d_fmt_want.p;
    t.n_ctg = (int32_t)st->ctg.size();
    t.n_ids = (int32_t)st->ids.size();
    t.n_info = (int32_t)st->info.size();
    t.n_fmt = (int32_t)st->fmt.size();
    t.n_samples = (uint32_t)st->samples.size();
    t.core_mask = st->core_mask;
    t.gt_id = -1;
    if (st->ver_major < 4 || (st->ver_major == 4 && st->ver_minor < 4))   // bcf_record_check: updatephasing only below VCFv4.4
        for (size_t i = 0; i < st->ids.size(); ++i)
            if (st->ids[i].present && st->ids[i].name == "GT") t.gt_id = (int32_t)i;
    t.any_fmt = 0;
    for (uint8_t w : st->fmt_want) t.any_fmt |= w;
    return DH_OK;
}

uint32_t info_width(const Field &f) { return f.is_list ? 16u : f.type == HT_FLAG ? 1u : f.type == HT_STR ? 16u : 4u; }
uint32_t fmt_width(const Field &f) { return (f.is_list || f.type == HT_STR) ? 16u : 4u; }
bool fmt_is_gt(const Field &f) { return f.name == "GT" && f.type == HT_STR; }
bool fmt_is_numlist(const Field &f) { return f.is_list && (f.type == HT_INT || f.type == HT_REAL); }
// compact wire layout: GT 4 bytes, numeric lists a 2-byte count, the rest as in the DuckDB layout
uint32_t fmt_width_compact(const Field &f) { return fmt_is_gt(f) ? 4u : fmt_is_numlist(f) ? 2u : fmt_width(f); }
const uint32_t CORE_WIDTH[BCF_CORE_COLS] = {16, 8, 16, 16, 16, 8, 16};

int issue_bcf_batch(dh_bam *b) {
    dh_bcf_state *st = b->bcf;
    const Batch &bb = b->batches[b->next_batch_i++];
    SegSlot &s = b->ss[b->cur_seg % NSEG];
    const int slot = scan_take_slot(b);
    if (slot < 0) { set_error("internal: no free batch slot"); return DH_EINVAL; }
    BatchSlot &o = b->bs[slot];
    const uint64_t n_rec = bb.rows, ns = st->samples.size(), rows = n_rec * ns;
    const size_t ni = st->info.size(), nf = st->fmt.size();
    auto al = [](uint64_t x) { return (x + A256) & ~A256; };
    uint64_t off = 256;   // [0, 8): bump counter
    uint64_t core_off[BCF_CORE_COLS], core_voff[BCF_CORE_COLS];
    std::vector<uint64_t> info_off(ni, UINT64_MAX), info_voff(ni, UINT64_MAX), fmt_off(nf, UINT64_MAX), fmt_voff(nf, UINT64_MAX);
    for (int c = 0; c < BCF_CORE_COLS; ++c) {
        core_off[c] = core_voff[c] = UINT64_MAX;
        if (!((st->core_mask >> c) & 1u)) continue;
        core_off[c] = off;
        off = al(off + CORE_WIDTH[c] * n_rec);
        if (c == BC_ID || c == BC_QUAL) { core_voff[c] = off; off = al(off + ((n_rec + 63) / 64) * 8); }
    }
    for (size_t c = 0; c < ni; ++c) {
        if (!st->info_want[c]) continue;
        info_off[c] = off;
        off = al(off + info_width(st->info[c]) * n_rec);
        if (st->info[c].type != HT_FLAG) { info_voff[c] = off; off = al(off + ((n_rec + 63) / 64) * 8); }
    }
    const uint64_t site_end = off;
    const bool compact = (b->mode & DH_BAM_COMPACT_OUTPUT) != 0;
    std::vector<uint64_t> fmt_roff(nf, UINT64_MAX);
    for (size_t c = 0; c < nf; ++c) {
        if (!st->fmt_want[c]) continue;
        fmt_off[c] = off;
        off = al(off + (compact ? fmt_width_compact(st->fmt[c]) : fmt_width(st->fmt[c])) * rows);
        fmt_voff[c] = off;
        off = al(off + ((rows + 63) / 64) * 8 + 8);
        if (compact && fmt_is_numlist(st->fmt[c])) { fmt_roff[c] = off; off = al(off + 8 * n_rec); }
    }
    const uint64_t heap_off = off, fixed = (bb.heap + A256) & ~A256, main_bytes = heap_off + fixed, total = main_bytes + st->bump_cap;
    int rc;
    if ((rc = o.arena.reserve(total + 256)) != DH_OK || (rc = o.h_arena.reserve(total + 256)) != DH_OK) return rc;
    const size_t tbl = (2 * (ni + nf) + nf) * sizeof(void *) + 64;
    if ((rc = st->h_ptrs[slot].reserve(tbl)) != DH_OK) return rc;
    uint8_t *dev = (uint8_t *)o.arena.p, *host = (uint8_t *)o.h_arena.p;
    void **tp = (void **)st->h_ptrs[slot].p;
    void **d_info = tp, **d_info_v = tp + ni, **d_fmt = tp + 2 * ni, **d_fmt_v = tp + 2 * ni + nf, **d_fmt_rec = tp + 2 * ni + 2 * nf;
    st->out_fmt_rec[slot].assign(nf, nullptr);
    st->out_info[slot].assign(ni, nullptr);
    st->out_info_v[slot].assign(ni, nullptr);
    st->out_fmt[slot].assign(nf, nullptr);
    st->out_fmt_v[slot].assign(nf, nullptr);
    cudaStream_t sk = b->ctx->stream, sd = b->ctx->copy_streams[1];
    DH_CUDA(cudaMemsetAsync(dev, 0, site_end, sk));
    BcfDecodeParams dp;
    memset(&dp, 0, sizeof dp);
    dh_bcf_batch &d = st->desc[slot];
    memset(&d, 0, sizeof d);
    for (int c = 0; c < BCF_CORE_COLS; ++c) {
        if (core_off[c] == UINT64_MAX) continue;
        dp.core[c] = dev + core_off[c];
        d.core[c] = host + core_off[c];
        if (core_voff[c] != UINT64_MAX) { dp.core_valid[c] = (unsigned long long *)(dev + core_voff[c]); d.core_validity[c] = (const uint64_t *)(host + core_voff[c]); }
    }
    for (size_t c = 0; c < ni; ++c) {
        d_info[c] = d_info_v[c] = nullptr;
        if (info_off[c] == UINT64_MAX) continue;
        d_info[c] = dev + info_off[c];
        st->out_info[slot][c] = host + info_off[c];
        if (info_voff[c] != UINT64_MAX) { d_info_v[c] = dev + info_voff[c]; st->out_info_v[slot][c] = (const uint64_t *)(host + info_voff[c]); }
    }
    for (size_t c = 0; c < nf; ++c) {
        d_fmt[c] = d_fmt_v[c] = d_fmt_rec[c] = nullptr;
        if (fmt_off[c] == UINT64_MAX) continue;
        if (fmt_roff[c] != UINT64_MAX) { d_fmt_rec[c] = dev + fmt_roff[c]; st->out_fmt_rec[slot][c] = host + fmt_roff[c]; }
        d_fmt[c] = dev + fmt_off[c];
        d_fmt_v[c] = dev + fmt_voff[c];
        st->out_fmt[slot][c] = host + fmt_off[c];
        st->out_fmt_v[slot][c] = (const uint64_t *)(host + fmt_voff[c]);
        DH_CUDA(cudaMemsetAsync(dev + fmt_voff[c], 0, ((rows + 63) / 64) * 8 + 8, sk));
    }
    dp.u = (const uint8_t *)s.out.p + b->carry_cap;
    dp.walk = (const BlockWalk *)s.h_walk.p;
    dp.blk_start = (const int64_t *)s.h_blk_start.p;
    dp.rec = (const uint16_t *)s.rec.p;
    dp.rec_bias = s.rec_bias;
    dp.row_base = (const uint32_t *)s.h_row_base.p;
    dp.heap_base = (const uint64_t *)s.h_heap_base.p;
    dp.b_lo = bb.b_lo;
    dp.b_hi = bb.b_hi;
    dp.info = (uint8_t *const *)d_info;
    dp.info_valid = (unsigned long long *const *)d_info_v;
    dp.fmt = (uint8_t *const *)d_fmt;
    dp.fmt_valid = (unsigned long long *const *)d_fmt_v;
    dp.fmt_rec = (uint8_t *const *)d_fmt_rec;
    dp.compact = compact ? 1u : 0u;
    dp.heap = dev + heap_off;
    dp.heap_host = (uint64_t)(uintptr_t)(host + heap_off);
    dp.fixed_bytes = fixed;
    dp.bump_cap = st->bump_cap;
    dp.bump_used = (unsigned long long *)dev;
    DH_CUDA(cudaEventRecord(o.ev_k0, sk));
    DH_CUDA(launch_bcf_decode(dp, st->tabs, sk));
    b->ctx->launches++;
    DH_CUDA(cudaEventRecord(o.ev_k1, sk));
    DH_CUDA(cudaStreamWaitEvent(sd, o.ev_k1, 0));
    DH_CUDA(cudaEventRecord(o.ev_c0, sd));
    const bool keep = (b->mode & DH_BAM_DEVICE_OUTPUT) != 0;
    if (!keep) DH_CUDA(cudaMemcpyAsync(host, dev, main_bytes, cudaMemcpyDeviceToHost, sd));
    else DH_CUDA(cudaMemcpyAsync(host, dev, 256, cudaMemcpyDeviceToHost, sd));   // the string-region counter is always checked
    DH_CUDA(cudaEventRecord(o.ev_done, sd));
    o.timed = true;
    d.n_records = n_rec;
    d.n_samples = ns;
    d.rec0 = bb.row0;
    d.info = st->out_info[slot].data();
    d.info_validity = st->out_info_v[slot].data();
    d.format = st->out_fmt[slot].data();
    d.format_validity = st->out_fmt_v[slot].data();
    d.format_rec = st->out_fmt_rec[slot].data();
    d.layout = compact ? DH_BAM_COMPACT_OUTPUT : 0u;
    d.heap = host + heap_off;
    d.heap_bytes = fixed;
    d.bytes = keep ? 0 : main_bytes;
    d.slot = slot;
    st->bump_off[slot] = main_bytes;
    o.on_device = keep;
    o.state = SLOT_INFLIGHT;
    b->st.d2h_bytes += d.bytes;
    b->inflight.push_back(slot);
    b->issued++;
    return DH_OK;
}

}  // namespace

namespace dh {
cudaError_t bcf_launch_walk(dh_bam *b, const WalkParams &wp, cudaStream_t stream) { return launch_bcf_walk(wp, b->bcf->tabs, stream); }
void bcf_state_free(dh_bcf_state *st) {
    if (!st) return;
    st->d_ctg.release(); st->d_ids.release(); st->d_info_col.release(); st->d_fmt_col.release(); st->d_info_meta.release();
    st->d_fmt_meta.release(); st->d_info_want.release(); st->d_fmt_want.release();
    for (auto &p : st->h_ptrs) p.release();
    delete st;
}
}  // namespace dh

extern "C" {

dh_bcf *dh_bcf_open(dh_ctx *ctx, const uint8_t *data, uint64_t size) {
    if (!ctx || !data) { set_error("dh_bcf_open: bad arguments"); return nullptr; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) { set_error("cudaSetDevice failed"); return nullptr; }
    dh_bam *b = new dh_bam();
    b->kind = 1;
    b->ctx = ctx;
    b->data = data;
    b->size = size;
    b->batch_rows = 2048;           // records per batch (x n_samples FORMAT cells each)
    b->carry_cap = 64ull << 20;     // a record straddling two segments: sites x samples x fields can be megabytes
    memset(&b->st, 0, sizeof b->st);
    dh_bcf_state *st = new dh_bcf_state();
    b->bcf = st;
    auto fail = [&](const char *msg) -> dh_bcf * { if (msg) set_error("%s", msg); dh_bam_close(b); return nullptr; };
    uint64_t bad = 0;
    if (walk_blocks(data, size, 0, b->bt, &bad) != DH_OK) return fail("not a BGZF file (read_bcf serves BCF; VCF text is not available in this build)");
    if (b->bt.blocks.empty()) return fail("empty file");
    std::vector<uint8_t> u;
    if (scan_inflate_head(b, 9, u) != DH_OK) return fail(nullptr);
    if (u.size() < 9 || memcmp(u.data(), "BCF\2", 4) != 0 || u[4] > 2) return fail("invalid BCF2 magic string (only BCF is served by this build)");
    const uint32_t l_text = (uint32_t)u[5] | ((uint32_t)u[6] << 8) | ((uint32_t)u[7] << 16) | ((uint32_t)u[8] << 24);
    if (scan_inflate_head(b, 9ull + l_text, u) != DH_OK) return fail(nullptr);
    if (u.size() < 9ull + l_text) return fail("truncated BCF header");
    b->text.assign((const char *)u.data() + 9, l_text);
    const std::string ctext(b->text.c_str());   // htslib parses the header as a C string
    if (!parse_header_text(st, ctext)) return fail("could not parse the BCF header (no #CHROM line)");
    // first record: inflated offset 9 + l_text
    const uint64_t p = 9ull + l_text;
    const size_t nb = b->bt.blocks.size();
    size_t blk = std::upper_bound(b->bt.uoff.begin(), b->bt.uoff.end(), p) - b->bt.uoff.begin() - 1;
    if (blk >= nb) { blk = nb; b->first_off = 0; } else b->first_off = (uint32_t)(p - b->bt.uoff[blk]);
    b->first_blk = blk;
    // schema: INFO_* then FORMAT_* columns in dictionary order (bcf_reader.c:612-713)
    for (size_t i = 0; i < st->ids.size(); ++i) {
        const DictId &d = st->ids[i];
        if (!d.present) continue;
        if (d.has[CL_INFO]) {
            st->info.push_back(Field{d.name, (int)i, d.type[CL_INFO], field_is_list(false, d.name, d.vl[CL_INFO])});
            if (d.name == "CSQ" || d.name == "BCSQ" || d.name == "ANN") st->has_vep = true;
        }
        if (d.has[CL_FMT] && !st->samples.empty()) st->fmt.push_back(Field{d.name, (int)i, d.type[CL_FMT], field_is_list(true, d.name, d.vl[CL_FMT])});
    }
    if (st->info.size() > 1024 || st->fmt.size() > 256) return fail("more INFO (1024) or FORMAT (256) fields than this build serves");
    st->info_want.assign(st->info.size(), 1);
    st->fmt_want.assign(st->fmt.size(), 1);
    if (scan_ensure_events(b) != DH_OK || upload_tables(b) != DH_OK) return fail(nullptr);
    dh_bcf *h = new dh_bcf();
    h->scan = b;
    return h;
}

void dh_bcf_close(dh_bcf *h) {
    if (!h) return;
    dh_bam_close(h->scan);
    delete h;
}

int32_t dh_bcf_n_samples(const dh_bcf *h) { return h ? (int32_t)h->scan->bcf->samples.size() : 0; }
const char *dh_bcf_sample_name(const dh_bcf *h, int32_t i) {
    return (h && i >= 0 && (size_t)i < h->scan->bcf->samples.size()) ? h->scan->bcf->samples[(size_t)i].c_str() : nullptr;
}
int32_t dh_bcf_n_contigs(const dh_bcf *h) { return h ? (int32_t)h->scan->bcf->ctg.size() : 0; }
const char *dh_bcf_contig_name(const dh_bcf *h, int32_t rid) {
    if (!h || rid < 0 || (size_t)rid >= h->scan->bcf->ctg.size() || !h->scan->bcf->ctg_present[(size_t)rid]) return nullptr;
    return h->scan->bcf->ctg[(size_t)rid].c_str();
}
int32_t dh_bcf_n_info(const dh_bcf *h) { return h ? (int32_t)h->scan->bcf->info.size() : 0; }
int32_t dh_bcf_n_format(const dh_bcf *h) { return h ? (int32_t)h->scan->bcf->fmt.size() : 0; }
static int field_out(const std::vector<Field> &v, int32_t i, dh_bcf_field *out) {
    if (i < 0 || (size_t)i >= v.size() || !out) { set_error("field index out of range"); return DH_EINVAL; }
    out->name = v[(size_t)i].name.c_str();
    out->type = v[(size_t)i].type;
    out->is_list = v[(size_t)i].is_list;
    return DH_OK;
}
int dh_bcf_info_field(const dh_bcf *h, int32_t i, dh_bcf_field *out) { return h ? field_out(h->scan->bcf->info, i, out) : DH_EINVAL; }
int dh_bcf_format_field(const dh_bcf *h, int32_t i, dh_bcf_field *out) { return h ? field_out(h->scan->bcf->fmt, i, out) : DH_EINVAL; }
int dh_bcf_has_vep(const dh_bcf *h) { return h && h->scan->bcf->has_vep; }
const char *dh_bcf_header_text(const dh_bcf *h, uint32_t *len) {
    if (!h) return nullptr;
    if (len) *len = (uint32_t)h->scan->text.size();
    return h->scan->text.c_str();
}

int dh_bcf_set_projection(dh_bcf *h, uint32_t core_mask, const uint8_t *info_want, const uint8_t *format_want) {
    if (!h) { set_error("dh_bcf_set_projection: null handle"); return DH_EINVAL; }
    dh_bam *b = h->scan;
    dh_bcf_state *st = b->bcf;
    if (b->scanning && !b->inflight.empty()) { set_error("dh_bcf_set_projection: scan in progress"); return DH_EINVAL; }
    st->core_mask = core_mask & ((1u << BCF_CORE_COLS) - 1);
    for (size_t i = 0; i < st->info.size(); ++i) st->info_want[i] = info_want ? (info_want[i] != 0) : 1;
    for (size_t i = 0; i < st->fmt.size(); ++i) st->fmt_want[i] = format_want ? (format_want[i] != 0) : 1;
    DH_CUDA(cudaSetDevice(b->ctx->device));
    return upload_tables(b);
}

int dh_bcf_configure(dh_bcf *h, uint64_t seg_comp_bytes, uint32_t batch_records, uint64_t string_bytes) {
    if (!h) return DH_EINVAL;
    if (string_bytes) h->scan->bcf->bump_cap = (string_bytes + 255) & ~255ull;
    return dh_bam_configure(h->scan, seg_comp_bytes, batch_records, 0);
}
int dh_bcf_set_mode(dh_bcf *h, uint32_t flags) { return h ? dh_bam_set_mode(h->scan, flags) : DH_EINVAL; }

int dh_bcf_scan_begin(dh_bcf *h) {
    if (!h) { set_error("dh_bcf_scan_begin: null handle"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(h->scan->mu);
    h->scan->filt_mode = 0;
    h->scan->pending.clear();
    return scan_begin_range(h->scan, 0, UINT64_MAX, true);
}

int dh_bcf_scan_region(dh_bcf *h, const dh_index *idx, const char *region) {
    if (!h || !idx || !region) { set_error("dh_bcf_scan_region: bad arguments"); return DH_EINVAL; }
    dh_bam *b = h->scan;
    std::lock_guard<std::mutex> lk(b->mu);
    std::string err;
    const char *one[1] = {region};
    const int n = regions_to_intervals(one, 1, b->bcf->ctg, b->filt, err);
    if (n < 0) { set_error("%s", err.c_str()); return DH_EINVAL; }
    if (n == 0) { set_error("region '%s' names no contig of this file", region); return DH_ENOREGION; }
    return scan_start_filtered(b, idx, 2);
}

static int bcf_acquire_locked(dh_bam *b, dh_bcf_batch *out) {
    dh_bcf_state *st = b->bcf;
    DH_CUDA(cudaSetDevice(b->ctx->device));
    // refill the free slots; `exhausted` = the current range has nothing left to issue (else the slots ran out first)
    auto top_up = [&](bool &exhausted) -> int {
        exhausted = false;
        while (scan_free_slots(b) > 0) {
            int w = scan_have_work(b);
            if (w < 0) return w;
            if (w == 0) { exhausted = true; break; }
            int rc = issue_bcf_batch(b);
            if (rc != DH_OK) return rc;
        }
        return DH_OK;
    };
    bool exhausted = false;
    int rc = top_up(exhausted);
    if (rc != DH_OK) { b->scanning = false; return rc; }
    while (b->inflight.empty() && exhausted && !b->pending.empty() && b->final_status == 0) {
        // this range is exhausted: go on with the next one of the query (rows keep counting)
        const IdxChunk c = b->pending.front();
        b->pending.pop_front();
        if ((rc = scan_begin_range(b, c.u, c.v, false)) != DH_OK || (rc = top_up(exhausted)) != DH_OK) { b->scanning = false; return rc; }
    }
    if (b->inflight.empty()) {
        b->st.rows = b->rows_total;
        b->st.status = b->final_status;
        if (!exhausted) {      // every slot is held by the caller and batches remain: not the end of the scan
            set_error("every batch slot is held by the caller: release one or raise the number of slots");
            return DH_EBUSY;
        }
        return 0;
    }
    const int slot = b->inflight.front();
    BatchSlot &o = b->bs[slot];
    b->inflight.pop_front();
    DH_CUDA(cudaEventSynchronize(o.ev_done));
    o.state = SLOT_HELD;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, o.ev_k0, o.ev_k1) == cudaSuccess) b->st.decode_ms += ms;
    // strings longer than 12 bytes / string-list cells live in a region whose use is only known now
    const uint64_t used = *(const uint64_t *)o.h_arena.p;
    if (used > st->bump_cap) {
        b->scanning = false;
        set_error("a batch needs %llu bytes for long strings but %llu are reserved: raise string_bytes (dh_bcf_configure)",
                  (unsigned long long)used, (unsigned long long)st->bump_cap);
        return DH_ENOMEM;
    }
    if (used && !(b->mode & DH_BAM_DEVICE_OUTPUT)) {
        cudaStream_t sd = b->ctx->copy_streams[1];
        DH_CUDA(cudaMemcpyAsync((uint8_t *)o.h_arena.p + st->bump_off[slot], (const uint8_t *)o.arena.p + st->bump_off[slot], used, cudaMemcpyDeviceToHost, sd));
        DH_CUDA(cudaStreamSynchronize(sd));
        b->st.d2h_bytes += used;
        st->desc[slot].bytes += used;
    }
    st->desc[slot].heap_bytes += used;
    rc = top_up(exhausted);
    if (rc != DH_OK) { b->scanning = false; return rc; }
    *out = st->desc[slot];
    if (b->inflight.empty() && b->chain_done) out->status = b->final_status;
    b->returned++;
    b->st.rows = b->rows_total;
    b->st.status = b->final_status;
    return 1;
}

int dh_bcf_acquire_batch(dh_bcf *h, dh_bcf_batch *out) {
    if (!h || !out) { set_error("dh_bcf_acquire_batch: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(h->scan->mu);
    if (!h->scan->scanning) { set_error("dh_bcf_acquire_batch: no scan in progress"); return DH_EINVAL; }
    return bcf_acquire_locked(h->scan, out);
}

int dh_bcf_release_batch(dh_bcf *h, int32_t slot) {
    if (!h || slot < 0 || slot >= h->scan->n_slots) { set_error("dh_bcf_release_batch: bad slot"); return DH_EINVAL; }
    dh_bam *b = h->scan;
    std::lock_guard<std::mutex> lk(b->mu);
    if (b->bs[slot].state != SLOT_HELD) { set_error("dh_bcf_release_batch: slot %d is not held", slot); return DH_EINVAL; }
    scan_release_slot(b, slot);
    return DH_OK;
}

int dh_bcf_next_batch(dh_bcf *h, dh_bcf_batch *out) {
    if (!h || !out) { set_error("dh_bcf_next_batch: bad arguments"); return DH_EINVAL; }
    dh_bam *b = h->scan;
    std::lock_guard<std::mutex> lk(b->mu);
    if (!b->scanning) { set_error("dh_bcf_next_batch: no scan in progress"); return DH_EINVAL; }
    if (b->auto_slot >= 0 && b->bs[b->auto_slot].state == SLOT_HELD) scan_release_slot(b, b->auto_slot);
    b->auto_slot = -1;
    const int rc = bcf_acquire_locked(b, out);
    if (rc == 1) b->auto_slot = out->slot;
    return rc;
}

int dh_bcf_set_slots(dh_bcf *h, int n_slots) { return h ? dh_bam_set_slots(h->scan, n_slots) : DH_EINVAL; }

int dh_bcf_scan_stats_get(const dh_bcf *h, dh_bam_scan_stats *out) { return h ? dh_bam_scan_stats_get(h->scan, out) : DH_EINVAL; }

}  // extern "C"
