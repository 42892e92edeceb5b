// dh_bam.cu — host pipeline of the read_bam scan behind the C-ABI (include/duckhts_cuda.h, dh_bam_*).
//
// Per segment of BGZF blocks:   H2D (pinned, copy stream)  ->  bgzf_inflate_kernel  ->  bam_walk_kernel
//   -> host validates the boundary chain (entry[i] == exit[i-1]) and cuts the segment into row batches
// per batch:                    bam_decode_kernel -> one D2H copy of the batch arena (columns + validity + heap)
// Four streams: H2D copies, inflate kernels, walk + decode kernels, D2H copies.  Three segments (the one being
// decoded and two being uploaded/inflated ahead) and up to five batches are in flight, so the D2H copy engine --
// the binding resource when every column returns to the host -- never waits for a kernel.  Everything the host
// does per segment is O(#BGZF blocks) bookkeeping; no record or byte of payload is touched on the CPU.
#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <deque>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include <ctype.h>

#include "dh_scan.h"
#include "inflate_core.cuh"

using namespace dh;

namespace {

uint32_t col_width(const dh_bam *b, int c) {
    switch (c) {
    case COL_FLAG: return 2;
    case COL_MAPQ: return 4;
    case COL_POS: case COL_PNEXT: case COL_TLEN: return 8;
    default: break;
    }
    if (c < BAM_CORE_COLS) return 16;
    char t = (char)b->tags[c - BAM_CORE_COLS].type;
    return (t == 'i' || t == 'f' || t == 'c') ? 8 : 16;
}
bool col_nullable(int c) { return c == COL_RG || c == COL_SAMPLE || c >= BAM_CORE_COLS; }
// compact wire layout (DH_BAM_COMPACT_OUTPUT, include/duckhts_cuda.h)
uint32_t col_width_compact(const dh_bam *b, int c) {
    switch (c) {
    case COL_FLAG: return 2;
    case COL_MAPQ: return 1;
    default: break;
    }
    if (c < BAM_CORE_COLS) return 4;   // POS, PNEXT, TLEN int32; RNAME, RNEXT, SAMPLE_ID codes; string lengths
    const char t = (char)b->tags[c - BAM_CORE_COLS].type;
    return (t == 'i' || t == 'f' || t == 'c') ? 8 : t == 'B' ? 16 : t == 'A' ? 1 : 4;
}
bool is_compact(const dh_bam *b) { return (b->mode & DH_BAM_COMPACT_OUTPUT) != 0; }

StrCell make_cell(dh_bam *b, const std::string &s) {
    StrCell c;
    memset(&c, 0, sizeof c);
    c.len = (uint32_t)s.size();
    if (s.size() <= 12) memcpy(c.w, s.data(), s.size());
    else {
        b->dict.push_back(s);
        const char *p = b->dict.back().data();
        memcpy(&c.w[0], p, 4);
        uint64_t a = (uint64_t)(uintptr_t)p;
        c.w[1] = (uint32_t)a;
        c.w[2] = (uint32_t)(a >> 32);
    }
    return c;
}

size_t block_of(const dh_bam *b, uint64_t coffset) {
    size_t lo = 0, hi = b->bt.blocks.size();
    while (lo < hi) {
        size_t mid = (lo + hi) / 2;
        if (b->bt.blocks[mid].coffset < coffset) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// ---- header (bam_hdr_read, htslib/sam.c:229-342) on the host, from GPU-inflated bytes ----
int parse_header(dh_bam *b) {
    const size_t nb = b->bt.blocks.size();
    size_t take = std::min<size_t>(nb, 4);
    std::vector<uint8_t> u;
    for (;;) {
        uint64_t cend = take ? b->bt.blocks[take - 1].coffset + b->bt.blocks[take - 1].csize : 0;
        u.resize((size_t)b->bt.uoff[take] + 8);
        uint64_t n = 0;
        int rc = dh_bgzf_inflate_host(b->ctx, b->data, cend, u.data(), u.size(), &n, nullptr);
        if (rc != DH_OK) return rc;
        const uint64_t have = n;
        auto need_more = [&](uint64_t want) { return want > have && take < nb; };
        auto le32 = [&](uint64_t p) { return (uint32_t)u[p] | ((uint32_t)u[p + 1] << 8) | ((uint32_t)u[p + 2] << 16) | ((uint32_t)u[p + 3] << 24); };
        bool again = false, bad = false;
        uint64_t p = 0;
        do {
            if (need_more(12)) { again = true; break; }
            if (have < 12 || memcmp(u.data(), "BAM\1", 4)) { bad = true; break; }
            uint32_t l_text = le32(4);
            p = 8;
            if (need_more(p + l_text + 4)) { again = true; break; }
            if (p + l_text + 4 > have) { bad = true; break; }
            b->text.assign((const char *)u.data() + p, l_text);
            p += l_text;
            int32_t n_ref = (int32_t)le32(p);
            p += 4;
            if (n_ref < 0) { bad = true; break; }
            b->ref_names.clear();
            b->ref_lens.clear();
            for (int32_t i = 0; i < n_ref && !again && !bad; ++i) {
                if (need_more(p + 4)) { again = true; break; }
                if (p + 4 > have) { bad = true; break; }
                int32_t l = (int32_t)le32(p);
                p += 4;
                if (l <= 0) { bad = true; break; }
                if (need_more(p + (uint64_t)l + 4)) { again = true; break; }
                if (p + (uint64_t)l + 4 > have) { bad = true; break; }
                std::string name((const char *)u.data() + p, (size_t)l);
                name = std::string(name.c_str());  // C-string semantics (tid2name), incl. the missing-NUL fix-up
                b->ref_names.push_back(name);
                p += (uint64_t)l;
                b->ref_lens.push_back(le32(p));
                p += 4;
            }
        } while (0);
        if (bad) { set_error("invalid BAM binary header"); return DH_EFORMAT; }
        if (again) { take = std::min(nb, take * 4); continue; }
        // first alignment record: inflated offset p -> (block, in-block offset); skip to the next non-empty block
        size_t blk = std::upper_bound(b->bt.uoff.begin(), b->bt.uoff.end(), p) - b->bt.uoff.begin() - 1;
        if (blk >= nb) { blk = nb; b->first_off = 0; }
        else b->first_off = (uint32_t)(p - b->bt.uoff[blk]);
        b->first_blk = blk;
        return DH_OK;
    }
}

// @RG ID -> SM dictionary in header order (first line with an ID wins in the kernel's lookup)
void build_dictionaries(dh_bam *b) {
    b->ref_cells.clear();
    for (auto &n : b->ref_names) b->ref_cells.push_back(make_cell(b, n));
    b->ref_cells.push_back(make_cell(b, "*"));
    b->rg.clear();
    b->rg_ids.clear();
    b->rg_samples.clear();
    const std::string &t = b->text;
    size_t p = 0;
    while (p < t.size()) {
        size_t eol = t.find('\n', p);
        if (eol == std::string::npos) eol = t.size();
        if (eol - p >= 4 && !t.compare(p, 4, "@RG\t")) {
            std::string id, sm;
            bool has_id = false, has_sm = false;
            size_t f = p + 4;
            while (f < eol) {
                size_t fe = t.find('\t', f);
                if (fe == std::string::npos || fe > eol) fe = eol;
                if (fe - f >= 3 && t[f + 2] == ':') {
                    if (t[f] == 'I' && t[f + 1] == 'D' && !has_id) { id = t.substr(f + 3, fe - f - 3); has_id = true; }
                    if (t[f] == 'S' && t[f + 1] == 'M' && !has_sm) { sm = t.substr(f + 3, fe - f - 3); has_sm = true; }
                }
                f = fe + 1;
            }
            if (has_id) {
                RgEntry e;
                memset(&e, 0, sizeof e);
                e.id_off = (uint32_t)b->rg_ids.size();
                e.id_len = (uint32_t)id.size();
                b->rg_ids += id;
                sm = std::string(sm.c_str());
                e.has_sample = has_sm && !sm.empty();
                if (e.has_sample) e.sample = make_cell(b, sm);
                b->rg.push_back(e);
                b->rg_samples.push_back(e.has_sample ? sm : std::string());
            }
        }
        p = eol + 1;
    }
}

// bit c of col_mask = column c projected; tag columns past bit 63 (more than 51 tag requests) are always materialised
static inline bool col_on(const dh_bam *b, int c) { return c >= 64 || ((b->col_mask >> c) & 1ull); }

int upload_tables(dh_bam *b) {
    int rc;
    if ((rc = b->d_ref_cells.reserve(b->ref_cells.size() * sizeof(StrCell))) != DH_OK) return rc;
    DH_CUDA(cudaMemcpy(b->d_ref_cells.p, b->ref_cells.data(), b->ref_cells.size() * sizeof(StrCell), cudaMemcpyHostToDevice));
    if ((rc = b->d_rg.reserve(std::max<size_t>(1, b->rg.size()) * sizeof(RgEntry))) != DH_OK) return rc;
    if (!b->rg.empty()) DH_CUDA(cudaMemcpy(b->d_rg.p, b->rg.data(), b->rg.size() * sizeof(RgEntry), cudaMemcpyHostToDevice));
    if ((rc = b->d_rg_ids.reserve(std::max<size_t>(1, b->rg_ids.size()))) != DH_OK) return rc;
    if (!b->rg_ids.empty()) DH_CUDA(cudaMemcpy(b->d_rg_ids.p, b->rg_ids.data(), b->rg_ids.size(), cudaMemcpyHostToDevice));
    uint32_t any_tag = 0;
    for (size_t k = 0; k < b->tags.size(); ++k) any_tag |= (b->tags[k].want = col_on(b, BAM_CORE_COLS + (int)k));
    if ((rc = b->d_tags.reserve(std::max<size_t>(1, b->tags.size()) * sizeof(TagReq))) != DH_OK) return rc;
    if (!b->tags.empty()) DH_CUDA(cudaMemcpy(b->d_tags.p, b->tags.data(), b->tags.size() * sizeof(TagReq), cudaMemcpyHostToDevice));
    ScanTables &t = b->tabs;
    t.ref_cells = (const StrCell *)b->d_ref_cells.p;
    t.rg = (const RgEntry *)b->d_rg.p;
    t.rg_ids = (const uint8_t *)b->d_rg_ids.p;
    t.tags = (const TagReq *)b->d_tags.p;
    t.n_ref = (int32_t)b->ref_names.size();
    t.n_rg = (uint32_t)b->rg.size();
    t.n_tags = (uint32_t)b->tags.size();
    t.any_tag_cols = any_tag;
    t.aux_map = b->aux_map ? 1u : 0u;
    t.compact = is_compact(b) ? 1u : 0u;
    t.flag_require = b->flag_require;
    t.flag_exclude = b->flag_exclude;
    t.min_mapq = b->min_mapq;
    t.pad2 = 0;
    t.aux_excl = nullptr;
    if (b->aux_map && !b->aux_excl.empty()) {
        if ((rc = b->d_aux_excl.reserve(b->aux_excl.size() * 4)) != DH_OK) return rc;
        DH_CUDA(cudaMemcpy(b->d_aux_excl.p, b->aux_excl.data(), b->aux_excl.size() * 4, cudaMemcpyHostToDevice));
        t.aux_excl = (const uint32_t *)b->d_aux_excl.p;
    }
    t.col_mask = b->col_mask;
    return DH_OK;
}

cudaStream_t s_comp(dh_bam *b) { return b->ctx->stream; }
cudaStream_t s_copy(dh_bam *b) { return b->ctx->copy_streams[0]; }
cudaStream_t s_d2h(dh_bam *b) { return b->ctx->copy_streams[1]; }
cudaStream_t s_inf(dh_bam *b) { return b->ctx->copy_streams[2]; }
#ifndef DH_SM_RESERVE_DEVICE
#define DH_SM_RESERVE_DEVICE 36
#endif
// Phase 1 of the inflate is a persistent kernel that takes whole SMs (227 KB of shared memory each).  When batches go
// back to the host, a few SMs stay free so that the decode kernels feeding the D2H copy engine are never locked out.
// With column batches left on the device the walk/decode kernels of the previous segment are the second half of the work:
// measured on the 100 M-record scan (r1k, DH_INFLATE_SM_RESERVE = 0 / 8 / 16 / 24 / 32 / 48): 226 / 251 / 270 / 284 / 301 / 285
// Mrecords/s, so 32 SMs were left to them; with the round-2 phase 2 and whole-round segments (r2x, r3g: 24 / 28 / 32 / 36 / 40 = 301 / 316 /
// 321 / 325 / 296) it is 36.
int inflate_sms(const dh_bam *b) {
    static const int reserve_dev = getenv("DH_INFLATE_SM_RESERVE") ? atoi(getenv("DH_INFLATE_SM_RESERVE")) : DH_SM_RESERVE_DEVICE;
    // BCF: the FORMAT decode of a segment is short next to its inflate (15 vs 16 ms per 177 MB): no reserve (2.8 vs 2.2 M sites/s)
    const int reserve = (b->mode & DH_BAM_DEVICE_OUTPUT) ? (b->kind == 1 ? 0 : reserve_dev) : 8;
    return reserve < b->ctx->sm_count ? b->ctx->sm_count - reserve : b->ctx->sm_count;
}

int ensure_events(dh_bam *b) {
    for (auto &s : b->ss)
        if (!s.ev_h2d) {
            DH_CUDA(cudaEventCreate(&s.ev_h2d));
            DH_CUDA(cudaEventCreate(&s.ev_inf0));
            DH_CUDA(cudaEventCreate(&s.ev_inf1));
            DH_CUDA(cudaEventCreate(&s.ev_inf_done));
            DH_CUDA(cudaEventCreateWithFlags(&s.ev_mid, cudaEventDisableTiming));
            DH_CUDA(cudaEventCreate(&s.ev_free));
            DH_CUDA(cudaEventCreate(&s.ev_walk0));
            DH_CUDA(cudaEventCreate(&s.ev_walk1));
            DH_CUDA(cudaEventCreate(&s.ev_walk2));
        }
    for (int i = 0; i < b->n_slots; ++i) {
        BatchSlot &s = b->bs[i];
        if (s.ev_done) continue;
        {
            DH_CUDA(cudaEventCreate(&s.ev_k0));
            DH_CUDA(cudaEventCreate(&s.ev_k1));
            DH_CUDA(cudaEventCreate(&s.ev_c0));
            DH_CUDA(cudaEventCreate(&s.ev_done));
        }
    }
    return DH_OK;
}

int take_slot(dh_bam *b) {
    for (int i = 0; i < b->n_slots; ++i)
        if (b->bs[i].state == SLOT_FREE) return i;
    return -1;
}
int free_slots(const dh_bam *b) {
    int n = 0;
    for (int i = 0; i < b->n_slots; ++i) n += b->bs[i].state == SLOT_FREE;
    return n;
}

// pread of one big range by a few threads (one thread moves page-cache bytes at a fraction of the PCIe rate)
bool read_range(int fd, uint8_t *buf, uint64_t n, uint64_t off) {
    if (n == 0) return true;
    const int nt = (int)std::min<uint64_t>(8, (n + (32ull << 20) - 1) / (32ull << 20));
    std::vector<int> ok((size_t)nt, 1);
    auto work = [&](int t) {
        uint64_t lo = n * (uint64_t)t / (uint64_t)nt, hi = n * (uint64_t)(t + 1) / (uint64_t)nt;
        while (lo < hi) {
            const ssize_t k = pread(fd, buf + lo, (size_t)std::min<uint64_t>(hi - lo, 64ull << 20), (off_t)(off + lo));
            if (k < 0 && errno == EINTR) continue;
            if (k <= 0) { ok[(size_t)t] = 0; return; }
            lo += (uint64_t)k;
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    for (int v : ok) if (!v) return false;
    return true;
}

// File mode: read and frame the next segment (rows A1 / A3 as bytes arrive: htslib's framing-only reader thread, bgzf.c:1485-1539,
// does the same ahead of its inflate workers).  Appends the segment's blocks to the table and the segment to b->segs; sets
// b->frame_done when nothing is left for this scan.  The compressed bytes stay in a pinned staging buffer until issue_inflate copies
// them to the device; two buffers alternate, so reading segment k + 1 overlaps the copy of segment k.
int frame_segment(dh_bam *b) {
    if (b->frame_done) return DH_OK;
    BlockTable &bt = b->bt;
    const int k = (int)b->segs.size(), st = k & 1;
    if (b->frame_pos >= b->fsize) { b->frame_done = true; return DH_OK; }
    if (!b->ev_stage[st]) DH_CUDA(cudaEventCreate(&b->ev_stage[st]));
    if (b->stage_used[st]) DH_CUDA(cudaEventSynchronize(b->ev_stage[st]));   // segment k - 2 has left this buffer
    const uint64_t cbase = b->frame_pos & ~15ull;
    const uint64_t want = std::min<uint64_t>(b->seg_comp + (b->frame_pos - cbase), b->fsize - cbase);
    int rc = b->stage[st].reserve(b->seg_comp + 64);
    if (rc != DH_OK) return rc;
    uint8_t *buf = (uint8_t *)b->stage[st].p;
    if (!read_range(b->fd, buf, want, cbase)) { set_error("read: %s", strerror(errno)); return DH_EIO; }
    b->stage_used[st] = true;
    const size_t b0 = bt.blocks.size();
    uint64_t off = b->frame_pos - cbase, u = bt.uoff.back();
    const uint64_t u0 = u;
    bool broken = false;
    while (off < want) {
        if (want - off < 18) break;
        const uint8_t *h = buf + off;
        const bool ok = h[0] == 31 && h[1] == 139 && h[2] == 8 && (h[3] & 4) && (h[10] | (h[11] << 8)) == 6 && h[12] == 'B' && h[13] == 'C' && (h[14] | (h[15] << 8)) == 2;
        if (!ok) { broken = true; break; }
        const uint32_t bsize = (uint32_t)(h[16] | (h[17] << 8)) + 1;
        if (bsize < 18 + 8) { broken = true; break; }
        if (want - off < bsize) break;                       // continues in the next segment (or is cut by the end of the file)
        dh_bgzf_block B;
        B.coffset = cbase + off;
        B.csize = bsize;
        const uint8_t *f = h + bsize - 8;
        B.crc = (uint32_t)f[0] | ((uint32_t)f[1] << 8) | ((uint32_t)f[2] << 16) | ((uint32_t)f[3] << 24);
        B.isize = (uint32_t)f[4] | ((uint32_t)f[5] << 8) | ((uint32_t)f[6] << 16) | ((uint32_t)f[7] << 24);
        B.reserved = 0;
        if (B.isize > BGZF_MAX_OUT) { broken = true; break; }
        if (bt.blocks.size() - b0 >= (1u << 17) || u + B.isize - u0 > (3ull << 30)) break;
        // range scans stop framing once the look-ahead behind the range end is in the table
        if (b->limit_blk != SIZE_MAX && bt.blocks.size() >= b->end_blk && u - bt.uoff[b->limit_blk] >= b->carry_cap) { b->frame_done = true; break; }
        bt.blocks.push_back(B);
        u += B.isize;
        bt.uoff.push_back(u);
        bt.eof_marker = (bsize == 28 && B.isize == 0);
        if (b->limit_coff != UINT64_MAX && b->limit_blk == SIZE_MAX && B.coffset >= b->limit_coff) {
            // the block the range ends in (bgzf_tell's "end of block k" = start of block k + 1; empty blocks are skipped the same way)
            size_t lb = bt.blocks.size() - 1;
            uint32_t lo = B.coffset == b->limit_coff ? b->limit_uoff : 0;
            if (lo >= B.isize) { lb += 1; lo = 0; }           // resolved when that block is framed: limit_off 0 of the next block
            b->limit_blk = lb;
            b->limit_off = lo;
            b->end_blk = lb + (lo ? 1 : 0) + 2;
        }
        off += bsize;
    }
    const size_t b1 = bt.blocks.size();
    bt.end_off = cbase + off;
    if (b1 == b0) {
        // nothing whole: a broken header, or a block cut by the end of the file (htslib reads what lies in front of it, then fails)
        if (!b->frame_done && (broken || cbase + want >= b->fsize)) b->truncated = true;
        else if (!b->frame_done) { set_error("internal: no BGZF block fits a segment"); return DH_EINVAL; }
        b->frame_done = true;
        return DH_OK;
    }
    Seg sg{b0, b1, st};
    b->segs.push_back(sg);
    b->frame_pos = cbase + off;
    if (broken) { b->truncated = true; b->frame_done = true; }
    if (b->frame_pos >= b->fsize) b->frame_done = true;
    return DH_OK;
}

// does segment k exist?  (memory image: all segments are planned by begin_range; file mode: they are framed as the scan gets there)
int ensure_seg(dh_bam *b, int k, bool &exists) {
    while (b->fd >= 0 && (size_t)k >= b->segs.size() && !b->frame_done) {
        const int rc = frame_segment(b);
        if (rc != DH_OK) return rc;
    }
    exists = (size_t)k < b->segs.size();
    return DH_OK;
}

// H2D + inflate of segment k into slot k % NSEG (asynchronous)
int issue_inflate(dh_bam *b, int k) {
    SegSlot &s = b->ss[k % NSEG];
    if (s.issued_for == k) return DH_OK;
    const Seg &sg = b->segs[k];
    const BlockTable &bt = b->bt;
    s.b0 = sg.b0;
    s.b1 = sg.b1;
    const size_t nblk = sg.b1 - sg.b0;
    s.cbase = bt.blocks[sg.b0].coffset & ~15ull;
    const uint64_t cend = bt.blocks[sg.b1 - 1].coffset + bt.blocks[sg.b1 - 1].csize;
    const uint64_t cbytes = cend - s.cbase;
    s.obytes = bt.uoff[sg.b1] - bt.uoff[sg.b0];
    int rc;
    // the slot's previous tenant (segment k - NSEG) has finished on the compute stream before anything here runs:
    // its kernels were all enqueued on s_comp earlier, and the copy stream waits for its last inflate below.
    const bool staged = (b->mode & DH_BAM_DEVICE_INPUT) && b->d_image.p;
    if (!staged && (rc = s.comp.reserve(cbytes + DH_COMP_PAD)) != DH_OK) return rc;
    if ((rc = s.out.reserve(b->carry_cap + s.obytes + 2 * BGZF_MAX_OUT)) != DH_OK) return rc;
    if ((rc = s.desc.reserve(nblk * sizeof(BlockDesc))) != DH_OK) return rc;
    if ((rc = s.status.reserve(nblk * sizeof(int))) != DH_OK) return rc;
    if ((rc = s.h_desc.reserve(nblk * sizeof(BlockDesc))) != DH_OK) return rc;
    if ((rc = s.h_status.reserve(64)) != DH_OK) return rc;
    const uint64_t cw = fill_descs(bt, sg.b0, sg.b1, s.cbase, sg.b0, b->carry_cap, (BlockDesc *)s.h_desc.p);
    cudaStream_t sc = s_copy(b), si = s_inf(b);
    // DH_PHASE_STREAMS=2: phase 2 of segment k runs on a second stream, so phase 1 of segment k + 1 (next on `si`) overlaps it
    static const bool two = getenv("DH_PHASE_STREAMS") && atoi(getenv("DH_PHASE_STREAMS")) == 2;
    DH_CUDA(cudaStreamWaitEvent(sc, s.ev_inf1, 0));  // previous inflate from this comp buffer is done (no-op if never recorded)
    if (!staged) {
        // memory image: straight from the caller's (pinned) buffer; file mode: from the staging buffer frame_segment() filled
        const uint8_t *src = b->fd >= 0 ? (const uint8_t *)b->stage[sg.stage].p : b->data + s.cbase;
        DH_CUDA(cudaMemcpyAsync(s.comp.p, src, cbytes, cudaMemcpyHostToDevice, sc));
        DH_CUDA(cudaMemsetAsync((uint8_t *)s.comp.p + cbytes, 0, DH_COMP_PAD, sc));
        if (b->fd >= 0) DH_CUDA(cudaEventRecord(b->ev_stage[sg.stage], sc));   // the staging buffer may be refilled once this copy is done
    }
    DH_CUDA(cudaMemcpyAsync(s.desc.p, s.h_desc.p, nblk * sizeof(BlockDesc), cudaMemcpyHostToDevice, sc));
    DH_CUDA(cudaEventRecord(s.ev_h2d, sc));
    DH_CUDA(cudaStreamWaitEvent(si, s.ev_h2d, 0));
    DH_CUDA(cudaStreamWaitEvent(si, s.ev_free, 0));  // the slot's previous tenant has been decoded (and its tail carried over)
    DH_CUDA(cudaEventRecord(s.ev_inf0, si));
    const uint8_t *comp_ptr = staged ? (const uint8_t *)b->d_image.p + s.cbase : (const uint8_t *)s.comp.p;
    if ((rc = inflate_blocks(b->ctx, comp_ptr, (const BlockDesc *)s.desc.p, (uint32_t)nblk, cw, (uint8_t *)s.out.p, (int *)s.status.p,
                             s.work, (unsigned int *)b->ctx->counter.p + 16 + 2 * (k % NSEG), si, inflate_sms(b), two ? b->ctx->inf2_stream : nullptr,
                             two ? s.ev_mid : nullptr)) != DH_OK)
        return rc;
    cudaStream_t st = two ? b->ctx->inf2_stream : si;      // the stream phase 2 ended on
    DH_CUDA(cudaEventRecord(s.ev_inf1, st));
    DH_CUDA(launch_status_first_bad((const int *)s.status.p, (uint32_t)nblk, (int *)s.h_status.p, st));
    b->ctx->launches++;
    DH_CUDA(cudaEventRecord(s.ev_inf_done, st));
    s.issued_for = k;
    b->st.h2d_bytes += (staged ? 0 : cbytes) + nblk * sizeof(BlockDesc);
    b->st.inflated_bytes += s.obytes;
    b->st.n_blocks += nblk;
    return DH_OK;
}

// Boundary walk of segment k + host validation; fills b->batches.  Synchronises with the walk results only.
int walk_segment(dh_bam *b, int k) {
    SegSlot &s = b->ss[k % NSEG];
    const BlockTable &bt = b->bt;
    const size_t nblk = s.b1 - s.b0;
    const uint32_t n = (uint32_t)nblk + 1;  // + pseudo-block 0 = the carried-over partial record
    cudaStream_t sk = s_comp(b);
    int rc;
    // control arrays (block starts, entries, walk results, row/heap bases) live in pinned host memory that the
    // kernels read and write directly: a few bytes per BGZF block, and they never queue behind the bulk copies
    const uint64_t min_rec = b->kind == 1 ? 32 : 36;   // smallest record: bounds the number of record offsets a block can hold
    if ((rc = s.rec.reserve(((s.obytes + b->carry_cap) / min_rec + n + 8) * 2)) != DH_OK ||
        (rc = s.h_blk_start.reserve((n + 1) * 8)) != DH_OK || (rc = s.h_entry.reserve(n * 8)) != DH_OK ||
        (rc = s.h_todo.reserve(n)) != DH_OK || (rc = s.h_walk.reserve(n * sizeof(BlockWalk))) != DH_OK ||
        (rc = s.h_row_base.reserve(n * 4)) != DH_OK || (rc = s.h_heap_base.reserve(n * 8)) != DH_OK ||
        (rc = s.h_map_base.reserve(n * 4)) != DH_OK)
        return rc;
    DH_CUDA(cudaStreamWaitEvent(sk, s.ev_inf_done, 0));
    uint8_t *ubase = (uint8_t *)s.out.p + b->carry_cap;
    // carry: the incomplete record at the end of the previous segment goes in front of this segment's bytes
    if (b->carry > 0) {
        SegSlot &ps = b->ss[(k - 1) % NSEG];
        const uint8_t *src = (const uint8_t *)ps.out.p + b->carry_cap + ps.obytes - b->carry;
        DH_CUDA(cudaMemcpyAsync(ubase - b->carry, src, (size_t)b->carry, cudaMemcpyDeviceToDevice, sk));
    }
    int64_t *hs = (int64_t *)s.h_blk_start.p, *he = (int64_t *)s.h_entry.p;
    uint8_t *ht = (uint8_t *)s.h_todo.p;
    hs[0] = -b->carry;
    for (size_t i = 0; i < nblk; ++i) hs[i + 1] = (int64_t)(bt.uoff[s.b0 + i] - bt.uoff[s.b0]);
    hs[n] = (int64_t)s.obytes;
    const bool first_seg = k == 0;
    for (uint32_t i = 0; i < n; ++i) { he[i] = INT64_MIN; ht[i] = 1; }
    he[0] = -b->carry;                       // the carried record starts exactly there
    if (first_seg) { he[0] = b->first_entry; he[1] = b->first_entry; }   // scan start is a known record boundary
    else if (b->carry == 0) he[1] = 0;       // previous segment ended exactly on a record boundary
    bool has_next = false;
    if ((rc = ensure_seg(b, k + 1, has_next)) != DH_OK) return rc;
    const bool is_final = b->to_eof && !has_next && (b->fd >= 0 || s.b1 == bt.blocks.size());
    int64_t limit_end = INT64_MAX;
    if (!b->to_eof && b->limit_blk < s.b1) {
        limit_end = b->limit_blk >= s.b0 ? (int64_t)(bt.uoff[b->limit_blk] - bt.uoff[s.b0]) + b->limit_off : 0;
    }
    WalkParams wp;
    wp.u = ubase;
    wp.blk_start = hs;
    wp.entry_in = he;
    wp.todo = ht;
    wp.out = (BlockWalk *)s.h_walk.p;
    wp.n = n;
    wp.is_final = is_final ? 1 : 0;
    wp.seg_end = (int64_t)s.obytes;
    wp.limit_end = limit_end;
    wp.rec = (uint16_t *)s.rec.p;
    wp.rec_bias = s.rec_bias = -b->carry;
    wp.filt_mode = b->filt_mode;
    wp.filt_n = (uint32_t)b->filt.size();
    wp.filt = (const RegIntv *)b->d_filt.p;
    BlockWalk *hw = (BlockWalk *)s.h_walk.p;
    uint32_t validated = 0;  // blocks [0, validated) have entry == predecessor's exit
    bool stopped = false;
    uint32_t stop_blk = 0;
    bool first_round = true;
    for (;;) {
        if (first_round) DH_CUDA(cudaEventRecord(s.ev_walk0, sk));
        if (b->kind == 1) DH_CUDA(bcf_launch_walk(b, wp, sk));
        else DH_CUDA(launch_bam_walk(wp, b->tabs, sk));
        b->ctx->launches++;
        DH_CUDA(cudaEventRecord(first_round ? s.ev_walk1 : s.ev_walk2, sk));
        DH_CUDA(cudaEventSynchronize(first_round ? s.ev_walk1 : s.ev_walk2));
        if (first_round) {
            // verdicts of the inflate kernel (fused CRC32/ISIZE): stop at the first bad block like bgzf_read does
            const int *hst = (const int *)s.h_status.p;   // {first failing block or -1, its code}
            if (hst[0] >= 0) {
                set_error("BGZF block %zu failed inflate/CRC check (code %d)", s.b0 + (size_t)hst[0], hst[1]);
                return DH_EDATA;
            }
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, s.ev_inf0, s.ev_inf1) == cudaSuccess) b->st.inflate_ms += ms;
            if (cudaEventElapsedTime(&ms, s.ev_walk0, s.ev_walk1) == cudaSuccess) b->st.walk_ms += ms;
        }
        first_round = false;
        memset(ht, 0, n);
        // pointer-jumping check: entry[i] must equal exit[i-1]
        if (validated == 0) validated = 1;  // block 0's entry is given
        bool need_round = false;
        while (validated < n) {
            const BlockWalk &pw = hw[validated - 1];
            if (pw.stop != 0) { stopped = true; stop_blk = validated - 1; break; }
            if (pw.exit_off >= limit_end) { stopped = true; stop_blk = validated - 1; break; }
            int64_t expect = pw.exit_off;
            BlockWalk &w = hw[validated];
            if (w.entry != expect) {
                if (expect >= hs[validated + 1]) {
                    // a long record spans this whole block: nothing starts here (no GPU work needed)
                    w.entry = w.exit_off = expect;
                    w.count = 0;
                    w.heap_bytes = 0;
                    w.map_entries = w.map_heap = 0;
                    w.stop = 0;
                } else {
                    he[validated] = expect;
                    ht[validated] = 1;
                    need_round = true;
                    b->st.rewalk_rounds++;
                    break;
                }
            }
            ++validated;
        }
        if (!need_round) {
            if (!stopped && hw[n - 1].stop != 0) { stopped = true; stop_blk = n - 1; }
            if (!stopped && hw[n - 1].exit_off >= limit_end) { stopped = true; stop_blk = n - 1; }
            break;
        }
    }
    // rows / heap prefix sums, batches
    const uint32_t last = stopped ? stop_blk : n - 1;
    b->batches.clear();
    b->next_batch_i = 0;
    uint32_t *hrb = (uint32_t *)s.h_row_base.p;
    uint64_t *hhb = (uint64_t *)s.h_heap_base.p;
    uint32_t *hmb = (uint32_t *)s.h_map_base.p;
    Batch cur{0, 0, 0, 0, b->rows_total};
    for (uint32_t i = 0; i <= last; ++i) {
        const BlockWalk &w = hw[i];
        // a batch ends at batch_rows rows or 2 GiB of string heap (32-bit heap offsets of the compact layout; bounded pinned arenas)
        if (cur.rows > 0 && (cur.rows + w.count > b->batch_rows || cur.heap + w.heap_bytes > (2ull << 30))) {
            cur.b_hi = i;
            b->batches.push_back(cur);
            cur = Batch{i, i, 0, 0, cur.row0 + cur.rows};
        }
        hrb[i] = (uint32_t)cur.rows;
        hhb[i] = cur.heap;
        hmb[i] = (uint32_t)cur.map;
        cur.rows += w.count;
        cur.heap += w.heap_bytes;
        cur.map += w.map_entries;
    }
    cur.b_hi = last + 1;
    if (cur.rows > 0) b->batches.push_back(cur);
    for (auto &bb : b->batches) b->rows_total += bb.rows;
    // how the chain continues
    const BlockWalk &lw = hw[last];
    b->carry = 0;
    auto voff_of = [&](int64_t off) -> uint64_t {  // segment-relative inflated offset -> virtual offset
        if (off < 0) return 0;
        uint64_t abs_u = bt.uoff[s.b0] + (uint64_t)off;
        size_t blk = std::upper_bound(bt.uoff.begin() + s.b0, bt.uoff.end(), abs_u) - bt.uoff.begin() - 1;
        if (blk >= bt.blocks.size()) return ((bt.end_off) << 16);
        return (bt.blocks[blk].coffset << 16) | (abs_u - bt.uoff[blk]);
    };
    if (stopped && lw.stop < 0) {
        b->final_status = lw.stop;   // bam_read1 / sam_read1_bam error: the reference ends the scan silently here
        b->chain_done = true;
        b->st.end_voffset = voff_of(lw.exit_off);
    } else if (stopped && lw.stop == 0) {
        b->chain_done = true;        // reached voff_end
        b->st.end_voffset = voff_of(lw.exit_off);
    } else if (stopped && lw.stop == 1) {
        b->carry = (int64_t)s.obytes - lw.exit_off;
        if ((uint64_t)b->carry > b->carry_cap) {
            // a limit of this pipeline, not a read error of the file: reported, never a silent end of scan
            set_error("a record of more than %llu bytes straddles two inflate segments: raise the carry buffer (dh_bam_configure)",
                      (unsigned long long)b->carry_cap);
            return DH_ENOMEM;
        }
        if (!has_next) {
            // range scan whose last record runs past the inflated look-ahead (carry_cap bytes behind the range end)
            set_error("the record at the end of the scan range is longer than the %llu-byte look-ahead: raise the carry buffer (dh_bam_configure)",
                      (unsigned long long)b->carry_cap);
            return DH_ENOMEM;
        }
    } else {
        b->st.end_voffset = voff_of(lw.exit_off);
        if (!has_next) {
            b->chain_done = true;
            // behind the last good block of a truncated image htslib's next read fails (bam_read1: bgzf_read != 4 and != 0 -> -2)
            if (b->truncated && is_final && b->final_status == 0) b->final_status = -2;
        }
    }
    return DH_OK;
}

int issue_batch(dh_bam *b) {
    const int slot = take_slot(b);
    if (slot < 0) { set_error("internal: no free batch slot"); return DH_EINVAL; }
    const Batch &bb = b->batches[b->next_batch_i++];
    SegSlot &s = b->ss[b->cur_seg % NSEG];
    BatchSlot &o = b->bs[slot];
    const bool compact = is_compact(b);
    // arena layout: [col data | validity]* aligned to 256, (compact: rec_off,) heap last
    uint64_t off = 0, coff[BAM_MAX_COLS], voff[BAM_MAX_COLS];
    const int ncols = BAM_CORE_COLS + (int)b->tags.size();
    for (int c = 0; c < ncols; ++c) {
        coff[c] = voff[c] = UINT64_MAX;
        if (!col_on(b, c)) continue;
        coff[c] = off;
        off = (off + (uint64_t)(compact ? col_width_compact(b, c) : col_width(b, c)) * bb.rows + A256) & ~A256;
        if (col_nullable(c)) {
            voff[c] = off;
            off = (off + ((bb.rows + 63) / 64) * 8 + A256) & ~A256;
        }
    }
    uint64_t map_off[4] = {0, 0, 0, 0};   // list entries, validity, key cells, value cells
    if (b->aux_map) {
        const uint64_t sz[4] = {16 * bb.rows, ((bb.rows + 63) / 64) * 8, 16 * bb.map, 16 * bb.map};
        for (int q = 0; q < 4; ++q) { map_off[q] = off; off = (off + sz[q] + A256) & ~A256; }
    }
    uint64_t recoff_off = 0;
    if (compact) { recoff_off = off; off = (off + 4 * bb.rows + A256) & ~A256; }
    const uint64_t heap_off = off, total = off + bb.heap;
    int rc;
    if ((rc = o.arena.reserve(total + 256)) != DH_OK) return rc;
    if ((rc = o.h_arena.reserve(total + 256)) != DH_OK) return rc;
    cudaStream_t sk = s_comp(b), sd = s_d2h(b);
    DecodeParams dp;
    memset(&dp, 0, sizeof dp);
    dp.u = (const uint8_t *)s.out.p + b->carry_cap;
    dp.walk = (const BlockWalk *)s.h_walk.p;
    dp.blk_start = (const int64_t *)s.h_blk_start.p;
    dp.rec = (const uint16_t *)s.rec.p;
    dp.rec_bias = s.rec_bias;
    dp.row_base = (const uint32_t *)s.h_row_base.p;
    dp.heap_base = (const uint64_t *)s.h_heap_base.p;
    dp.b_lo = bb.b_lo;
    dp.b_hi = bb.b_hi;
    dp.heap = (uint8_t *)o.arena.p + heap_off;
    dp.heap_host = (uint64_t)(uintptr_t)((uint8_t *)o.h_arena.p + heap_off);
    dp.rec_off = compact ? (uint32_t *)((uint8_t *)o.arena.p + recoff_off) : nullptr;
    memset(&o.desc, 0, sizeof o.desc);
    for (int c = 0; c < ncols; ++c) {
        if (coff[c] == UINT64_MAX) continue;
        dp.col[c] = (uint8_t *)o.arena.p + coff[c];
        o.desc.data[c] = (uint8_t *)o.h_arena.p + coff[c];
        if (voff[c] != UINT64_MAX) {
            dp.valid[c] = (unsigned long long *)((uint8_t *)o.arena.p + voff[c]);
            o.desc.validity[c] = (const uint64_t *)((uint8_t *)o.h_arena.p + voff[c]);
            DH_CUDA(cudaMemsetAsync(dp.valid[c], 0, ((bb.rows + 63) / 64) * 8, sk));
        }
    }
    DH_CUDA(cudaEventRecord(o.ev_k0, sk));
    DH_CUDA(compact ? launch_bam_decode_compact(dp, b->tabs, sk) : launch_bam_decode(dp, b->tabs, sk));
    b->ctx->launches++;
    o.dp = dp;
    o.dev_recoff = recoff_off;
    if (b->aux_map) {
        AuxMapParams mp;
        memset(&mp, 0, sizeof mp);
        mp.u = dp.u; mp.walk = dp.walk; mp.blk_start = dp.blk_start; mp.rec = dp.rec; mp.rec_bias = dp.rec_bias;
        mp.row_base = dp.row_base; mp.heap_base = dp.heap_base; mp.map_base = (const uint32_t *)s.h_map_base.p;
        mp.b_lo = bb.b_lo; mp.b_hi = bb.b_hi;
        uint8_t *da = (uint8_t *)o.arena.p, *ha = (uint8_t *)o.h_arena.p;
        mp.entries = da + map_off[0];
        mp.valid = (unsigned long long *)(da + map_off[1]);
        mp.keys = da + map_off[2];
        mp.values = da + map_off[3];
        mp.heap = dp.heap;
        mp.heap_host = dp.heap_host;
        DH_CUDA(cudaMemsetAsync(mp.valid, 0, ((bb.rows + 63) / 64) * 8, sk));
        DH_CUDA(launch_bam_auxmap(mp, b->tabs, sk));
        b->ctx->launches++;
        o.desc.aux_map = ha + map_off[0];
        o.desc.aux_map_validity = (const uint64_t *)(ha + map_off[1]);
        o.desc.aux_keys = ha + map_off[2];
        o.desc.aux_values = ha + map_off[3];
        o.desc.n_aux = bb.map;
    }
    DH_CUDA(cudaEventRecord(o.ev_k1, sk));
    DH_CUDA(cudaStreamWaitEvent(sd, o.ev_k1, 0));
    DH_CUDA(cudaEventRecord(o.ev_c0, sd));
    const bool keep = (b->mode & DH_BAM_DEVICE_OUTPUT) != 0;
    if (total && !keep) DH_CUDA(cudaMemcpyAsync(o.h_arena.p, o.arena.p, total, cudaMemcpyDeviceToHost, sd));
    DH_CUDA(cudaEventRecord(o.ev_done, sd));
    o.timed = true;
    o.desc.n_rows = bb.rows;
    o.desc.row0 = bb.row0;
    o.desc.status = 0;
    o.desc.n_cols = ncols;
    o.desc.heap = (const uint8_t *)o.h_arena.p + heap_off;
    o.desc.heap_bytes = bb.heap;
    o.desc.bytes = total;
    o.desc.slot = slot;
    o.desc.layout = compact ? DH_BAM_COMPACT_OUTPUT : 0u;
    o.desc.rec_off = compact ? (const uint32_t *)((const uint8_t *)o.h_arena.p + recoff_off) : nullptr;
    o.on_device = keep;
    o.last = false;
    o.state = SLOT_INFLIGHT;
    b->st.d2h_bytes += keep ? 0 : total;
    b->inflight.push_back(slot);
    b->issued++;
    return DH_OK;
}

// make sure there is a next batch to issue: advance through segments.  Returns 1 if a batch is available,
// 0 if the scan is exhausted, < 0 on error.
int have_work(dh_bam *b) {
    for (;;) {
        if (b->cur_seg >= 0 && b->next_batch_i < b->batches.size()) return 1;
        if (b->chain_done) return 0;
        int k = b->cur_seg + 1;
        bool exists = false;
        int rc = ensure_seg(b, k, exists);
        if (rc != DH_OK) return rc;
        if (!exists) { b->chain_done = true; return 0; }
        rc = issue_inflate(b, k);  // no-op when prefetched
        if (rc != DH_OK) return rc;
        b->cur_seg = k;
        if ((rc = walk_segment(b, k)) != DH_OK) return rc;
        // Segment k-1 is no longer needed: its decode kernels and the carry copy out of its tail are all enqueued on
        // the decode stream.  Its slot is the one segment k+2 will be inflated into.
        if (k >= 1) DH_CUDA(cudaEventRecord(b->ss[(k - 1) % NSEG].ev_free, s_comp(b)));
        // prefetch: H2D + inflate of the next two segments overlap this segment's decode batches and their D2H copies
        // (file mode reads ahead through two staging buffers: one segment beyond the one being decoded)
        for (int a = 1; a < (b->fd >= 0 ? 2 : NSEG) && !b->chain_done; ++a) {
            if ((rc = ensure_seg(b, k + a, exists)) != DH_OK) return rc;
            if (exists && (rc = issue_inflate(b, k + a)) != DH_OK) return rc;
        }
    }
}

}  // namespace

extern "C" {

dh_bam *dh_bam_open(dh_ctx *ctx, const uint8_t *data, uint64_t size) {
    if (!ctx || !data) { set_error("dh_bam_open: bad arguments"); return nullptr; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) { set_error("cudaSetDevice failed"); return nullptr; }
    dh_bam *b = new dh_bam();
    b->ctx = ctx;
    b->data = data;
    b->size = size;
    memset(&b->st, 0, sizeof b->st);
    uint64_t bad = 0;
    if (walk_blocks(data, size, 0, b->bt, &bad) != DH_OK) {
        // htslib opens a truncated BAM or one with garbage behind a block and returns every record in front of the broken block; the
        // read then fails (bgzf_read -> -1, bam_read1 -> -2 / -3 / -4) and the reference ends the scan silently (bam_reader.c:754-765)
        if (b->bt.blocks.empty()) {
            set_error("invalid BGZF header at offset %llu", (unsigned long long)bad);
            delete b;
            return nullptr;
        }
        b->bt.end_off = bad;
        b->truncated = true;
    }
    if (b->bt.blocks.empty()) { set_error("empty file"); delete b; return nullptr; }
    if (parse_header(b) != DH_OK) { delete b; return nullptr; }
    build_dictionaries(b);
    if (ensure_events(b) != DH_OK || upload_tables(b) != DH_OK) { dh_bam_close(b); return nullptr; }
    return b;
}

dh_bam *dh_bam_open_file(dh_ctx *ctx, const char *path) {
    if (!ctx || !path) { set_error("dh_bam_open_file: bad arguments"); return nullptr; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) { set_error("cudaSetDevice failed"); return nullptr; }
    const int fd = open(path, O_RDONLY);
    struct stat sb;
    if (fd < 0 || fstat(fd, &sb) != 0 || !S_ISREG(sb.st_mode)) {
        set_error("cannot open %s: %s", path, strerror(errno));
        if (fd >= 0) close(fd);
        return nullptr;
    }
    dh_bam *b = new dh_bam();
    b->ctx = ctx;
    b->fd = fd;
    b->fsize = (uint64_t)sb.st_size;
    b->size = b->fsize;
    b->seg_comp = 512ull << 20;      // two pinned staging buffers of this size; the file itself is never pinned
    memset(&b->st, 0, sizeof b->st);
    // the header: the leading bytes of the file, more of them until the header blocks are all there (a 20,000-contig header is ~1 MB)
    uint64_t take = std::min<uint64_t>(b->fsize, 4ull << 20);
    for (;;) {
        b->head.resize((size_t)take);
        if (!read_range(fd, b->head.data(), take, 0)) { set_error("read: %s", strerror(errno)); dh_bam_close(b); return nullptr; }
        b->bt = BlockTable();
        uint64_t bad = 0;
        const int wrc = walk_blocks(b->head.data(), take, 0, b->bt, &bad);
        if (wrc != DH_OK) b->bt.end_off = bad;              // a block cut by the end of what was read (or a broken one): the blocks in front count
        if (b->bt.blocks.empty()) {
            set_error(b->fsize ? "invalid BGZF header at offset 0" : "empty file");
            dh_bam_close(b);
            return nullptr;
        }
        b->data = b->head.data();
        {   // not a BAM at all?  say so now instead of reading more and more of a file that will never parse
            const dh_bgzf_block &B0 = b->bt.blocks[0];
            std::vector<uint8_t> u0((size_t)B0.isize + 8);
            uint64_t n0 = 0;
            if (dh_bgzf_inflate_host(ctx, b->head.data(), B0.csize, u0.data(), u0.size(), &n0, nullptr) != DH_OK || (n0 >= 4 && memcmp(u0.data(), "BAM\1", 4) != 0)) {
                set_error("invalid BAM binary header");
                b->data = nullptr;
                dh_bam_close(b);
                return nullptr;
            }
        }
        const int prc = parse_header(b);
        b->data = nullptr;
        if (prc == DH_OK && b->first_blk < b->bt.blocks.size()) break;
        if (prc == DH_OK && take >= b->fsize) break;        // header only, no alignment block behind it
        if (take >= b->fsize) { if (prc == DH_OK) break; dh_bam_close(b); return nullptr; }
        take = std::min<uint64_t>(b->fsize, take * 4);      // the header (or the block of the first record) continues behind what was read
    }
    b->first_voff = b->first_blk < b->bt.blocks.size() ? ((b->bt.blocks[b->first_blk].coffset << 16) | b->first_off) : (b->bt.end_off << 16);
    build_dictionaries(b);
    if (ensure_events(b) != DH_OK || upload_tables(b) != DH_OK) { dh_bam_close(b); return nullptr; }
    return b;
}

void dh_bam_close(dh_bam *b) {
    if (!b) return;
    cudaSetDevice(b->ctx->device);
    cudaDeviceSynchronize();
    if (b->fd >= 0) close(b->fd);
    for (int i = 0; i < 2; ++i) { b->stage[i].release(); if (b->ev_stage[i]) cudaEventDestroy(b->ev_stage[i]); }
    for (auto &s : b->ss) {
        s.comp.release(); s.out.release(); s.desc.release(); s.status.release(); s.work.release(); s.rec.release();
        s.h_desc.release(); s.h_status.release(); s.h_blk_start.release(); s.h_entry.release(); s.h_todo.release();
        s.h_walk.release(); s.h_row_base.release(); s.h_heap_base.release(); s.h_map_base.release();
        for (cudaEvent_t e : {s.ev_mid, s.ev_h2d, s.ev_inf0, s.ev_inf1, s.ev_inf_done, s.ev_free, s.ev_walk0, s.ev_walk1, s.ev_walk2}) if (e) cudaEventDestroy(e);
    }
    for (auto &s : b->bs) {
        s.arena.release(); s.h_arena.release();
        for (cudaEvent_t e : {s.ev_k0, s.ev_k1, s.ev_c0, s.ev_done}) if (e) cudaEventDestroy(e);
    }
    b->d_dict_ptr.release(); b->d_dict_hash.release(); b->d_ref_hash.release(); b->d_sample_hash.release(); b->d_sums.release(); b->h_sums.release();
    b->d_aux_excl.release();
    if (b->ev_scan0) cudaEventDestroy(b->ev_scan0);
    b->d_ref_cells.release(); b->d_rg.release(); b->d_rg_ids.release(); b->d_tags.release(); b->d_image.release(); b->d_filt.release();
    if (b->bcf) bcf_state_free(b->bcf);
    delete b;
}

int32_t dh_bam_n_ref(const dh_bam *b) { return b ? (int32_t)b->ref_names.size() : 0; }
const char *dh_bam_ref_name(const dh_bam *b, int32_t tid) {
    return (b && tid >= 0 && (size_t)tid < b->ref_names.size()) ? b->ref_names[tid].c_str() : nullptr;
}
uint32_t dh_bam_ref_len(const dh_bam *b, int32_t tid) {
    return (b && tid >= 0 && (size_t)tid < b->ref_lens.size()) ? b->ref_lens[tid] : 0;
}
const char *dh_bam_header_text(const dh_bam *b, uint32_t *len) {
    if (!b) return nullptr;
    if (len) *len = (uint32_t)b->text.size();
    return b->text.c_str();
}
uint64_t dh_bam_first_voffset(const dh_bam *b) {
    if (!b) return 0;
    if (b->fd >= 0) return b->first_voff;
    if (b->first_blk >= b->bt.blocks.size()) return b->bt.end_off << 16;
    return (b->bt.blocks[b->first_blk].coffset << 16) | b->first_off;
}
uint64_t dh_bam_n_blocks(const dh_bam *b) { return b ? b->bt.blocks.size() : 0; }

int dh_bam_set_projection(dh_bam *b, uint64_t col_mask, const dh_tag_req *tags, int n_tags) {
    if (!b || n_tags < 0 || n_tags > BAM_MAX_TAGS || (n_tags && !tags)) { set_error("dh_bam_set_projection: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (b->scanning && !b->inflight.empty()) { set_error("dh_bam_set_projection: scan in progress"); return DH_EINVAL; }
    b->tags.clear();
    for (int i = 0; i < n_tags; ++i) {
        TagReq r;
        r.t0 = (uint8_t)tags[i].tag[0];
        r.t1 = (uint8_t)tags[i].tag[1];
        char ty = tags[i].type == 'H' ? 'Z' : tags[i].type;
        if (ty == 'c') {   // N3: a CIGAR metric as a derived BIGINT column (tag[0] = metric, tag[1] = the operator of cigar_has_op)
            if (!strchr("SHlrqaRo", (char)r.t0) || !r.t0) { set_error("unknown CIGAR metric '%c'", tags[i].tag[0]); return DH_EINVAL; }
            if (r.t0 == 'o') {
                r.t1 = (uint8_t)toupper(r.t1);
                if (!r.t1 || !strchr("MIDNSHP=X", (char)r.t1)) { set_error("cigar_has_op: '%c' is not a CIGAR operator", tags[i].tag[1]); return DH_EINVAL; }
            }
        } else
        if (ty != 'i' && ty != 'f' && ty != 'Z' && ty != 'A' && ty != 'B') { set_error("unsupported tag column type '%c'", tags[i].type); return DH_EINVAL; }
        r.type = (uint8_t)ty;
        r.want = 1;
        b->tags.push_back(r);
    }
    const int ncols = BAM_CORE_COLS + n_tags;
    b->col_mask = ncols >= 64 ? col_mask : (col_mask & ((1ull << ncols) - 1));
    DH_CUDA(cudaSetDevice(b->ctx->device));
    return upload_tables(b);
}

int dh_bam_set_aux_map(dh_bam *b, int enable, const char *exclude, int n_exclude) {
    if (!b || n_exclude < 0 || (n_exclude && !exclude)) { set_error("dh_bam_set_aux_map: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (b->scanning && !b->inflight.empty()) { set_error("dh_bam_set_aux_map: scan in progress"); return DH_EINVAL; }
    b->aux_map = enable != 0;
    b->aux_excl.clear();
    if (enable && n_exclude) {
        b->aux_excl.assign(2048, 0u);
        for (int i = 0; i < n_exclude; ++i) {
            const uint32_t key = (uint8_t)exclude[2 * i] | ((uint32_t)(uint8_t)exclude[2 * i + 1] << 8);
            b->aux_excl[key >> 5] |= 1u << (key & 31u);
        }
    }
    DH_CUDA(cudaSetDevice(b->ctx->device));
    return upload_tables(b);
}

int dh_bam_configure(dh_bam *b, uint64_t seg_comp_bytes, uint32_t batch_rows, uint64_t carry_bytes) {
    if (!b) return DH_EINVAL;
    if (seg_comp_bytes) b->seg_comp = std::min<uint64_t>(std::max<uint64_t>(seg_comp_bytes, 1 << 16), 2ull << 30);
    if (batch_rows) b->batch_rows = batch_rows;
    if (carry_bytes) {
        b->carry_cap = (carry_bytes + 255) & ~255ull;
        for (auto &s : b->ss) { s.out.release(); s.issued_for = -1; }
    }
    return DH_OK;
}

int dh_bam_set_mode(dh_bam *b, uint32_t flags) {
    if (!b) return DH_EINVAL;
    DH_CUDA(cudaSetDevice(b->ctx->device));
    if ((flags & DH_BAM_DEVICE_INPUT) && b->fd >= 0) { set_error("DH_BAM_DEVICE_INPUT needs a memory image (dh_bam_open), not a file handle"); return DH_EINVAL; }
    if ((flags & DH_BAM_DEVICE_INPUT) && !b->d_image.p) {
        int rc = b->d_image.reserve(b->bt.end_off + DH_COMP_PAD + 16);
        if (rc != DH_OK) return rc;
        DH_CUDA(cudaMemcpy(b->d_image.p, b->data, b->bt.end_off, cudaMemcpyHostToDevice));
        DH_CUDA(cudaMemset((uint8_t *)b->d_image.p + b->bt.end_off, 0, DH_COMP_PAD));
    }
    if (!(flags & DH_BAM_DEVICE_INPUT)) b->d_image.release();
    b->mode = flags;
    b->tabs.compact = is_compact(b) ? 1u : 0u;
    return DH_OK;
}

int dh_bam_set_filter(dh_bam *b, uint32_t flag_require, uint32_t flag_exclude, uint32_t min_mapq) {
    if (!b) { set_error("dh_bam_set_filter: null handle"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (b->scanning && !b->inflight.empty()) { set_error("dh_bam_set_filter: scan in progress"); return DH_EINVAL; }
    b->flag_require = flag_require & 0xffffu;
    b->flag_exclude = flag_exclude & 0xffffu;
    b->min_mapq = min_mapq;
    b->tabs.flag_require = b->flag_require;
    b->tabs.flag_exclude = b->flag_exclude;
    b->tabs.min_mapq = b->min_mapq;
    return DH_OK;
}

int32_t dh_bam_n_rg(const dh_bam *b) { return b ? (int32_t)b->rg.size() : 0; }
const char *dh_bam_rg_id(const dh_bam *b, int32_t i, uint32_t *len) {
    if (!b || i < 0 || (size_t)i >= b->rg.size()) return nullptr;
    if (len) *len = b->rg[(size_t)i].id_len;
    return b->rg_ids.data() + b->rg[(size_t)i].id_off;
}
const char *dh_bam_rg_sample(const dh_bam *b, int32_t i) {
    if (!b || i < 0 || (size_t)i >= b->rg.size() || !b->rg[(size_t)i].has_sample) return nullptr;
    return b->rg_samples[(size_t)i].c_str();
}

dh_bam *dh_bam_clone(const dh_bam *src, dh_ctx *ctx) {
    if (!src || !ctx || src->kind != 0) { set_error("dh_bam_clone: bad arguments"); return nullptr; }
    if (cudaSetDevice(ctx->device) != cudaSuccess) { set_error("cudaSetDevice failed"); return nullptr; }
    dh_bam *b = new dh_bam();
    b->ctx = ctx;
    b->data = src->data;
    b->size = src->size;
    if (src->fd >= 0) {      // file mode: an own descriptor (own file position is not needed: pread), the header as parsed
        b->fd = dup(src->fd);
        if (b->fd < 0) { set_error("dup: %s", strerror(errno)); delete b; return nullptr; }
        b->fsize = src->fsize;
        b->first_voff = src->first_voff;
    }
    b->bt = src->bt;
    b->text = src->text;
    b->ref_names = src->ref_names;
    b->ref_lens = src->ref_lens;
    b->first_blk = src->first_blk;
    b->first_off = src->first_off;
    b->truncated = src->truncated;
    b->seg_comp = src->seg_comp;
    b->batch_rows = src->batch_rows;
    b->carry_cap = src->carry_cap;
    b->n_slots = src->n_slots;
    memset(&b->st, 0, sizeof b->st);
    build_dictionaries(b);
    if (ensure_events(b) != DH_OK || upload_tables(b) != DH_OK) { dh_bam_close(b); return nullptr; }
    return b;
}

// (block, in-block offset) of a virtual offset.  htslib's bgzf_tell reports "end of block k" as (k, isize) while the next
// record really starts at (k + 1, 0); empty blocks are skipped the same way (htslib/bgzf.c:1145-1228).
static int locate_voff(const dh_bam *b, uint64_t voff, const char *what, size_t &blk, uint32_t &off) {
    const BlockTable &bt = b->bt;
    const size_t nb = bt.blocks.size();
    blk = block_of(b, voff >> 16);
    off = (uint32_t)(voff & 0xffff);
    if (blk < nb && bt.blocks[blk].coffset != (voff >> 16)) { set_error("%s is not a BGZF block address", what); return DH_EINVAL; }
    while (blk < nb && off >= bt.blocks[blk].isize) {
        if (off > bt.blocks[blk].isize) { set_error("%s points beyond the end of its BGZF block", what); return DH_EINVAL; }
        ++blk;
        off = 0;
    }
    return DH_OK;
}

// Start scanning one virtual-offset range.  `fresh` = first range of a scan: counters and statistics restart.
static int begin_range(dh_bam *b, uint64_t voff_beg, uint64_t voff_end, bool fresh) {
    DH_CUDA(cudaSetDevice(b->ctx->device));
    // drop whatever an abandoned scan left in flight (this context's streams only: other contexts may share the device)
    DH_CUDA(cudaStreamSynchronize(b->ctx->stream));
    for (int i = 0; i < 3; ++i) DH_CUDA(cudaStreamSynchronize(b->ctx->copy_streams[i]));
    DH_CUDA(cudaStreamSynchronize(b->ctx->inf2_stream));
    if (b->fd >= 0) {
        // file mode: the block table restarts at the block the range begins in and grows as segments are read (frame_segment)
        const uint64_t vb = voff_beg ? voff_beg : b->first_voff;
        BlockTable &t = b->bt;
        t.blocks.clear();
        t.uoff.assign(1, 0);
        t.end_off = vb >> 16;
        t.eof_marker = false;
        b->frame_pos = vb >> 16;
        b->frame_done = b->frame_pos >= b->fsize;
        b->truncated = false;
        b->to_eof = voff_end == UINT64_MAX || (voff_end >> 16) >= b->fsize;
        b->limit_coff = b->to_eof ? UINT64_MAX : (voff_end >> 16);
        b->limit_uoff = (uint32_t)(voff_end & 0xffff);
        b->limit_blk = SIZE_MAX;
        b->limit_off = 0;
        b->end_blk = SIZE_MAX;
        b->segs.clear();
        for (auto &s : b->ss) s.issued_for = -1;
        b->first_entry = (int64_t)(vb & 0xffff);     // "end of block k" chains into block k + 1 by itself (the walk starts nothing in k)
        b->cur_seg = -1;
        b->batches.clear();
        b->next_batch_i = 0;
        b->inflight.clear();
        b->issued = b->returned = 0;
        b->carry = 0;
        b->chain_done = b->frame_done;
        if (fresh) {
            for (auto &o : b->bs) o.state = SLOT_FREE;
            b->auto_slot = -1;
            b->rows_total = 0;
            b->final_status = 0;
            memset(&b->st, 0, sizeof b->st);
            b->st.end_voffset = vb;
            if (!b->ev_scan0) DH_CUDA(cudaEventCreate(&b->ev_scan0));
            DH_CUDA(cudaEventRecord(b->ev_scan0, s_comp(b)));
        }
        b->scanning = true;
        return DH_OK;
    }
    const BlockTable &bt = b->bt;
    const size_t nb = bt.blocks.size();
    size_t b_first = b->first_blk;
    uint32_t off_first = b->first_off;
    int rc;
    if (voff_beg != 0 && (rc = locate_voff(b, voff_beg, "voff_beg", b_first, off_first)) != DH_OK) return rc;
    b->to_eof = voff_end == UINT64_MAX || (voff_end >> 16) >= bt.end_off;
    size_t b_stop = nb;
    if (!b->to_eof) {
        if ((rc = locate_voff(b, voff_end, "voff_end", b->limit_blk, b->limit_off)) != DH_OK) return rc;
        if (b->limit_blk >= nb) b->to_eof = true;
        // inflate up to the limit block plus look-ahead for a record that straddles the range end: at least two blocks, and as many
        // as hold carry_cap inflated bytes (a 150 kb long read is several blocks; the look-ahead costs a few MB per range)
        else {
            b_stop = std::min(nb, b->limit_blk + (b->limit_off ? 1 : 0) + 2);
            while (b_stop < nb && bt.uoff[b_stop] - bt.uoff[b->limit_blk] < b->carry_cap) ++b_stop;
        }
    }
    b->end_blk = b_stop;
    b->segs.clear();
    size_t s0 = std::min(b_first, b_stop);
    while (s0 < b_stop) {
        size_t e = s0;
        const uint64_t cbase = bt.blocks[s0].coffset & ~15ull;
        while (e < b_stop && e - s0 < (1u << 17) && bt.blocks[e].coffset + bt.blocks[e].csize - cbase <= b->seg_comp &&
               bt.uoff[e + 1] - bt.uoff[s0] <= (3ull << 30))
            ++e;
        if (e == s0) e = s0 + 1;
        // Phase 1 of the inflate decodes `round` streams at a time and a warp takes its next 32 only when all of its lanes are done: a
        // segment of 3.4 rounds costs 4 (BCF: 95,000 blocks per GiB).  Cut it to whole rounds where that wastes more than 7 %.
        const size_t round = bgzf_inflate_v3_round(inflate_sms(b));
        if (e < b_stop && e - s0 > round) {
            const size_t k = (e - s0) / round, up = k + ((e - s0) % round ? 1 : 0);
            if ((double)(e - s0) < 0.93 * (double)(up * round)) e = s0 + k * round;
        }
        b->segs.push_back(Seg{s0, e});
        s0 = e;
    }
    for (auto &s : b->ss) s.issued_for = -1;
    b->first_entry = off_first;
    b->cur_seg = -1;
    b->batches.clear();
    b->next_batch_i = 0;
    b->inflight.clear();
    b->issued = b->returned = 0;
    b->carry = 0;
    b->chain_done = b->segs.empty();
    if (fresh) {
        for (auto &o : b->bs) o.state = SLOT_FREE;   // whatever the caller still held belongs to the abandoned scan
        b->auto_slot = -1;
        b->rows_total = 0;
        b->final_status = 0;
        memset(&b->st, 0, sizeof b->st);
        b->st.end_voffset = voff_beg ? voff_beg : dh_bam_first_voffset(b);
        if (!b->ev_scan0) DH_CUDA(cudaEventCreate(&b->ev_scan0));
        DH_CUDA(cudaEventRecord(b->ev_scan0, s_comp(b)));
    }
    b->scanning = true;
    return DH_OK;
}

static int check_layout(const dh_bam *b) {
    if (is_compact(b) && b->aux_map) { set_error("the compact layout does not carry the AUXILIARY_TAGS column: scan in the DuckDB layout"); return DH_EINVAL; }
    return DH_OK;
}

int dh_bam_scan_begin(dh_bam *b, uint64_t voff_beg, uint64_t voff_end) {
    if (!b) { set_error("dh_bam_scan_begin: null handle"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (check_layout(b) != DH_OK) return DH_EINVAL;
    b->filt_mode = 0;
    b->filt.clear();
    b->pending.clear();
    return begin_range(b, voff_beg, voff_end, true);
}

// ---- A8: indexed scans ------------------------------------------------------------------------------------
dh_index *dh_index_load(dh_ctx *ctx, const uint8_t *data, uint64_t size) {
    if (!data || size < 8) { set_error("dh_index_load: bad arguments"); return nullptr; }
    std::vector<uint8_t> plain;
    if (data[0] == 0x1f && data[1] == 0x8b) {   // CSI files are BGZF-compressed: inflate on the GPU like any other BGZF stream
        if (!ctx) { set_error("dh_index_load: a context is needed to inflate a BGZF-compressed index"); return nullptr; }
        uint64_t n = 0;
        if (dh_bgzf_inflate_host(ctx, data, size, nullptr, 0, &n, nullptr) != DH_OK) return nullptr;
        plain.resize((size_t)n + 8);
        if (dh_bgzf_inflate_host(ctx, data, size, plain.data(), plain.size(), &n, nullptr) != DH_OK) return nullptr;
        data = plain.data();
        size = n;
    }
    dh_index *ix = new dh_index();
    std::string err;
    if (!index_parse(data, size, ix->ix, err)) { set_error("%s", err.c_str()); delete ix; return nullptr; }
    return ix;
}
void dh_index_free(dh_index *ix) { delete ix; }
int32_t dh_index_n_ref(const dh_index *ix) { return ix ? (int32_t)ix->ix.refs.size() : 0; }
int dh_index_is_csi(const dh_index *ix) { return ix && ix->ix.csi; }
uint64_t dh_index_n_no_coor(const dh_index *ix) { return ix ? ix->ix.n_no_coor : 0; }

static int start_filtered(dh_bam *b, const dh_index *idx, int mode) {
    std::vector<IdxChunk> ranges;
    plan_ranges(idx->ix, b->filt, ranges);
    b->pending.assign(ranges.begin(), ranges.end());
    b->filt_mode = mode;
    DH_CUDA(cudaSetDevice(b->ctx->device));
    int rc = b->d_filt.reserve(std::max<size_t>(1, b->filt.size()) * sizeof(RegIntv));
    if (rc != DH_OK) return rc;
    if (!b->filt.empty()) DH_CUDA(cudaMemcpy(b->d_filt.p, b->filt.data(), b->filt.size() * sizeof(RegIntv), cudaMemcpyHostToDevice));
    if (b->pending.empty()) {
        // nothing indexed under the query: an iterator that is already finished (hts_itr_query :3467-3470, :3533-3537)
        rc = begin_range(b, 0, UINT64_MAX, true);
        b->segs.clear();
        b->chain_done = true;
        return rc;
    }
    const IdxChunk c = b->pending.front();
    b->pending.pop_front();
    return begin_range(b, c.u, c.v, true);
}

int dh_bam_scan_regions(dh_bam *b, const dh_index *idx, const char *const *regions, int n_regions) {
    if (!b || !idx || (n_regions > 0 && !regions)) { set_error("dh_bam_scan_regions: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (check_layout(b) != DH_OK) return DH_EINVAL;
    std::string err;
    const int n = regions_to_intervals(regions, n_regions, b->ref_names, b->filt, err);
    if (n < 0) { set_error("%s", err.c_str()); return DH_EINVAL; }
    if (n == 0) { set_error("no region names a reference sequence of this file"); return DH_ENOREGION; }
    return start_filtered(b, idx, 2);
}

int dh_bam_scan_contigs(dh_bam *b, const dh_index *idx, int32_t tid_lo, int32_t tid_hi) {
    if (!b || !idx) { set_error("dh_bam_scan_contigs: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (check_layout(b) != DH_OK) return DH_EINVAL;
    const int32_t n_ref = (int32_t)b->ref_names.size();
    tid_lo = std::max(tid_lo, 0);
    tid_hi = std::min(tid_hi, n_ref);
    b->filt.clear();
    for (int32_t t = tid_lo; t < tid_hi; ++t) {
        RegIntv v;
        v.tid = t; v.beg = 0; v.end = INT64_MAX; v.pad = 0;
        b->filt.push_back(v);
    }
    return start_filtered(b, idx, (tid_lo == 0 && tid_hi == n_ref) ? 1 : 2);
}

int dh_bam_split_contigs(const dh_bam *b, const dh_index *idx, int n_parts, int32_t *tid_bounds) {
    if (!b || !idx || n_parts < 1 || !tid_bounds) { set_error("dh_bam_split_contigs: bad arguments"); return DH_EINVAL; }
    // the index may cover fewer or more references than the header names: bounds are over the header's contigs
    const int32_t n_ref = (int32_t)b->ref_names.size();
    int rc = dh_index_split_contigs(idx, n_parts, tid_bounds);
    if (rc != DH_OK) return rc;
    for (int p = 0; p <= n_parts; ++p) tid_bounds[p] = std::min(tid_bounds[p], n_ref);
    tid_bounds[n_parts] = n_ref;
    return DH_OK;
}

// Pipeline step shared by the three entry points below (caller holds b->mu): refill the free slots, wait for the oldest
// batch in flight and hand its slot over as HELD.
static int acquire_locked(dh_bam *b, dh_bam_batch *out) {
    DH_CUDA(cudaSetDevice(b->ctx->device));
    // refill the free slots; `exhausted` = the current range has nothing left to issue (else the slots ran out first)
    auto top_up = [&](bool &exhausted) -> int {
        exhausted = false;
        while (free_slots(b) > 0) {
            int w = have_work(b);
            if (w < 0) return w;
            if (w == 0) { exhausted = true; break; }
            int rc = issue_batch(b);
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
        if ((rc = begin_range(b, c.u, c.v, false)) != DH_OK || (rc = top_up(exhausted)) != DH_OK) { b->scanning = false; return rc; }
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
    static const bool trace = getenv("DH_TRACE") != nullptr;
    if (trace && b->ev_scan0) {   // per-batch timeline relative to scan_begin: decode start/end, D2H start/end (ms)
        float t0 = 0, t1 = 0, t2 = 0, t3 = 0;
        cudaEventElapsedTime(&t0, b->ev_scan0, o.ev_k0);
        cudaEventElapsedTime(&t1, b->ev_scan0, o.ev_k1);
        cudaEventElapsedTime(&t2, b->ev_scan0, o.ev_c0);
        cudaEventElapsedTime(&t3, b->ev_scan0, o.ev_done);
        fprintf(stderr, "[dh_trace] batch %llu rows %llu decode %.1f-%.1f d2h %.1f-%.1f (%.1f GB/s)\n", (unsigned long long)b->returned,
                (unsigned long long)o.desc.n_rows, t0, t1, t2, t3, o.desc.bytes / 1e6 / (t3 - t2 > 0 ? t3 - t2 : 1));
    }
    rc = top_up(exhausted);  // keep the pipeline full while the caller consumes this one
    if (rc != DH_OK) { b->scanning = false; return rc; }
    *out = o.desc;
    if (b->inflight.empty() && b->chain_done) out->status = b->final_status;
    b->returned++;
    b->st.rows = b->rows_total;
    b->st.status = b->final_status;
    return 1;
}

int dh_bam_prefetch(dh_bam *b) {
    if (!b) { set_error("dh_bam_prefetch: null handle"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (!b->scanning) { set_error("dh_bam_prefetch: no scan in progress"); return DH_EINVAL; }
    DH_CUDA(cudaSetDevice(b->ctx->device));
    // uploads + inflate of the first segments are enqueued (asynchronous); the boundary walk, which needs their result on the
    // host, is left to the first acquire
    if (b->cur_seg != -1 || b->chain_done) return DH_OK;
    for (int k = 0; k < (b->fd >= 0 ? 2 : NSEG); ++k) {
        bool exists = false;
        int rc = ensure_seg(b, k, exists);
        if (rc == DH_OK && exists) rc = issue_inflate(b, k);
        if (rc != DH_OK) { b->scanning = false; return rc; }
        if (!exists) break;
    }
    return DH_OK;
}

int dh_bam_acquire_batch(dh_bam *b, dh_bam_batch *out) {
    if (!b || !out) { set_error("dh_bam_acquire_batch: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (!b->scanning) { set_error("dh_bam_acquire_batch: no scan in progress"); return DH_EINVAL; }
    return acquire_locked(b, out);
}

int dh_bam_release_batch(dh_bam *b, int32_t slot) {
    if (!b || slot < 0 || slot >= b->n_slots) { set_error("dh_bam_release_batch: bad slot"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (b->bs[slot].state != SLOT_HELD) { set_error("dh_bam_release_batch: slot %d is not held", slot); return DH_EINVAL; }
    b->bs[slot].state = SLOT_FREE;
    if (b->auto_slot == slot) b->auto_slot = -1;
    return DH_OK;
}

int dh_bam_next_batch(dh_bam *b, dh_bam_batch *out) {
    if (!b || !out) { set_error("dh_bam_next_batch: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (!b->scanning) { set_error("dh_bam_next_batch: no scan in progress"); return DH_EINVAL; }
    if (b->auto_slot >= 0 && b->bs[b->auto_slot].state == SLOT_HELD) b->bs[b->auto_slot].state = SLOT_FREE;   // "valid until the next call"
    b->auto_slot = -1;
    const int rc = acquire_locked(b, out);
    if (rc == 1) b->auto_slot = out->slot;
    return rc;
}

int dh_bam_set_slots(dh_bam *b, int n_slots) {
    if (!b || n_slots < 3 || n_slots > MAX_SLOTS) { set_error("dh_bam_set_slots: 3..%d slots", MAX_SLOTS); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (n_slots == b->n_slots) return DH_OK;
    DH_CUDA(cudaSetDevice(b->ctx->device));
    // whatever scan was under way is abandoned: its batches are forgotten once the streams are idle
    DH_CUDA(cudaStreamSynchronize(b->ctx->stream));
    for (int i = 0; i < 3; ++i) DH_CUDA(cudaStreamSynchronize(b->ctx->copy_streams[i]));
    DH_CUDA(cudaStreamSynchronize(b->ctx->inf2_stream));
    for (auto &o : b->bs) o.state = SLOT_FREE;
    b->inflight.clear();
    b->auto_slot = -1;
    b->scanning = false;
    b->n_slots = n_slots;
    return ensure_events(b);
}

// ---- per-column checksums of a held batch (parity at sizes no cell-by-cell comparison reaches) -----------------------------
static int checksum_tables(dh_bam *b) {
    if (b->checksum_tables) return DH_OK;
    b->h_ref_hash.clear(); b->h_sample_hash.clear(); b->h_dict_ptr.clear(); b->h_dict_hash.clear();
    for (auto &n : b->ref_names) b->h_ref_hash.push_back(ck_bytes((const uint8_t *)n.data(), (uint32_t)n.size()));
    b->h_ref_hash.push_back(ck_bytes((const uint8_t *)"*", 1));
    for (auto &sm : b->rg_samples) b->h_sample_hash.push_back(ck_bytes((const uint8_t *)sm.data(), (uint32_t)sm.size()));
    std::vector<std::pair<uint64_t, uint64_t>> d;
    for (auto &str : b->dict) d.emplace_back((uint64_t)(uintptr_t)str.data(), ck_bytes((const uint8_t *)str.data(), (uint32_t)str.size()));
    std::sort(d.begin(), d.end());
    for (auto &e : d) { b->h_dict_ptr.push_back(e.first); b->h_dict_hash.push_back(e.second); }
    auto up = [&](DevBuf &buf, const std::vector<uint64_t> &v) -> int {
        int rc = buf.reserve(std::max<size_t>(1, v.size()) * 8);
        if (rc != DH_OK) return rc;
        if (!v.empty()) DH_CUDA(cudaMemcpy(buf.p, v.data(), v.size() * 8, cudaMemcpyHostToDevice));
        return DH_OK;
    };
    int rc;
    if ((rc = up(b->d_ref_hash, b->h_ref_hash)) != DH_OK || (rc = up(b->d_sample_hash, b->h_sample_hash)) != DH_OK ||
        (rc = up(b->d_dict_ptr, b->h_dict_ptr)) != DH_OK || (rc = up(b->d_dict_hash, b->h_dict_hash)) != DH_OK ||
        (rc = b->d_sums.reserve(BAM_MAX_COLS * 8)) != DH_OK || (rc = b->h_sums.reserve(BAM_MAX_COLS * 8)) != DH_OK)
        return rc;
    b->checksum_tables = true;
    return DH_OK;
}

int dh_bam_batch_checksum(dh_bam *b, const dh_bam_batch *batch, uint64_t *sums) {
    if (!b || !batch || !sums || batch->slot < 0 || batch->slot >= b->n_slots || b->kind != 0) { set_error("dh_bam_batch_checksum: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    BatchSlot &o = b->bs[batch->slot];
    if (o.state != SLOT_HELD) { set_error("dh_bam_batch_checksum: the batch is not held by the caller"); return DH_EINVAL; }
    DH_CUDA(cudaSetDevice(b->ctx->device));
    int rc = checksum_tables(b);
    if (rc != DH_OK) return rc;
    ChecksumParams cp;
    memset(&cp, 0, sizeof cp);
    const int ncols = BAM_CORE_COLS + (int)b->tags.size();
    cp.n_rows = o.desc.n_rows;
    cp.heap_bytes = o.desc.heap_bytes;
    cp.compact = o.desc.layout == DH_BAM_COMPACT_OUTPUT ? 1u : 0u;
    cp.n_cols = (uint32_t)ncols;
    cp.n_ref = (int32_t)b->ref_names.size();
    cp.n_dict = (uint32_t)b->h_dict_ptr.size();
    static const uint8_t CORE_KIND[BAM_CORE_COLS] = {CK_STR, CK_U16, CK_DICT, CK_POS, CK_MAPQ, CK_STR, CK_DICT, CK_POS, CK_TLEN, CK_STR, CK_STR, CK_STR, CK_SAMPLE};
    for (int c = 0; c < ncols; ++c) {
        if (c < BAM_CORE_COLS) cp.kind[c] = CORE_KIND[c];
        else {
            const char t = (char)b->tags[(size_t)(c - BAM_CORE_COLS)].type;
            cp.kind[c] = (t == 'i' || t == 'c') ? CK_TAG_I : t == 'f' ? CK_TAG_F : t == 'A' ? CK_TAG_A : t == 'B' ? CK_TAG_B : CK_STR;
        }
    }
    if (o.on_device) {
        cp.dp = o.dp;
        cp.ref_hash = (const uint64_t *)b->d_ref_hash.p;
        cp.sample_hash = (const uint64_t *)b->d_sample_hash.p;
        cp.dict_ptr = (const uint64_t *)b->d_dict_ptr.p;
        cp.dict_hash = (const uint64_t *)b->d_dict_hash.p;
        cp.sums = (uint64_t *)b->d_sums.p;
        cudaStream_t sk = s_comp(b);
        DH_CUDA(cudaMemsetAsync(b->d_sums.p, 0, BAM_MAX_COLS * 8, sk));
        DH_CUDA(launch_bam_checksum(cp, sk));
        b->ctx->launches++;
        DH_CUDA(cudaMemcpyAsync(b->h_sums.p, b->d_sums.p, BAM_MAX_COLS * 8, cudaMemcpyDeviceToHost, sk));
        DH_CUDA(cudaStreamSynchronize(sk));
        for (int c = 0; c < ncols; ++c) sums[c] += ((const uint64_t *)b->h_sums.p)[c];
    } else {
        for (int c = 0; c < ncols; ++c) {
            cp.dp.col[c] = (uint8_t *)o.desc.data[c];
            cp.dp.valid[c] = (unsigned long long *)o.desc.validity[c];
        }
        cp.dp.heap = (uint8_t *)o.desc.heap;
        cp.dp.rec_off = (uint32_t *)o.desc.rec_off;
        cp.ref_hash = b->h_ref_hash.data();
        cp.sample_hash = b->h_sample_hash.data();
        cp.dict_ptr = b->h_dict_ptr.data();
        cp.dict_hash = b->h_dict_hash.data();
        uint64_t local[BAM_MAX_COLS] = {};
        cp.sums = local;
        static const int nthr = getenv("DH_HOST_THREADS") ? atoi(getenv("DH_HOST_THREADS")) : (int)std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
        bam_checksum_host(cp, nthr);
        for (int c = 0; c < ncols; ++c) sums[c] += local[c];
    }
    return DH_OK;
}

// ---- N2: bam_index -----------------------------------------------------------------------------------------------------------
void dh_free(void *p) { free(p); }

int dh_bam_index_build(dh_bam *b, uint8_t **bai, uint64_t *bai_size) {
    if (!b || !bai || !bai_size || b->kind != 0) { set_error("dh_bam_index_build: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    *bai = nullptr;
    *bai_size = 0;
    // the scan: every record, no column (nothing is decoded, only the boundary walk runs)
    const uint64_t keep_mask = b->col_mask;
    const bool keep_aux = b->aux_map;
    const uint32_t fr = b->flag_require, fe = b->flag_exclude, mq = b->min_mapq;
    b->col_mask = 0;
    b->aux_map = false;
    b->flag_require = b->flag_exclude = b->min_mapq = 0;
    b->filt_mode = 0;
    b->filt.clear();
    b->pending.clear();
    int rc = upload_tables(b);
    auto restore = [&]() { b->col_mask = keep_mask; b->aux_map = keep_aux; b->flag_require = fr; b->flag_exclude = fe; b->min_mapq = mq; upload_tables(b); b->scanning = false; };
    if (rc == DH_OK) rc = begin_range(b, 0, UINT64_MAX, true);
    if (rc != DH_OK) { restore(); return rc; }
    const int32_t n_ref = (int32_t)b->ref_names.size();
    // linear index on the device: per contig the windows its length needs, plus slack for reads that run over the end
    std::vector<uint32_t> lin_base((size_t)n_ref + 1, 0), lin_n((size_t)std::max(n_ref, 1), 0);
    for (int32_t t = 0; t < n_ref; ++t) {
        lin_n[(size_t)t] = (uint32_t)std::min<uint64_t>(32769, ((uint64_t)b->ref_lens[(size_t)t] >> 14) + 64);
        lin_base[(size_t)t + 1] = lin_base[(size_t)t] + lin_n[(size_t)t];
    }
    const size_t lin_total = lin_base[(size_t)n_ref];
    DevBuf d_lin, d_lin_base, d_lin_n, d_runs, d_blocks, d_counter;
    PinBuf h_runs, h_blocks;
    std::vector<IdxRun> runs;                 // the file as runs of equal (tid, bin), merged across blocks
    int64_t prev_tid = -2, prev_beg = 0;
    int err = 0;
    auto cleanup = [&]() { d_lin.release(); d_lin_base.release(); d_lin_n.release(); d_runs.release(); d_blocks.release(); d_counter.release(); h_runs.release(); h_blocks.release(); };
    if ((rc = d_lin.reserve(std::max<size_t>(1, lin_total) * 8)) != DH_OK || (rc = d_lin_base.reserve(((size_t)n_ref + 1) * 4)) != DH_OK ||
        (rc = d_lin_n.reserve(std::max<size_t>(1, (size_t)n_ref) * 4)) != DH_OK || (rc = d_counter.reserve(64)) != DH_OK) { cleanup(); restore(); return rc; }
    cudaStream_t sk = s_comp(b);
    cudaMemsetAsync(d_lin.p, 0xff, std::max<size_t>(1, lin_total) * 8, sk);
    cudaMemcpyAsync(d_lin_base.p, lin_base.data(), ((size_t)n_ref + 1) * 4, cudaMemcpyHostToDevice, sk);
    if (n_ref) cudaMemcpyAsync(d_lin_n.p, lin_n.data(), (size_t)n_ref * 4, cudaMemcpyHostToDevice, sk);
    for (;;) {
        const int w = have_work(b);            // inflates and walks the next segment (and starts the ones behind it)
        if (w < 0) { rc = w; break; }
        if (w == 0) break;
        SegSlot &s = b->ss[b->cur_seg % NSEG];
        const uint32_t n = (uint32_t)(s.b1 - s.b0) + 1;
        uint64_t n_rec = 0;
        const BlockWalk *hw = (const BlockWalk *)s.h_walk.p;
        for (uint32_t i = 0; i < n; ++i) n_rec += hw[i].count;
        if ((rc = d_runs.reserve(std::max<uint64_t>(1, n_rec) * sizeof(IdxRun))) != DH_OK || (rc = d_blocks.reserve((size_t)n * sizeof(IdxBlock))) != DH_OK ||
            (rc = h_runs.reserve(std::max<uint64_t>(1, n_rec) * sizeof(IdxRun))) != DH_OK || (rc = h_blocks.reserve((size_t)n * sizeof(IdxBlock))) != DH_OK) break;
        IndexParams ip;
        memset(&ip, 0, sizeof ip);
        ip.u = (const uint8_t *)s.out.p + b->carry_cap;
        ip.walk = hw;
        ip.blk_start = (const int64_t *)s.h_blk_start.p;
        ip.rec = (const uint16_t *)s.rec.p;
        ip.rec_bias = s.rec_bias;
        ip.n = n;
        ip.u_base = (int64_t)b->bt.uoff[s.b0];
        ip.runs = (IdxRun *)d_runs.p;
        ip.run_cap = (uint32_t)std::min<uint64_t>(n_rec, 0xffffffffu);
        ip.counter = (unsigned int *)d_counter.p;
        ip.blocks = (IdxBlock *)d_blocks.p;
        ip.lin = (unsigned long long *)d_lin.p;
        ip.lin_base = (const uint32_t *)d_lin_base.p;
        ip.lin_n = (const uint32_t *)d_lin_n.p;
        ip.n_ref = n_ref;
        // only the blocks the walk accepted count: a stop (invalid record) ends the index like it ends sam_index
        cudaMemsetAsync(d_counter.p, 0, 4, sk);
        cudaMemsetAsync(d_blocks.p, 0, (size_t)n * sizeof(IdxBlock), sk);
        cudaError_t ce = launch_bam_index(ip, b->tabs, sk);
        b->ctx->launches++;
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(h_blocks.p, d_blocks.p, (size_t)n * sizeof(IdxBlock), cudaMemcpyDeviceToHost, sk);
        if (ce == cudaSuccess && n_rec) ce = cudaMemcpyAsync(h_runs.p, d_runs.p, n_rec * sizeof(IdxRun), cudaMemcpyDeviceToHost, sk);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(sk);
        if (ce != cudaSuccess) { rc = cuda_fail(ce, "bam_index_kernel"); break; }
        // blocks in order, up to the batches the walk planned (rows behind an invalid record are not part of the file for htslib)
        uint32_t last_blk = 0;
        for (auto &bb : b->batches) last_blk = std::max(last_blk, bb.b_hi);
        const IdxBlock *hb = (const IdxBlock *)h_blocks.p;
        const IdxRun *hr = (const IdxRun *)h_runs.p;
        for (uint32_t i = 0; i < last_blk && !err; ++i) {
            const IdxBlock &ib = hb[i];
            if (!ib.count) continue;
            if (ib.err) { err = ib.err; break; }
            // sortedness across blocks (hts_idx_push :2597-2616)
            if (prev_tid != -2 && ib.first_tid == prev_tid && ib.first_tid >= 0 && prev_beg > ib.first_beg) { err = 1; break; }
            prev_tid = ib.last_tid;
            prev_beg = ib.last_beg;
            for (uint32_t q = 0; q < ib.n_runs; ++q) {
                const IdxRun &r = hr[ib.base + q];
                if (!runs.empty() && runs.back().tid == r.tid && runs.back().bin == r.bin) { runs.back().n_mapped += r.n_mapped; runs.back().n_unmapped += r.n_unmapped; }
                else runs.push_back(r);
            }
        }
        b->next_batch_i = b->batches.size();   // nothing to decode: on to the next segment
        if (err) break;
    }
    // the linear index comes back once
    std::vector<uint64_t> lin(std::max<size_t>(1, lin_total));
    if (rc == DH_OK && !err && lin_total) {
        cudaError_t ce = cudaMemcpy(lin.data(), d_lin.p, lin_total * 8, cudaMemcpyDeviceToHost);
        if (ce != cudaSuccess) rc = cuda_fail(ce, "linear index copy");
    }
    const int final_status = b->final_status;
    const uint64_t file_end = b->fd >= 0 ? b->fsize : b->bt.end_off;
    cleanup();
    restore();
    if (rc != DH_OK) return rc;
    if (err) {
        set_error(err == 1 ? "unsorted positions: bam_index needs a coordinate-sorted file" : err == 2 ? "a coordinate beyond 2^29 cannot be stored in a BAI index (use CSI)"
                  : err == 3 ? "a read lies far beyond the end of its reference sequence" : "internal: run buffer overflow");
        return DH_EFORMAT;
    }
    if (final_status < -1 || b->truncated) { set_error("corrupted or truncated BAM file (record status %d)", final_status); return DH_EFORMAT; }   // sam_index: ret < -1 -> error
    // ---- hts_idx_push / hts_idx_finish on the run sequence ----
    const BlockTable &bt = b->bt;
    // bgzf_tell with the stream read up to inflated offset u.  A read that ends exactly at a block end leaves the stream at (address of
    // the block that follows, 0) (bgzf_read, htslib/bgzf.c: block_address = bgzf_htell() once block_offset == block_length), so the
    // header's last block never appears when the header fills it, and the offset after the last record is the EOF marker's address.
    auto tell = [&](uint64_t u) -> uint64_t {
        const size_t nb = bt.blocks.size();
        if (nb == 0) return file_end << 16;
        if (u == 0) return bt.blocks[0].coffset << 16;
        const size_t blk = (size_t)(std::upper_bound(bt.uoff.begin(), bt.uoff.end(), u - 1) - bt.uoff.begin()) - 1;   // holds byte u - 1
        if (blk >= nb) return file_end << 16;
        if (u < bt.uoff[blk + 1]) return (bt.blocks[blk].coffset << 16) | (u - bt.uoff[blk]);
        return (blk + 1 < nb ? bt.blocks[blk + 1].coffset : file_end) << 16;
    };
    // contigs must form one group each; placed reads may not follow reads without a contig (:2599-2610)
    {
        std::vector<uint8_t> seen((size_t)std::max(n_ref, 1), 0);
        int32_t cur = -2;
        bool no_coor = false;
        for (auto &r : runs) {
            if (r.tid == cur) continue;
            if (r.tid >= 0 && (no_coor || seen[(size_t)r.tid])) { set_error(no_coor ? "reads without a contig are not in a single block at the end" : "chromosome blocks not continuous"); return DH_EFORMAT; }
            if (r.tid >= 0) seen[(size_t)r.tid] = 1; else no_coor = true;
            cur = r.tid;
        }
    }
    const uint64_t final_offset = tell(bt.uoff.back());     // bgzf_tell after the read that met the end of the file (hts_idx_finish)
    struct Chunk { uint64_t u, v; };
    std::vector<std::map<uint32_t, std::vector<Chunk>>> bins((size_t)std::max(n_ref, 1));
    struct Meta { uint64_t off_beg = 0, off_end = 0, n_mapped = 0, n_unmapped = 0; bool any = false; };
    std::vector<Meta> meta((size_t)std::max(n_ref, 1));
    uint64_t n_no_coor = 0;
    for (size_t i = 0; i < runs.size(); ++i) {
        const IdxRun &r = runs[i];
        const uint64_t beg = tell(r.start_u), end = i + 1 < runs.size() ? tell(runs[i + 1].start_u) : final_offset;
        if (r.tid < 0) { n_no_coor += (uint64_t)r.n_mapped + r.n_unmapped; continue; }
        bins[(size_t)r.tid][r.bin].push_back(Chunk{beg, end});
        Meta &m = meta[(size_t)r.tid];
        if (!m.any) { m.any = true; m.off_beg = beg; }
        m.off_end = end;
        m.n_mapped += r.n_mapped;
        m.n_unmapped += r.n_unmapped;
    }
    // hts_idx_finish: update_loff's backward fill of the linear index (:2433-2436), compress_binning (:2457-2508)
    std::vector<std::vector<uint64_t>> lidx((size_t)std::max(n_ref, 1));
    for (int32_t t = 0; t < n_ref; ++t) {
        const uint64_t *src = lin.data() + lin_base[(size_t)t];
        int64_t nwin = 0;
        for (int64_t wv = (int64_t)lin_n[(size_t)t] - 1; wv >= 0; --wv) if (src[wv] != UINT64_MAX) { nwin = wv + 1; break; }
        auto &L = lidx[(size_t)t];
        L.resize((size_t)nwin);
        for (int64_t wv = 0; wv < nwin; ++wv) L[(size_t)wv] = src[wv] == UINT64_MAX ? UINT64_MAX : tell(src[wv]);
        for (int64_t wv = nwin - 2; wv >= 0; --wv) if (L[(size_t)wv] == UINT64_MAX) L[(size_t)wv] = L[(size_t)wv + 1];
        auto &B = bins[(size_t)t];
        static const uint32_t LEVEL_FIRST[6] = {0, 1, 9, 73, 585, 4681};
        for (int l = 5; l > 0; --l) {
            const uint32_t start = LEVEL_FIRST[l];
            std::vector<uint32_t> keys;
            for (auto &kv : B) if (kv.first >= start && kv.first < 37449) keys.push_back(kv.first);
            for (uint32_t k : keys) {
                auto it = B.find(k);
                if (it == B.end()) continue;
                auto &pl = it->second;
                if (l < 5 && pl.size() > 1) std::sort(pl.begin(), pl.end(), [](const Chunk &a, const Chunk &c) { return a.u < c.u; });
                if ((pl.back().v >> 16) - (pl.front().u >> 16) < 0x10000ull) {      // HTS_MIN_MARKER_DIST
                    const uint32_t parent = (k - 1) >> 3;
                    auto pit = B.find(parent);
                    if (pit == B.end()) continue;
                    pit->second.insert(pit->second.end(), pl.begin(), pl.end());
                    B.erase(it);
                }
            }
        }
        auto z = B.find(0);
        if (z != B.end()) std::sort(z->second.begin(), z->second.end(), [](const Chunk &a, const Chunk &c) { return a.u < c.u; });
        for (auto &kv : B) {       // merge adjacent chunks that start from the same BGZF block
            auto &pl = kv.second;
            size_t m = 0;
            for (size_t q = 1; q < pl.size(); ++q) {
                if ((pl[m].v >> 16) >= (pl[q].u >> 16)) { if (pl[m].v < pl[q].v) pl[m].v = pl[q].v; }
                else pl[++m] = pl[q];
            }
            pl.resize(m + 1);
        }
    }
    // ---- hts_idx_save: "BAI\1", n_ref, per contig {n_bin, {bin, n_chunk, chunks}, n_intv, ioffsets}, n_no_coor ----
    std::vector<uint8_t> out;
    auto put32 = [&](uint32_t v) { for (int q = 0; q < 4; ++q) out.push_back((uint8_t)(v >> (8 * q))); };
    auto put64 = [&](uint64_t v) { for (int q = 0; q < 8; ++q) out.push_back((uint8_t)(v >> (8 * q))); };
    out.insert(out.end(), {'B', 'A', 'I', 1});
    put32((uint32_t)n_ref);
    for (int32_t t = 0; t < n_ref; ++t) {
        auto &B = bins[(size_t)t];
        const Meta &m = meta[(size_t)t];
        put32((uint32_t)(B.size() + (m.any ? 1 : 0)));
        for (auto &kv : B) {
            put32(kv.first);
            put32((uint32_t)kv.second.size());
            for (auto &c : kv.second) { put64(c.u); put64(c.v); }
        }
        if (m.any) { put32(37450u); put32(2u); put64(m.off_beg); put64(m.off_end); put64(m.n_mapped); put64(m.n_unmapped); }
        put32((uint32_t)lidx[(size_t)t].size());
        for (uint64_t v : lidx[(size_t)t]) put64(v);
    }
    put64(n_no_coor);
    uint8_t *res = (uint8_t *)malloc(out.size());
    if (!res) { set_error("out of memory"); return DH_ENOMEM; }
    memcpy(res, out.data(), out.size());
    *bai = res;
    *bai_size = out.size();
    return DH_OK;
}

int dh_index_split_contigs(const dh_index *idx, int n_parts, int32_t *tid_bounds) {
    if (!idx || n_parts < 1 || !tid_bounds) { set_error("dh_index_split_contigs: bad arguments"); return DH_EINVAL; }
    const int32_t n_ref = (int32_t)idx->ix.refs.size();
    std::vector<uint64_t> pre((size_t)n_ref + 1, 0);   // compressed bytes of contigs [0, t)
    for (int32_t t = 0; t < n_ref; ++t) {
        uint64_t u, v, w = 0;
        if (contig_span(idx->ix, t, u, v) && (v >> 16) >= (u >> 16)) w = (v >> 16) - (u >> 16) + 1;
        pre[(size_t)t + 1] = pre[(size_t)t] + w;
    }
    tid_bounds[0] = 0;
    for (int p = 1; p < n_parts; ++p) {
        const uint64_t target = pre[(size_t)n_ref] / (uint64_t)n_parts * (uint64_t)p;
        int32_t t = (int32_t)(std::lower_bound(pre.begin(), pre.end(), target) - pre.begin());
        // cut at the contig boundary nearest to the target
        if (t > 0 && target - pre[(size_t)t - 1] < pre[(size_t)std::min(t, n_ref)] - target) --t;
        tid_bounds[p] = std::min(std::max(t, tid_bounds[p - 1]), n_ref);
    }
    tid_bounds[n_parts] = n_ref;
    return DH_OK;
}

int dh_bam_scan_stats_get(const dh_bam *b, dh_bam_scan_stats *out) {
    if (!b || !out) return DH_EINVAL;
    std::lock_guard<std::mutex> lk(b->mu);
    *out = b->st;
    out->rows = b->rows_total;
    out->status = b->final_status;
    return DH_OK;
}

// compressed bytes [off, off + n) of the image, wherever it lives
static bool fetch_bytes(const dh_bam *b, uint64_t off, uint64_t n, std::vector<uint8_t> &out) {
    if (off > b->size || n > b->size - off) return false;
    out.resize((size_t)n);
    if (b->fd < 0) { memcpy(out.data(), b->data + off, (size_t)n); return true; }
    return read_range(b->fd, out.data(), n, off);
}
static inline bool hdr_ok(const uint8_t *h) {
    return h[0] == 31 && h[1] == 139 && h[2] == 8 && (h[3] & 4) && (h[10] | (h[11] << 8)) == 6 && h[12] == 'B' && h[13] == 'C' && (h[14] | (h[15] << 8)) == 2;
}

// Does a record start at offset 0 of the BGZF block at file offset `coff`?  htslib starts a block at a record boundary whenever the
// next record does not fit the current one (bgzf_flush_try, htslib/sam.c:883), but the blocks behind a record longer than a block
// start in the middle of it, and foreign writers cut anywhere (htslib/test/bgzf_boundaries).  The block is inflated (GPU) and its
// records are chained on the host: every header must be plausible and the chain must end exactly at the block end or run past it
// with its last record.  *csize_out = the block's size.
static bool block_starts_a_record(const dh_bam *b, uint64_t coff, uint32_t *csize_out) {
    std::vector<uint8_t> c;
    if (!fetch_bytes(b, coff, std::min<uint64_t>(65536, b->size - coff), c) || c.size() < 26 || !hdr_ok(c.data())) return false;
    const uint32_t csize = (uint32_t)(c[16] | (c[17] << 8)) + 1;
    if (csize < 26 || csize > c.size()) return false;
    if (csize_out) *csize_out = csize;
    const uint32_t isize = (uint32_t)c[csize - 4] | ((uint32_t)c[csize - 3] << 8) | ((uint32_t)c[csize - 2] << 16) | ((uint32_t)c[csize - 1] << 24);
    if (isize < 36 || isize > BGZF_MAX_OUT) return false;
    std::vector<uint8_t> u((size_t)isize + 8);
    uint64_t n = 0;
    if (dh_bgzf_inflate_host(b->ctx, c.data(), csize, u.data(), u.size(), &n, nullptr) != DH_OK || n != isize) return false;
    const int32_t n_ref = (int32_t)b->ref_names.size();
    auto le32 = [&](uint64_t p) { return (uint32_t)u[p] | ((uint32_t)u[p + 1] << 8) | ((uint32_t)u[p + 2] << 16) | ((uint32_t)u[p + 3] << 24); };
    uint64_t off = 0;
    int seen = 0;
    while (off + 36 <= n) {
        const uint32_t bl = le32(off);
        if (bl < 32u || bl >= (1u << 28)) return false;
        const int32_t tid = (int32_t)le32(off + 4), pos = (int32_t)le32(off + 8), l_qseq = (int32_t)le32(off + 20), mtid = (int32_t)le32(off + 24),
                      mpos = (int32_t)le32(off + 28);
        const uint32_t l_qname = u[off + 12], n_cigar = le32(off + 16) & 0xffffu;
        if (tid < -1 || tid >= n_ref || mtid < -1 || mtid >= n_ref || pos < -1 || mpos < -1 || l_qname < 1 || l_qseq < 0) return false;
        if (32ull + l_qname + 4ull * n_cigar + (((uint64_t)l_qseq + 1) >> 1) + (uint64_t)l_qseq > bl) return false;
        const uint64_t qend = off + 36 + l_qname - 1;
        if (qend < n && u[qend] != 0) return false;
        off += 4ull + bl;
        ++seen;
    }
    return seen > 0 && off >= n;      // off < n: fewer than 36 bytes left that are no record header
}

// file offset of the first BGZF block at or behind `target` (north-star item (a): the BSIZE chain cannot be entered in the middle,
// so a candidate is every offset that carries the 16 fixed header bytes and whose chain of BSIZE fields meets two more such headers
// or the end of the file); UINT64_MAX if there is none in the next 256 KB
static uint64_t find_block_at_or_after(const dh_bam *b, uint64_t target) {
    if (b->fd < 0) {
        const size_t blk = block_of(b, target);
        return blk < b->bt.blocks.size() ? b->bt.blocks[blk].coffset : UINT64_MAX;
    }
    std::vector<uint8_t> w;
    const uint64_t n = std::min<uint64_t>(256u << 10, b->size - std::min(target, b->size));
    if (n < 28 || !fetch_bytes(b, target, n, w)) return UINT64_MAX;
    for (uint64_t o = 0; o + 28 <= n; ++o) {
        if (!hdr_ok(w.data() + o)) continue;
        uint64_t q = o;
        int ok = 0;
        for (int hop = 0; hop < 3; ++hop) {
            const uint64_t bs = (uint64_t)(w[q + 16] | (w[q + 17] << 8)) + 1;
            if (bs < 26) break;
            q += bs;
            if (target + q == b->size) { ok = 3; break; }              // the chain ends exactly at the end of the file
            if (q + 18 > n) { ok = hop + 1 >= 2 ? 3 : ok; break; }      // ran out of window after at least two good hops
            if (!hdr_ok(w.data() + q)) break;
            ok = hop + 1;
        }
        if (ok >= 2) return target + o;
    }
    return UINT64_MAX;
}

int dh_bam_split(const dh_bam *b, int n_parts, uint64_t *voffs) {
    if (!b || n_parts < 1 || !voffs) { set_error("dh_bam_split: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(b->mu);
    if (cudaSetDevice(b->ctx->device) != cudaSuccess) { set_error("cudaSetDevice failed"); return DH_ECUDA; }
    voffs[0] = dh_bam_first_voffset(b);
    const uint64_t end = b->fd >= 0 ? b->fsize : b->bt.end_off;
    // the block holding the first record stays whole in part 0; cut the rest by compressed bytes
    uint64_t c0 = std::min(end, voffs[0] >> 16);
    if (voffs[0] & 0xffff) {      // first record in the middle of a block: cuts start behind that block
        uint32_t cs = 0;
        std::vector<uint8_t> h;
        if (fetch_bytes(b, c0, std::min<uint64_t>(18, end - c0), h) && h.size() == 18 && hdr_ok(h.data())) cs = (uint32_t)(h[16] | (h[17] << 8)) + 1;
        c0 = std::min(end, c0 + cs);
    }
    uint64_t prev = c0;
    bool collapsed = false;
    for (int p = 1; p < n_parts; ++p) {
        const uint64_t target = std::max(prev, c0 + (end - c0) * (uint64_t)p / (uint64_t)n_parts);
        // a cut must be a record boundary: move on to the next block that verifiably starts one (a few tries: the blocks behind a long
        // read); a file whose blocks do not start records at all (foreign writer) is not cut: part 0 gets everything
        uint64_t coff = collapsed ? UINT64_MAX : find_block_at_or_after(b, target);
        for (int tries = 0; coff != UINT64_MAX && coff < end; ++tries) {
            uint32_t cs = 0;
            if (block_starts_a_record(b, coff, &cs)) break;
            if (cs == 0 || tries >= 64) { collapsed = true; coff = UINT64_MAX; break; }
            coff += cs;                                                // the next block of the chain
            if (coff + 28 > end) coff = UINT64_MAX;                    // only the EOF marker (or nothing) is left
        }
        voffs[p] = (collapsed || coff == UINT64_MAX || coff >= end) ? UINT64_MAX : (coff << 16);
        if (voffs[p] != UINT64_MAX) prev = coff;
    }
    voffs[n_parts] = UINT64_MAX;
    return DH_OK;
}

}  // extern "C"

// ---- pipeline steps shared with dh_bcf.cu ----------------------------------------------------------------------
namespace dh {
int scan_ensure_events(dh_bam *b) { return ensure_events(b); }
int scan_begin_range(dh_bam *b, uint64_t voff_beg, uint64_t voff_end, bool fresh) { return begin_range(b, voff_beg, voff_end, fresh); }
int scan_have_work(dh_bam *b) { return have_work(b); }
int scan_take_slot(dh_bam *b) { return take_slot(b); }
void scan_release_slot(dh_bam *b, int slot) { b->bs[slot].state = SLOT_FREE; if (b->auto_slot == slot) b->auto_slot = -1; }
int scan_free_slots(const dh_bam *b) { return free_slots(b); }
int scan_start_filtered(dh_bam *b, const dh_index *idx, int mode) { return start_filtered(b, idx, mode); }
int scan_inflate_head(dh_bam *b, uint64_t want, std::vector<uint8_t> &u) {
    const size_t nb = b->bt.blocks.size();
    size_t take = 0;
    while (take < nb && b->bt.uoff[take] < want) ++take;
    const uint64_t cend = take ? b->bt.blocks[take - 1].coffset + b->bt.blocks[take - 1].csize : 0;
    u.resize((size_t)b->bt.uoff[take] + 8);
    uint64_t n = 0;
    int rc = take ? dh_bgzf_inflate_host(b->ctx, b->data, cend, u.data(), u.size(), &n, nullptr) : DH_OK;
    if (rc != DH_OK) return rc;
    u.resize((size_t)n);
    return DH_OK;
}
}  // namespace dh
