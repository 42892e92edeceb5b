// dh_text.cu — C-ABI of the text path: a BGZF-compressed, tab-separated text file read as batches of lines
// with every projected field located on the GPU.
//
// Reference: read_tabix's sequential scan (src/tabix_reader.c:852-1030): hts_getline per line, get_field per cell.  Here the file is
// read in segments of whole BGZF blocks: pread -> H2D -> inflate (the kernels of A2) behind the unfinished last line of the previous
// segment -> text_count / text_lines / text_fields (text_split.cu) -> D2H of the finished lines and their cell offsets.  The host
// side of the table function (duckhts_ext.c) applies what depends on a line's number (line_skip, the header line) and converts typed
// columns with strtoll / strtod like the reference; the GPU finds every line and every field.
#include <errno.h>
#include <fcntl.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include "bgzf_inflate.cuh"
#include "dh_internal.h"
#include "text_split.cuh"

using namespace dh;

struct dh_text {
    dh_ctx *ctx = nullptr;
    int fd = -1;
    uint64_t fsize = 0, foff = 0;
    bool file_done = false, finished = false, scanning = false;
    int status = 0;
    uint32_t n_cols = 1, n_proj = 1;
    std::vector<uint16_t> slot;
    int meta = '#';
    uint64_t seg_comp = 64ull << 20;
    uint64_t carry = 0, line0 = 0;
    PinBuf h_comp, h_text, h_cells, h_flags, h_desc, h_status, h_misc;
    DevBuf d_comp, d_out, d_desc, d_status, d_work, d_counts, d_nl, d_cells, d_flags, d_slot, d_carry;
    std::mutex mu;
};

static inline bool hdr_ok(const uint8_t *h) {
    return h[0] == 31 && h[1] == 139 && h[2] == 8 && (h[3] & 4) && (h[10] | (h[11] << 8)) == 6 && h[12] == 'B' && h[13] == 'C' && (h[14] | (h[15] << 8)) == 2;
}
static inline uint32_t rd32(const uint8_t *p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

// whole BGZF blocks at the front of buf[0, n): table + how many compressed bytes they take; stops at `max_out` inflated bytes
static bool frame(const uint8_t *buf, uint64_t n, uint64_t max_out, BlockTable &bt, uint64_t &used, bool &bad) {
    bt.blocks.clear();
    bt.uoff.assign(1, 0);
    uint64_t off = 0, u = 0;
    bad = false;
    while (n - off >= 18) {
        if (!hdr_ok(buf + off)) { bad = true; break; }
        const uint32_t bsize = (uint32_t)(buf[off + 16] | (buf[off + 17] << 8)) + 1;
        if (bsize < 26) { bad = true; break; }
        if (n - off < bsize) break;
        dh_bgzf_block b;
        b.coffset = off; b.csize = bsize; b.crc = rd32(buf + off + bsize - 8); b.isize = rd32(buf + off + bsize - 4); b.reserved = 0;
        if (b.isize > 65536u) { bad = true; break; }
        if (u + b.isize > max_out && !bt.blocks.empty()) break;
        bt.blocks.push_back(b);
        u += b.isize;
        bt.uoff.push_back(u);
        off += bsize;
    }
    used = off;
    bt.end_off = off;
    return !bt.blocks.empty();
}

extern "C" {

dh_text *dh_text_open_file(dh_ctx *c, const char *path) {
    if (!c || !path) { set_error("dh_text_open_file: bad arguments"); return nullptr; }
    const int fd = open(path, O_RDONLY);
    struct stat sb;
    if (fd < 0 || fstat(fd, &sb) != 0 || !S_ISREG(sb.st_mode)) { set_error("cannot open %s: %s", path, strerror(errno)); if (fd >= 0) close(fd); return nullptr; }
    uint8_t h[18];
    if (sb.st_size < 18 || pread(fd, h, 18, 0) != 18 || !hdr_ok(h)) { set_error("%s is not BGZF-compressed: the text path reads bgzip'ed files only", path); close(fd); return nullptr; }
    dh_text *t = new dh_text();
    t->ctx = c;
    t->fd = fd;
    t->fsize = (uint64_t)sb.st_size;
    if (getenv("DH_TEXT_SEG_KB")) t->seg_comp = std::max<uint64_t>(64, (uint64_t)atoll(getenv("DH_TEXT_SEG_KB"))) << 10;
    t->slot.assign(1, 0);
    return t;
}

void dh_text_close(dh_text *t) {
    if (!t) return;
    cudaSetDevice(t->ctx->device);
    cudaStreamSynchronize(t->ctx->stream);
    for (PinBuf *b : {&t->h_comp, &t->h_text, &t->h_cells, &t->h_flags, &t->h_desc, &t->h_status, &t->h_misc}) b->release();
    for (DevBuf *b : {&t->d_comp, &t->d_out, &t->d_desc, &t->d_status, &t->d_work, &t->d_counts, &t->d_nl, &t->d_cells, &t->d_flags, &t->d_slot, &t->d_carry}) b->release();
    if (t->fd >= 0) close(t->fd);
    delete t;
}

// the first inflated bytes of the file (whole blocks, at most `cap`): what bind needs to see the header and the first lines
int dh_text_head(dh_text *t, uint8_t *buf, uint64_t cap, uint64_t *n) {
    if (!t || !buf || !n) { set_error("dh_text_head: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(t->mu);
    *n = 0;
    const uint64_t want = std::min<uint64_t>(t->fsize, std::max<uint64_t>(cap, 1 << 16) + 65536);
    std::vector<uint8_t> comp((size_t)want);
    if (!dh_parallel_io(t->fd, comp.data(), want, 0, false)) { set_error("read: %s", strerror(errno)); return DH_EIO; }
    BlockTable bt;
    uint64_t used = 0;
    bool bad = false;
    if (!frame(comp.data(), want, cap, bt, used, bad)) return DH_OK;
    if (bt.uoff.back() > cap) {                         // a single block larger than cap: hand out what fits
        std::vector<uint8_t> tmp((size_t)bt.uoff.back());
        uint64_t got = 0;
        const int rc = dh_bgzf_inflate_host(t->ctx, comp.data(), used, tmp.data(), tmp.size(), &got, nullptr);
        if (rc != DH_OK) return rc;
        *n = std::min<uint64_t>(got, cap);
        memcpy(buf, tmp.data(), (size_t)*n);
        return DH_OK;
    }
    return dh_bgzf_inflate_host(t->ctx, comp.data(), used, buf, cap, n, nullptr);
}

int dh_text_configure(dh_text *t, uint32_t n_cols, const uint32_t *proj, uint32_t n_proj, int meta_char) {
    if (!t || n_cols == 0 || n_cols > 65534u || (n_proj && !proj)) { set_error("dh_text_configure: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(t->mu);
    t->n_cols = n_cols;
    t->slot.assign(n_cols, 0xffff);
    for (uint32_t i = 0; i < n_proj; ++i) {
        if (proj[i] >= n_cols || t->slot[proj[i]] != 0xffff) { set_error("dh_text_configure: projected column %u out of range or named twice", proj[i]); return DH_EINVAL; }
        t->slot[proj[i]] = (uint16_t)i;
    }
    t->n_proj = n_proj;
    t->meta = meta_char;
    return DH_OK;
}

int dh_text_scan_begin(dh_text *t) {
    if (!t) { set_error("dh_text_scan_begin: null handle"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(t->mu);
    DH_CUDA(cudaSetDevice(t->ctx->device));
    DH_CUDA(cudaStreamSynchronize(t->ctx->stream));
    int rc = t->d_slot.reserve((size_t)t->n_cols * 2);
    if (rc != DH_OK) return rc;
    DH_CUDA(cudaMemcpy(t->d_slot.p, t->slot.data(), (size_t)t->n_cols * 2, cudaMemcpyHostToDevice));
    t->foff = 0;
    t->file_done = t->finished = false;
    t->status = 0;
    t->carry = 0;
    t->line0 = 0;
    t->scanning = true;
    return DH_OK;
}

int dh_text_next_batch(dh_text *t, dh_text_batch *out) {
    if (!t || !out) { set_error("dh_text_next_batch: bad arguments"); return DH_EINVAL; }
    std::lock_guard<std::mutex> lk(t->mu);
    if (!t->scanning) { set_error("dh_text_next_batch: no scan in progress"); return DH_EINVAL; }
    memset(out, 0, sizeof *out);
    dh_ctx *c = t->ctx;
    DH_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;
    int rc;
    for (;;) {
        if (t->finished) { out->status = t->status; return 0; }
        // ---- the next segment of whole blocks ----
        uint64_t u = 0;
        size_t nb = 0;
        bool carry_placed = false;
        if (!t->file_done) {
            const uint64_t want = std::min<uint64_t>(t->seg_comp + 65536, t->fsize - t->foff);
            if ((rc = t->h_comp.reserve(t->seg_comp + 65536 + 64)) != DH_OK) return rc;
            if (want && !dh_parallel_io(t->fd, (uint8_t *)t->h_comp.p, want, t->foff, false)) { set_error("read: %s", strerror(errno)); return DH_EIO; }
            BlockTable bt;
            uint64_t used = 0;
            bool bad = false;
            const bool any = want && frame((const uint8_t *)t->h_comp.p, want, 1ull << 31, bt, used, bad);
            if (!any) {
                // the end of the file, a cut block or a broken header: like a failing bgzf_read, the lines in front are all there is
                if (t->foff < t->fsize) t->status = -2;
                t->file_done = true;
            } else {
                nb = bt.blocks.size();
                u = bt.uoff.back();
                if ((rc = t->d_comp.reserve(used + DH_COMP_PAD)) != DH_OK || (rc = t->d_out.reserve(t->carry + u + 4096)) != DH_OK ||
                    (rc = t->d_desc.reserve(nb * sizeof(BlockDesc))) != DH_OK || (rc = t->d_status.reserve(nb * sizeof(int))) != DH_OK ||
                    (rc = t->h_desc.reserve(nb * sizeof(BlockDesc))) != DH_OK || (rc = t->h_status.reserve(nb * sizeof(int))) != DH_OK)
                    return rc;
                const uint64_t cw = fill_descs(bt, 0, nb, 0, 0, t->carry, (BlockDesc *)t->h_desc.p);
                DH_CUDA(cudaMemcpyAsync(t->d_comp.p, t->h_comp.p, used, cudaMemcpyHostToDevice, s));
                DH_CUDA(cudaMemsetAsync((uint8_t *)t->d_comp.p + used, 0, DH_COMP_PAD, s));
                DH_CUDA(cudaMemcpyAsync(t->d_desc.p, t->h_desc.p, nb * sizeof(BlockDesc), cudaMemcpyHostToDevice, s));
                if (t->carry) DH_CUDA(cudaMemcpyAsync(t->d_out.p, t->d_carry.p, t->carry, cudaMemcpyDeviceToDevice, s));
                carry_placed = true;
                if ((rc = inflate_blocks(c, (const uint8_t *)t->d_comp.p, (const BlockDesc *)t->d_desc.p, (uint32_t)nb, cw, (uint8_t *)t->d_out.p,
                                         (int *)t->d_status.p, t->d_work, (unsigned int *)c->counter.p, s)) != DH_OK)
                    return rc;
                DH_CUDA(cudaMemcpyAsync(t->h_status.p, t->d_status.p, nb * sizeof(int), cudaMemcpyDeviceToHost, s));
                DH_CUDA(cudaStreamSynchronize(s));
                const int *hs = (const int *)t->h_status.p;
                for (size_t i = 0; i < nb; ++i)
                    if (hs[i] != 0) {                       // a block that fails inflate / CRC: the bytes in front of it count, the scan ends there
                        u = bt.uoff[i];
                        t->status = -2;
                        t->file_done = true;
                        break;
                    }
                t->foff += used;
                if (t->foff >= t->fsize) t->file_done = true;
            }
        }
        if (!carry_placed && t->carry) {
            // nothing (more) to read: the unfinished last line is all there is
            if ((rc = t->d_out.reserve(t->carry + 4096)) != DH_OK) return rc;
            DH_CUDA(cudaMemcpyAsync(t->d_out.p, t->d_carry.p, t->carry, cudaMemcpyDeviceToDevice, s));
        }
        uint64_t n = t->carry + u;
        if (n == 0) { if (t->file_done) t->finished = true; continue; }
        if (n >= (1ull << 32) - 64) { set_error("a text segment of %llu bytes: lines this long are not served", (unsigned long long)n); return DH_ENOMEM; }
        uint8_t *text = (uint8_t *)t->d_out.p;
        if ((rc = t->h_misc.reserve(64)) != DH_OK) return rc;
        if (t->file_done) {
            // a last line without '\n' is a line (hts_getline returns it): give it one
            DH_CUDA(cudaMemcpyAsync(t->h_misc.p, text + n - 1, 1, cudaMemcpyDeviceToHost, s));
            DH_CUDA(cudaStreamSynchronize(s));
            if (*(const uint8_t *)t->h_misc.p != '\n') { DH_CUDA(cudaMemsetAsync(text + n, '\n', 1, s)); ++n; }
        }
        DH_CUDA(cudaMemsetAsync(text + n, 0, 64, s));           // the 16-byte loads behind the end read zeros
        // ---- lines ----
        const uint64_t chunks = text_chunks(n);
        if ((rc = t->d_counts.reserve((chunks + 1) * 4)) != DH_OK) return rc;
        DH_CUDA(launch_text_count(text, n, (uint32_t *)t->d_counts.p, s));
        c->launches += 2;
        DH_CUDA(cudaMemcpyAsync(t->h_misc.p, (const uint32_t *)t->d_counts.p + chunks, 4, cudaMemcpyDeviceToHost, s));
        DH_CUDA(cudaStreamSynchronize(s));
        const uint64_t n_lines = *(const uint32_t *)t->h_misc.p;
        if (n_lines == 0) {
            // not one complete line in this segment: all of it waits for the next one
            if ((rc = t->d_carry.reserve(n + 64)) != DH_OK) return rc;
            DH_CUDA(cudaMemcpyAsync(t->d_carry.p, text, n, cudaMemcpyDeviceToDevice, s));
            DH_CUDA(cudaStreamSynchronize(s));
            t->carry = n;
            if (t->file_done) t->finished = true;               // cannot happen (a '\n' was appended), but never loop
            continue;
        }
        if ((rc = t->d_nl.reserve(n_lines * 4)) != DH_OK || (rc = t->d_cells.reserve(std::max<uint64_t>(1, n_lines * t->n_proj * 8))) != DH_OK ||
            (rc = t->d_flags.reserve(n_lines)) != DH_OK)
            return rc;
        DH_CUDA(launch_text_lines(text, n, (const uint32_t *)t->d_counts.p, (uint32_t *)t->d_nl.p, s));
        FieldParams fp;
        fp.text = text; fp.nl = (const uint32_t *)t->d_nl.p; fp.n_lines = n_lines; fp.n_cols = t->n_cols; fp.n_proj = t->n_proj;
        fp.slot = (const uint16_t *)t->d_slot.p; fp.cells = (uint32_t *)t->d_cells.p; fp.flags = (uint8_t *)t->d_flags.p; fp.meta = t->meta;
        DH_CUDA(launch_text_fields(fp, s));
        c->launches += 2;
        DH_CUDA(cudaMemcpyAsync(t->h_misc.p, (const uint32_t *)t->d_nl.p + (n_lines - 1), 4, cudaMemcpyDeviceToHost, s));
        DH_CUDA(cudaStreamSynchronize(s));
        const uint64_t text_bytes = (uint64_t)*(const uint32_t *)t->h_misc.p + 1;        // up to and including the last '\n'
        if ((rc = t->h_text.reserve(text_bytes + 64)) != DH_OK || (rc = t->h_cells.reserve(std::max<uint64_t>(1, n_lines * t->n_proj * 8))) != DH_OK ||
            (rc = t->h_flags.reserve(n_lines)) != DH_OK)
            return rc;
        DH_CUDA(cudaMemcpyAsync(t->h_text.p, text, text_bytes, cudaMemcpyDeviceToHost, s));
        if (t->n_proj) DH_CUDA(cudaMemcpyAsync(t->h_cells.p, t->d_cells.p, n_lines * t->n_proj * 8, cudaMemcpyDeviceToHost, s));
        DH_CUDA(cudaMemcpyAsync(t->h_flags.p, t->d_flags.p, n_lines, cudaMemcpyDeviceToHost, s));
        const uint64_t tail = n - text_bytes;
        if (tail) {
            if ((rc = t->d_carry.reserve(tail + 64)) != DH_OK) return rc;
            DH_CUDA(cudaMemcpyAsync(t->d_carry.p, text + text_bytes, tail, cudaMemcpyDeviceToDevice, s));
        }
        DH_CUDA(cudaStreamSynchronize(s));
        t->carry = tail;
        if (t->file_done && tail == 0) t->finished = true;
        out->n_lines = n_lines;
        out->line0 = t->line0;
        out->text = (const uint8_t *)t->h_text.p;
        out->text_bytes = text_bytes;
        out->cells = (const uint32_t *)t->h_cells.p;
        out->flags = (const uint8_t *)t->h_flags.p;
        out->n_proj = t->n_proj;
        out->status = 0;
        t->line0 += n_lines;
        return 1;
    }
}

}  // extern "C"
