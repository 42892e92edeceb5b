/* gen_synth.c — seeded synthetic-input generator that writes THROUGH THE VENDORED htslib (BGZF level 6).
 *
 * TEST/BENCH INFRASTRUCTURE (lives under oracle/ because it links the reference's libhts.a from
 * oracle/_ref; it is never on the product path).  BASELINE.json: "All inputs are synthetic BAM/BCF
 * files from a seeded generator written by the vendored htslib at BGZF level 6".  SURVEY.md §8(d)
 * defines the record mix.
 *
 *   gen_synth bam  OUT N SEED [--paired] [--contigs C] [--part P/K] [--no-header] [--no-eof] [--index]
 *                       [--unplaced M] [--threads T] [--btags]
 *       coordinate-sorted BAM, C contigs chr1..chrC of 50 Mb, N records in total (split evenly over
 *       contigs; every contig gets >= 1 read).  --part P/K writes only the contigs of part P
 *       of K (contig c belongs to part c*K/C), so K processes can write in parallel; the caller
 *       concatenates the parts (BGZF members concatenate) — part 0 carries the header, only the last
 *       part keeps the EOF block.  Record content depends only on (SEED, contig), never on K.
 *   gen_synth bgzf OUT N SEED [--paired]
 *       the same record stream (header included) written with plain bgzf_write() in 64 KiB calls, so
 *       BGZF block edges fall anywhere inside records (config 2 payload; also exercises the
 *       record-straddles-block path of the boundary resolver).
 */
#include <inttypes.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "htslib/bgzf.h"
#include "htslib/hts.h"
#include "htslib/kstring.h"
#include "htslib/sam.h"

#define CONTIG_LEN 50000000
#define READ_LEN 150

typedef struct { uint64_t s; } rng_t;
static inline uint64_t rng_next(rng_t *r) {
    uint64_t x = r->s;
    x ^= x << 13; x ^= x >> 7; x ^= x << 17;
    r->s = x;
    return x;
}
static inline uint32_t rng_below(rng_t *r, uint32_t n) { return (uint32_t)((rng_next(r) >> 11) % n); }
static void rng_seed(rng_t *r, uint64_t seed, uint64_t stream) {
    r->s = seed ^ (0x9E3779B97F4A7C15ull * (stream + 1));
    if (!r->s) r->s = 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < 8; ++i) rng_next(r);
}

typedef struct {
    int64_t pos;
    int64_t mpos;
    int32_t isize;
    uint16_t flag;
    uint64_t pair_id;
    int rg;
} pending_t;

static const char *RG_ID[4] = {"rg1", "rg2", "rg3", "rg4"};
static const char *RG_SM[4] = {"sampleA", "sampleA", "sampleB", "sampleB"};

static sam_hdr_t *make_header(int n_contigs) {
    sam_hdr_t *h = sam_hdr_init();
    kstring_t ks = {0, 0, NULL};
    ksprintf(&ks, "@HD\tVN:1.6\tSO:coordinate\n");
    for (int c = 0; c < n_contigs; ++c) ksprintf(&ks, "@SQ\tSN:chr%d\tLN:%d\n", c + 1, CONTIG_LEN);
    for (int g = 0; g < 4; ++g) ksprintf(&ks, "@RG\tID:%s\tSM:%s\tPL:SYNTH\n", RG_ID[g], RG_SM[g]);
    ksprintf(&ks, "@PG\tID:gen_synth\tPN:gen_synth\tVN:1\n");
    if (sam_hdr_add_lines(h, ks.s, ks.l) < 0) { fprintf(stderr, "header build failed\n"); exit(1); }
    free(ks.s);
    return h;
}

/* fill one record.  unmapped: FUNMAP set, placed at mate's coordinates, no CIGAR, no NM/MD/AS */
static int g_btags = 0;   /* --btags: B-array tags (CG, FZ, ML of the SAMtags table and others) + tags of every other aux type */

/* B arrays of every sub-type under the three LIST columns of standard_tags (CG, FZ, ML), their declared sub-type or not
 * (a float array under an integer LIST column), empty arrays, and non-standard tags for auxiliary_tags */
static void add_btags(bam1_t *b, rng_t *r) {
    uint8_t buf[4 * 64];
    const uint32_t k = rng_below(r, 16);
    if (k < 9) {   /* ML: base-modification likelihoods, B:C, 0..40 values */
        const uint32_t n = rng_below(r, 8) == 0 ? 0 : 1 + rng_below(r, 40);
        for (uint32_t i = 0; i < n; ++i) buf[i] = (uint8_t)rng_below(r, 256);
        bam_aux_update_array(b, "ML", 'C', n, buf);
    } else if (k == 9) {   /* ML as floats: the writer then stores doubles in the BIGINT child */
        float f[5];
        const uint32_t n = 1 + rng_below(r, 5);
        for (uint32_t i = 0; i < n; ++i) f[i] = (float)rng_below(r, 1000) / 8.0f - 20.0f;
        bam_aux_update_array(b, "ML", 'f', n, f);
    } else if (k == 10) {   /* ML as signed bytes / shorts */
        int16_t h[6];
        int8_t c[6];
        const uint32_t n = 1 + rng_below(r, 6);
        for (uint32_t i = 0; i < n; ++i) { h[i] = (int16_t)(rng_below(r, 65536) - 32768); c[i] = (int8_t)(rng_below(r, 256) - 128); }
        if (rng_below(r, 2)) bam_aux_update_array(b, "ML", 's', n, h); else bam_aux_update_array(b, "ML", 'c', n, c);
    }
    if (rng_below(r, 4) == 0) {   /* CG: BAM-only long CIGAR, B:I */
        uint32_t cg[12];
        const uint32_t n = 1 + rng_below(r, 12);
        for (uint32_t i = 0; i < n; ++i) cg[i] = (uint32_t)rng_next(r);
        bam_aux_update_array(b, "CG", 'I', n, cg);
    }
    if (rng_below(r, 5) == 0) {   /* FZ: flow signal intensities, B:S; sometimes B:i with negative values */
        uint16_t fz[20];
        int32_t fi[4];
        const uint32_t n = 1 + rng_below(r, 20);
        for (uint32_t i = 0; i < n; ++i) fz[i] = (uint16_t)rng_below(r, 65536);
        for (uint32_t i = 0; i < 4; ++i) fi[i] = (int32_t)rng_next(r);
        if (rng_below(r, 6)) bam_aux_update_array(b, "FZ", 'S', n, fz); else bam_aux_update_array(b, "FZ", 'i', 4, fi);
    }
    /* non-standard tags of every aux type (for the auxiliary_tags MAP): XA:A XC:c XS:s XI:I Xf:f XZ:Z XH:H XB:B:i XD:B:f */
    const uint32_t m = rng_below(r, 64);
    if (m & 1) { char a = (char)('!' + rng_below(r, 90)); bam_aux_append(b, "XA", 'A', 1, (uint8_t *)&a); }
    if (m & 2) bam_aux_update_int(b, "XC", (int64_t)rng_below(r, 200) - 100);
    if (m & 4) bam_aux_update_int(b, "XS", (int64_t)rng_below(r, 60000) - 30000);
    if (m & 8) bam_aux_update_int(b, "XI", (int64_t)(uint32_t)rng_next(r));
    if (m & 16) {
        static const float FV[] = {0.0f, 1.0f, -1.5f, 3.14159274f, 1e-5f, 1e-4f, 123456.0f, 1234567.0f, 1e20f, -2.5e-10f, 0.1f, 100000.0f, 999999.5f, 0.000123456f};
        const uint32_t pick = rng_below(r, 20);
        bam_aux_update_float(b, "Xf", pick < 14 ? FV[pick] : (float)rng_below(r, 1000000) / 997.0f);
    }
    if (m & 32) {
        char z[40];
        const uint32_t n = rng_below(r, 30);
        for (uint32_t i = 0; i < n; ++i) z[i] = (char)('a' + rng_below(r, 26));
        z[n] = 0;
        bam_aux_update_str(b, "XZ", -1, z);
    }
    if (rng_below(r, 8) == 0) { const char *hx = "1AE301"; bam_aux_append(b, "XH", 'H', 7, (const uint8_t *)hx); }
    if (rng_below(r, 8) == 0) {
        int32_t v[3] = {(int32_t)rng_next(r), -1, 7};
        bam_aux_update_array(b, "XB", 'i', 1 + rng_below(r, 3), v);
    }
    if (rng_below(r, 8) == 0) {
        float v[3] = {(float)rng_below(r, 100000) / 64.0f, 1e-7f, -3.0e9f};
        bam_aux_update_array(b, "XD", 'f', 1 + rng_below(r, 3), v);
    }
}

static void fill_record(bam1_t *b, rng_t *r, int tid, int64_t pos, uint16_t flag, int mtid, int64_t mpos,
                        int32_t isize, uint64_t id, int rg, int ctg_label) {
    char qname[64];
    int lq = snprintf(qname, sizeof qname, "SYN:%d:%" PRIu64 ":%u", ctg_label, id, (unsigned)(rng_next(r) & 0xfff));
    char seq[READ_LEN], qual[READ_LEN];
    static const char ACGT[4] = {'A', 'C', 'G', 'T'};
    for (int i = 0; i < READ_LEN; i += 16) {
        uint64_t x = rng_next(r), y = rng_next(r);
        for (int j = 0; j < 16 && i + j < READ_LEN; ++j) {
            seq[i + j] = ((y >> (j * 4)) & 0x3ff) == 0 ? 'N' : ACGT[(x >> (2 * j)) & 3];
            /* 4-level qualities 37/25/11/2 with 70/15/10/5 % */
            uint32_t u = (uint32_t)((x >> (32 + 2 * (j & 7))) ^ (y >> (40 + j))) % 100;
            qual[i + j] = u < 70 ? 37 : u < 85 ? 25 : u < 95 ? 11 : 2;
        }
    }
    uint32_t cigar[2];
    int n_cigar = 0;
    int unmapped = (flag & BAM_FUNMAP) != 0;
    if (!unmapped) {
        if (rng_below(r, 10) == 0) {
            uint32_t k = 1 + rng_below(r, 40);
            cigar[0] = bam_cigar_gen(k, BAM_CSOFT_CLIP);
            cigar[1] = bam_cigar_gen(READ_LEN - k, BAM_CMATCH);
            n_cigar = 2;
        } else {
            cigar[0] = bam_cigar_gen(READ_LEN, BAM_CMATCH);
            n_cigar = 1;
        }
    }
    int mapq = unmapped ? 0 : (int)rng_below(r, 61);
    if (bam_set1(b, lq, qname, flag, tid, pos, mapq, n_cigar, cigar, mtid, mpos, isize, READ_LEN, seq, qual, 64) < 0) {
        fprintf(stderr, "bam_set1 failed\n");
        exit(1);
    }
    if (!unmapped) {
        int nm = rng_below(r, 4) == 0 ? (int)rng_below(r, 6) : 0;
        bam_aux_update_int(b, "NM", nm);
        char md[32];
        if (nm == 0) snprintf(md, sizeof md, "%d", READ_LEN);
        else { int k = 1 + rng_below(r, READ_LEN - 2); snprintf(md, sizeof md, "%d%c%d", k, ACGT[rng_below(r, 4)], READ_LEN - 1 - k); }
        bam_aux_update_str(b, "MD", -1, md);
        bam_aux_update_int(b, "AS", (int)(READ_LEN - 5 * nm - rng_below(r, 10) + (rng_below(r, 50) == 0 ? 200 : 0)));
    }
    bam_aux_update_str(b, "RG", -1, RG_ID[rg]);
    if (g_btags) add_btags(b, r);
}

typedef int (*emit_fn)(void *ctx, bam1_t *b);

/* generate every record of contig `c` (n_ctg records), in coordinate order */
static void gen_contig(int c, uint64_t n_ctg, uint64_t seed, int paired, emit_fn emit, void *ctx) {
    rng_t r;
    rng_seed(&r, seed, (uint64_t)c);
    bam1_t *b = bam_init1();
    pending_t pend[256];
    int npend = 0;
    uint64_t emitted = 0, id = 0;
    /* mean step so the reads span the contig */
    uint64_t starts = paired ? (n_ctg + 1) / 2 : n_ctg;
    uint32_t step = (uint32_t)((CONTIG_LEN - 2000) / (starts ? starts : 1));
    if (step < 1) step = 1;
    int64_t pos = 0;
    while (emitted < n_ctg) {
        pos += (int64_t)rng_below(&r, 2 * step) ;
        if (pos > CONTIG_LEN - 1000) pos = CONTIG_LEN - 1000;
        /* flush pending mates that sort before this start */
        for (;;) {
            int best = -1;
            for (int i = 0; i < npend; ++i)
                if (pend[i].pos <= pos && (best < 0 || pend[i].pos < pend[best].pos)) best = i;
            if (best < 0 || emitted >= n_ctg) break;
            pending_t p = pend[best];
            pend[best] = pend[--npend];
            fill_record(b, &r, c, p.pos, p.flag, c, p.mpos, p.isize, p.pair_id, p.rg, c + 1);
            emit(ctx, b);
            ++emitted;
        }
        if (emitted >= n_ctg) break;
        int rg = (int)rng_below(&r, 4);
        if (!paired) {
            uint16_t flag = rng_below(&r, 2) ? BAM_FREVERSE : 0;
            fill_record(b, &r, c, pos, flag, -1, -1, 0, id++, rg, c + 1);
            emit(ctx, b);
            ++emitted;
        } else {
            int32_t frag = READ_LEN + 150 + (int32_t)rng_below(&r, 351); /* mate offset 150..500 */
            int64_t mpos = pos + frag - READ_LEN;
            int mate_unmapped = rng_below(&r, 100) == 0;
            uint16_t f1 = BAM_FPAIRED | BAM_FREAD1 | BAM_FMREVERSE | (mate_unmapped ? BAM_FMUNMAP : BAM_FPROPER_PAIR);
            uint16_t f2 = BAM_FPAIRED | BAM_FREAD2 | (mate_unmapped ? BAM_FUNMAP : (BAM_FPROPER_PAIR | BAM_FREVERSE));
            if (mate_unmapped) { f1 &= ~BAM_FMREVERSE; }
            fill_record(b, &r, c, pos, f1, c, mate_unmapped ? pos : mpos, mate_unmapped ? 0 : frag, id, rg, c + 1);
            emit(ctx, b);
            ++emitted;
            if (npend < 256) {
                pending_t p;
                p.pos = mate_unmapped ? pos : mpos; /* unmapped mate is placed at its mate's position */
                p.mpos = pos;
                p.isize = mate_unmapped ? 0 : -frag;
                p.flag = f2;
                p.pair_id = id;
                p.rg = rg;
                pend[npend++] = p;
            }
            ++id;
        }
    }
    bam_destroy1(b);
}

static int emit_bam(void *ctx, bam1_t *b) {
    if (bam_write1((BGZF *)ctx, b) < 0) { fprintf(stderr, "bam_write1 failed\n"); exit(1); }
    return 0;
}

/* bgzf mode: serialise records ourselves and push them with plain bgzf_write (no per-record flush) */
typedef struct { BGZF *fp; uint8_t *buf; size_t n, cap; } raw_ctx;
static void raw_flush(raw_ctx *rc) {
    size_t off = 0;
    while (off < rc->n) {
        size_t k = rc->n - off > 65536 ? 65536 : rc->n - off;
        if (bgzf_write(rc->fp, rc->buf + off, k) < 0) { fprintf(stderr, "bgzf_write failed\n"); exit(1); }
        off += k;
    }
    rc->n = 0;
}
static int emit_raw(void *ctx, bam1_t *b) {
    raw_ctx *rc = (raw_ctx *)ctx;
    const bam1_core_t *c = &b->core;
    uint32_t block_len = b->l_data - c->l_extranul + 32;
    if (rc->n + block_len + 4 > rc->cap) raw_flush(rc);
    uint8_t *p = rc->buf + rc->n;
    uint32_t x[9];
    x[0] = block_len;
    x[1] = (uint32_t)c->tid;
    x[2] = (uint32_t)c->pos;
    x[3] = (uint32_t)c->bin << 16 | c->qual << 8 | (c->l_qname - c->l_extranul);
    x[4] = (uint32_t)c->flag << 16 | (c->n_cigar & 0xffff);
    x[5] = (uint32_t)c->l_qseq;
    x[6] = (uint32_t)c->mtid;
    x[7] = (uint32_t)c->mpos;
    x[8] = (uint32_t)c->isize;
    memcpy(p, x, 36);
    memcpy(p + 36, b->data, c->l_qname - c->l_extranul);
    memcpy(p + 36 + c->l_qname - c->l_extranul, b->data + c->l_qname, b->l_data - c->l_qname);
    rc->n += 4 + block_len;
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 5) {
        fprintf(stderr, "usage: gen_synth bam|bgzf OUT N SEED [--paired] [--contigs C] [--part P/K] [--no-header] "
                        "[--no-eof] [--index] [--unplaced M] [--threads T] [--btags]\n");
        return 2;
    }
    const char *mode = argv[1], *out = argv[2];
    uint64_t n = strtoull(argv[3], NULL, 10), seed = strtoull(argv[4], NULL, 0);
    int paired = 0, n_contigs = 22, part = 0, nparts = 1, no_header = 0, no_eof = 0, do_index = 0, threads = 0;
    uint64_t unplaced = 0;
    for (int i = 5; i < argc; ++i) {
        if (!strcmp(argv[i], "--paired")) paired = 1;
        else if (!strcmp(argv[i], "--contigs") && i + 1 < argc) n_contigs = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--part") && i + 1 < argc) sscanf(argv[++i], "%d/%d", &part, &nparts);
        else if (!strcmp(argv[i], "--no-header")) no_header = 1;
        else if (!strcmp(argv[i], "--no-eof")) no_eof = 1;
        else if (!strcmp(argv[i], "--index")) do_index = 1;
        else if (!strcmp(argv[i], "--unplaced") && i + 1 < argc) unplaced = strtoull(argv[++i], NULL, 10);
        else if (!strcmp(argv[i], "--threads") && i + 1 < argc) threads = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--btags")) g_btags = 1;
        else { fprintf(stderr, "unknown option %s\n", argv[i]); return 2; }
    }
    sam_hdr_t *hdr = make_header(n_contigs);
    BGZF *fp = bgzf_open(out, "w6");
    if (!fp) { fprintf(stderr, "cannot open %s\n", out); return 1; }
    if (threads > 1) bgzf_mt(fp, threads, 256);
    int raw = !strcmp(mode, "bgzf");
    raw_ctx rc = {fp, NULL, 0, 1 << 20};
    if (raw) rc.buf = (uint8_t *)malloc(rc.cap);
    if (!no_header) {
        if (raw) {
            /* header bytes through the same unaligned writer: magic, l_text, text, n_ref, refs */
            const char *text = sam_hdr_str(hdr);
            uint32_t l_text = (uint32_t)sam_hdr_length(hdr), n_ref = (uint32_t)sam_hdr_nref(hdr);
            bgzf_write(fp, "BAM\1", 4);
            bgzf_write(fp, &l_text, 4);
            bgzf_write(fp, text, l_text);
            bgzf_write(fp, &n_ref, 4);
            for (uint32_t i = 0; i < n_ref; ++i) {
                const char *nm = sam_hdr_tid2name(hdr, (int)i);
                uint32_t l = (uint32_t)strlen(nm) + 1, ln = (uint32_t)sam_hdr_tid2len(hdr, (int)i);
                bgzf_write(fp, &l, 4);
                bgzf_write(fp, nm, l);
                bgzf_write(fp, &ln, 4);
            }
        } else if (bam_hdr_write(fp, hdr) < 0) { fprintf(stderr, "bam_hdr_write failed\n"); return 1; }
    }
    for (int c = 0; c < n_contigs; ++c) {
        if ((int)((int64_t)c * nparts / n_contigs) != part) continue;
        uint64_t n_ctg = n / n_contigs + ((uint64_t)c < n % n_contigs ? 1 : 0);
        if (n_ctg == 0) n_ctg = 1;
        gen_contig(c, n_ctg, seed, paired, raw ? emit_raw : emit_bam, raw ? (void *)&rc : (void *)fp);
    }
    if (unplaced && part == nparts - 1) {
        /* unplaced unmapped reads (tid = -1) at the end of the file: exercises SURVEY Q1 */
        rng_t r;
        rng_seed(&r, seed, 1000003);
        bam1_t *b = bam_init1();
        for (uint64_t i = 0; i < unplaced; ++i) {
            fill_record(b, &r, -1, -1, BAM_FUNMAP, -1, -1, 0, i, (int)rng_below(&r, 4), 0);
            if (raw) emit_raw(&rc, b); else emit_bam(fp, b);
        }
        bam_destroy1(b);
    }
    if (raw) raw_flush(&rc);
    if (bgzf_close(fp) < 0) { fprintf(stderr, "bgzf_close failed\n"); return 1; }
    if (no_eof) {
        /* drop the 28-byte EOF marker so parts can be concatenated */
        FILE *f = fopen(out, "rb+");
        fseek(f, 0, SEEK_END);
        long sz = ftell(f);
        fclose(f);
        if (truncate(out, sz - 28) != 0) { perror("truncate"); return 1; }
    }
    if (do_index && !raw) {
        if (sam_index_build3(out, NULL, 0, threads > 1 ? threads : 0) < 0) { fprintf(stderr, "index build failed\n"); return 1; }
    }
    sam_hdr_destroy(hdr);
    return 0;
}
