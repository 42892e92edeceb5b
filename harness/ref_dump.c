/* ref_dump.c — runs the REFERENCE's own code (libhts.a from /root/reference, built by oracle/Makefile) over a
 * BAM/SAM file and dumps, per record, exactly the cells src/bam_reader.c:779-966 would store, through the same
 * htslib calls (sam_read1, bam_get_*, seq_nt16_str, bam_aux_get, bam_aux2i/2f/2A/2Z, sam_hdr_find_tag_id,
 * kputw/bam_cigar_opchr).  TEST INFRASTRUCTURE: generates tests/golden/*.tsv and pins oracle/dh_oracle.c.
 *
 *   ref_dump dump FILE [TAG:type ...]     one line per record, tab separated, NULL = \N, bytes outside
 *                                         [0x21,0x7e] and '\' as \xHH; last line "#status <code> <n>"
 *   ref_dump sam2bam IN OUT               convert (for turning htslib's .sam fixtures into BAM)
 *   ref_dump bgunzip FILE                 raw bgzf_read() stream to stdout
 *   ref_dump dumpregion FILE REGIONS [TAG:type ...]   like dump, but through sam_index_load3 + sam_itr_regarray +
 *                                         sam_itr_next exactly as src/bam_reader.c:642-671,749-752 (REGIONS = the
 *                                         `region` parameter: comma separated); "#no-iterator" if regarray fails
 *   ref_dump dumpcontigs FILE [TAG:type ...]          the index-driven scan without a region: sam_itr_queryi per
 *                                         contig in order (bam_reader.c:689-716), empty contigs skipped
 *   ref_dump index FILE MIN_SHIFT         sam_index_build (0 = BAI, 14 = CSI)
 */
#include <inttypes.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "htslib/bgzf.h"
#include "htslib/hts.h"
#include "htslib/kstring.h"
#include "htslib/sam.h"

static void put_esc(const char *s, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        unsigned char c = (unsigned char)s[i];
        if (c < 0x21 || c > 0x7e || c == '\\') printf("\\x%02x", c);
        else putchar(c);
    }
}
static void put_cstr(const char *s) { put_esc(s, strlen(s)); }

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: ref_dump dump|sam2bam|bgunzip ...\n"); return 2; }
    if (!strcmp(argv[1], "bgunzip")) {
        BGZF *fp = bgzf_open(argv[2], "r");
        if (!fp) return 1;
        static char buf[65536];
        ssize_t n;
        while ((n = bgzf_read(fp, buf, sizeof buf)) > 0) fwrite(buf, 1, (size_t)n, stdout);
        int bad = n < 0;
        bgzf_close(fp);
        return bad;
    }
    if (!strcmp(argv[1], "sam2bam")) {
        samFile *in = sam_open(argv[2], "r"), *out = sam_open(argv[3], "wb");
        if (!in || !out) return 1;
        sam_hdr_t *h = sam_hdr_read(in);
        if (!h || sam_hdr_write(out, h) < 0) return 1;
        bam1_t *b = bam_init1();
        int r;
        while ((r = sam_read1(in, h, b)) >= 0)
            if (sam_write1(out, h, b) < 0) return 1;
        sam_close(in);
        return sam_close(out) < 0 || r < -1;
    }
    if (!strcmp(argv[1], "index")) return sam_index_build(argv[2], argc > 3 ? atoi(argv[3]) : 0) < 0;
    const int by_region = !strcmp(argv[1], "dumpregion"), by_contig = !strcmp(argv[1], "dumpcontigs");
    const int tag0 = by_region ? 4 : 3;
    samFile *fp = sam_open(argv[2], "r");
    if (!fp) { fprintf(stderr, "cannot open %s\n", argv[2]); return 1; }
    sam_hdr_t *hdr = sam_hdr_read(fp);
    if (!hdr) { printf("#header-invalid\n"); return 0; }
    bam1_t *b = bam_init1();
    kstring_t ks = {0, 0, NULL}, sm = {0, 0, NULL};
    char *buf = malloc(1 << 16);
    size_t cap = 1 << 16;
    uint64_t n = 0;
    int ret;
    printf("#n_ref %d\n", sam_hdr_nref(hdr));
    for (int i = 0; i < sam_hdr_nref(hdr); ++i) { printf("#ref "); put_cstr(sam_hdr_tid2name(hdr, i)); printf(" %" PRId64 "\n", (int64_t)sam_hdr_tid2len(hdr, i)); }
    hts_idx_t *idx = NULL;
    hts_itr_t *itr = NULL;
    int contig = -1;
    if (by_region || by_contig) {
        idx = sam_index_load3(fp, argv[2], NULL, HTS_IDX_SILENT_FAIL);
        if (!idx) { printf("#no-index\n"); return 0; }
    }
    if (by_region) {
        char *dup = strdup(argv[3]), *regs[256];
        unsigned nreg = 0;
        for (char *tok = strtok(dup, ","); tok && nreg < 256; tok = strtok(NULL, ",")) regs[nreg++] = tok;   /* parse_regions, bam_reader.c:318-348 */
        itr = sam_itr_regarray(idx, hdr, regs, nreg);
        if (!itr) { printf("#no-iterator\n"); return 0; }
    }
    for (;;) {
        if (by_contig && !itr) {
            if (++contig >= sam_hdr_nref(hdr)) { ret = -1; break; }
            itr = sam_itr_queryi(idx, contig, 0, HTS_POS_MAX);
            if (!itr) continue;
        }
        ret = itr ? sam_itr_next(fp, itr, b) : sam_read1(fp, hdr, b);
        if (ret < 0) {
            if (by_contig && ret == -1) { hts_itr_destroy(itr); itr = NULL; continue; }
            break;
        }
        int l = b->core.l_qseq;
        if ((size_t)l + 1 > cap) { cap = (size_t)l + 1; buf = realloc(buf, cap); }
        put_cstr(bam_get_qname(b));
        printf("\t%u\t", (unsigned)b->core.flag);
        put_cstr(b->core.tid >= 0 ? sam_hdr_tid2name(hdr, b->core.tid) : "*");
        printf("\t%" PRId64 "\t%d\t", (int64_t)b->core.pos + 1, (int)b->core.qual);
        if (b->core.n_cigar > 0) {
            ks.l = 0;
            const uint32_t *cig = bam_get_cigar(b);
            for (uint32_t i = 0; i < b->core.n_cigar; ++i) { kputw((int)bam_cigar_oplen(cig[i]), &ks); kputc(bam_cigar_opchr(bam_cigar_op(cig[i])), &ks); }
            put_esc(ks.s, ks.l);
        } else putchar('*');
        putchar('\t');
        put_cstr(b->core.mtid >= 0 ? sam_hdr_tid2name(hdr, b->core.mtid) : "*");
        printf("\t%" PRId64 "\t%" PRId64 "\t", (int64_t)b->core.mpos + 1, (int64_t)b->core.isize);
        if (l > 0) { for (int i = 0; i < l; ++i) buf[i] = seq_nt16_str[bam_seqi(bam_get_seq(b), i)]; buf[l] = 0; put_cstr(buf); } else putchar('*');
        putchar('\t');
        if (l > 0 && bam_get_qual(b)[0] != 255) { for (int i = 0; i < l; ++i) buf[i] = (char)(bam_get_qual(b)[i] + 33); buf[l] = 0; put_cstr(buf); } else putchar('*');
        putchar('\t');
        uint8_t *aux = bam_aux_get(b, "RG");
        const char *rg = aux ? bam_aux2Z(aux) : NULL;
        if (rg) put_cstr(rg); else printf("\\N");
        putchar('\t');
        sm.l = 0;
        if (rg && sam_hdr_find_tag_id(hdr, "RG", "ID", rg, "SM", &sm) == 0 && sm.s && sm.l > 0) put_cstr(sm.s); else printf("\\N");
        for (int t = tag0; t < argc; ++t) {
            const char *tag = argv[t];
            char type = tag[3];
            putchar('\t');
            uint8_t *a = bam_aux_get(b, tag);
            if (!a) { printf("\\N"); continue; }
            if (type == 'i') printf("%" PRId64, bam_aux2i(a));
            else if (type == 'f') { double d = bam_aux2f(a); uint64_t u; memcpy(&u, &d, 8); printf("%016" PRIx64, u); }
            else if (type == 'A') { char tmp[2] = {bam_aux2A(a), 0}; put_cstr(tmp); if (!tmp[0]) printf("\\e"); }
            else { const char *z = bam_aux2Z(a); if (z) { put_cstr(z); if (!z[0]) printf("\\e"); } else printf("\\N"); }
        }
        putchar('\n');
        ++n;
    }
    printf("#status %d %" PRIu64 "\n", ret < -1 ? ret : 0, n);
    return 0;
}
