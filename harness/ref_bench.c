/* ref_bench.c — times the REFERENCE's own CPU code (libhts.a compiled from /root/reference by
 * oracle/Makefile) for the hot path.  BENCH INFRASTRUCTURE: bench.py's cpu_baseline / --impl reference legs.
 *
 *   ref_bench bgunzip FILE THREADS [MAX_OUT_BYTES]
 *       the bgzf_read(64 KiB) loop of src/bgzip.c:154-200 with bgzf_mt(fp, THREADS, 256) when THREADS > 1,
 *       output discarded; stops after MAX_OUT_BYTES inflated bytes if given.
 *   ref_bench sam_read1 FILE THREADS [MAX_RECORDS]
 *       sam_open + hts_set_threads(fp, THREADS) + sam_read1 loop with the column conversions of
 *       src/bam_reader.c:779-966 done through the same htslib calls (second CPU baseline beside the
 *       reference extension under harness/fake_duckdb).
 * Prints one JSON line: {"mode":..,"threads":..,"seconds":..,"bytes_out":..,"records":..}
 */
#include <inttypes.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "htslib/bgzf.h"
#include "htslib/hts.h"
#include "htslib/kstring.h"
#include "htslib/sam.h"

static double now(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + ts.tv_nsec * 1e-9;
}

int main(int argc, char **argv) {
    if (argc < 4) { fprintf(stderr, "usage: ref_bench bgunzip|sam_read1 FILE THREADS [LIMIT]\n"); return 2; }
    const char *mode = argv[1], *path = argv[2];
    int threads = atoi(argv[3]);
    uint64_t limit = argc > 4 ? strtoull(argv[4], NULL, 10) : 0;
    if (!strcmp(mode, "bgunzip")) {
        BGZF *fp = bgzf_open(path, "r");
        if (!fp) { fprintf(stderr, "cannot open %s\n", path); return 1; }
        if (threads > 1) bgzf_mt(fp, threads, 256);
        static uint8_t buf[65536];
        uint64_t total = 0, sum = 0;
        double t0 = now();
        for (;;) {
            ssize_t n = bgzf_read(fp, buf, sizeof buf);
            if (n < 0) { fprintf(stderr, "bgzf_read error\n"); return 1; }
            if (n == 0) break;
            total += (uint64_t)n;
            sum += buf[0] + buf[n - 1];
            if (limit && total >= limit) break;
        }
        double t1 = now();
        bgzf_close(fp);
        printf("{\"mode\":\"bgunzip\",\"threads\":%d,\"seconds\":%.6f,\"bytes_out\":%" PRIu64 ",\"records\":0,\"sink\":%" PRIu64 "}\n",
               threads, t1 - t0, total, sum & 1);
        return 0;
    }
    if (!strcmp(mode, "sam_read1")) {
        samFile *fp = sam_open(path, "r");
        if (!fp) { fprintf(stderr, "cannot open %s\n", path); return 1; }
        if (threads > 0) hts_set_threads(fp, threads);
        sam_hdr_t *hdr = sam_hdr_read(fp);
        if (!hdr) return 1;
        bam1_t *b = bam_init1();
        kstring_t ks = {0, 0, NULL};
        char *seq = malloc(1 << 20), *qual = malloc(1 << 20);
        uint64_t n = 0, sink = 0;
        double t0 = now();
        while (sam_read1(fp, hdr, b) >= 0) {
            int l = b->core.l_qseq;
            ks.l = 0;
            const uint32_t *cig = bam_get_cigar(b);
            for (uint32_t i = 0; i < b->core.n_cigar; ++i) { kputw(bam_cigar_oplen(cig[i]), &ks); kputc(bam_cigar_opchr(cig[i]), &ks); }
            const uint8_t *s = bam_get_seq(b), *q = bam_get_qual(b);
            for (int i = 0; i < l; ++i) seq[i] = seq_nt16_str[bam_seqi(s, i)];
            for (int i = 0; i < l; ++i) qual[i] = (char)(q[i] + 33);
            const char *rn = b->core.tid >= 0 ? sam_hdr_tid2name(hdr, b->core.tid) : "*";
            uint8_t *a;
            int64_t nm = (a = bam_aux_get(b, "NM")) ? bam_aux2i(a) : -1, as = (a = bam_aux_get(b, "AS")) ? bam_aux2i(a) : -1;
            const char *md = (a = bam_aux_get(b, "MD")) ? bam_aux2Z(a) : NULL, *rg = (a = bam_aux_get(b, "RG")) ? bam_aux2Z(a) : NULL;
            sink += (uint64_t)(ks.l + (l ? seq[0] + qual[l - 1] : 0) + rn[0] + nm + as + (md ? md[0] : 0) + (rg ? rg[0] : 0) + strlen(bam_get_qname(b)));
            ++n;
            if (limit && n >= limit) break;
        }
        double t1 = now();
        printf("{\"mode\":\"sam_read1\",\"threads\":%d,\"seconds\":%.6f,\"bytes_out\":0,\"records\":%" PRIu64 ",\"sink\":%" PRIu64 "}\n",
               threads, t1 - t0, n, sink & 1);
        return 0;
    }
    fprintf(stderr, "unknown mode %s\n", mode);
    return 2;
}
