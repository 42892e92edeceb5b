/* gen_bcf.c — seeded synthetic BCF generator that writes THROUGH THE VENDORED htslib (BGZF level 6).
 * TEST/BENCH INFRASTRUCTURE (links the reference's libhts.a from oracle/_ref; never on the product path).
 * Shape = BASELINE.json configs[4] / SURVEY.md §8(d) config 5: biallelic SNV sites over C contigs x S samples,
 * FORMAT GT (diploid, mostly phased) : DP : GQ : AD (Number=R); INFO AC (A) AN (1) AF (A, Float) DP (1) DB (Flag);
 * header ##fileformat=VCFv4.2.
 *
 *   gen_bcf OUT N_SITES N_SAMPLES SEED [--contigs C] [--edge] [--index]
 *     --edge   adds what the plain shape never produces: multi-allelic and indel sites (long REF/ALT strings),
 *              IDs, failing FILTERs, missing QUAL, missing / partially missing FORMAT values, haploid and missing
 *              genotypes, records that omit FORMAT fields, INFO string / string-list / fixed-count fields,
 *              FORMAT float, string and Number=2 fields.
 *     --index  also writes OUT.csi (bcf_index_build)
 */
#include <inttypes.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "htslib/hts.h"
#include "htslib/kstring.h"
#include "htslib/vcf.h"

typedef struct { uint64_t s; } rng_t;
static inline uint64_t rng_next(rng_t *r) { uint64_t x = r->s; x ^= x << 13; x ^= x >> 7; x ^= x << 17; r->s = x; return x; }
static inline uint32_t rng_below(rng_t *r, uint32_t n) { return (uint32_t)((rng_next(r) >> 11) % n); }
static inline double rng_unit(rng_t *r) { return (double)(rng_next(r) >> 11) / 9007199254740992.0; }

int main(int argc, char **argv) {
    if (argc < 5) { fprintf(stderr, "usage: gen_bcf OUT N_SITES N_SAMPLES SEED [--contigs C] [--edge] [--index]\n"); return 2; }
    const char *out = argv[1];
    uint64_t n_sites = strtoull(argv[2], NULL, 10);
    int n_samples = atoi(argv[3]);
    uint64_t seed = strtoull(argv[4], NULL, 0);
    int n_contigs = 22, edge = 0, do_index = 0;
    for (int i = 5; i < argc; ++i) {
        if (!strcmp(argv[i], "--contigs") && i + 1 < argc) n_contigs = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--edge")) edge = 1;
        else if (!strcmp(argv[i], "--index")) do_index = 1;
        else { fprintf(stderr, "unknown option %s\n", argv[i]); return 2; }
    }
    htsFile *fp = hts_open(out, "wb6");
    if (!fp) { fprintf(stderr, "cannot open %s\n", out); return 1; }
    bcf_hdr_t *h = bcf_hdr_init("w");
    bcf_hdr_set_version(h, "VCFv4.2");
    char line[512];
    for (int c = 0; c < n_contigs; ++c) { snprintf(line, sizeof line, "##contig=<ID=chr%d,length=50000000>", c + 1); bcf_hdr_append(h, line); }
    bcf_hdr_append(h, "##FILTER=<ID=q10,Description=\"Quality below 10\">");
    bcf_hdr_append(h, "##FILTER=<ID=s50,Description=\"Less than 50% of samples have data\">");
    bcf_hdr_append(h, "##INFO=<ID=AC,Number=A,Type=Integer,Description=\"Allele count\">");
    bcf_hdr_append(h, "##INFO=<ID=AN,Number=1,Type=Integer,Description=\"Allele number\">");
    bcf_hdr_append(h, "##INFO=<ID=AF,Number=A,Type=Float,Description=\"Allele frequency\">");
    bcf_hdr_append(h, "##INFO=<ID=DP,Number=1,Type=Integer,Description=\"Total depth\">");
    bcf_hdr_append(h, "##INFO=<ID=DB,Number=0,Type=Flag,Description=\"dbSNP membership\">");
    if (edge) {
        bcf_hdr_append(h, "##INFO=<ID=AA,Number=1,Type=String,Description=\"Ancestral allele\">");
        bcf_hdr_append(h, "##INFO=<ID=ANNO,Number=.,Type=String,Description=\"Free annotations, comma separated\">");
        bcf_hdr_append(h, "##INFO=<ID=SB,Number=4,Type=Integer,Description=\"Strand bias counts\">");
        bcf_hdr_append(h, "##INFO=<ID=MQ,Number=1,Type=Float,Description=\"Mapping quality\">");
    }
    bcf_hdr_append(h, "##FORMAT=<ID=GT,Number=1,Type=String,Description=\"Genotype\">");
    bcf_hdr_append(h, "##FORMAT=<ID=DP,Number=1,Type=Integer,Description=\"Read depth\">");
    bcf_hdr_append(h, "##FORMAT=<ID=GQ,Number=1,Type=Integer,Description=\"Genotype quality\">");
    bcf_hdr_append(h, "##FORMAT=<ID=AD,Number=R,Type=Integer,Description=\"Allelic depths\">");
    if (edge) {
        bcf_hdr_append(h, "##FORMAT=<ID=GL,Number=G,Type=Float,Description=\"Genotype likelihoods\">");
        bcf_hdr_append(h, "##FORMAT=<ID=HQ,Number=2,Type=Integer,Description=\"Haplotype quality\">");
        bcf_hdr_append(h, "##FORMAT=<ID=FT,Number=1,Type=String,Description=\"Sample filter\">");
        bcf_hdr_append(h, "##FORMAT=<ID=PQ,Number=1,Type=Float,Description=\"Phasing quality (float on purpose)\">");
    }
    for (int s = 0; s < n_samples; ++s) { snprintf(line, sizeof line, "S%04d", s + 1); bcf_hdr_add_sample(h, line); }
    bcf_hdr_add_sample(h, NULL);
    if (bcf_hdr_write(fp, h) < 0) { fprintf(stderr, "header write failed\n"); return 1; }

    bcf1_t *rec = bcf_init();
    const int ns = n_samples > 0 ? n_samples : 1;
    int32_t *gt = malloc(sizeof(int32_t) * ns * 2), *dp = malloc(sizeof(int32_t) * ns), *gq = malloc(sizeof(int32_t) * ns),
            *ad = malloc(sizeof(int32_t) * ns * 4), *hq = malloc(sizeof(int32_t) * ns * 2);
    float *gl = malloc(sizeof(float) * ns * 10), *pq = malloc(sizeof(float) * ns);
    const char **ft = malloc(sizeof(char *) * ns);
    static const char BASES[] = "ACGT";
    for (int c = 0; c < n_contigs; ++c) {
        uint64_t n_ctg = n_sites / n_contigs + ((uint64_t)c < n_sites % n_contigs ? 1 : 0);
        rng_t r;
        r.s = seed ^ (0x9E3779B97F4A7C15ull * (uint64_t)(c + 1));
        if (!r.s) r.s = 1;
        for (int i = 0; i < 8; ++i) rng_next(&r);
        int64_t pos = 10000;
        for (uint64_t k = 0; k < n_ctg; ++k) {
            bcf_clear(rec);
            pos += 1 + rng_below(&r, (uint32_t)(45000000 / (n_ctg ? n_ctg : 1)) + 1);
            rec->rid = c;
            rec->pos = pos;
            const int kind = edge ? (int)rng_below(&r, 16) : 0;   /* 0..9 plain SNV; the rest: edge shapes */
            int n_allele = 2;
            char als[256];
            char refb = BASES[rng_below(&r, 4)], altb = BASES[rng_below(&r, 4)];
            if (altb == refb) altb = BASES[(strchr(BASES, refb) - BASES + 1) & 3];
            if (kind == 10) { n_allele = 3; snprintf(als, sizeof als, "%c,%c,%c", refb, altb, BASES[(strchr(BASES, altb) - BASES + 1) & 3]); }
            else if (kind == 11) { snprintf(als, sizeof als, "%cACGTACGTACGTACGTAC,%c", refb, refb); }          /* long deletion */
            else if (kind == 12) { n_allele = 4; snprintf(als, sizeof als, "%c,%cTTTTTTTTTTTTTTTTTTTTTT,%cG,<DEL>", refb, refb, refb); }
            else if (kind == 13) { n_allele = 1; snprintf(als, sizeof als, "%c", refb); }                           /* no ALT */
            else snprintf(als, sizeof als, "%c,%c", refb, altb);
            bcf_update_alleles_str(h, rec, als);
            if (edge && rng_below(&r, 3) == 0) { snprintf(line, sizeof line, "rs%u%s", rng_below(&r, 100000000), kind == 11 ? ";alt_identifier_x" : ""); bcf_update_id(h, rec, line); }
            if (edge && rng_below(&r, 10) == 0) bcf_float_set_missing(rec->qual); else rec->qual = (float)(rng_below(&r, 10000) / 10.0);
            if (edge && rng_below(&r, 5) == 0) {
                int32_t flt[2] = {bcf_hdr_id2int(h, BCF_DT_ID, "q10"), bcf_hdr_id2int(h, BCF_DT_ID, "s50")};
                bcf_update_filter(h, rec, flt, 1 + (int)rng_below(&r, 2));
            } else if (!edge || rng_below(&r, 2)) { int32_t pass = bcf_hdr_id2int(h, BCF_DT_ID, "PASS"); bcf_update_filter(h, rec, &pass, 1); }
            /* genotypes: AF ~ skewed to rare */
            const double u = rng_unit(&r), af = u * u * u * 0.5;
            int32_t ac[3] = {0, 0, 0}, an = 0, tot_dp = 0;
            const int ploidy = (kind == 14) ? 1 : 2;
            for (int s = 0; s < n_samples; ++s) {
                int a[2];
                for (int p = 0; p < 2; ++p) a[p] = rng_unit(&r) < af ? 1 + (n_allele > 2 ? (int)rng_below(&r, (uint32_t)n_allele - 1) : 0) : 0;
                if (n_allele == 1) a[0] = a[1] = 0;
                const int phased = rng_below(&r, 100) < 95, missing = rng_below(&r, 200) == 0;
                if (ploidy == 1) gt[s] = missing ? bcf_gt_missing : bcf_gt_unphased(a[0]);
                else if (missing) { gt[2 * s] = bcf_gt_missing; gt[2 * s + 1] = (edge && (s & 1)) ? bcf_int32_vector_end : bcf_gt_missing; }
                else { gt[2 * s] = bcf_gt_unphased(a[0]); gt[2 * s + 1] = phased ? bcf_gt_phased(a[1]) : bcf_gt_unphased(a[1]); }
                if (!missing) { an += ploidy; for (int p = 0; p < ploidy; ++p) if (a[p]) ac[a[p] - 1 < 3 ? a[p] - 1 : 2]++; }
                dp[s] = (edge && rng_below(&r, 50) == 0) ? bcf_int32_missing : (int32_t)(5 + rng_below(&r, kind == 15 ? 70000 : 60));
                gq[s] = (edge && rng_below(&r, 50) == 0) ? bcf_int32_missing : (int32_t)rng_below(&r, 100);
                if (dp[s] != bcf_int32_missing) tot_dp += dp[s] & 0xffff;
                for (int j = 0; j < n_allele; ++j) ad[s * n_allele + j] = (int32_t)rng_below(&r, 40);
                if (edge && rng_below(&r, 40) == 0) { ad[s * n_allele] = bcf_int32_missing; if (n_allele > 1) ad[s * n_allele + n_allele - 1] = bcf_int32_vector_end; }
                if (edge) {
                    const int ng = n_allele * (n_allele + 1) / 2;
                    for (int j = 0; j < ng; ++j) gl[s * ng + j] = -(float)(rng_below(&r, 1000) / 10.0);
                    if (rng_below(&r, 30) == 0) bcf_float_set_missing(gl[s * ng]);
                    if (rng_below(&r, 30) == 0 && ng > 1) bcf_float_set_vector_end(gl[s * ng + ng - 1]);
                    hq[2 * s] = (int32_t)rng_below(&r, 60); hq[2 * s + 1] = rng_below(&r, 20) ? (int32_t)rng_below(&r, 60) : bcf_int32_missing;
                    if (rng_below(&r, 25) == 0) hq[2 * s] = bcf_int32_missing;
                    ft[s] = rng_below(&r, 10) ? "PASS" : (rng_below(&r, 2) ? "LowGQ;LowDP;StrandBias" : ".");
                    pq[s] = (float)(rng_below(&r, 500) / 10.0);
                    if (rng_below(&r, 20) == 0) bcf_float_set_missing(pq[s]);
                }
            }
            if (n_allele > 1) {
                bcf_update_info_int32(h, rec, "AC", ac, n_allele - 1);
                float afv[3];
                for (int j = 0; j < n_allele - 1; ++j) afv[j] = an ? (float)ac[j < 3 ? j : 2] / (float)an : 0.f;
                bcf_update_info_float(h, rec, "AF", afv, n_allele - 1);
            }
            if (!(edge && kind == 13)) bcf_update_info_int32(h, rec, "AN", &an, 1);
            bcf_update_info_int32(h, rec, "DP", &tot_dp, 1);
            if (rng_below(&r, 4) == 0) bcf_update_info_flag(h, rec, "DB", NULL, 1);
            if (edge) {
                if (rng_below(&r, 2)) bcf_update_info_string(h, rec, "AA", rng_below(&r, 8) ? "A" : "a_rather_long_ancestral_allele_string");
                const uint32_t an_kind = rng_below(&r, 6);
                static const char *ANN[5] = {"x", "alpha,beta", "one,,three", "a_very_long_annotation_token,short,", ",lead"};
                if (an_kind < 5) bcf_update_info_string(h, rec, "ANNO", ANN[an_kind]);
                if (rng_below(&r, 3) == 0) { int32_t sb[4] = {(int32_t)rng_below(&r, 100), (int32_t)rng_below(&r, 100), 70000, bcf_int32_missing}; bcf_update_info_int32(h, rec, "SB", sb, 4); }
                if (rng_below(&r, 3) == 0) { float mq = (float)(rng_below(&r, 600) / 10.0); if (rng_below(&r, 5) == 0) bcf_float_set_missing(mq); bcf_update_info_float(h, rec, "MQ", &mq, 1); }
            }
            if (n_samples > 0) {
                bcf_update_genotypes(h, rec, gt, n_samples * ploidy);
                if (edge && kind == 9) goto write_it;                         /* kind 9: a site with GT only */
                if (!(edge && kind == 8)) bcf_update_format_int32(h, rec, "DP", dp, n_samples);   /* kind 8 omits DP */
                bcf_update_format_int32(h, rec, "GQ", gq, n_samples);
                bcf_update_format_int32(h, rec, "AD", ad, n_samples * n_allele);
                if (edge && kind != 7) {
                    bcf_update_format_float(h, rec, "GL", gl, n_samples * (n_allele * (n_allele + 1) / 2));
                    bcf_update_format_int32(h, rec, "HQ", hq, n_samples * 2);
                    bcf_update_format_string(h, rec, "FT", ft, n_samples);
                    bcf_update_format_float(h, rec, "PQ", pq, n_samples);
                }
            }
        write_it:
            if (bcf_write(fp, h, rec) < 0) { fprintf(stderr, "bcf_write failed\n"); return 1; }
        }
    }
    if (hts_close(fp) < 0) { fprintf(stderr, "close failed\n"); return 1; }
    if (do_index && bcf_index_build(out, 14) < 0) { fprintf(stderr, "index build failed\n"); return 1; }
    return 0;
}
