/* dh_oracle.h — CPU restatement ("port") of the reference's read_bam / bgunzip hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under duckhts_b200/ or include/ may include, link or call this;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it, and
 * only as the checker or as the timed CPU baseline.
 *
 * Parity pinned: tests/test_oracle.py checks every function below against (a) the reference's own
 * fixtures (tests/golden/, produced by running the reference — libhts.a compiled from
 * /root/reference by oracle/Makefile — through oracle/ref_dump.c and the reference extension under
 * harness/fake_duckdb.c) and (b) the known answers of test/sql/duckhts.test.
 *
 * Each function cites the reference file:line it follows (paths relative to the reference tree,
 * "htslib/" = third_party/htslib/).  zlib (system libz; the reference's only third-party dependency on
 * this path — htslib/bgzf.c:775-793) is called exactly where the reference calls it.
 */
#ifndef DH_ORACLE_H
#define DH_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    uint64_t coffset;
    uint32_t csize, isize, crc, reserved;
} dho_block;

/* htslib/bgzf.c:896-903 (check_header) + :1147-1216 (BSIZE walk of bgzf_read_block).
 * Returns number of blocks (written up to cap) or -1 on a bad header; *end_off = offset after the last block. */
int64_t dho_bgzf_walk(const uint8_t *data, uint64_t size, dho_block *out, int64_t cap, uint64_t *end_off);

/* htslib/bgzf.c:762-804 (bgzf_uncompress, zlib variant) + :808-824 (inflate_block) applied to every
 * block in file order, i.e. the bgzf_read() loop of src/bgzip.c:180-200.
 * Returns 0, or -1 (header), -2 (zlib error -> BGZF_ERR_ZLIB), -3 (CRC mismatch -> BGZF_ERR_CRC);
 * *bad_block = index of the failing block.  out may be NULL (size query / verify only). */
int dho_bgzf_inflate(const uint8_t *data, uint64_t size, uint8_t *out, uint64_t cap, uint64_t *out_len,
                     int64_t *bad_block);

/* One block only: dst must hold 65536 bytes.  Returns inflated length, -2 zlib error, -3 CRC mismatch. */
int dho_bgzf_inflate_block(const uint8_t *block, uint32_t csize, uint8_t *dst);

/* ---- BAM -------------------------------------------------------------------------------------------
 * Columnar result of the reference's read_bam scan in sequential (no index, no region) mode over an
 * inflated BAM byte stream.  Mirrors, cell for cell, what src/bam_reader.c:779-966 stores. */
typedef struct {
    uint64_t *off; /* n+1 offsets into heap */
    char *heap;
    uint8_t *valid; /* 1 = not NULL (NULL only where the reference calls set_null) */
    uint64_t heap_len, heap_cap;
} dho_strcol;

typedef struct {
    char tag[2];
    char type; /* 'i' BIGINT (bam_aux2i), 'f' DOUBLE (bam_aux2f), 'Z'/'H' VARCHAR (bam_aux2Z), 'A' VARCHAR(1),
                  'B' LIST (children as 8-byte words in the string column, 8*len bytes per row) */
} dho_tag_req;

typedef struct {
    int64_t *i;    /* type 'i' */
    double *f;     /* type 'f' */
    dho_strcol s;  /* type 'Z','H','A' */
    uint8_t *valid;
} dho_tagcol;

typedef struct {
    uint64_t n;          /* records */
    uint16_t *flag;      /* FLAG USMALLINT  bam_reader.c:790-794 */
    int64_t *pos;        /* POS = pos+1     :806-810 */
    int32_t *mapq;       /* MAPQ INTEGER    :812-816 */
    int64_t *pnext;      /* PNEXT = mpos+1  :845-849 */
    int64_t *tlen;       /* TLEN = isize    :851-855 */
    int32_t *tid, *mtid; /* as read; RNAME/RNEXT below are the strings the reference emits */
    /* VARCHAR columns, strlen-truncated exactly like duckdb_vector_assign_string_element;
     * CIGAR uses the _len variant (:819-834) */
    dho_strcol qname, rname, cigar, rnext, seq, qual, rg, sample;
    int n_tags;
    dho_tagcol *tags;
    /* header (htslib/sam.c:229-342) */
    int32_t n_ref;
    char **ref_name;
    uint32_t *ref_len;
    char *text;
    uint32_t l_text;
    uint64_t first_record; /* inflated offset of the first alignment record */
    uint64_t *rec_off;     /* inflated offset of every record's block_size word */
    /* 0 = clean EOF; <0 = the bam_read1/sam_read1_bam error code that ended the scan after n records
     * (the reference stops silently) */
    int status;
    int64_t *endpos;       /* bam_endpos (htslib/sam.c:519-524) of every record: what index iterators compare */
} dho_bam_columns;

/* htslib/sam.c:229-342 (bam_hdr_read), :779-855 (bam_read1), :4124-4134 (sam_read1_bam range check),
 * :4785-4855,5093-5171 (aux walk + conversions), src/bam_reader.c:375-404,779-966 (column conversions),
 * htslib/header.c sam_hdr_find_tag_id (RG ID -> SM).  `u` = inflated byte stream of a whole BAM file.
 * Returns 0, or -1 if the header itself is invalid (the reference fails in bind). */
int dho_bam_decode(const uint8_t *u, uint64_t n, const dho_tag_req *tags, int n_tags, dho_bam_columns *out);
/* A16: tag request type 'M' = the AUXILIARY_TAGS MAP of the row (bam_reader.c:967-1031), serialised in the string column as
 * key NUL value NUL ...; dho_set_aux_exclude names the n two-byte tags left out (standard_tags := true: BAM_STD_TAGS). */
void dho_set_aux_exclude(const char *tags, int n);
void dho_bam_free(dho_bam_columns *c);

/* ---- A8: BAI/CSI index + iterators (dh_oracle_idx.inc) ------------------------------------------------
 * idx_read_core / update_loff (htslib/hts.c:2920-2983, 2426-2455); hts_itr_query + hts_itr_next (:3421-3595,
 * 4266-4318) = sam_itr_queryi; hts_itr_multi_bam + hts_itr_multi_next (:3597-3745, 4320-4604) = sam_itr_regarray.
 * Records are described by arrays: virtual offset of each record and of the record after it (bgzf_tell before /
 * after the read), tid, pos, bam_endpos.  The functions write the indices of the records the iterator returns. */
typedef struct dho_index dho_index;
dho_index *dho_index_parse(const uint8_t *data, uint64_t size);   /* uncompressed BAI / CSI image */
void dho_index_free(dho_index *ix);
int32_t dho_index_n_ref(const dho_index *ix);
int64_t dho_itr_query_rows(const dho_index *ix, uint64_t n_rec, const uint64_t *voff, const uint64_t *voff_next, const int32_t *rtid,
                           const int64_t *rbeg, const int64_t *rend, int tid, int64_t beg, int64_t end, uint64_t *out);
int64_t dho_itr_regions_rows(const dho_index *ix, uint64_t n_rec, const uint64_t *voff, const uint64_t *voff_next, const int32_t *rtid,
                             const int64_t *rbeg, const int64_t *rend, int n_iv, const int32_t *iv_tid, const int64_t *iv_beg,
                             const int64_t *iv_end, uint64_t *out);

#ifdef __cplusplus
}
#endif
#endif
