#!/bin/bash
# Regenerates tests/golden/ from the REFERENCE (run in the build container, where /root/reference exists):
#   * small binary fixtures of the reference's own test suites (test/data, third_party/htslib/test), the .sam ones
#     converted to BAM by the reference's htslib (harness/_bin/ref_dump sam2bam);
#   * for each, the cells the reference scan stores (harness/_bin/ref_dump dump = htslib calls of
#     src/bam_reader.c:779-966), and the raw bgzf_read() stream sha256.
# The outputs are committed; the GPU box never needs /root/reference.
set -e
cd "$(dirname "$0")"
REF=/root/reference
HT=$REF/third_party/htslib/test
D=../../harness/_bin/ref_dump
TAGS="NM:i MD:Z AS:i XS:i XA:A XZ:Z Xf:f F0:f I1:i ZA:A H0:Z"
cp $REF/test/data/range.bam range.bam
cp $HT/bgzf_boundaries/bgzf_boundaries1.bam $HT/bgzf_boundaries/bgzf_boundaries2.bam $HT/bgzf_boundaries/bgzf_boundaries3.bam .
cp $HT/colons.bam colons.bam
cp $HT/bgziptest.txt.gz bgziptest.txt.gz
for s in 'auxf#values' 'ce#unmap' 'ce#unmap1' 'ce#unmap2' 'c1#noseq' 'ce#large_seq' 'ce#5b' 'c1#clip' 'ce#1' 'ce#supp' 'c1#bounds' 'xx#rg'; do
  o=$(echo "$s" | tr '#' '_')
  $D sam2bam "$HT/$s.sam" "$o.bam"
done
zcat $REF/test/data/rg.sam.gz > /tmp/rg.sam && $D sam2bam /tmp/rg.sam rg.bam
zcat $REF/test/data/aux_tags.sam.gz > /tmp/aux_tags.sam && $D sam2bam /tmp/aux_tags.sam aux_tags.bam
chmod u+w *.bam *.gz
for f in *.bam; do
  $D dump "$f" $TAGS > "${f%.bam}.tsv"
  $D bgunzip "$f" | sha256sum | cut -d' ' -f1 > "${f%.bam}.sha256"
done
$D bgunzip bgziptest.txt.gz | sha256sum | cut -d' ' -f1 > bgziptest.txt.sha256
ls -la
# ---- A8: index-driven scans of range.bam (.bai from the reference's test/data) through sam_itr_regarray / sam_itr_queryi
cp $REF/test/data/range.bam.bai range.bam.bai; chmod u+w range.bam.bai
k=0
while IFS= read -r reg; do
  $D dumpregion range.bam "$reg" $TAGS > "range.region$k.tsv"; echo "$reg" > "range.region$k.txt"; k=$((k+1))
done <<'REGIONS'
CHROMOSOME_I
CHROMOSOME_I:1-1000
CHROMOSOME_I:1-1000,CHROMOSOME_I:1-1000
CHROMOSOME_II:2000-2500,CHROMOSOME_IV:1000-1100,CHROMOSOME_II:1-1200
CHROMOSOME_III:2,900-
CHROMOSOME_V
nosuchcontig,CHROMOSOME_IV:-1500
nosuchcontig
CHROMOSOME_II:1136-1136
{CHROMOSOME_X}:1-10,CHROMOSOME_III:1k-1.1k
REGIONS
$D dumpcontigs range.bam $TAGS > range.contigs.tsv
# ---- B1-B10: read_bcf.  The reference EXTENSION itself (oracle/_ref/libduckhts_ref.so, unchanged sources) under the
# fake DuckDB host prints every cell; the files are the reference's own fixture and a small seeded --edge file.
H=../../harness/_bin/fake_duckdb; RX=../../oracle/_ref/libduckhts_ref.so
cp $REF/test/data/vcf_file.bcf vcf_file.bcf; chmod u+w vcf_file.bcf
cp $REF/test/data/vcf_file.bcf.csi vcf_file.bcf.csi; chmod u+w vcf_file.bcf.csi   # the index the reference's region tests use (test/sql/duckhts.test:86-105)
../../harness/_bin/gen_bcf edge_small.bcf 60 5 0x51 --edge --contigs 3
../../harness/_bin/gen_bcf plain_small.bcf 40 9 0x52 --contigs 2
for f in vcf_file edge_small plain_small; do
  $H $RX read_bcf $f.bcf tidy_format=true --dump > $f.bcf.tidy.txt 2>/dev/null
  $H $RX read_bcf $f.bcf --dump > $f.bcf.wide.txt 2>/dev/null
done
# ---- N4: the reference's text fixtures and what its read_tabix returns for them (the unchanged extension under harness/fake_duckdb)
H=../../harness/_bin/fake_duckdb
R=../../oracle/_ref/libduckhts_ref.so
for f in meta_tabix.tsv.gz header_tabix.tsv.gz gff_file.gff.gz formatcols.vcf.gz no_contig.vcf.gz; do
  cp $REF/test/data/$f $f; chmod u+w $f
  for x in tbi csi; do if [ -f $REF/test/data/$f.$x ]; then cp $REF/test/data/$f.$x $f.$x; chmod u+w $f.$x; fi; done
done
$H $R read_tabix meta_tabix.tsv.gz --dump --types > tabix.meta.tsv
$H $R read_tabix meta_tabix.tsv.gz header_names=chrom,pos,value --dump --types > tabix.meta_names.tsv
$H $R read_tabix meta_tabix.tsv.gz auto_detect=true --dump --types > tabix.meta_auto.tsv
$H $R read_tabix header_tabix.tsv.gz header=true --dump --types > tabix.header.tsv
$H $R read_tabix header_tabix.tsv.gz header=true column_types=VARCHAR,BIGINT,VARCHAR --dump --types > tabix.header_types.tsv
$H $R read_tabix gff_file.gff.gz --dump --types > tabix.gff.tsv
$H $R read_tabix gff_file.gff.gz --cols 0,2 --dump --types > tabix.gff_proj.tsv
$H $R read_tabix formatcols.vcf.gz --dump --types > tabix.vcf.tsv
$H $R read_tabix no_contig.vcf.gz auto_detect=true --dump --types > tabix.no_contig_auto.tsv
