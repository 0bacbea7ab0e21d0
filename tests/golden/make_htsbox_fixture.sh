#!/bin/bash
# BAM fixtures written by an INDEPENDENT implementation: the htsbox binary vendored in the reference
# (/root/reference/bin/VolcanoSV-vc/Small_INDEL/htsbox/htsbox, htslib lite-r346) converts the SAM text of the committed
# fixtures and indexes the result.  tests/test_bam_cpu.py::test_htsbox_written_bam reads them through the native feeder
# (vsv_bam_*) and compares with the SAM-text loader.  Run in the build container (needs /root/reference).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
HB=/root/reference/bin/VolcanoSV-vc/Small_INDEL/htsbox/htsbox
T=$(mktemp -d)
zcat "$HERE/e2e_wgs/input/reads.sam.gz" > "$T/reads_wgs.sam"
zcat "$HERE/e2e/input/contigs.sam.gz" > "$T/contigs_chr1.sam"
for n in reads_wgs contigs_chr1; do
  "$HB" samview -bS "$T/$n.sam" > "$HERE/htsbox/$n.bam"
  "$HB" bamidx "$HERE/htsbox/$n.bam"
done
rm -rf "$T"
ls -la "$HERE/htsbox"
