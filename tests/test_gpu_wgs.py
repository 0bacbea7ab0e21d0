"""Whole-genome mode (LI/volcanosv-vc-large-indel.py without -chr, MAIN:266-319) against the files
the UNMODIFIED reference wrote for 22 small synthetic chromosomes (tests/golden/e2e_wgs, generated
by tests/golden/make_e2e_wgs_golden.py).  Byte for byte.  Pins what a single chromosome cannot:
concatenation order, vcf-sort's dictionary chromosome order, the genome-wide coverage median and
the last_match_idx carry across chromosomes (GTD:94)."""
import gzip
import os
import shutil
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
WGS = os.path.join(HERE, "golden", "e2e_wgs")
EXP = os.path.join(WGS, "expected")
FILES = ["VCF_header", "raw_variants_wgs.vcf", "cute_sig/DEL.sigs", "cute_sig/INS.sigs", "cute_sig/INS.sigs.gte30auto",
         "raw_variants_wgs_cutesv_sig_support_mins30_fl1000.csv", "raw_variants_wgs_filter_DEL.vcf",
         "GT_Correction/bnd_del_real.tsv", "GT_Correction/bnd_del_real.tsv.newgt", "GT_Correction/bnd_ins_real.tsv",
         "GT_Correction/bnd_ins_real.tsv.newgt", "raw_variants_wgs_filter_DEL.vcf.newgt.DEL",
         "raw_variants_wgs_filter_DEL.vcf.newgt.INS", "variants_filtered_GT_corrected.vcf", "Sample_volcanosv_large_indel.vcf"]


@pytest.fixture(scope="module")
def work(tmp_path_factory):
    w = str(tmp_path_factory.mktemp("e2e_wgs"))
    for root, _, files in os.walk(os.path.join(WGS, "input")):
        for fn in files:
            src = os.path.join(root, fn)
            dst = os.path.join(w, os.path.relpath(src, os.path.join(WGS, "input")))
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            if fn.endswith(".gz"):
                with gzip.open(src, "rb") as fi, open(dst[:-3], "wb") as fo:
                    shutil.copyfileobj(fi, fo)
            else:
                shutil.copy(src, dst)
    return w


def check(out, work):
    for c in range(1, 23):
        rel = "chr%d/final_vcf/volcano_variant_no_redundancy.vcf" % c
        assert open(os.path.join(out, rel)).read() == open(os.path.join(EXP, rel)).read(), rel
    for rel in FILES:
        got = open(os.path.join(out, rel)).read().replace(work, "$WORK")
        assert got == open(os.path.join(EXP, rel)).read(), rel


def argv(work, out):
    return ["volcanosv-vc-large-indel.py", "-i", os.path.join(work, "asm"), "-o", out, "-dtype", "Hifi", "-bam",
            os.path.join(work, "reads.sam"), "-ref", os.path.join(work, "ref.fa"), "-t", "4", "-px", "Sample",
            "-cbam", os.path.join(work, "contigs_chr{chr}.sam")]


def test_whole_genome_mode_end_to_end(work):
    from volcanosv_b200 import cli
    out = os.path.join(work, "out")
    cli.main(argv(work, out))
    check(out, work)


def test_whole_genome_mode_two_ranks(work):
    """Two processes (one per GPU in production; both on GPU 0 here) split the chromosomes by LPT,
    meet in the output directory, rank 0 finishes: same bytes."""
    out = os.path.join(work, "out2")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533", WORLD_SIZE="2", LOCAL_RANK="0",
               PYTHONPATH=os.path.dirname(HERE) + os.pathsep + os.environ.get("PYTHONPATH", ""))
    procs = [subprocess.Popen([sys.executable, "-m", "volcanosv_b200.cli"] + argv(work, out), env=dict(env, RANK=str(r)),
                              cwd=os.path.dirname(HERE)) for r in range(2)]
    for p in procs:
        assert p.wait(timeout=900) == 0
    check(out, work)


def _whole_genome_contigs(work):
    """One contig FASTA and one coordinate-sorted contig alignment file for the whole genome."""
    allfa, allsam = os.path.join(work, "all_contigs.fa"), os.path.join(work, "all_contigs.sam")
    with open(allfa, "w") as fo:
        for c in range(1, 23):
            fo.write(open(os.path.join(work, "asm", "chr%d" % c, "assembly", "final_contigs", "Sample_final_contigs.fa")).read())
    with open(allsam, "w") as fo:
        fo.write("@HD\tVN:1.6\tSO:coordinate\n")
        body = []
        for c in range(1, 23):
            for line in open(os.path.join(work, "contigs_chr%d.sam" % c)):
                if line.startswith("@SQ"):
                    fo.write(line)
                elif not line.startswith("@"):
                    body.append(line)
        fo.writelines(body)
    return allfa, allsam


def test_otherasm_entry_point(work):
    """LI/volcanosv-vc-large-indel-otherasm.py: one whole-genome contig FASTA + one contig alignment file,
    contigs split by chromosome on the fly; against the unmodified reference's run of that script."""
    from volcanosv_b200 import cli
    allfa, allsam = _whole_genome_contigs(work)
    out = os.path.join(work, "out_otherasm")
    cli.main(["volcanosv-vc-large-indel-otherasm.py", "-i", allfa, "-o", out, "-dtype", "Hifi", "-bam",
              os.path.join(work, "reads.sam"), "-ref", os.path.join(work, "ref.fa"), "-t", "4", "-px", "Sample", "-cbam", allsam])
    exp = os.path.join(WGS, "expected_otherasm")
    for rel in os.listdir(exp):
        got = open(os.path.join(out, rel)).read().replace(work, "$WORK")
        assert got == open(os.path.join(exp, rel)).read(), rel


def test_otherasm_2hap_entry_point(work):
    """LI/volcanosv-vc-large-indel-otherasm-2hap.py: one FASTA per haplotype, the merged sorted contig
    alignments (2hap:250-258) supplied; against the unmodified reference's run of that script
    (tests/golden/make_e2e_2hap_golden.py)."""
    from volcanosv_b200 import cli
    allfa, allsam = _whole_genome_contigs(work)
    hp = {k: open(os.path.join(work, "%s.fa" % k), "w") for k in ("hp1", "hp2")}
    cur = None
    for line in open(allfa):
        if line.startswith(">"):
            cur = hp["hp1" if "hp1" in line else "hp2"]
        cur.write(line)
    for f in hp.values():
        f.close()
    out = os.path.join(work, "out_2hap")
    cli.main(["volcanosv-vc-large-indel-otherasm-2hap.py", "-hp1", os.path.join(work, "hp1.fa"), "-hp2",
              os.path.join(work, "hp2.fa"), "-o", out, "-dtype", "Hifi", "-bam", os.path.join(work, "reads.sam"),
              "-ref", os.path.join(work, "ref.fa"), "-t", "4", "-px", "Sample", "-cbam", allsam])
    assert open(os.path.join(out, "contig_by_chr", "merged.fa")).read() == \
        open(os.path.join(work, "hp1.fa")).read() + open(os.path.join(work, "hp2.fa")).read()
    exp = os.path.join(WGS, "expected_otherasm_2hap")
    for rel in os.listdir(exp):
        got = open(os.path.join(out, rel)).read().replace(work, "$WORK")
        assert got == open(os.path.join(exp, rel)).read(), rel


def test_whole_genome_mode_from_bam(work):
    """Same run with BAM inputs (22 references in one reads BAM, per-chromosome contig BAMs) through the
    native feeder: same bytes as from SAM text."""
    sys.path.insert(0, HERE)
    import bamio
    from volcanosv_b200 import cli
    bamio.sam_to_bam(os.path.join(work, "reads.sam"), os.path.join(work, "reads.bam"), block=40000)
    for c in range(1, 23):
        bamio.sam_to_bam(os.path.join(work, "contigs_chr%d.sam" % c), os.path.join(work, "contigs_chr%d.bam" % c))
    out = os.path.join(work, "out_bam")
    a = argv(work, out)
    a[a.index("-bam") + 1] = os.path.join(work, "reads.bam")
    a[a.index("-cbam") + 1] = os.path.join(work, "contigs_chr{chr}.bam")
    cli.main(a)
    check(out, work)
