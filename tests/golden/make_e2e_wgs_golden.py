#!/usr/bin/env python
"""Whole-genome golden fixture: runs the reference's large-indel stage UNMODIFIED in WGS mode
(`python3 LI/volcanosv-vc-large-indel.py ...` WITHOUT -chr: chr1..chr22 through
Raw_variant_call.py, the concatenated raw VCF through filter_GT_correction.py, MAIN:266-319) on 22
small synthetic diploid chromosomes and stores inputs + the outputs that matter under
tests/golden/e2e_wgs/.  What it pins beyond the single-chromosome fixture: chromosome order of
the concatenation and of vcf-sort, the genome-wide coverage median (filter_vcf_by_sig_cov), the
`last_match_idx` carry across chromosomes (GTD:94) and the whole-BAM cuteSV-style scan.

Same test-only stand-ins as make_e2e_golden.py (tests/golden/shims, tests/golden/stubs).
Run in the build container (needs /root/reference):  python tests/golden/make_e2e_wgs_golden.py
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_e2e_golden as G  # noqa: E402

OUT = os.path.join(HERE, "e2e_wgs")
KEEP = ("raw_variants_wgs.vcf", "variants_filtered_GT_corrected.vcf", "Sample_volcanosv_large_indel.vcf", "VCF_header",
        "cute_sig/DEL.sigs", "cute_sig/INS.sigs", "cute_sig/INS.sigs.gte30auto", "cute_sig/reads.sigs",
        "GT_Correction/bnd_del_real.tsv", "GT_Correction/bnd_del_real.tsv.newgt", "GT_Correction/bnd_ins_real.tsv",
        "GT_Correction/bnd_ins_real.tsv.newgt", "raw_variants_wgs_filter_DEL.vcf", "raw_variants_wgs_filter_DEL.vcf.newgt.DEL",
        "raw_variants_wgs_filter_DEL.vcf.newgt.INS", "raw_variants_wgs_cutesv_sig_support_mins30_fl1000.csv")


def write_sam(path, sq, recs):
    with open(path, "w") as f:
        f.write("@HD\tVN:1.6\tSO:coordinate\n")
        for name, ln in sq:
            f.write("@SQ\tSN:%s\tLN:%d\n" % (name, ln))
        for chrom, r in recs:
            name, flag, pos, mapq, cig, seq = r[:6]
            tags = r[6] if len(r) > 6 else []
            f.write("\t".join([name, str(flag), chrom, str(pos + 1), str(mapq), G.cigar_str(cig), "*", "0", "0", seq, "*"] +
                              list(tags)) + "\n")


def main():
    if not os.path.isdir(G.LI):
        sys.exit("needs /root/reference (build container only)")
    if os.path.exists(OUT):
        shutil.rmtree(OUT)
    inp = os.path.join(OUT, "input")
    os.makedirs(inp)
    sq, all_reads = [], []
    with open(os.path.join(inp, "ref.fa"), "w") as fref, open(os.path.join(inp, "planted_svs.tsv"), "w") as fsv:
        for c in range(1, 23):
            chrom = "chr%d" % c
            L = 30_000 + 700 * ((c * 7) % 11)
            ref, svs, contig_recs, contig_fa, reads = G.make_dataset(seed=7000 + c, L=L, chrom=chrom, lite=True,
                                                                      name_prefix="c%d" % c, coverage=26,
                                                                      ps_offset=c * 1_000_000)   # contig names unique genome-wide
            sq.append((chrom, L))
            fref.write(">%s\n" % chrom)
            for i in range(0, L, 60):
                fref.write(ref[i:i + 60] + "\n")
            asm = os.path.join(inp, "asm", chrom, "assembly", "final_contigs")
            os.makedirs(asm)
            with open(os.path.join(asm, "Sample_final_contigs.fa"), "w") as f:
                for n, s in contig_fa:
                    f.write(">%s\n%s\n" % (n, s))
            write_sam(os.path.join(inp, "contigs_%s.sam" % chrom), [(chrom, L)], [(chrom, r) for r in contig_recs])
            all_reads += [(chrom, r) for r in reads]
            for s in svs:
                fsv.write("%s\t%d\t%s\t%d\t%s\n" % (chrom, s["pos"], s["type"], s["len"], s["zyg"]))
    write_sam(os.path.join(inp, "reads.sam"), sq, all_reads)
    work = "/tmp/vsv_e2e_wgs_ref"
    if os.path.exists(work):
        shutil.rmtree(work)
    shutil.copytree(inp, work)
    env = dict(os.environ)
    env.update(PYTHONPATH=os.path.join(HERE, "shims"), PATH=os.path.join(HERE, "stubs") + ":" + env["PATH"],
               LC_ALL="C", PYTHONHASHSEED="0", VSV_CONTIG_SAM_DIR=work)
    cmd = ["python3", G.LI + "volcanosv-vc-large-indel.py", "-i", os.path.join(work, "asm"), "-o", os.path.join(work, "out"),
           "-dtype", "Hifi", "-bam", os.path.join(work, "reads.sam"), "-ref", os.path.join(work, "ref.fa"),
           "-t", "4", "-px", "Sample"]
    with open(os.path.join(OUT, "reference_run.log"), "w") as log:
        log.write(" ".join(cmd) + "\n")
        log.flush()
        rc = subprocess.call(cmd, env=env, stdout=log, stderr=subprocess.STDOUT, cwd=work)
    print("reference exit code", rc)
    exp = os.path.join(OUT, "expected")
    n = 0
    rels = list(KEEP) + ["chr%d/final_vcf/volcano_variant_no_redundancy.vcf" % c for c in range(1, 23)]
    for rel in rels:
        src = os.path.join(work, "out", rel)
        if not os.path.exists(src):
            print("missing in the reference run:", rel)
            continue
        dst = os.path.join(exp, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        with open(src, "rb") as fi:
            data = fi.read().replace(work.encode(), b"$WORK")
        with open(dst, "wb") as fo:
            fo.write(data)
        n += 1
    print("kept %d reference output files under %s" % (n, exp))
    run_otherasm(work, env)
    for root, _, files in os.walk(inp):
        for fn in files:
            if fn.endswith((".sam", ".fa")):
                subprocess.check_call(["gzip", "-9", "-n", "-f", os.path.join(root, fn)])


def run_otherasm(work, env):
    """LI/volcanosv-vc-large-indel-otherasm.py on the same data: ONE haplotype-resolved contig FASTA
    and ONE contig alignment file for the whole genome (the script splits contigs by chromosome)."""
    allfa = os.path.join(work, "all_contigs.fa")
    allsam = os.path.join(work, "all_contigs.sam")
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
    env = dict(env, VSV_CONTIG_SAM=allsam)
    env.pop("VSV_CONTIG_SAM_DIR", None)
    out = os.path.join(work, "out_otherasm")
    cmd = ["python3", G.LI + "volcanosv-vc-large-indel-otherasm.py", "-i", allfa, "-o", out, "-dtype", "Hifi", "-bam",
           os.path.join(work, "reads.sam"), "-ref", os.path.join(work, "ref.fa"), "-t", "4", "-px", "Sample"]
    with open(os.path.join(OUT, "reference_run_otherasm.log"), "w") as log:
        log.write(" ".join(cmd) + "\n")
        log.flush()
        rc = subprocess.call(cmd, env=env, stdout=log, stderr=subprocess.STDOUT, cwd=work)
    print("otherasm reference exit code", rc)
    exp = os.path.join(OUT, "expected_otherasm")
    os.makedirs(exp, exist_ok=True)
    for rel in ("raw_variants_wgs.vcf", "variants_filtered_GT_corrected.vcf", "Sample_volcanosv_large_indel.vcf"):
        src = os.path.join(out, rel)
        if not os.path.exists(src):
            print("missing in the otherasm run:", rel)
            continue
        with open(src, "rb") as fi, open(os.path.join(exp, rel), "wb") as fo:
            fo.write(fi.read().replace(work.encode(), b"$WORK"))
        same = open(src, "rb").read() == open(os.path.join(work, "out", rel), "rb").read()
        print(rel, "identical to the per-chromosome-assembly run:", same)


if __name__ == "__main__":
    main()
