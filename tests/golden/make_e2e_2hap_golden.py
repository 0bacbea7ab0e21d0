"""Golden for LI/volcanosv-vc-large-indel-otherasm-2hap.py: the UNMODIFIED reference script run here on the
whole-genome fixture of make_e2e_wgs_golden.py (tests/golden/e2e_wgs/input), the two haplotypes given as two
FASTA files.  Test-only stand-ins as for the other file-level goldens (shims/ for pysam / cigar / edlib /
matplotlib, stubs/ for minimap2 / samtools / vcf-sort): `minimap2` prints the fixture's alignments of the
contigs of the FASTA it was given, `samtools merge | samtools sort -o` concatenates and coordinate-sorts them.
Writes tests/golden/e2e_wgs/expected_otherasm_2hap/ and reference_run_otherasm_2hap.log.

    python tests/golden/make_e2e_2hap_golden.py        (needs /root/reference; a few minutes)
"""
import gzip
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_e2e_golden as G  # noqa: E402  (LI = the reference's Large_INDEL directory)

OUT = os.path.join(HERE, "e2e_wgs")


def stage_inputs(work):
    """Fixture files -> work (gunzipped), plus the whole-genome contig SAM and one FASTA per haplotype."""
    if os.path.exists(work):
        shutil.rmtree(work)
    src_root = os.path.join(OUT, "input")
    for root, _, files in os.walk(src_root):
        for fn in files:
            src = os.path.join(root, fn)
            dst = os.path.join(work, os.path.relpath(src, src_root))
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            if fn.endswith(".gz"):
                with gzip.open(src, "rb") as fi, open(dst[:-3], "wb") as fo:
                    shutil.copyfileobj(fi, fo)
            else:
                shutil.copy(src, dst)
    allsam = os.path.join(work, "all_contigs.sam")
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
    hp = {k: open(os.path.join(work, "%s.fa" % k), "w") for k in ("hp1", "hp2")}
    for c in range(1, 23):
        cur = None
        for line in open(os.path.join(work, "asm", "chr%d" % c, "assembly", "final_contigs", "Sample_final_contigs.fa")):
            if line.startswith(">"):
                cur = hp["hp1" if "hp1" in line else "hp2"]
            cur.write(line)
    for f in hp.values():
        f.close()
    return allsam


def main():
    work = "/tmp/vsv_e2e_2hap_ref"
    allsam = stage_inputs(work)
    env = dict(os.environ)
    env.update(PYTHONPATH=os.path.join(HERE, "shims"), PATH=os.path.join(HERE, "stubs") + ":" + env["PATH"],
               LC_ALL="C", PYTHONHASHSEED="0", VSV_CONTIG_SAM_BY_FASTA=allsam)
    out = os.path.join(work, "out_2hap")
    cmd = ["python3", G.LI + "volcanosv-vc-large-indel-otherasm-2hap.py", "-hp1", os.path.join(work, "hp1.fa"),
           "-hp2", os.path.join(work, "hp2.fa"), "-o", out, "-dtype", "Hifi", "-bam", os.path.join(work, "reads.sam"),
           "-ref", os.path.join(work, "ref.fa"), "-t", "4", "-px", "Sample"]
    with open(os.path.join(OUT, "reference_run_otherasm_2hap.log"), "w") as log:
        log.write(" ".join(cmd) + "\n")
        log.flush()
        rc = subprocess.call(cmd, env=env, stdout=log, stderr=subprocess.STDOUT, cwd=work)
    print("2hap reference exit code", rc)
    merged = os.path.join(out, "Sample.sorted.bam")
    print("merged contig alignments identical to the one-file fixture:",
          open(merged).read() == open(allsam).read())
    exp = os.path.join(OUT, "expected_otherasm_2hap")
    os.makedirs(exp, exist_ok=True)
    for rel in ("raw_variants_wgs.vcf", "variants_filtered_GT_corrected.vcf", "Sample_volcanosv_large_indel.vcf"):
        src = os.path.join(out, rel)
        if not os.path.exists(src):
            print("missing in the 2hap run:", rel)
            continue
        with open(src, "rb") as fi, open(os.path.join(exp, rel), "wb") as fo:
            fo.write(fi.read().replace(work.encode(), b"$WORK"))
        one = os.path.join(OUT, "expected_otherasm", rel)
        print(rel, "identical to the one-FASTA otherasm run:", open(os.path.join(exp, rel), "rb").read() == open(one, "rb").read())


if __name__ == "__main__":
    main()
