#!/usr/bin/env python
"""ONT and CLR variants of the end-to-end golden: the SAME inputs as tests/golden/e2e/input, the
reference's stage run UNMODIFIED with `-dtype ONT` / `-dtype CLR` (different contig-signature
scripts ECS_O / ECS_C, split-alignment rules, CLR gates and genotype thresholds).  Keeps the files
that differ by data type under tests/golden/e2e/expected_<dtype>/chr1/.
Run in the build container (needs /root/reference):  python tests/golden/make_e2e_dtype_golden.py
"""
import gzip
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LI = "/root/reference/bin/VolcanoSV-vc/Large_INDEL/"
E2E = os.path.join(HERE, "e2e")
KEEP = ("Sample_volcanosv_large_indel_chr1.vcf", "volcano_raw_variant.vcf", "volcano_variant_filtered.vcf",
        "final_vcf/volcano_variant_no_redundancy.vcf", "final_vcf/variants_filtered_GT_corrected.vcf",
        "final_vcf/GT_Correction/bnd_del_real.tsv", "final_vcf/GT_Correction/bnd_ins_real.tsv",
        "final_vcf/volcano_variant_no_redundancy_filter_DEL.vcf",
        "final_vcf/volcano_variant_no_redundancy_cutesv_sig_support_mins30_fl1000.csv")


def main():
    if not os.path.isdir(LI):
        sys.exit("needs /root/reference (build container only)")
    for dtype in ("ONT", "CLR"):
        work = "/tmp/vsv_e2e_ref_%s" % dtype
        if os.path.exists(work):
            shutil.rmtree(work)
        for root, _, files in os.walk(os.path.join(E2E, "input")):
            for fn in files:
                src = os.path.join(root, fn)
                dst = os.path.join(work, os.path.relpath(src, os.path.join(E2E, "input")))
                os.makedirs(os.path.dirname(dst), exist_ok=True)
                if fn.endswith(".gz"):
                    with gzip.open(src, "rb") as fi, open(dst[:-3], "wb") as fo:
                        shutil.copyfileobj(fi, fo)
                else:
                    shutil.copy(src, dst)
        env = dict(os.environ)
        env.update(PYTHONPATH=os.path.join(HERE, "shims"), PATH=os.path.join(HERE, "stubs") + ":" + env["PATH"],
                   LC_ALL="C", PYTHONHASHSEED="0", VSV_CONTIG_SAM=os.path.join(work, "contigs.sam"))
        cmd = ["python3", LI + "volcanosv-vc-large-indel.py", "-i", os.path.join(work, "asm"), "-o", os.path.join(work, "out"),
               "-dtype", dtype, "-bam", os.path.join(work, "reads.sam"), "-ref", os.path.join(work, "ref.fa"), "-chr", "1",
               "-t", "2", "-px", "Sample"]
        exp = os.path.join(E2E, "expected_%s" % dtype, "chr1")
        if os.path.exists(os.path.dirname(exp)):
            shutil.rmtree(os.path.dirname(exp))
        os.makedirs(exp)
        with open(os.path.join(os.path.dirname(exp), "reference_run.log"), "w") as log:
            log.write(" ".join(cmd) + "\n")
            log.flush()
            rc = subprocess.call(cmd, env=env, stdout=log, stderr=subprocess.STDOUT, cwd=work)
        print(dtype, "reference exit code", rc)
        for rel in KEEP:
            src = os.path.join(work, "out", "chr1", rel)
            if not os.path.exists(src):
                print("missing:", rel)
                continue
            dst = os.path.join(exp, rel)
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            with open(src, "rb") as fi:
                data = fi.read().replace(work.encode(), b"$WORK")
            with open(dst, "wb") as fo:
                fo.write(data)


if __name__ == "__main__":
    main()
