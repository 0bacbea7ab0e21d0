#!/usr/bin/env python
"""Mid-scale end-to-end golden ("widen the parity pins"): the reference's stage UNMODIFIED
(`python3 LI/volcanosv-vc-large-indel.py ... -dtype ONT -chr 1`) on a 1 Mb synthetic chromosome with ONT-shaped
reads - a small indel every ~33 bp (hundreds of CIGAR ops per read, ~10^5 `.sigs` rows) - and the cases the 140 kb
fixture does not reach:
  * 8 % of the reads split with SA tags, some with more than 7 segments (SE:370 drops them),
  * exact duplicate `.sigs` rows (a secondary alignment of the same read with the same CIGAR: `sort -u`, SE:637-638),
  * many signatures at the same position (ties of `sort -k2,2 -k3,3n` and of sort_sig),
  * an INS call in a read-coverage gap (depth 0: pandas inf / nan path of GTI:228),
  * read names shared by several records (ERS split pairs).
Read sequences are runs of one letter per read (the SAM text must stay small enough to commit); the 140 kb fixture
keeps real sequences.  Same test-only stand-ins as make_e2e_golden.py (shims/, stubs/).
Run in the build container (needs /root/reference):  python tests/golden/make_e2e_mid_golden.py
"""
import gzip
import os
import random
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_e2e_golden as G  # noqa: E402

LI = G.LI
OUT = os.path.join(HERE, "e2e_mid")
L = 1_000_000
COVERAGE = 15
GAP = (400_000, 404_000)      # no read overlaps this window: the INS planted inside has depth 0


def ont_events(hsv, rng, a, b, noisy=True):
    """events of one read: the planted SVs it spans (imprecise now and then) + a small indel every ~33 bp; 96 % of
    them 1-6 bp, 3 % 10-29 bp, 1 % 30-44 bp"""
    ev = []
    for s in hsv:
        if not (a + 300 < s["pos"] < b - 300):
            continue
        ln, sq, p = s["len"], s["seq"], s["pos"]
        if noisy and rng.random() < 0.25:
            p += rng.randint(-3, 3)
            if s["type"] == "INS":
                k = rng.randint(0, max(0, min(6, ln - 31)))
                sq, ln = sq[: ln - k], ln - k
            else:
                ln += rng.randint(0, 5)
        ev.append((p, "I" if s["type"] == "INS" else "D", ln, sq))
    for _ in range(int((b - a) * 0.03)):
        p = rng.randint(a + 50, b - 50)
        u = rng.random()
        ln = rng.randint(1, 6) if u < 0.96 else (rng.randint(10, 29) if u < 0.99 else rng.randint(30, 44))
        t = "I" if rng.random() < 0.5 else "D"
        ev.append((p, t, ln, "A" * ln if t == "I" else ""))
    ev.sort(key=lambda e: e[0])
    out, last_end = [], a + 20
    for (p, t, ln, sq) in ev:
        if p <= last_end + 3:
            continue
        out.append((p, t, ln, sq if t == "D" or sq else "A" * ln))
        last_end = p + (ln if t == "D" else 0)
    return out


def make_dataset(seed=20261105):
    rng = random.Random(seed)
    ref = G.rand_seq(rng, L)
    svs = [s for s in G.make_svs(rng, L) if not (GAP[0] - 2500 < s["pos"] < GAP[1] + 2500)]
    svs.append(dict(pos=(GAP[0] + GAP[1]) // 2, type="INS", len=420, seq=G.rand_seq(rng, 420), zyg="hom"))   # > 250 bp: passes the FP filter
    svs.sort(key=lambda x: x["pos"])
    haps = {1: G.hap_svs(svs, 1, rng), 2: G.hap_svs(svs, 2, rng)}
    contig_recs, contig_fa = [], []
    for hap in (1, 2):
        for k, (a, b) in enumerate([(150, 520_000), (500_000, L - 120)]):
            name = "PS%d_%d_hp%d_%d" % (a + 1, b, hap, k)
            ev = G.events_for(haps[hap], rng, a, b, 0.00012, False)
            ev = [e for e in ev if e[2] < 30 or any(e[0] == s["pos"] for s in haps[hap])]
            rev = (hap == 2 and k == 0)
            c, q = G.build_alignment(ref, ev, a, b)
            contig_recs.append((name, 16 if rev else 0, a, 60, c, q))
            contig_fa.append((name, G.revcomp(q) if rev else q))
    contig_recs.sort(key=lambda r: r[2])
    reads = []
    n_reads = int(L * COVERAGE / 9000)
    for i in range(n_reads):
        hap = rng.choice((1, 2))
        ln = max(600, int(rng.gauss(9000, 3000)))
        a = rng.randint(0, L - 700)
        b = min(L, a + ln)
        if b - a < 500 or (a < GAP[1] and b > GAP[0]):
            continue
        rev = rng.random() < 0.5
        mapq = 60 if rng.random() < 0.88 else rng.choice([0, 5, 19, 20, 33, 49, 50, 55])
        name = "read%05d" % i
        letter = "ACGT"[i % 4]
        ev = ont_events(haps[hap], rng, a, b)
        big = [e for e in ev if e[2] >= 100 and a + 800 < e[0] < b - 800 - (e[2] if e[1] == "D" else 0)]
        u = rng.random()
        if big and u < 0.5:
            # split read: the primary keeps the left part and soft-clips the rest, supplementary alignments (hard clipped) carry
            # the right part; now and then the right part is cut into many pieces (> 7 segments)
            p, t, sl, sq = big[0]
            r2s = p + sl if t == "D" else p
            c1, q1 = G.build_alignment(ref, [e for e in ev if e[0] < p], a, p)
            mid_len = sl if t == "I" else 0
            pieces = 1 if rng.random() < 0.8 else rng.randint(7, 9)
            cuts = sorted(rng.sample(range(r2s + 200, b - 200), pieces - 1)) if pieces > 1 and b - r2s > 400 + pieces else []
            bounds = [r2s] + cuts + [b]
            segs, sa, qpos = [], [], len(q1) + mid_len
            total_q2 = 0
            parts = []
            for s0, s1 in zip(bounds[:-1], bounds[1:]):
                c2, q2 = G.build_alignment(ref, [e for e in ev if s0 < e[0] < s1 - (e[2] if e[1] == "D" else 0) - 1], s0, s1)
                parts.append((s0, c2, len(q2)))
                total_q2 += len(q2)
            full_len = len(q1) + mid_len + total_q2
            st = "-" if rev else "+"
            for (s0, c2, l2) in parts:
                tail = full_len - qpos - l2
                sa_cig = "%dS%s%s" % (qpos, G.cigar_str(c2), ("%dS" % tail) if tail else "")
                sa.append("%s,%d,%s,%s,%d,%d" % ("chr1", s0 + 1, st, sa_cig, mapq, rng.randint(0, 30)))
                segs.append((s0, [(5, qpos)] + c2 + ([(5, tail)] if tail else []), l2))
                qpos += l2
            reads.append((name, 16 if rev else 0, a, mapq, c1 + [(4, full_len - len(q1))], letter * full_len,
                          ["SA:Z:" + ";".join(sa) + ";"]))
            for (s0, cg, l2) in segs:
                reads.append((name, 2064 if rev else 2048, s0, mapq, cg, letter * l2, []))
        else:
            c, q = G.build_alignment(ref, ev, a, b)
            lead = rng.choice([0, 0, 0, rng.randint(5, 300)])
            tail = rng.choice([0, 0, 0, rng.randint(5, 300)])
            cig = ([(4, lead)] if lead else []) + c + ([(4, tail)] if tail else [])
            reads.append((name, 16 if rev else 0, a, mapq, cig, letter * (lead + len(q) + tail), []))
            if u > 0.97:    # a secondary alignment identical to the primary: every `.sigs` row of it is an exact duplicate
                reads.append((name, (16 if rev else 0) | 256, a, mapq, cig, letter * (lead + len(q) + tail), []))
    reads.sort(key=lambda r: r[2])
    return ref, svs, contig_recs, contig_fa, reads


def main():
    if not os.path.isdir(LI):
        sys.exit("needs /root/reference (build container only)")
    if os.path.exists(OUT):
        shutil.rmtree(OUT)
    inp = os.path.join(OUT, "input")
    asm = os.path.join(inp, "asm", "chr1", "assembly", "final_contigs")
    os.makedirs(asm)
    ref, svs, contig_recs, contig_fa, reads = make_dataset()
    with open(os.path.join(inp, "ref.fa"), "w") as f:
        f.write(">chr1\n")
        for i in range(0, L, 60):
            f.write(ref[i:i + 60] + "\n")
    with open(os.path.join(asm, "Sample_final_contigs.fa"), "w") as f:
        for n, s in contig_fa:
            f.write(">%s\n%s\n" % (n, s))
    G.write_sam(os.path.join(inp, "contigs.sam"), L, contig_recs)
    G.write_sam(os.path.join(inp, "reads.sam"), L, reads)
    work = "/tmp/vsv_e2e_mid_ref"
    if os.path.exists(work):
        shutil.rmtree(work)
    shutil.copytree(inp, work)
    env = dict(os.environ)
    env.update(PYTHONPATH=os.path.join(HERE, "shims"), PATH=os.path.join(HERE, "stubs") + ":" + env["PATH"],
               LC_ALL="C", PYTHONHASHSEED="0", VSV_CONTIG_SAM=os.path.join(work, "contigs.sam"))
    cmd = ["python3", LI + "volcanosv-vc-large-indel.py", "-i", os.path.join(work, "asm"), "-o", os.path.join(work, "out"),
           "-dtype", "ONT", "-bam", os.path.join(work, "reads.sam"), "-ref", os.path.join(work, "ref.fa"), "-chr", "1",
           "-t", "2", "-px", "Sample"]
    with open(os.path.join(OUT, "reference_run.log"), "w") as log:
        log.write(" ".join(cmd) + "\n")
        log.flush()
        rc = subprocess.call(cmd, env=env, stdout=log, stderr=subprocess.STDOUT, cwd=work)
    print("reference exit code", rc)
    exp = os.path.join(OUT, "expected")
    keep = ("Sample_volcanosv_large_indel_chr1.vcf", "volcano_raw_variant.vcf", "volcano_variant_filtered.vcf",
            "reads_signature/chr1_reads_sig.txt", "reads_signature/chr1_DEL_reads_split.txt", "reads_signature/chr1_INS_reads_split.txt",
            "final_vcf/volcano_variant_no_redundancy.vcf", "final_vcf/variants_filtered_GT_corrected.vcf",
            "final_vcf/cute_sig/DEL.sigs", "final_vcf/cute_sig/INS.sigs", "final_vcf/cute_sig/INS.sigs.gte30auto",
            "final_vcf/GT_Correction/bnd_del_real.tsv", "final_vcf/GT_Correction/bnd_ins_real.tsv",
            "final_vcf/volcano_variant_no_redundancy_filter_DEL.vcf",
            "final_vcf/volcano_variant_no_redundancy_cutesv_sig_support_mins30_fl1000.csv")
    for rel in keep:
        src = os.path.join(work, "out", "chr1", rel)
        if not os.path.exists(src):
            print("missing:", rel)
            continue
        dst = os.path.join(exp, rel) + ".gz"
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        data = open(src, "rb").read().replace(work.encode(), b"$WORK")
        with open(dst, "wb") as raw, gzip.GzipFile(fileobj=raw, mode="wb", compresslevel=9, mtime=0) as fo:
            fo.write(data)
    for root, _, files in os.walk(inp):
        for fn in files:
            if fn.endswith((".sam", ".fa")):
                subprocess.check_call(["gzip", "-9", "-n", "-f", os.path.join(root, fn)])
    sz = sum(os.path.getsize(os.path.join(r, f)) for r, _, fs in os.walk(OUT) for f in fs)
    print("fixture size: %.2f MB" % (sz / 1e6))


if __name__ == "__main__":
    main()
