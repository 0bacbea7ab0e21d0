#!/usr/bin/env python
"""End-to-end golden fixture: runs the reference's large-indel stage UNMODIFIED
(`python3 LI/volcanosv-vc-large-indel.py ... -chr 1`, every operator script as its own
process, exactly as the reference chains them) on a small synthetic diploid chromosome and
stores inputs + every output file under tests/golden/e2e/.

Only test-side stand-ins are used for what this image lacks:
  tests/golden/shims/   pysam (SAM text), cigar, edlib, matplotlib; sitecustomize pins
                        np.argsort to kind='stable'
  tests/golden/stubs/   minimap2 (prints the pre-made contig alignments), samtools
                        (faidx / sort=cat / index=noop), vcf-sort (`sort -k1,1d -k2,2n`)
Run in the build container (needs /root/reference):  python tests/golden/make_e2e_golden.py
"""
import os
import random
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LI = "/root/reference/bin/VolcanoSV-vc/Large_INDEL/"
OUT = os.path.join(HERE, "e2e")
COMP = {"A": "T", "C": "G", "G": "C", "T": "A", "N": "N"}


def revcomp(s):
    return "".join(COMP[c] for c in reversed(s))


def rand_seq(rng, n):
    return "".join(rng.choice("ACGT") for _ in range(n))


def make_svs(rng, L):
    """Planted SVs: (pos, type, len, inserted sequence, zygosity)."""
    svs, pos = [], 5000
    sizes = [34, 41, 55, 72, 96, 130, 180, 240, 310, 420, 650, 900, 1300, 1900]
    while pos < L - 9000:
        t = "INS" if rng.random() < 0.55 else "DEL"
        ln = rng.choice(sizes) + rng.randint(0, 9)
        zyg = rng.choice(["hom", "hom", "h1", "h2"])
        svs.append(dict(pos=pos, type=t, len=ln, seq=rand_seq(rng, ln) if t == "INS" else "", zyg=zyg))
        pos += (ln if t == "DEL" else 0) + rng.randint(3500, 7000)
    return svs


def hap_svs(svs, hap, rng):
    """SVs carried by one haplotype; homozygous ones are jittered a little on hap 2."""
    out = []
    for s in svs:
        if s["zyg"] == "hom" or s["zyg"] == "h%d" % hap:
            s2 = dict(s)
            if hap == 2 and s["zyg"] == "hom" and rng.random() < 0.5:
                s2["pos"] = s["pos"] + rng.randint(-40, 40)
                if s["type"] == "INS":
                    k = rng.randint(0, min(8, s["len"] - 31))
                    s2["seq"] = s["seq"][: s["len"] - k]
                    s2["len"] = len(s2["seq"])
                else:
                    s2["len"] = s["len"] + rng.randint(0, 12)
            out.append(s2)
    return sorted(out, key=lambda x: x["pos"])


def build_alignment(ref, events, a, b):
    """CIGAR (list of (op,len)) and query sequence of an alignment of a haplotype segment to
    ref[a:b) carrying `events` = sorted [(pos, 'I'|'D', len, seq)] with a < pos, non-overlapping."""
    cig, q, cur = [], [], a
    for (p, t, ln, sq) in events:
        if p <= cur or p >= b - 1 or (t == "D" and p + ln >= b - 1):
            continue
        cig.append((0, p - cur))
        q.append(ref[cur:p])
        if t == "D":
            cig.append((2, ln))
            cur = p + ln
        else:
            cig.append((1, ln))
            q.append(sq)
            cur = p
    cig.append((0, b - cur))
    q.append(ref[cur:b])
    # merge nothing: M runs are separated by construction
    return cig, "".join(q)


def cigar_str(cig):
    return "".join("%d%s" % (l, "MIDNSHP=X"[o]) for o, l in cig)


def events_for(hsv, rng, a, b, small_rate, noisy):
    ev = []
    for s in hsv:
        if not (a + 300 < s["pos"] < b - 300):
            continue
        ln, sq, p = s["len"], s["seq"], s["pos"]
        if noisy and rng.random() < 0.12:                      # imprecise carriage of the SV
            p += rng.randint(-3, 3)
            if s["type"] == "INS":
                k = rng.randint(0, max(0, min(6, ln - 31)))
                sq, ln = sq[: ln - k], ln - k
            else:
                ln += rng.randint(0, 5)
        ev.append((p, "I" if s["type"] == "INS" else "D", ln, sq))
    n_small = int((b - a) * small_rate)
    for _ in range(n_small):
        p = rng.randint(a + 50, b - 50)
        u = rng.random()
        ln = rng.randint(1, 6) if u < 0.8 else (rng.randint(10, 29) if u < 0.97 else rng.randint(30, 44))
        t = "I" if rng.random() < 0.5 else "D"
        ev.append((p, t, ln, rand_seq(rng, ln) if t == "I" else ""))
        if rng.random() < 0.08:                                # a second nearby INS: exercises the SE merge
            ev.append((p + rng.randint(20, 90), "I", rng.randint(10, 25), None))
    ev = [(p, t, ln, sq if sq is not None or t == "D" else rand_seq(rng, ln)) for (p, t, ln, sq) in ev]
    ev.sort(key=lambda e: e[0])
    out, last_end = [], a + 20
    for (p, t, ln, sq) in ev:                                  # drop events that touch the previous one
        if p <= last_end + 8:
            continue
        out.append((p, t, ln, sq))
        last_end = p + (ln if t == "D" else 0)
    return out


def make_dataset(seed=20240611, L=140_000, chrom="chr1", lite=False, name_prefix="", coverage=32, ps_offset=0):
    """`lite` (whole-genome fixture: many small chromosomes) leaves out the hand-placed special cases,
    keeps read sequences compressible and forces one homozygous INS per chromosome (CSS:87 needs one)."""
    rng = random.Random(seed)
    ref = rand_seq(rng, L)
    svs = make_svs(rng, L)
    if lite:
        svs[0].update(type="INS", zyg="hom", seq=rand_seq(rng, svs[0]["len"]))
        return _finish_dataset(rng, ref, svs, L, chrom, True, name_prefix, coverage, ps_offset)
    return _finish_dataset(rng, ref, svs, L, chrom, False, name_prefix, coverage, ps_offset)


def _finish_dataset(rng, ref, svs, L, chrom, lite, name_prefix, coverage, ps_offset):
    # near-duplicate calls for remove_redundancy: the two haplotypes carry similar SVs too far apart
    # to be paired (> 200 bp) but close enough to be linked (INS: <= 500 bp + edit similarity;
    # DEL: <= 3000 bp + overlap)
    def mid(i):
        a = svs[i]["pos"] + (svs[i]["len"] if svs[i]["type"] == "DEL" else 0)
        return (a + svs[i + 1]["pos"]) // 2
    if not lite:
        s1 = rand_seq(rng, 300)
        s2 = "".join(c if rng.random() > 0.08 else rng.choice("ACGT") for c in s1) + rand_seq(rng, 11)
        m = mid(2)
        svs.append(dict(pos=m - 160, type="INS", len=len(s1), seq=s1, zyg="h1"))
        svs.append(dict(pos=m + 160, type="INS", len=len(s2), seq=s2, zyg="h2"))
        m = mid(6)
        svs.append(dict(pos=m - 300, type="DEL", len=500, seq="", zyg="h1"))
        svs.append(dict(pos=m + 50, type="DEL", len=523, seq="", zyg="h2"))
    svs.sort(key=lambda x: x["pos"])
    haps = {1: hap_svs(svs, 1, rng), 2: hap_svs(svs, 2, rng)}
    # contig-only artefacts (no read carries them): the read-support filter must drop them
    if lite:
        ghosts = {1: [(mid(1), "I", 61, rand_seq(rng, 61))] if len(svs) > 2 and rng.random() < 0.4 else [], 2: []}
    else:
        ghosts = {1: [(mid(9), "I", 61, rand_seq(rng, 61)), (mid(12), "D", 47, "")], 2: [(mid(15), "I", 120, rand_seq(rng, 120))]}
    # ---------------- contigs: two per haplotype, overlapping by ~8 kb; one on the minus strand,
    # one split (primary + supplementary) at a big deletion
    contig_recs, contig_fa = [], []
    for hap in (1, 2):
        spans = [(150, int(L * 0.56)), (int(L * 0.5), L - 120)] if lite else [(150, 76_000), (68_000, L - 120)]
        for k, (a, b) in enumerate(spans):
            name = "PS%d_%d_hp%d_%d" % (a + 1 + ps_offset, b + ps_offset, hap, k)
            ev = events_for(haps[hap], rng, a, b, 0.00012, False)
            ev = [e for e in ev if e[2] < 30 or any(e[0] == s["pos"] for s in haps[hap])]
            ev = sorted(ev + [g for g in ghosts[hap] if a + 500 < g[0] < b - 500], key=lambda e: e[0])
            rev = (hap == 2 and k == 0)
            big = [e for e in ev if e[1] == "D" and e[2] >= 600]
            if hap == 1 and k == 1 and big:
                # split at the first big deletion: rec1 = [a, p) + soft clip, rec2 = hard clip + [p+len, b)
                p, _, ln, _ = big[0]
                c1, q1 = build_alignment(ref, [e for e in ev if e[0] < p], a, p)
                c2, q2 = build_alignment(ref, [e for e in ev if e[0] > p + ln], p + ln, b)
                full = q1 + q2
                contig_recs.append((name, 0, a, 60, c1 + [(4, len(q2))], full))
                contig_recs.append((name, 2048, p + ln, 60, [(5, len(q1))] + c2, q2))
                contig_fa.append((name, full))
            else:
                c, q = build_alignment(ref, ev, a, b)
                contig_recs.append((name, 16 if rev else 0, a, 60, c, q))
                contig_fa.append((name, revcomp(q) if rev else q))
    contig_recs.sort(key=lambda r: r[2])
    # ---------------- reads: HiFi-like, 32x
    reads = []
    n_reads = int(L * coverage / 9000)
    rseq = (lambda n: "A" * n) if lite else (lambda n: rand_seq(rng, n))
    for i in range(n_reads):
        hap = rng.choice((1, 2))
        ln = max(300, int(rng.gauss(9000, 2500))) if rng.random() > 0.03 else rng.randint(200, 480)
        a = rng.randint(0, L - 400)
        b = min(L, a + ln)
        if b - a < 150:
            continue
        rev = rng.random() < 0.5
        mapq = 60 if rng.random() < 0.88 else rng.choice([0, 5, 19, 20, 33, 49, 50, 55])
        name = name_prefix + "read%05d" % i
        ev = events_for(haps[hap], rng, a, b, 0.0009, True)
        big = [e for e in ev if e[2] >= 600 and a + 800 < e[0] < b - 800 - (e[2] if e[1] == "D" else 0)]
        if big and rng.random() < 0.6:
            # split read: primary keeps the left part and soft-clips the rest, the supplementary
            # alignment (hard-clipped) carries the right part; SA tag on the primary
            p, t, sl, sq = big[0]
            r2s = p + sl if t == "D" else p
            c1, q1 = build_alignment(ref, [e for e in ev if e[0] < p], a, p)
            c2, q2 = build_alignment(ref, [e for e in ev if e[0] > r2s], r2s, b)
            mid = sq if t == "I" else ""
            full = q1 + mid + q2
            qseq = rseq(len(full))
            sa_cig = "%dS%s" % (len(q1) + len(mid), cigar_str(c2))
            st = "-" if rev else "+"
            sa = "SA:Z:%s,%d,%s,%s,%d,%d;" % (chrom, r2s + 1, st, sa_cig, mapq, rng.randint(0, 30))
            reads.append((name, 16 if rev else 0, a, mapq, c1 + [(4, len(mid) + len(q2))], qseq, [sa]))
            reads.append((name, 2064 if rev else 2048, r2s, mapq, [(5, len(q1) + len(mid))] + c2,
                          qseq[len(q1) + len(mid):], []))
        else:
            c, q = build_alignment(ref, ev, a, b)
            lead = rng.choice([0, 0, 0, rng.randint(5, 300)])
            tail = rng.choice([0, 0, 0, rng.randint(5, 300)])
            cig = ([(4, lead)] if lead else []) + c + ([(4, tail)] if tail else [])
            reads.append((name, 16 if rev else 0, a, mapq, cig, rseq(lead + len(q) + tail), []))
    reads.sort(key=lambda r: r[2])
    return ref, svs, contig_recs, contig_fa, reads


def write_sam(path, L, recs):
    with open(path, "w") as f:
        f.write("@HD\tVN:1.6\tSO:coordinate\n@SQ\tSN:chr1\tLN:%d\n" % L)
        for r in recs:
            name, flag, pos, mapq, cig, seq = r[:6]
            tags = r[6] if len(r) > 6 else []
            f.write("\t".join([name, str(flag), "chr1", str(pos + 1), str(mapq), cigar_str(cig), "*", "0", "0", seq, "*"] +
                              list(tags)) + "\n")


def main():
    if not os.path.isdir(LI):
        sys.exit("needs /root/reference (build container only)")
    if os.path.exists(OUT):
        shutil.rmtree(OUT)
    inp = os.path.join(OUT, "input")
    asm = os.path.join(inp, "asm", "chr1", "assembly", "final_contigs")
    os.makedirs(asm)
    ref, svs, contig_recs, contig_fa, reads = make_dataset()
    L = len(ref)
    with open(os.path.join(inp, "ref.fa"), "w") as f:
        f.write(">chr1\n")
        for i in range(0, L, 60):
            f.write(ref[i:i + 60] + "\n")
    with open(os.path.join(asm, "Sample_final_contigs.fa"), "w") as f:
        for n, s in contig_fa:
            f.write(">%s\n%s\n" % (n, s))
    write_sam(os.path.join(inp, "contigs.sam"), L, contig_recs)
    write_sam(os.path.join(inp, "reads.sam"), L, reads)
    with open(os.path.join(inp, "planted_svs.tsv"), "w") as f:
        for s in svs:
            f.write("%d\t%s\t%d\t%s\n" % (s["pos"], s["type"], s["len"], s["zyg"]))
    work = "/tmp/vsv_e2e_ref"
    if os.path.exists(work):
        shutil.rmtree(work)
    shutil.copytree(inp, work)
    env = dict(os.environ)
    env.update(PYTHONPATH=os.path.join(HERE, "shims"), PATH=os.path.join(HERE, "stubs") + ":" + env["PATH"],
               LC_ALL="C", PYTHONHASHSEED="0", VSV_CONTIG_SAM=os.path.join(work, "contigs.sam"))
    cmd = ["python3", LI + "volcanosv-vc-large-indel.py", "-i", os.path.join(work, "asm"), "-o", os.path.join(work, "out"),
           "-dtype", "Hifi", "-bam", os.path.join(work, "reads.sam"), "-ref", os.path.join(work, "ref.fa"), "-chr", "1",
           "-t", "2", "-px", "Sample"]
    with open(os.path.join(OUT, "reference_run.log"), "w") as log:
        log.write(" ".join(cmd) + "\n")
        log.flush()
        rc = subprocess.call(cmd, env=env, stdout=log, stderr=subprocess.STDOUT, cwd=work)
    print("reference exit code", rc)
    # keep every output file of the reference run (text, small); paths relative to out/
    exp = os.path.join(OUT, "expected")
    n = 0
    for root, _, files in os.walk(os.path.join(work, "out")):
        for fn in files:
            src = os.path.join(root, fn)
            rel = os.path.relpath(src, os.path.join(work, "out"))
            if fn.endswith(".sam") or fn.endswith(".bam") or "ref_by_chr" in rel or "/signatures/" in rel \
                    or os.path.getsize(src) > 3_000_000:
                continue
            dst = os.path.join(exp, rel)
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            with open(src, "rb") as fi:
                data = fi.read().replace(work.encode(), b"$WORK")
            with open(dst, "wb") as fo:
                fo.write(data)
            n += 1
    print("kept %d reference output files under %s" % (n, exp))
    # the committed inputs are gzip'ed (tests gunzip them into a temp dir)
    for root, _, files in os.walk(inp):
        for fn in files:
            if fn.endswith((".sam", ".fa")):
                subprocess.check_call(["gzip", "-9", "-n", "-f", os.path.join(root, fn)])


if __name__ == "__main__":
    main()
