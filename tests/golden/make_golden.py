#!/usr/bin/env python
"""Generate tests/golden/*.json by EXECUTING THE REFERENCE's own functions
(/root/reference, read-only) on seeded synthetic inputs.  Run in the build container:

    python tests/golden/make_golden.py

The fixtures hold inputs and the reference's outputs; tests compare the oracle and the
CUDA path against them on machines where /root/reference does not exist.
Pins applied: np.argsort -> kind='stable'.
"""
import json
import os
import random
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness as H  # noqa: E402

OPC = "MIDNSHP=X"


def rand_cigar(rng, alphabet, n_inner, clip=True):
    """alphabet: list of match-like op codes to draw from (0 / 7 / 8 / 3)."""
    c = []
    if clip and rng.random() < 0.5:
        c.append((rng.choice([4, 5]), rng.randint(1, 3000)))
    for i in range(n_inner):
        c.append((rng.choice(alphabet), rng.randint(1, 900)))
        if i != n_inner - 1:
            op = rng.choice([1, 2])
            u = rng.random()
            if u < 0.45:
                ln = rng.randint(1, 9)
            elif u < 0.6:
                ln = rng.randint(10, 29)
            elif u < 0.9:
                ln = rng.randint(30, 400)
            else:
                ln = rng.randint(400, 5000)
            c.append((op, ln))
            # sometimes an adjacent second variant op with no match between
            if rng.random() < 0.15:
                c.append((3 - op if rng.random() < 0.5 else op, rng.choice([5, 12, 35, 180, 400])))
    if clip and rng.random() < 0.5:
        c.append((rng.choice([4, 5]), rng.randint(1, 3000)))
    return c


def seq_for(rng, cigar):
    n = sum(l for o, l in cigar if o in (0, 1, 4, 7, 8))
    return "".join(rng.choice("ACGT") for _ in range(n))


def sig_rows(lst):
    return [[s[2], s[3], s[5], s[6]] for s in lst]


def gold_contig_scan():
    ns = H.load_defs("extract_contig_signature_Hifi.py")
    rng = random.Random(101)
    cases = []
    for i in range(60):
        cg = rand_cigar(rng, [0], rng.randint(1, 40))
        rd = H.Read("PS1_9_hp1_%d" % i, rng.randint(0, 10**6), cg, mapq=60, is_reverse=rng.random() < 0.5)
        d, s, oref, oq = ns["extract_sig_from_cigar"](rd, 30)
        cases.append(dict(pos=rd.pos, cigar=cg, dels=sig_rows(d), inss=sig_rows(s), oref=oref, oq=oq))
    return cases


def gold_reads_scan():
    ns = H.load_defs("extract_reads_signature.py")
    rng = random.Random(202)
    cases = []
    for i in range(60):
        cg = rand_cigar(rng, [0, 7, 8, 0, 3], rng.randint(1, 40))
        rd = H.Read("r%d" % i, rng.randint(0, 10**6), cg)
        d, s, oref, oq = ns["extract_sig_from_cigar"](rd, 30)
        cases.append(dict(pos=rd.pos, cigar=cg, dels=[[x[2], x[3], x[5]] for x in d],
                          inss=[[x[2], x[3], x[5]] for x in s], oref=oref, oq=oq))
    return cases


def gold_parse_read():
    ns = H.load_defs("sig_extract.py", keep_assign=True)
    rng = random.Random(303)
    cases = []
    for i in range(80):
        cg = rand_cigar(rng, [0, 7, 8, 0], rng.randint(1, 30))
        seq = seq_for(rng, cg)
        rev = rng.random() < 0.5
        flag = rng.choice([0, 16, 0, 16, 2048, 2064, 256]) if i % 5 == 0 else (16 if rev else 0)
        mapq = rng.choice([60, 60, 60, 19, 20, 5])
        tags = []
        sa = []
        if rng.random() < 0.5:
            L = len(seq)
            rd_end = None
            for _ in range(rng.randint(1, 3)):
                fs = rng.randint(0, max(1, L - 50))
                m = rng.randint(20, max(21, L - fs))
                ls = max(0, L - fs - m)
                dlen = rng.choice([0, 0, 40, 700])
                cg_sa = ("%dS" % fs if fs else "") + "%dM" % m + ("%dD%dM" % (dlen, 1) if dlen else "") + ("%dS" % ls if ls else "")
                chrom = "chr1" if rng.random() < 0.85 else "chr2"
                strand = "+" if (rng.random() < 0.8) != rev else "-"
                sa.append("%s,%d,%s,%s,%d,%d" % (chrom, rng.randint(1, 10**6), strand, cg_sa,
                                                 rng.choice([60, 60, 10]), rng.randint(0, 50)))
            tags.append(("SA", ";".join(sa) + ";"))
        rd = H.Read("q%d" % i, rng.randint(1000, 10**6), cg, mapq=mapq, is_reverse=rev, chrom="chr1",
                    seq=seq, flag=flag, tags=tags)
        cand = ns["parse_read"](rd, "chr1", 30, 20, 7, 500, 10, 0, 100, 100000)
        rows = []
        for e in cand:
            if e[-2] == "DEL":
                rows.append(["DEL", int(e[0]), int(e[1])])
            elif e[-2] == "INS":
                rows.append(["INS", int(e[0]), int(e[1]), e[3]])
        cases.append(dict(pos=rd.pos, cigar=cg, seq=seq, flag=flag, mapq=mapq, sa=sa, rows=rows))
    # the SURVEY App. B-8 known-answer reads (true 500 bp deletion / 300 bp insertion via SA)
    for nm, pos, cg, sa in [("readB", 50000, [(0, 3000), (4, 4000)], ["chr1,53501,+,3000S4000M,60,10"]),
                            ("readC", 80000, [(0, 3000), (4, 4300)], ["chr1,83001,+,3300S4000M,60,10"])]:
        seq = seq_for(rng, cg)
        rd = H.Read(nm, pos, cg, chrom="chr1", seq=seq, flag=0, tags=[("SA", ";".join(sa) + ";")])
        cand = ns["parse_read"](rd, "chr1", 30, 20, 7, 500, 10, 0, 100, 100000)
        rows = [[e[-2], int(e[0]), int(e[1])] + ([e[3]] if e[-2] == "INS" else []) for e in cand if e[-2] in ("DEL", "INS")]
        cases.append(dict(pos=pos, cigar=cg, seq=seq, flag=0, mapq=60, sa=sa, rows=rows))
    return cases


def gold_split_pair():
    out = {}
    for key, script in [("hifi", "extract_contig_signature_Hifi.py"), ("ont", "extract_contig_signature_ONT.py"),
                        ("clr", "extract_contig_signature_CLR.py"), ("reads", "extract_reads_signature.py")]:
        ns = H.load_defs(script)
        rng = random.Random(404)
        cases = []
        for i in range(300):
            rl = rng.randint(5000, 40000)
            cut = rng.randint(1000, rl - 1000)
            ovl = rng.choice([0, 0, rng.randint(-600, 600), rng.randint(-4000, 4000)])
            a_q = cut
            b_q = cut + ovl
            b_q = min(max(b_q, 1), rl - 1)
            c1 = [(0, a_q), (rng.choice([4, 5]), rl - a_q)]
            if rng.random() < 0.1:
                c1 = [(4, 100), (0, a_q - 100), (0 if rng.random() < 0.3 else 4, rl - a_q)]
            c2 = [(rng.choice([4, 5]), b_q), (0, rl - b_q)]
            if rng.random() < 0.05:
                c2 = [(0, b_q), (0, rl - b_q)]
            p1 = rng.randint(1000, 10**6)
            gap = rng.choice([0, rng.randint(-500, 500), rng.randint(30, 3000), rng.randint(-3000, -30),
                              rng.randint(-60000, 60000)])
            p2 = max(p1, p1 + a_q + gap)
            rev = rng.random() < 0.5
            rev2 = rev if rng.random() < 0.9 else (not rev)
            m1, m2 = rng.choice([60, 60, 55, 20]), rng.choice([60, 60, 50, 3])
            r1 = H.Read("n", p1, c1, mapq=m1, is_reverse=rev)
            r2 = H.Read("n", p2, c2, mapq=m2, is_reverse=rev2)
            mm = 0 if key == "reads" else 50
            d, s = ns["extract_sig_from_split"](r1, r2, min_mapq=mm, max_svlen=50000)
            res = None
            if d:
                res = ["DEL", d[0][2], d[0][3], d[0][5], d[0][6]]
            if s:
                assert res is None
                res = ["INS", s[0][2], s[0][3], s[0][5], s[0][6]]
            cases.append(dict(p1=p1, c1=c1, m1=m1, rev1=rev, p2=p2, c2=c2, m2=m2, rev2=rev2, min_mapq=mm, res=res))
        out[key] = cases
    return out


def gold_cluster_pair():
    ns = H.load_defs("extract_contig_signature_Hifi.py")
    rng = random.Random(505)
    out = dict(cluster=[], pair=[])

    def rand_sigs(n, tp, name):
        p = 1000
        lst = []
        for i in range(n):
            p += rng.choice([0, 3, 20, 60, 99, 100, 101, 150, 400, 2000])
            ln = rng.choice([30, 45, 60, 61, 90, 119, 120, 121, 300, 1000])
            q = rng.randint(0, 10**5)
            lst.append(["chr10", tp, p, ln, name + "_%d" % rng.randint(0, 3), q, q + (ln if tp == "INS" else 1),
                        rng.choice("+-"), rng.choice(["cigar", "split-alignment"]), rng.choice([60, 55, "60-52"])])
        return lst

    for t in range(30):
        for tp in ("DEL", "INS"):
            lst = rand_sigs(rng.randint(0, 40), tp, "PS1_2_hp1")
            if not lst:
                continue
            fn = ns["cluster_del"] if tp == "DEL" else ns["cluster_ins"]
            kept = fn(lst)
            out["cluster"].append(dict(type=tp, pos=[s[2] for s in lst], len=[s[3] for s in lst],
                                       kept=[lst.index(k) if lst.count(k) == 1 else [i for i, s in enumerate(lst) if s is k][0] for k in kept]))
    for t in range(30):
        h1 = rand_sigs(rng.randint(0, 25), "DEL", "PS1_2_hp1") + rand_sigs(rng.randint(0, 25), "INS", "PS1_2_hp1")
        h2 = rand_sigs(rng.randint(0, 25), "DEL", "PS1_2_hp2") + rand_sigs(rng.randint(0, 25), "INS", "PS1_2_hp2")
        h1 = ns["sort_sig"](h1)
        h2 = ns["sort_sig"](h2)
        paired = ns["pair_sig"](h1, h2, 1000, 200, 0.5, 0.5)
        out["pair"].append(dict(h1=h1, h2=h2, paired=paired))
    return out


def gold_window():
    out = {}
    rng = random.Random(606)
    # FP filter
    ns = H.load_defs("FP_filter_v1.py")
    cases = []
    for t in range(20):
        sig2 = sorted([["chr10", rng.choice(["DEL", "INS"]), rng.randint(0, 60000), rng.choice([0, 30, 50, 80, 120, 300])]
                       for _ in range(rng.randint(0, 200))], key=lambda s: s[2])
        calls = [["chr10", "DEL", rng.randint(0, 60000), rng.choice([30, 60, 100, 250, 251, 800])] for _ in range(40)]
        sup = ns["eval_sig"](calls, sig2, 1000, 250, 500, 0.5)
        cases.append(dict(call_pos=[c[2] for c in calls], call_len=[c[3] for c in calls],
                          sig_pos=[s[2] for s in sig2], sig_len=[s[3] for s in sig2], support=[int(x) for x in sup]))
    out["fp"] = cases
    # coverage
    ns = H.load_defs("calculate_signature_support.py", extra=dict(flanking=1000, min_size=30))
    cases = []
    for t in range(20):
        n = rng.randint(1, 150)
        isig = sorted([(p, p + 1, rng.randint(10, 500)) for p in (rng.randint(0, 80000) for _ in range(n))])
        cpos = sorted(rng.randint(0, 80000) for _ in range(30))
        icall = [(p, p + 1, 50, "id", "0/1", "INS") for p in cpos]
        dci = ns["calc_ins_call_cov"](icall, isig)
        dsig = sorted([(p, p + l, -l) for p, l in ((rng.randint(0, 80000), rng.choice([10, 40, 100, 900, 6000])) for _ in range(n))])
        dcall = sorted([(p, p + l, -l, "id", "0/1", "DEL") for p, l in ((rng.randint(1000, 80000), rng.choice([30, 100, 1500])) for _ in range(30))])
        dcd = ns["calc_del_call_cov"](dcall, dsig)
        cases.append(dict(ins_sig_pos=[s[0] for s in isig], ins_sig_len=[s[2] for s in isig], ins_call_pos=cpos,
                          ins_cov=[float(dci.get(p, 0)) for p in cpos],
                          del_sig_pos=[s[0] for s in dsig], del_sig_len=[-s[2] for s in dsig],
                          del_call_start=[c[0] for c in dcall], del_call_end=[c[1] for c in dcall],
                          del_cov=[float(dcd.get((c[0], c[1]), 0)) for c in dcall]))
    out["cov"] = cases
    return out


def gold_genotype(tmpdir="/tmp/vsv_golden"):
    os.makedirs(tmpdir, exist_ok=True)
    out = {}
    rng = random.Random(707)
    gtd = H.load_defs("correct_gt_del_real_data.py", extra=dict(vtype="DEL", n_thread=1))
    gti = H.load_defs("correct_gt_ins_real_data.py", extra=dict(vtype="INS", n_thread=1))
    cases = []
    for t in range(12):
        chroms = ["chr1", "chr10", "chr2"] if t % 2 else ["chr1"]
        rows = []
        for c in chroms:
            for _ in range(rng.randint(5, 120)):
                rows.append((c, rng.randint(100, 40000), rng.choice([30, 35, 60, 61, 100, 167, 300, 12, 2000]), "r%d" % rng.randint(0, 50)))
        rows += rows[: len(rows) // 3]
        rows.sort(key=lambda r: (r[0], r[1], str(r[2]), r[3]))
        for vt, ns in (("DEL", gtd), ("INS", gti)):
            path = os.path.join(tmpdir, "%s_%d.sigs" % (vt, t))
            with open(path, "w") as f:
                for r in rows:
                    f.write("%s\t%s\t%d\t%d\t%s%s\n" % (vt, r[0], r[1], r[2], r[3], "\tACGT" if vt == "INS" else ""))
            cands = []
            for c in sorted(chroms, key=lambda x: int(x[3:])):
                for p in sorted(rng.randint(300, 40000) for _ in range(25)):
                    cands.append((c, p, rng.choice([30, 50, 100, 101, 300, 1001, 2500])))
            if vt == "DEL":
                sig_list = ns["load_sig"](path)
                var_list = [("0/1", sv, "%s\t%d\tid%d\tN\tN\t20\tPASS\tSVLEN=-%d;SVTYPE=DEL\tGT\t0/1" % (c, p, i, sv))
                            for i, (c, p, sv) in enumerate(cands)]
                sup = ns["match_varlist_siglist"](sig_list, var_list, 0.6, 2.3)
                uniq = [[s[0], s[1], s[2], s[3]] for s in sig_list]
                mids = None
            else:
                sv_list = [[int(c[3:]), p, sv, "0/1", "id%d" % i] for i, (c, p, sv) in enumerate(cands)]
                sup, mids = ns["extract_sig_support"](sv_list, path, 2.3, 0.6)
                dc = ns["load_sig_file"](path)
                uniq = [["chr%d" % k[0], k[1], k[2], v] for k, v in dc.items()]
            cases.append(dict(vtype=vt, rows=[[r[0], r[1], r[2]] for r in rows], uniq=uniq,
                              cands=[[c, p, sv] for c, p, sv in cands], support=[int(x) for x in sup], match_id=mids))
    out["support"] = cases
    # depth
    H.FakePysam.reads = []
    rs = sorted(rng.randint(0, 50000) for _ in range(400))
    for i, s in enumerate(rs):
        H.FakePysam.reads.append(H.Read("d%d" % i, s, [(0, rng.choice([200, 1500, 5000, 20000]))], chrom="chr1"))
    dcases = []
    for _ in range(60):
        p, sv = rng.randint(300, 50000), rng.choice([30, 100, 1000, 1001, 4000])
        dcases.append(dict(pos=p, svlen=sv, del_depth=float(gtd["check_full_cover_reads"]("x", "chr1", p, sv)),
                           ins_depth=int(gti["check_full_cover_reads"]("x", "chr1", p, 100))))
    out["depth"] = dict(read_start=[r.reference_start for r in H.FakePysam.reads],
                        read_end=[r.reference_end for r in H.FakePysam.reads], cases=dcases)
    # thresholds
    import pandas as pd
    tcases = []
    for dtype in ("Hifi", "ONT", "CLR"):
        for vt, ns in (("DEL", gtd), ("INS", gti)):
            para = ns["read_para"](H.LI + "para/GT_correction_para_%s_%s.txt" % (dtype, vt))
            thr = [para["t_large_11"], para["t_small_11"], para["t_large_01"], para["t_small_01"]]
            rowsd = []
            for _ in range(40):
                sv = rng.choice([30, 1000, 1001, 5000])
                gt = rng.choice(["0/1", "1/1", "./."])
                nsup = rng.choice([0, 1, 3, 6, 59, 60, 61, 69, 71, 79, 91, 100])
                d2 = rng.choice([0, 2, 10, 200, 200, 201])
                if vt == "DEL":
                    ratio = 1 if d2 == 0 else nsup / (d2 / 2)
                else:
                    with np.errstate(all="ignore"):
                        ratio = float(np.float64(nsup) / np.float64(d2 / 2))
                rowsd.append((sv, gt, nsup, d2, ratio))
            df = pd.DataFrame(dict(svlen=[r[0] for r in rowsd], call_gt=[r[1] for r in rowsd],
                                   n_ratio=[r[4] for r in rowsd]))
            # the reference round-trips the frame through a TSV before thresholding (GTD:231,379)
            pth = os.path.join(tmpdir, "df.tsv")
            df.to_csv(pth, sep="\t", index=False)
            df = pd.read_csv(pth, sep="\t")
            new = ns["correct_gt_eval"](df, *thr)
            tcases.append(dict(dtype=dtype, vtype=vt, thr=[None if x != x else x for x in thr],
                               rows=[[r[0], r[1], r[2], r[3]] for r in rowsd], new_gt=list(new)))
    out["thresholds"] = tcases
    return out


def gold_sa():
    """analysis_split_read (SE:193-326) on coherent random segment lists."""
    ns = H.load_defs("sig_extract.py", keep_assign=True)
    rng = random.Random(808)
    cases = []
    for t in range(300):
        L = rng.randint(1200, 4000)
        k = rng.choice([2, 2, 2, 3, 3, 4, 5])
        cuts = sorted(rng.sample(range(100, L - 100), k - 1))
        qb = [0] + cuts + [L]
        strand = rng.choice("+-")
        r = rng.randint(10000, 10**6)
        segs = []
        for i in range(k):
            qs, qe = qb[i], qb[i + 1]
            qs2 = qs + rng.choice([0, 0, 0, rng.randint(-50, 50), rng.randint(30, 600)]) if i else qs
            qs2 = min(max(qs2, 0), qe - 1)
            chrom = "chr1" if rng.random() < 0.9 else "chr2"
            st = strand if rng.random() < 0.9 else ("-" if strand == "+" else "+")
            segs.append([qs2, qe, r, r + (qe - qs2), chrom, st])
            r = r + (qe - qs2) + rng.choice([0, 0, rng.randint(-29, 100), rng.randint(30, 5000), rng.randint(-200, -30), 200000])
        if strand == "-":
            # minus-strand reads list query coordinates mirrored
            segs = [[L - s[1], L - s[0]] + s[2:] for s in segs]
        rng.shuffle(segs)
        query = "".join(rng.choice("ACGT") for _ in range(L))
        cand = []
        ns["analysis_split_read"]([list(s) for s in segs], 30, L, "rd", cand, 100000, query)
        rows = []
        for e in cand:
            if e[-2] == "DEL":
                rows.append(["DEL", int(e[0]), int(e[1])])
            elif e[-2] == "INS":
                rows.append(["INS", int(e[0]), int(e[1]), e[3]])
        cases.append(dict(L=L, segs=segs, query=query, rows=rows))
    return cases


def gold_clr_gate():
    """ECS_C:53-70,425-427: ins_pct / var_dist of the reference on contig CIGARs around both thresholds (0.13, 200), incl. exact
    ties (13/100, 200.0) and contigs the gate rejects."""
    ns = H.load_defs("extract_contig_signature_CLR.py")
    rng = random.Random(99)
    cases = []
    fixed = [[(0, 87), (1, 13)], [(0, 870), (1, 130), (0, 1)], [(0, 200)], [(0, 100), (2, 5), (0, 300)], [(0, 199), (1, 40), (0, 200)],
             [(5, 10), (0, 50), (1, 30), (0, 50), (4, 7)], [(0, 1), (1, 1)], [(0, 400), (1, 200)]]
    for cg in fixed + [None] * 60:
        if cg is None:
            k = rng.randint(1, 40)
            mrun = rng.choice([20, 60, 150, 200, 260, 900])
            cg = []
            for _ in range(k):
                cg.append((0, max(1, int(rng.gauss(mrun, mrun / 4)))))
                u = rng.random()
                cg.append((1, rng.randint(1, mrun // 2 + 2)) if u < 0.75 else (2, rng.randint(1, 60)))
            cg.append((0, rng.randint(1, 400)))
        ip, vd = ns["ins_pct"](cg), ns["var_dist"](cg)
        cases.append(dict(cigar=[list(x) for x in cg], ins_pct=ip, var_dist=vd, gate=bool((ip <= 0.13) or (vd >= 200))))
    return cases


def main():
    assert H.available(), "reference not mounted; goldens can only be regenerated in the build container"
    jobs = dict(contig_scan=gold_contig_scan, reads_scan=gold_reads_scan, parse_read=gold_parse_read,
                split_pair=gold_split_pair, cluster_pair=gold_cluster_pair, window=gold_window,
                genotype=gold_genotype, sa=gold_sa, clr_gate=gold_clr_gate)
    only = sys.argv[1:]
    for k, fn in jobs.items():
        if only and k not in only:
            continue
        with open(os.path.join(HERE, k + ".json"), "w") as f:
            json.dump(fn(), f, separators=(",", ":"))
        print("wrote", k, os.path.getsize(os.path.join(HERE, k + ".json")))


if __name__ == "__main__":
    main()
