"""The oracle (oracle/vsv_oracle.py) against fixtures produced by executing the
reference's own functions (tests/golden/make_golden.py).  CPU only."""
import math

import numpy as np

import oracle as O


def _scan_case(c, mode, min_len, min_mapq=0, mapq=60):
    ops, off = O.pack_cigar([c["cigar"]])
    return O.scan_cigar(ops, off, [c["pos"]], [mapq], mode, min_len, min_mapq)


def test_contig_scan_merge(golden):
    for c in golden("contig_scan"):
        sigs, rt, qt, xt = _scan_case(c, O.MODE_CONTIG, 30)
        d, s = O.merge_in_record(sigs, O.MODE_CONTIG)
        assert [[x[2], x[3], x[4], x[5]] for x in d] == c["dels"]
        assert [[x[2], x[3], x[4], x[5]] for x in s] == c["inss"]
        assert c["pos"] + rt[0] == c["oref"] and qt[0] == c["oq"]


def test_reads_scan(golden):
    for c in golden("reads_scan"):
        sigs, rt, qt, xt = _scan_case(c, O.MODE_READS, 30)
        assert [[x[2], x[3], x[4]] for x in sigs if x[1] == O.DEL] == c["dels"]
        assert [[x[2], x[3], x[4]] for x in sigs if x[1] == O.INS] == c["inss"]
        assert c["pos"] + rt[0] == c["oref"] and qt[0] == c["oq"]


def parse_sa(s):
    """Host-side SA pre-parse as the product does it (SE:328-343,352-358)."""
    import re
    chrom, pos, strand, cg, mq, _nm = s.split(",")
    items = [(int(n), ch) for n, ch in re.findall(r"(\d+)([MIDNSHP=X])", cg)]
    fs = items[0][0] if items[0][1] == "S" else 0
    ls = items[-1][0] if items[-1][1] == "S" else 0
    bias = sum(n for n, ch in items if ch in "MD=X")
    return (chrom, int(pos), 0 if strand == "+" else 1, fs, ls, bias, int(mq))


def cute_rows(c, chrom="chr1", min_mapq=20):
    """Oracle restatement of parse_read (SE:438-512) for one record."""
    cg = c["cigar"]
    seq = c["seq"]
    qlen = len(seq)
    rows = []
    if qlen < 500:
        return rows
    ops, off = O.pack_cigar([cg])
    sigs, rt, qt, xt = O.scan_cigar(ops, off, [c["pos"]], [c["mapq"]], O.MODE_CUTE, 10, min_mapq)
    d, s = O.merge_in_record(sigs, O.MODE_CUTE)
    inss = [x for x in sigs if x[1] == O.INS]
    for m in s:
        piece = "".join(seq[inss[k][4]:inss[k][5]] for k in range(m[6], m[6] + m[7]))
        rows.append(["INS", m[2], m[3], piece])
    for m in d:
        rows.append(["DEL", m[2], m[3]])
    if c["flag"] in (0, 16) and c["sa"]:
        prim = None
        if c["mapq"] >= min_mapq:
            f, l = cg[0], cg[-1]
            left = f[1] if f[0] in (4, 5) else 0
            right = l[1] if l[0] in (4, 5) else 0
            ref_end = c["pos"] + int(rt[0]) + int(xt[0])
            if c["flag"] == 0:
                prim = (left, qlen - right, c["pos"], ref_end, chrom, 0)
            else:
                prim = (right, qlen - left, c["pos"], ref_end, chrom, 1)
        seg = O.sa_segments(prim, [parse_sa(x) for x in c["sa"]], qlen, min_mapq)
        for (tp, p, ln, q0, q1, rev) in O.sa_split_analysis(seg, qlen):
            if tp == O.INS:
                q = seq[::-1] if rev else seq
                rows.append(["INS", p, ln, q[q0:q1]])
            else:
                rows.append(["DEL", p, ln])
    return rows


def test_parse_read(golden):
    for c in golden("parse_read"):
        assert cute_rows(c) == c["rows"], c["sa"]


def test_split_pair(golden):
    g = golden("split_pair")
    var = dict(hifi=O.SPLIT_HIFI, ont=O.SPLIT_ONT, clr=O.SPLIT_CLR, reads=O.SPLIT_READS)
    nhit = 0
    for key, cases in g.items():
        for c in cases:
            def info(cg, pos, readsmode):
                ref_end = pos + sum(l for o, l in cg if o in (0, 2, 3, 7, 8))
                rlset = (0, 7, 8, 1, 4, 5) if readsmode else (0, 1, 4, 5)
                rl = sum(l for o, l in cg if o in rlset)
                return ref_end, rl, (cg[0][1] << 4) | cg[0][0], (cg[-1][1] << 4) | cg[-1][0]
            e1, rl1, f1, l1 = info(c["c1"], c["p1"], key == "reads")
            e2, rl2, f2, l2 = info(c["c2"], c["p2"], key == "reads")
            r = O.split_pair(var[key], c["p1"], e1, rl1, l1, c["m1"], c["rev1"], c["p2"], f2, c["m2"],
                             c["rev2"], c["min_mapq"])
            exp = c["res"]
            if exp is None:
                assert r is None
            else:
                nhit += 1
                assert r is not None and [("DEL", "INS")[r[0]], r[1], r[2], r[3], r[4]] == exp
    assert nhit > 100


def test_cluster_and_pair(golden):
    g = golden("cluster_pair")
    for c in g["cluster"]:
        assert O.cluster_leader(c["pos"], c["len"], c["type"] == "DEL") == c["kept"]
    for c in g["pair"]:
        h1, h2 = c["h1"], c["h2"]
        t = {"DEL": 0, "INS": 1}
        p1, p2 = O.pair_haps([s[2] for s in h1], [s[3] for s in h1], [t[s[1]] for s in h1],
                             [s[2] for s in h2], [s[3] for s in h2], [t[s[1]] for s in h2])
        # rebuild the reference's paired list (ECS_H:570-598) from the partner arrays
        res = []
        for i, s in enumerate(h1):
            if p1[i] < 0:
                res.append((s[2], s[3], "0/1", s[4]))
            else:
                o = h2[p1[i]]
                w = s if s[3] > o[3] else o
                res.append((w[2], w[3], "1/1", w[4]))
        for j, s in enumerate(h2):
            if p2[j] < 0:
                res.append((s[2], s[3], "0/1", s[4]))
        order = O.stable_order_by_pos([r[0] for r in res])
        res = [res[k] for k in order]
        assert res == [(s[2], s[3], s[10], s[4]) for s in c["paired"]]


def test_window_joins(golden):
    g = golden("window")
    for c in g["fp"]:
        assert O.fp_support(c["call_pos"], c["call_len"], c["sig_pos"], c["sig_len"]) == c["support"]
    for c in g["cov"]:
        assert O.ins_call_cov(c["ins_call_pos"], c["ins_sig_pos"], c["ins_sig_len"]) == c["ins_cov"]
        assert [float(x) for x in O.del_call_cov(c["del_call_start"], c["del_call_end"], c["del_sig_pos"],
                                                 c["del_sig_len"])] == c["del_cov"]


def test_support_depth_thresholds(golden):
    g = golden("genotype")
    for c in g["support"]:
        cid = lambda s: int(s[3:])
        uc, up, ul, un = O.group_sigs([cid(r[0]) for r in c["rows"] if c["vtype"] == "DEL" or r[2] >= 30],
                                      [r[1] for r in c["rows"] if c["vtype"] == "DEL" or r[2] >= 30],
                                      [r[2] for r in c["rows"] if c["vtype"] == "DEL" or r[2] >= 30])
        assert [[cid(u[0]), u[1], u[2], u[3]] for u in c["uniq"]] == [list(x) for x in zip(uc, up, ul, un)]
        sup, mids = O.sig_support([cid(x[0]) for x in c["cands"]], [x[1] for x in c["cands"]],
                                  [x[2] for x in c["cands"]], uc, up, ul, un)
        assert sup == c["support"]
        if c["match_id"] is not None:
            assert mids == c["match_id"]
    d = g["depth"]
    for c in d["cases"]:
        assert O.depth_del(d["read_start"], d["read_end"], c["pos"], c["svlen"]) == int(round(2 * c["del_depth"]))
        assert O.depth_ins(d["read_start"], d["read_end"], c["pos"]) == 2 * c["ins_depth"]
    gts = {"0/1": 0, "1/1": 1}
    for c in g["thresholds"]:
        thr = [float("nan") if x is None else x for x in c["thr"]]
        vt = O.DEL if c["vtype"] == "DEL" else O.INS
        for (sv, gt, nsup, d2), exp in zip(c["rows"], c["new_gt"]):
            _, new = O.genotype(vt, sv, gts.get(gt, 2), nsup, d2, thr)
            assert {0: "0/1", 1: "1/1", 2: gt}[new] == exp


def test_sa_split_analysis(golden):
    nrow = 0
    for c in golden("sa"):
        seg = [(s[0], s[1], s[2], s[3], s[4], 0 if s[5] == "+" else 1) for s in c["segs"]]
        rows = []
        for (tp, p, ln, q0, q1, rev) in O.sa_split_analysis(seg, c["L"]):
            if tp == O.INS:
                q = c["query"][::-1] if rev else c["query"]
                rows.append(["INS", p, ln, q[q0:q1]])
            else:
                rows.append(["DEL", p, ln])
        assert rows == c["rows"], c["segs"]
        nrow += len(rows)
    assert nrow > 50
