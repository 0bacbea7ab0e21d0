"""CUDA stage 2 (split pairs, sort, cluster, pair, window joins) and stage 3 (group,
support, depth, genotype) against the reference-generated goldens and the oracle."""
import math
import os

import numpy as np
import pytest

import oracle as O
import volcanosv_b200 as V
from volcanosv_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = V.Context(0)
    yield c
    c.close()


def _info(cg, pos, readsmode):
    ref_end = pos + sum(l for o, l in cg if o in (0, 2, 3, 7, 8))
    rlset = (0, 7, 8, 1, 4, 5) if readsmode else (0, 1, 4, 5)
    rl = sum(l for o, l in cg if o in rlset)
    return ref_end, rl, (cg[0][1] << 4) | cg[0][0], (cg[-1][1] << 4) | cg[-1][0]


def test_split_pairs_golden(ctx, golden):
    g = golden("split_pair")
    var = dict(hifi=V.SPLIT_HIFI, ont=V.SPLIT_ONT, clr=V.SPLIT_CLR, reads=V.SPLIT_READS)
    for key, cases in g.items():
        pos, end, rl, fo, lo, mq, rv, nm = [], [], [], [], [], [], [], []
        for i, c in enumerate(cases):
            for cg, p, m, r in ((c["c1"], c["p1"], c["m1"], c["rev1"]), (c["c2"], c["p2"], c["m2"], c["rev2"])):
                e, l, f, la = _info(cg, p, key == "reads")
                pos.append(p); end.append(e); rl.append(l); fo.append(f); lo.append(la); mq.append(m); rv.append(r); nm.append(i)
        out = ctx.split_pairs(var[key], cases[0]["min_mapq"], pos, end, rl, fo, lo, mq, rv, nm)
        got = {int(s["rec"]) // 2: [("DEL", "INS")[s["type"]], int(s["pos"]), int(s["len"]), int(s["q0"]), int(s["q1"])] for s in out}
        assert all(int(s["first"]) == int(s["rec"]) + 1 for s in out)
        assert list(got) == sorted(got)
        for i, c in enumerate(cases):
            assert got.get(i) == c["res"], (key, i)


def test_split_pairs_driver_vs_oracle(ctx, golden):
    """Driver semantics (ECS_H:421-457): select filter, names in first-appearance order,
    consecutive records of a name in coordinate order -- records of the golden pairs are
    interleaved in BAM order, some names get a third record, some records are deselected."""
    rng = np.random.default_rng(5)
    cases = golden("split_pair")["hifi"]
    recs = []
    for i, c in enumerate(cases):
        nm = i if i % 7 else max(0, i - 1)          # every 7th pair shares the previous name
        for cg, p, m, r in ((c["c1"], c["p1"], c["m1"], c["rev1"]), (c["c2"], c["p2"], c["m2"], c["rev2"])):
            e, l, f, la = _info(cg, p, False)
            recs.append((p, e, l, f, la, m, r, nm))
    recs.sort(key=lambda t: t[0])
    pos, end, rl, fo, lo, mq, rv, raw = (np.array(x) for x in zip(*recs))
    n = len(pos)
    sel = (mq >= 50) & (rng.random(n) < 0.9)
    remap, nid = {}, np.full(n, -1, np.int32)
    for r in range(n):
        if sel[r]:
            nid[r] = remap.setdefault(int(raw[r]), len(remap))
    tot = 0
    for variant in (V.SPLIT_HIFI, V.SPLIT_ONT, V.SPLIT_CLR, V.SPLIT_READS):
        out = ctx.split_pairs(variant, 50, pos, end, rl, fo, lo, mq, rv, nid, sel)
        exp = O.split_pairs(variant, pos, end, rl, fo, lo, mq, rv, nid, sel, 50)
        got = [tuple(int(s[f]) for f in ("rec", "first", "type", "pos", "len", "q0", "q1")) for s in out]
        assert got == exp
        tot += len(exp)
    assert tot > 50


def test_argsort_stable(ctx):
    rng = np.random.default_rng(1)
    k = rng.integers(-50, 50, 5000).astype(np.int32)
    assert np.array_equal(ctx.argsort_stable(k), np.argsort(k, kind="stable"))


def test_cluster_pair_golden(ctx, golden):
    g = golden("cluster_pair")
    for c in g["cluster"]:
        assert ctx.cluster(c["type"] == "DEL", c["pos"], c["len"]).tolist() == c["kept"]
    t = {"DEL": 0, "INS": 1}
    for c in g["pair"]:
        h1, h2 = c["h1"], c["h2"]
        a = ([s[2] for s in h1], [s[3] for s in h1], [t[s[1]] for s in h1])
        b = ([s[2] for s in h2], [s[3] for s in h2], [t[s[1]] for s in h2])
        p1, p2 = ctx.pair_haps(*a, *b)
        e1, e2 = O.pair_haps(*a, *b)
        assert p1.tolist() == e1 and p2.tolist() == e2


def test_cluster_pair_large_vs_oracle(ctx):
    rng = np.random.default_rng(3)
    n = 3000
    pos = np.sort(rng.integers(0, 200000, n)).astype(np.int32)
    ln = rng.choice([30, 45, 60, 61, 90, 120, 121, 300, 1000], n).astype(np.int32)
    for is_del in (False, True):
        assert ctx.cluster(is_del, pos, ln).tolist() == O.cluster_leader(pos, ln, is_del)
    typ = rng.integers(0, 2, n).astype(np.int32)
    pos2 = np.sort(np.clip(pos + rng.integers(-250, 250, n), 0, None)).astype(np.int32)
    ln2 = rng.choice([30, 45, 60, 61, 90, 120, 121, 300, 1000], n).astype(np.int32)
    typ2 = rng.integers(0, 2, n).astype(np.int32)
    p1, p2 = ctx.pair_haps(pos, ln, typ, pos2, ln2, typ2)
    e1, e2 = O.pair_haps(pos, ln, typ, pos2, ln2, typ2)
    assert p1.tolist() == e1 and p2.tolist() == e2
    assert sum(1 for x in e1 if x >= 0) > 100


def test_window_joins_golden(ctx, golden):
    g = golden("window")
    for c in g["fp"]:
        assert ctx.fp_support(c["call_pos"], c["call_len"], c["sig_pos"], c["sig_len"]).tolist() == c["support"]
    for c in g["cov"]:
        assert [float(x) for x in ctx.ins_cov(c["ins_call_pos"], c["ins_sig_pos"], c["ins_sig_len"])] == c["ins_cov"]
        assert [float(x) for x in ctx.del_cov(c["del_call_start"], c["del_call_end"], c["del_sig_pos"], c["del_sig_len"])] == c["del_cov"]
    # empty signature table
    assert ctx.fp_support([5, 9], [30, 300], [], []).tolist() == [0, 60]
    assert ctx.ins_cov([5], [], []).tolist() == [0]
    assert ctx.del_cov([5], [9], [], []).tolist() == [0]


def test_support_depth_thresholds_golden(ctx, golden):
    g = golden("genotype")
    cid = lambda s: int(s[3:])
    for c in g["support"]:
        rows = [r for r in c["rows"] if c["vtype"] == "DEL" or r[2] >= 30]
        uc, up, ul, un = ctx.group_sigs([cid(r[0]) for r in rows], [r[1] for r in rows], [r[2] for r in rows])
        assert [[cid(u[0]), u[1], u[2], u[3]] for u in c["uniq"]] == [list(map(int, x)) for x in zip(uc, up, ul, un)]
        sup, mid, last = ctx.sig_support([cid(x[0]) for x in c["cands"]], [x[1] for x in c["cands"]],
                                         [x[2] for x in c["cands"]], uc, up, ul, un)
        assert sup.tolist() == c["support"]
        if c["match_id"] is not None:
            assert mid.tolist() == c["match_id"]
        # sharded by candidate halves with the carried last_match_idx == unsharded
        h = len(c["cands"]) // 2
        cc, cp, cl = [cid(x[0]) for x in c["cands"]], [x[1] for x in c["cands"]], [x[2] for x in c["cands"]]
        s1, _, l1 = ctx.sig_support(cc[:h], cp[:h], cl[:h], uc, up, ul, un)
        s2, _, l2 = ctx.sig_support(cc[h:], cp[h:], cl[h:], uc, up, ul, un, last_in=l1)
        assert s1.tolist() + s2.tolist() == c["support"] and l2 == last
    d = g["depth"]
    S, E, exp = [], [], []
    for c in d["cases"]:
        S.append(c["pos"] - 100); E.append(c["pos"] + 100); exp.append(c["ins_depth"])
    assert ctx.span_depth(d["read_start"], d["read_end"], S, E).tolist() == exp
    nan = float("nan")
    pos = [c["pos"] for c in d["cases"]]
    sv = [c["svlen"] for c in d["cases"]]
    z = [0] * len(pos)
    d2, ratio, ng = ctx.genotype(d["read_start"], d["read_end"], pos, sv, z, z, z, [nan] * 4, [nan] * 4)
    assert d2.tolist() == [int(round(2 * c["del_depth"])) for c in d["cases"]]
    d2, ratio, ng = ctx.genotype(d["read_start"], d["read_end"], pos, sv, [1] * len(pos), z, z, [nan] * 4, [nan] * 4)
    assert d2.tolist() == [2 * c["ins_depth"] for c in d["cases"]]


def test_genotype_thresholds_vs_oracle(ctx, golden):
    """Threshold decision incl. exact-tie ratios (3/5 vs 0.6, 69/100 vs 0.69), depth 0,
    inf/nan for INS; compared with the oracle that is pinned by the golden file."""
    g = golden("genotype")["thresholds"]
    rng = np.random.default_rng(9)
    # reads: k identical spanning reads over [0, 100000) -> depth k for any window
    for c in g:
        thr = [float("nan") if x is None else x for x in c["thr"]]
        vt = O.DEL if c["vtype"] == "DEL" else O.INS
        for depth in (0, 5, 100):
            rs = np.zeros(depth, np.int32)
            re = np.full(depth, 10**6, np.int32)
            n = 64
            sv = rng.choice([30, 1000, 1001, 5000], n).astype(np.int32)
            gt = rng.choice([0, 1, 2], n).astype(np.int32)
            ns = rng.choice([0, 1, 3, 6, 59, 60, 61, 69, 71, 79, 91, 100], n).astype(np.int32)
            pos = np.full(n, 5000, np.int32)
            tp = np.full(n, vt, np.int32)
            d2, ratio, ng = ctx.genotype(rs, re, pos, sv, tp, gt, ns, thr if vt == O.DEL else [float("nan")] * 4,
                                         thr if vt == O.INS else [float("nan")] * 4)
            for i in range(n):
                er, eg = O.genotype(vt, int(sv[i]), int(gt[i]), int(ns[i]), 2 * depth, thr)
                assert int(d2[i]) == 2 * depth and int(ng[i]) == eg
                assert (math.isnan(er) and math.isnan(ratio[i])) or er == ratio[i]


def test_depth_stress_properties(ctx):
    """10^6 reads x 10^5 windows: compare with a numpy evaluation on a sample and check
    monotonicity properties on all."""
    rs, re = synth.synth_intervals(3, 1_000_000)
    pos, svlen, typ, gt = synth.synth_candidates(4, 100_000)
    S, E = pos.copy(), (pos + np.minimum(svlen, 1000)).astype(np.int32)
    d = ctx.span_depth(rs, re, S, E)
    idx = np.random.default_rng(0).integers(0, len(S), 300)
    for i in idx:
        assert d[i] == O.span_count(rs, re, int(S[i]), int(E[i]))
    # widening the window can only lower the count
    d_wide = ctx.span_depth(rs, re, S - 500, E + 500)
    assert np.all(d_wide <= d)
    # dense windows (tiled path) == sparse evaluation of the same windows one by one
    Sd = np.arange(5_000_000, 5_000_000 + 4096, dtype=np.int32)
    dd = ctx.span_depth(rs, re, Sd, Sd + 50)
    for i in range(0, 4096, 257):
        assert dd[i] == O.span_count(rs, re, int(Sd[i]), int(Sd[i]) + 50)


def _sa_pack(cases):
    seg_off, cols, L = [0], [[] for _ in range(6)], []
    for segs, ln in cases:
        for s in segs:
            for k in range(6):
                cols[k].append(s[k])
        seg_off.append(seg_off[-1] + len(segs))
        L.append(ln)
    return seg_off, cols, L


def _sa_check(ctx, cases, max_parts=7):
    seg_off, cols, L = _sa_pack(cases)
    got = ctx.sa_analysis(seg_off, *cols, L, max_parts=max_parts)
    exp = []
    for r, (segs, ln) in enumerate(cases):
        for (tp, p, n, q0, q1, rev) in O.sa_split_analysis([tuple(s) for s in segs], ln, max_parts=max_parts):
            exp.append((r, tp, p, n, q0, q1, rev))
    g = [tuple(int(x[f]) for f in ("rec", "type", "pos", "len", "q0", "q1", "first")) for x in got]
    assert g == exp
    return len(exp)


def test_sa_analysis_golden(ctx, golden):
    """analysis_split_read (SE:193-326): reference-generated rows incl. the INS sequence slice."""
    g = golden("sa")
    cases = [([(s[0], s[1], s[2], s[3], 1 if s[4] == "chr1" else 2, 0 if s[5] == "+" else 1) for s in c["segs"]], c["L"])
             for c in g]
    seg_off, cols, L = _sa_pack(cases)
    got = ctx.sa_analysis(seg_off, *cols, L)
    rows = [[] for _ in g]
    for x in got:
        c = g[int(x["rec"])]
        if int(x["type"]) == V.INS:
            q = c["query"][::-1] if int(x["first"]) else c["query"]
            rows[int(x["rec"])].append(["INS", int(x["pos"]), int(x["len"]), q[int(x["q0"]):int(x["q1"])]])
        else:
            rows[int(x["rec"])].append(["DEL", int(x["pos"]), int(x["len"])])
    assert rows == [c["rows"] for c in g]
    assert sum(len(r) for r in rows) > 50


def test_sa_analysis_random_vs_oracle(ctx):
    rng = np.random.default_rng(99)
    cases = []
    for t in range(5000):
        k = int(rng.choice([0, 1, 2, 2, 2, 3, 3, 4, 5, 7, 8, 9]))
        ln = int(rng.integers(600, 30000))
        r = int(rng.integers(10000, 10**8))
        q = 0
        st0 = int(rng.integers(0, 2))
        segs = []
        for i in range(k):
            qe = q + int(rng.integers(50, max(60, ln // max(k, 1))))
            chrom = 1 if rng.random() < 0.85 else 2
            st = st0 if rng.random() < 0.9 else 1 - st0
            segs.append((q, qe, r, r + (qe - q) + int(rng.integers(-20, 20)), chrom, st))
            q = qe + int(rng.choice([0, 0, rng.integers(-60, 60), rng.integers(30, 2000)]))
            r = segs[-1][3] + int(rng.choice([0, rng.integers(-29, 101), rng.integers(30, 120000), rng.integers(-500, -30)]))
        if st0:
            segs = [(ln - s[1], ln - s[0]) + s[2:] for s in segs]
        order = rng.permutation(len(segs))
        cases.append(([segs[i] for i in order], ln))
    assert _sa_check(ctx, cases) > 500
    assert _sa_check(ctx, cases[:500], max_parts=16) > 50
    # empty input and the argument check
    assert len(ctx.sa_analysis([0], [], [], [], [], [], [], [])) == 0
    with pytest.raises(V.VsvError):
        ctx.sa_analysis([0, 2], [0, 5], [5, 9], [1, 2], [3, 4], [1, 1], [0, 0], [100], max_parts=40)


def test_edit_distance_batch(ctx):
    """vsv_edit_distance_batch == edlib NW distance (RR:75-81): known answers, random pairs against
    the oracle DP, block-edge lengths, long strings, empty strings."""
    import random
    rng = random.Random(7)
    strs = ["kitten", "sitting", "", "ACGT", "ACGTACGTNN", "A" * 64, "A" * 63 + "C", "C" * 65]
    pairs = [(0, 1), (2, 3), (3, 2), (2, 2), (3, 4), (5, 6), (5, 7), (6, 7), (4, 4)]
    for m in (1, 31, 63, 64, 65, 127, 128, 129, 500, 2100, 5000):
        a = "".join(rng.choice("ACGT") for _ in range(m))
        b = list(a)
        for _ in range(rng.randint(0, max(1, m // 10))):
            k = rng.randrange(len(b)) if b else 0
            u = rng.random()
            if u < 0.3 and len(b) > 1:
                b.pop(k)
            elif u < 0.6:
                b.insert(k, rng.choice("ACGTN"))
            elif b:
                b[k] = rng.choice("ACGT")
        c = "".join(rng.choice("ACGTN") for _ in range(rng.choice([m, m // 2 + 1, 2 * m + 3])))
        base = len(strs)
        strs += [a, "".join(b), c]
        pairs += [(base, base + 1), (base + 1, base), (base, base + 2), (base + 2, base + 1)]
    got = ctx.edit_distance_batch(strs, [p[0] for p in pairs], [p[1] for p in pairs])
    assert int(got[0]) == 3 and int(got[1]) == 4 and int(got[3]) == 0
    for (i, j), g in zip(pairs, got):
        assert int(g) == O.edit_distance(strs[i], strs[j]), (len(strs[i]), len(strs[j]))
    from oracle import edlib_ref as E
    if E.available():        # the reference's own vendored edlib, built into oracle/_ref by oracle/Makefile (travels with the snapshot)
        for (i, j), g in zip(pairs, got):
            assert int(g) == E.distance(strs[i], strs[j])
    # one symbol table per call: up to 32 distinct byte values anywhere in the table
    assert int(ctx.edit_distance_batch(["ABCDEFGHIJKLMNOPQRS", "ABCDEFGHIJKLMNOPQRSTUV"], [0], [1])[0]) == 3
    assert int(ctx.edit_distance_batch(["ABCDEFGHIJKLMNOPQRS", "ABC"], [0], [1])[0]) == 16
    with pytest.raises(V.VsvError):
        ctx.edit_distance_batch(["".join(chr(48 + i) for i in range(33)), "0"], [0], [1])
    with pytest.raises(V.VsvError):
        ctx.edit_distance_batch(["A", "C"], [0], [2])
    assert len(ctx.edit_distance_batch(["A"], [], [])) == 0


def test_edit_distance_banded_long(ctx):
    """The banded sweep of k_edit_banded: long pairs (one CTA per pair), pairs that finish in the first narrow round, pairs
    that need the band widened to the full matrix, very unequal lengths, a pair whose state does not fit shared memory
    (global scratch through the same pointers) -- all against the oracle DP and the reference's vendored edlib."""
    import random
    from oracle import edlib_ref as E
    rng = random.Random(19)

    def rand(n, alpha="ACGT"):
        return "".join(rng.choice(alpha) for _ in range(n))

    def mutate(a, frac):
        b = list(a)
        for _ in range(int(len(a) * frac)):
            k = rng.randrange(len(b))
            u = rng.random()
            if u < 0.33 and len(b) > 1:
                b.pop(k)
            elif u < 0.66:
                b.insert(k, rng.choice("ACGT"))
            else:
                b[k] = rng.choice("ACGT")
        return "".join(b)
    strs, pairs = [], []

    def add(a, b):
        strs.extend([a, b])
        pairs.append((len(strs) - 2, len(strs) - 1))
    a = rand(8193)
    add(a, a)                                        # first long size, identical
    add(a, mutate(a, 0.01))
    add(a, mutate(a, 0.2))                           # second round
    add(a, rand(8193))                               # unrelated: full matrix
    add(rand(9000), rand(17999))
    a = rand(20011)
    add(a, mutate(a, 0.004))
    add(a[:12000], a)                                # delta = 8011: the stripe is shifted
    add(rand(64), rand(30000))                       # one block against a long text
    add(rand(8192), rand(8192))                      # largest short pair
    a = rand(8200)
    add(a[:4100] + rand(300) + a[4100:], a)          # one long insertion
    got = ctx.edit_distance_batch(strs, [p[0] for p in pairs], [p[1] for p in pairs])
    wide = "ABCDEFGHIJKLMNOPQRSTUVWXYZ012345"
    a = rand(60000, wide)                            # 32 symbols x 938 blocks: 37 words per block > 200 KB -> global scratch
    w_strs = [a, mutate(a, 0.002), rand(500, wide), rand(700, wide)]
    w_got = ctx.edit_distance_batch(w_strs, [0, 2, 3], [1, 3, 0])
    for (i, j), g in zip([(0, 1), (2, 3), (3, 0)], w_got):
        exp = E.distance(w_strs[i], w_strs[j]) if E.available() else O.edit_distance(w_strs[i], w_strs[j])
        assert int(g) == exp, (len(w_strs[i]), len(w_strs[j]))
    for (i, j), g in zip(pairs, got):
        exp = E.distance(strs[i], strs[j]) if E.available() else O.edit_distance(strs[i], strs[j])
        assert int(g) == exp, (len(strs[i]), len(strs[j]))
    for i, j in pairs[:5]:                           # and the two checkers against each other where the DP is quick
        if E.available():
            assert E.distance(strs[i], strs[j]) == O.edit_distance(strs[i], strs[j])
    # the order of a pair does not matter
    rev = ctx.edit_distance_batch(strs, [p[1] for p in pairs], [p[0] for p in pairs])
    assert rev.tolist() == got.tolist()


def test_clr_contig_gate_golden(ctx):
    """a2: the CLR gate (ECS_C:53-70,425-427) from the device class sums against the reference's own ins_pct / var_dist
    (tests/golden/clr_gate.json, generated by executing the reference functions), incl. the exact ties 13/100 and 200.0 and
    contigs the gate REJECTS."""
    import json
    from volcanosv_b200 import operators as OP
    cases = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "clr_gate.json")))
    ops, off = [], [0]
    for c in cases:
        ops += [(ln << 4) | op for op, ln in c["cigar"]]
        off.append(len(ops))
    ops, off = np.asarray(ops, np.uint32), np.asarray(off, np.int64)
    i_sum, m_sum, m_cnt = ctx.cigar_class_sums(ops, off)
    for k, c in enumerate(cases):
        assert i_sum[k] / (m_sum[k] + i_sum[k]) == c["ins_pct"] and m_sum[k] / m_cnt[k] == c["var_dist"], k
    gate = OP.clr_gate(ctx, ops, off)
    assert gate.tolist() == [c["gate"] for c in cases]
    assert 10 <= sum(1 for c in cases if not c["gate"]) <= len(cases) - 10
