"""The C twin of the oracle against the golden-pinned Python oracle (CPU only)."""
import numpy as np

import oracle as O
from oracle import c_oracle as CO
from volcanosv_b200 import synth


def _t(a, f=("rec", "type", "pos", "len", "q0", "q1")):
    return [tuple(int(x[k]) for k in f) for x in a]


def test_c_scan_and_merge():
    b = synth.synth_records(7, 200, "hifi", eqx=True)
    for mode, ml, mq in ((0, 30, 50), (1, 30, 50), (2, 10, 20)):
        s, rt, qt, xt = CO.scan_cigar(b["ops"], b["op_off"], b["pos"], b["mapq"], mode, ml, mq, threads=3)
        es, ert, eqt, ext = O.scan_cigar(b["ops"], b["op_off"], b["pos"], b["mapq"], mode, ml, mq)
        assert _t(s) == es and np.array_equal(rt, ert) and np.array_equal(qt, eqt) and np.array_equal(xt, ext)
    d, i = CO.merge_cute(s)
    ed, ei = O.merge_in_record(es, O.MODE_CUTE)
    f8 = ("rec", "type", "pos", "len", "q0", "q1")
    assert _t(d, f8) == [x[:6] for x in ed] and _t(i, f8) == [x[:6] for x in ei]
    assert [int(x) for x in i["count"]] == [x[7] for x in ei]


def test_c_joins_and_genotype():
    rng = np.random.default_rng(2)
    rs, re = synth.synth_intervals(3, 5000, chrom_len=1_000_000)
    pos, svlen, typ, gt = synth.synth_candidates(4, 300, chrom_len=1_000_000)
    sp, sl, sn = synth.synth_sig_table(5, pos, svlen, per_cand=4)
    sup = CO.sig_support(pos, svlen, sp, sl, sn)
    esup, _ = O.sig_support([0] * len(pos), pos, svlen, [0] * len(sp), sp, sl, sn)
    assert sup.tolist() == esup and sum(esup) > 0
    assert CO.fp_support(pos, svlen, sp, sl).tolist() == O.fp_support(pos, svlen, sp, sl)
    assert [float(x) for x in CO.cov(False, pos, pos, sp, sl)] == O.ins_call_cov(pos, sp, sl)
    assert CO.cov(True, pos, pos + svlen, sp, sl).tolist() == O.del_call_cov(pos, pos + svlen, sp, sl)
    thr_d, thr_i = [0.6, 0.69, 0.71, 0.91], [float("nan"), 0.59, 0.63, 0.79]
    d2, ratio, ng = CO.genotype(rs, re, pos, svlen, typ, gt, sup, thr_d, thr_i, threads=2)
    for k in range(len(pos)):
        e2 = O.depth_del(rs, re, int(pos[k]), int(svlen[k])) if typ[k] == 0 else O.depth_ins(rs, re, int(pos[k]))
        er, eg = O.genotype(int(typ[k]), int(svlen[k]), int(gt[k]), int(sup[k]), e2, thr_d if typ[k] == 0 else thr_i)
        assert d2[k] == e2 and ng[k] == eg and (ratio[k] == er or (np.isnan(ratio[k]) and er != er))
    S = pos - 100
    assert CO.span_depth(rs, re, S, pos + 100, threads=2).tolist() == [O.span_count(rs, re, int(a), int(a) + 200) for a in S]


def test_sigs_order_is_c_locale_string_order():
    s = np.zeros(4, CO.SIG_DTYPE)
    s["pos"] = [5, 5, 5, 4]
    s["len"] = [95, 100, 10, 7]
    assert CO.sigs_order(s)["len"].tolist() == [7, 10, 100, 95]
