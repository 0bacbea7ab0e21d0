"""ctypes wrapper of oracle/libvsv_oracle_c.so (TEST INFRASTRUCTURE; see vsv_oracle_c.c)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "libvsv_oracle_c.so")
_lib = None

SIG_DTYPE = np.dtype([("rec", "<i4"), ("type", "<i4"), ("pos", "<i4"), ("len", "<i4"),
                      ("q0", "<i4"), ("q1", "<i4"), ("first", "<i4"), ("count", "<i4")])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            subprocess.check_call(["make", "-s", "-C", _HERE])
        _lib = C.CDLL(_PATH)
        _lib.orc_scan_cigar.restype = C.c_int64
        _lib.orc_merge_cute.restype = C.c_int64
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _a(x, dt):
    return None if x is None else np.ascontiguousarray(x, dtype=dt)


def num_threads():
    return int(lib().orc_num_threads())


def scan_cigar(ops, op_off, pos, mapq, mode, min_len, min_mapq, enable=None, threads=1, totals=True, cap=None):
    ops, op_off, pos, mapq, enable = _a(ops, np.uint32), _a(op_off, np.int64), _a(pos, np.int32), _a(mapq, np.uint8), _a(enable, np.uint8)
    n = len(pos)
    cap = cap or max(1024, len(ops) // 8)
    out = np.empty(cap, SIG_DTYPE)
    rt = np.zeros(n, np.int32) if totals else None
    qt = np.zeros(n, np.int32) if totals else None
    xt = np.zeros(n, np.int32) if totals else None
    k = lib().orc_scan_cigar(C.c_int(mode), C.c_int(min_len), C.c_int(min_mapq), _p(ops), _p(op_off), _p(pos), _p(mapq),
                             _p(enable), C.c_int64(n), _p(out), C.c_int64(cap), _p(rt), _p(qt), _p(xt), C.c_int(threads))
    if k > cap:
        return scan_cigar(ops, op_off, pos, mapq, mode, min_len, min_mapq, enable, threads, totals, cap=int(k))
    return out[:k], rt, qt, xt


def merge_cute(sigs):
    outs = []
    for tp in (0, 1):
        s = np.ascontiguousarray(sigs[sigs["type"] == tp])
        o = np.empty(len(s), SIG_DTYPE)
        k = lib().orc_merge_cute(_p(s), C.c_int64(len(s)), C.c_int(tp), _p(o)) if len(s) else 0
        outs.append(o[:k])
    return outs


def span_depth(rs, re, S, E, threads=1):
    rs, re, S, E = (_a(x, np.int32) for x in (rs, re, S, E))
    d = np.empty(len(S), np.int32)
    lib().orc_span_depth(_p(rs), _p(re), C.c_int64(len(rs)), _p(S), _p(E), C.c_int64(len(S)), _p(d), C.c_int(threads))
    return d


def genotype(rs, re, cpos, clen, ctype, cgt, nsup, thr_del, thr_ins, threads=1):
    rs, re, cpos, clen, ctype, cgt, nsup = (_a(x, np.int32) for x in (rs, re, cpos, clen, ctype, cgt, nsup))
    td, ti = _a(thr_del, np.float64), _a(thr_ins, np.float64)
    n = len(cpos)
    d2, ratio, ng = np.empty(n, np.int32), np.empty(n, np.float64), np.empty(n, np.int32)
    lib().orc_genotype(_p(rs), _p(re), C.c_int64(len(rs)), _p(cpos), _p(clen), _p(ctype), _p(cgt), _p(nsup), C.c_int64(n),
                       _p(td), _p(ti), _p(d2), _p(ratio), _p(ng), C.c_int(threads))
    return d2, ratio, ng


def sig_support(cpos, clen, sp, sl, sn):
    cpos, clen, sp, sl, sn = (_a(x, np.int32) for x in (cpos, clen, sp, sl, sn))
    out = np.zeros(len(cpos), np.int32)
    if len(sp):
        lib().orc_sig_support(_p(cpos), _p(clen), C.c_int64(len(cpos)), _p(sp), _p(sl), _p(sn), C.c_int64(len(sp)), _p(out))
    return out


def fp_support(cpos, clen, sp, sl):
    cpos, clen, sp, sl = (_a(x, np.int32) for x in (cpos, clen, sp, sl))
    out = np.zeros(len(cpos), np.int32)
    lib().orc_fp_support(_p(cpos), _p(clen), C.c_int64(len(cpos)), _p(sp), _p(sl), C.c_int64(len(sp)), _p(out))
    return out


def cov(is_del, cs, ce, sp, sl):
    cs, ce, sp, sl = (_a(x, np.int32) for x in (cs, ce, sp, sl))
    out = np.zeros(len(cs), np.int64)
    lib().orc_cov(C.c_int(1 if is_del else 0), _p(cs), _p(ce), C.c_int64(len(cs)), _p(sp), _p(sl), C.c_int64(len(sp)), _p(out))
    return out


def sigs_order(sigs):
    """`.sigs` order (SE:637-638, LC_ALL=C): pos numeric, then len as a decimal string; stable."""
    if len(sigs) == 0:
        return sigs
    ln = np.maximum(sigs["len"].astype(np.int64), 0)
    digits = np.ones(len(ln), np.int64)
    t = ln.copy()
    for _ in range(10):
        t //= 10
        digits += (t > 0)
    scaled = ln * (10 ** (10 - digits))          # left-justified to 10 digits == string order
    o = np.lexsort((digits, scaled, sigs["pos"]))  # lexsort is stable
    return sigs[o]


def group(pos, length, min_len=0):
    keep = length >= min_len
    pos, length = pos[keep], length[keep]
    if len(pos) == 0:
        z = np.zeros(0, np.int32)
        return z, z, z
    head = np.ones(len(pos), bool)
    head[1:] = (pos[1:] != pos[:-1]) | (length[1:] != length[:-1])
    idx = np.flatnonzero(head)
    cnt = np.diff(np.append(idx, len(pos))).astype(np.int32)
    return pos[idx], length[idx], cnt


def shard_step(batch, cands, thr_del, thr_ins, threads=1):
    """CPU restatement of volcanosv_b200.pipeline.ChromShard.run() (same call chain:
    RVC:83-96 ERS -> FPF; FGC:111-172 SE -> CSS -> GTD/GTI) for parity and baseline."""
    ops, off, pos, mapq, qlen = batch["ops"], batch["op_off"], batch["pos"], batch["mapq"], batch["qlen"]
    s_reads, rt, qt, xt = scan_cigar(ops, off, pos, mapq, 1, 30, 50, threads=threads)
    rend = (pos + rt).astype(np.int32)
    o = np.argsort(s_reads["pos"], kind="stable")
    s_reads = s_reads[o]
    enable = (qlen >= 500).astype(np.uint8)
    s_cute, *_ = scan_cigar(ops, off, pos, mapq, 2, 10, 20, enable=enable, threads=threads, totals=False)
    dels, inss = merge_cute(s_cute)
    dels, inss = sigs_order(dels), sigs_order(inss)
    cpos, clen, ctyp, cgt = cands["pos"], cands["svlen"], cands["type"], cands["gt"]
    fp = fp_support(cpos, clen, s_reads["pos"], s_reads["len"])
    cov_i = cov(False, cpos, cpos, inss["pos"], inss["len"])
    cov_d = cov(True, cpos, (cpos + clen).astype(np.int32), dels["pos"], dels["len"])
    ud = group(dels["pos"], dels["len"], 0)
    ui = group(inss["pos"], inss["len"], 30)
    sup_d = sig_support(cpos, clen, *ud)
    sup_i = sig_support(cpos, clen, *ui)
    nsup = np.where(ctyp == 0, sup_d, sup_i).astype(np.int32)
    d2, ratio, ng = genotype(pos, rend, cpos, clen, ctyp, cgt, nsup, thr_del, thr_ins, threads=threads)
    return dict(fp_support=fp, cov_ins=cov_i, cov_del=cov_d, n_support=nsup, depth2=d2, ratio=ratio, new_gt=ng,
                sig_reads=s_reads, sig_del=dels, sig_ins=inss)
