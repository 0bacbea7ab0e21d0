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


# --------------------------------------------------------------------------- resident step, round 2
def split_reads(pos, ref_end, readlen, first_op, last_op, is_rev, nxt):
    """extract_sig_from_split for READ records (ERS:147-197, min_mapq 0) on the consecutive same-name pairs
    (r, nxt[r]) (ERS:227-237), vectorised.  Returns SIG_DTYPE rows: rec = first record, first = second."""
    pos, ref_end, readlen = (np.asarray(x, np.int64) for x in (pos, ref_end, readlen))
    first_op, last_op = np.asarray(first_op, np.int64), np.asarray(last_op, np.int64)
    r = np.flatnonzero(np.asarray(nxt) >= 0)
    b = np.asarray(nxt)[r].astype(np.int64)
    clip = lambda w: ((w & 15) == 4) | ((w & 15) == 5)
    ok = (np.asarray(is_rev)[r] == np.asarray(is_rev)[b]) & clip(last_op[r]) & clip(first_op[b])      # ERS:160-166
    Ref1e, Ref2s = ref_end[r], pos[b]
    Read1e, Read2s = readlen[r] - (last_op[r] >> 4), first_op[b] >> 4
    D = (Ref2s - Ref1e) - (Read2s - Read1e)
    O = Ref1e - Ref2s
    ok &= np.abs(D) <= 50000
    is_del = ok & (O < 30) & (D >= 30)                                                                   # ERS:191-193
    is_ins = ok & (O < 30) & (D <= -30)                                                                  # ERS:194-196
    out = np.zeros(int(is_del.sum() + is_ins.sum()), SIG_DTYPE)
    k = np.flatnonzero(is_del | is_ins)
    out["rec"], out["first"], out["count"] = r[k], b[k], 1
    out["type"] = np.where(is_del[k], 0, 1)
    mid = np.trunc((Ref1e[k] + Ref2s[k]) / 2.0).astype(np.int64)                                         # int((a+b)/2)
    out["pos"] = np.where(is_del[k], Ref1e[k], mid)
    out["len"] = np.abs(D[k])
    out["q0"], out["q1"] = Read1e[k], Read2s[k]
    return out


def sa_rows(batch, ref_end, in_bed, min_mapq=20, min_read_len=500, sv_size=30, max_size=100000, max_parts=7):
    """SA-tag analysis rows (SE:345-371,193-326) from the pre-parsed SA columns of a batch (tests: small)."""
    from . import vsv_oracle as O
    dels, inss = [], []
    if "sa_rec" not in batch:
        return dels, inss
    ops, off = batch["ops"], batch["op_off"]
    for i, r in enumerate(batch["sa_rec"]):
        r = int(r)
        fl, qlen = int(batch["flag"][r]), int(batch["qlen"][r])
        if fl not in (0, 16) or qlen < min_read_len or not in_bed[r]:
            continue
        primary = None
        if int(batch["mapq"][r]) >= min_mapq:
            a, b = int(off[r]), int(off[r + 1])
            f, l = (int(ops[a]), int(ops[b - 1])) if b > a else (0, 0)
            left = (f >> 4) if (f & 15) in (4, 5) else 0
            right = (l >> 4) if (l & 15) in (4, 5) else 0
            primary = (left, qlen - right, int(batch["pos"][r]), int(ref_end[r]), int(batch.get("sa_self_chrom", 0)), 0) if fl == 0 else \
                (right, qlen - left, int(batch["pos"][r]), int(ref_end[r]), int(batch.get("sa_self_chrom", 0)), 1)
        ent = [(int(batch["sa_chrom"][e]), int(batch["sa_pos"][e]), int(batch["sa_strand"][e]), int(batch["sa_first_s"][e]),
                int(batch["sa_last_s"][e]), int(batch["sa_bias"][e]), int(batch["sa_mapq"][e]))
               for e in range(int(batch["sa_off"][i]), int(batch["sa_off"][i + 1]))]
        seg = O.sa_segments(primary, ent, qlen, min_mapq)
        if len(seg) < 2:
            continue
        for (tp, p, ln, q0, q1, rev) in O.sa_split_analysis(seg, qlen, sv_size, max_size, max_parts):
            (dels if tp == 0 else inss).append((r, p, ln, q0, q1, rev))
    return dels, inss


def shard_step2(batch, cands, thr_del, thr_ins, threads=1, bed=None, chrom_len=None, batch_size=10_000_000, drop_ins=()):
    """CPU restatement of the resident step of csrc/vsv_shard.cu (vsv_shard_scan + vsv_shard_candidates) with the
    reference's product semantics: RVC:83-96 (ERS cigar + split -> FPF) and FGC:111-172 (SE cigar + merge + SA,
    sort -u, .sigs order -> CSS; GTD on the DEL calls and GTI on the INS calls, each with its own last_match_idx)."""
    ops, off, pos, mapq, qlen = batch["ops"], batch["op_off"], batch["pos"], batch["mapq"], batch["qlen"]
    n = len(pos)
    s_reads, rt, qt, xt = scan_cigar(ops, off, pos, mapq, 1, 30, 50, threads=threads)
    rend = (pos + rt).astype(np.int32)
    readlen = (qt + xt).astype(np.int32)
    in_bed = np.ones(n, bool)
    if bed is not None:                                               # SE:58-80 + SE:524-535
        in_bed[:] = False
        t0 = (pos.astype(np.int64) // batch_size) * batch_size if chrom_len >= batch_size else np.zeros(n, np.int64)
        t1 = np.minimum(t0 + batch_size, chrom_len) if chrom_len >= batch_size else np.full(n, chrom_len, np.int64)
        for (r0, r1) in bed:
            att = ((t0 <= r0) & (t1 > r0)) | ((r0 <= t0) & (t0 < r1))
            in_bed |= att & ~((rend <= r0) | (pos >= r1))
    res = dict(ref_end=rend, readlen=readlen, in_bed=in_bed.astype(np.uint8), sig_reads_cigar=s_reads)
    split = np.zeros(0, SIG_DTYPE)
    if batch.get("next_same_name") is not None:
        has = off[1:] > off[:-1]
        last = max(len(ops) - 1, 0)
        f_op = np.where(has, ops[np.minimum(off[:-1], last)], 0)
        l_op = np.where(has, ops[np.maximum(off[1:] - 1, 0)], 0)
        is_rev = (batch["flag"] >> 4) & 1 if batch.get("flag") is not None else np.zeros(n, np.int64)
        split = split_reads(pos, rend, readlen, f_op, l_op, is_rev, batch["next_same_name"])
    res["sig_reads_split"] = split
    reads_all = np.concatenate([s_reads, split])
    reads_sorted = reads_all[np.argsort(reads_all["pos"], kind="stable")]
    enable = (qlen >= 500).astype(np.uint8)
    s_cute, *_ = scan_cigar(ops, off, pos, mapq, 2, 10, 20, enable=enable, threads=threads, totals=False)
    res["sig_cute"] = s_cute
    s_cute_in = s_cute[in_bed[s_cute["rec"]]] if bed is not None else s_cute
    dels, inss = merge_cute(s_cute_in)
    sd, si = sa_rows(batch, rend, in_bed)
    def with_sa(rows, sa, tp):
        if not sa:
            return rows
        extra = np.zeros(len(sa), SIG_DTYPE)
        for k, (r, p, ln, q0, q1, rev) in enumerate(sa):
            extra[k] = (r, tp | (256 if rev else 0), p, ln, q0, q1, -1, 0)
        return np.concatenate([rows, extra])
    dels, inss = with_sa(dels, sd, 0), with_sa(inss, si, 1)
    res["n_sa_rows"] = len(sd) + len(si)
    res["rows_del"], res["rows_ins"] = dels, inss
    name = batch.get("name_id")

    def dedup(rows, is_del):
        """sort -u: `.sigs` order, then rows equal to an earlier one in (pos, len, read) -> DEL dropped, INS reported"""
        rows = sigs_order(rows)
        keep = np.ones(len(rows), bool)
        maybe = 0
        if name is not None and len(rows):
            key = np.stack([rows["pos"].astype(np.int64), rows["len"].astype(np.int64), np.asarray(name)[rows["rec"]].astype(np.int64)], 1)
            _, first_idx, inv = np.unique(key, axis=0, return_index=True, return_inverse=True)
            later = first_idx[inv.reshape(-1)] != np.arange(len(rows))
            # np.unique returns the FIRST occurrence in array order, which is the `.sigs` order here
            if is_del:
                keep = ~later
            else:
                maybe = int(later.sum())
        return rows, keep, maybe
    dels_s, keep_d, _ = dedup(dels, True)
    inss_s, keep_i, maybe_i = dedup(inss, False)
    if len(drop_ins):
        raise NotImplementedError("confirmed INS duplicates are a host decision; tests drop them through the engine only")
    res["n_del_dup"], res["n_ins_maybe_dup"] = int((~keep_d).sum()), maybe_i
    ud = group(dels_s["pos"][keep_d], dels_s["len"][keep_d], 0)
    # GTI:81 keeps INS signatures >= 30 bp; a dropped duplicate never heads a group
    ui = group(inss_s["pos"][keep_i], inss_s["len"][keep_i], 30)
    res["udel"], res["uins"] = ud, ui
    if cands is None:
        return res
    cpos, clen, ctyp, cgt = cands["pos"], cands["svlen"], cands["type"], cands["gt"]
    res["fp_support"] = fp_support(cpos, clen, reads_sorted["pos"], reads_sorted["len"])
    cov_i = cov(False, cpos, cpos, inss_s["pos"][keep_i], inss_s["len"][keep_i])
    cov_d = cov(True, cpos, (cpos + clen).astype(np.int32), dels_s["pos"][keep_d], dels_s["len"][keep_d])
    res["cov"] = np.where(ctyp == 0, cov_d, cov_i)
    nsup = np.zeros(len(cpos), np.int32)
    for tp, u in ((0, ud), (1, ui)):
        m = ctyp == tp
        nsup[m] = sig_support(cpos[m], clen[m], *u)
    res["n_support"] = nsup
    d2, ratio, ng = genotype(pos, rend, cpos, clen, ctyp, cgt, nsup, thr_del, thr_ins, threads=threads)
    res.update(depth2=d2, ratio=ratio, new_gt=ng)
    return res
