"""CPU oracle (TEST INFRASTRUCTURE, not product code) for the large-indel hot path.

Every function restates one reference function on the packed struct-of-arrays layout
used at the C-ABI boundary (include/volcanosv_b200.h) and cites the reference lines it
follows.  Paths: LI/ = /root/reference/bin/VolcanoSV-vc/Large_INDEL/ ;
ECS_H = LI/extract_contig_signature_Hifi.py, ECS_O/_C = ..._ONT.py/_CLR.py,
ERS = LI/extract_reads_signature.py, SE = LI/sig_extract.py,
FPF = LI/FP_filter_v1.py, CSS = LI/calculate_signature_support.py,
GTD/GTI = LI/correct_gt_{del,ins}_real_data.py.

Pure-Python loops: meant for inputs the tests finish in seconds.  The C twin
(oracle/vsv_oracle_c.c) covers the two volume stages (scan, span depth) for large
sizes and for bench.py's cpu_baseline leg.

Parity status: pinned by tests/golden/*.json, produced by running the reference's own
functions in the build container (tests/golden/make_golden.py).
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import numpy as np

# CIGAR op codes (BAM): 0 M, 1 I, 2 D, 3 N, 4 S, 5 H, 6 P, 7 =, 8 X
OP_M, OP_I, OP_D, OP_N, OP_S, OP_H, OP_P, OP_EQ, OP_X = range(9)

MODE_CONTIG = 0   # ECS_H:53-85
MODE_READS = 1    # ERS:47-83
MODE_CUTE = 2     # SE:438-478

DEL, INS = 0, 1


def _mask(codes) -> int:
    m = 0
    for c in codes:
        m |= 1 << c
    return m


# (ref_mask, qry_mask, aux_mask, hard-clip sign) per mode.
#  contig: only M,S,D,I are handled (ECS_H:72-85); aux = H so rl = qry+aux (ECS_H:308-313)
#  reads : M,=,X both; S,I query; D,N ref (ERS:66-81); aux = H (ERS:148-153)
#  cute  : shift_del == shift_ins == sum over {M,D,=,X}; shift_ins_read = every op but D
#          (SE:461-478); aux = N so reference_end = pos + ref + aux.
MODE_TABLE = {
    MODE_CONTIG: (_mask([0, 2]), _mask([0, 1, 4]), _mask([5]), +1),
    MODE_READS: (_mask([0, 7, 8, 2, 3]), _mask([0, 7, 8, 4, 1]), _mask([5]), +1),
    MODE_CUTE: (_mask([0, 2, 7, 8]), 0xFFFF & ~_mask([2]), _mask([3]), -1),
}


def pack_cigar(cigars: Sequence[Sequence[Tuple[int, int]]]):
    """list of [(op,len),...] -> (ops u32 = len<<4|op, op_off i64[n+1])."""
    op_off = np.zeros(len(cigars) + 1, dtype=np.int64)
    for i, c in enumerate(cigars):
        op_off[i + 1] = op_off[i] + len(c)
    ops = np.zeros(int(op_off[-1]), dtype=np.uint32)
    k = 0
    for c in cigars:
        for op, ln in c:
            ops[k] = (int(ln) << 4) | int(op)
            k += 1
    return ops, op_off


def scan_cigar(ops, op_off, pos, mapq, mode, min_len, min_mapq, enable=None):
    """Stage-1 CIGAR walk.  ECS_H:53-85 / ERS:47-83 / SE:438-478 (scan part only).

    Returns (sigs, ref_total, qry_total, aux_total); sigs = list of
    (rec, type, pos, len, q0, q1) in (record, op) order.  Records with
    mapq < min_mapq or enable[r]==0 emit nothing (ECS_H:392-394, ERS:120, SE:439,447)
    but still get their totals (the reference needs reference_end for every record).
    """
    ref_mask, qry_mask, aux_mask, hsign = MODE_TABLE[mode]
    n = len(op_off) - 1
    sigs = []
    ref_t = np.zeros(n, dtype=np.int64)
    qry_t = np.zeros(n, dtype=np.int64)
    aux_t = np.zeros(n, dtype=np.int64)
    for r in range(n):
        a, b = int(op_off[r]), int(op_off[r + 1])
        emit = int(mapq[r]) >= min_mapq and (enable is None or enable[r])
        hc = 0
        if b > a and (int(ops[a]) & 15) == OP_H:
            hc = int(ops[a]) >> 4
        rs = qs = xs = 0
        p0 = int(pos[r])
        for k in range(a, b):
            w = int(ops[k])
            op, ln = w & 15, w >> 4
            if emit and ln >= min_len and (op == OP_D or op == OP_I):
                q0 = qs + hsign * hc
                if op == OP_D:
                    sigs.append((r, DEL, p0 + rs, ln, q0, q0 + 1))
                else:
                    sigs.append((r, INS, p0 + rs, ln, q0, q0 + ln))
            if (ref_mask >> op) & 1:
                rs += ln
            if (qry_mask >> op) & 1:
                qs += ln
            if (aux_mask >> op) & 1:
                xs += ln
        ref_t[r], qry_t[r], aux_t[r] = rs, qs, xs
    return sigs, ref_t, qry_t, aux_t


def merge_in_record(sigs, mode):
    """Per-record sequential merge of the scan output.

    contig: ECS_H:87-164 (cluster_ins_one_read / cluster_del_one_read);
    cute  : SE:373-435 (generate_combine_sigs, INS merge_dis=100, DEL merge_dis=0,
            including the end:=start quirk after the first flush, SE:429-430);
    reads : no per-read merge (ERS:47-83) -> identity.

    Returns (dels, inss); each a list of (rec, type, pos, len, q0, q1, first, count)
    where first/count index the consecutive pre-merge sigs OF THAT TYPE that were
    folded into the output (the host concatenates INS sequence pieces from them).
    """
    out = {DEL: [], INS: []}
    per_type = {DEL: [s for s in sigs if s[1] == DEL], INS: [s for s in sigs if s[1] == INS]}
    for tp in (DEL, INS):
        lst = per_type[tp]
        i = 0
        while i < len(lst):
            j = i
            while j < len(lst) and lst[j][0] == lst[i][0]:
                j += 1
            out[tp].extend(_merge_one_record(lst, i, j, tp, mode))
            i = j
    return out[DEL], out[INS]


def _merge_one_record(lst, i0, i1, tp, mode):
    res = []
    if mode == MODE_READS:
        for k in range(i0, i1):
            res.append(lst[k] + (k, 1))
        return res
    if mode == MODE_CONTIG:
        cur = list(lst[i0]) + [i0, 1]
        for k in range(i0 + 1, i1):
            s = lst[k]
            d = abs(s[2] - cur[2])
            if tp == INS:
                hit = ((cur[3] > 250 and s[3] > 250 and d < 250)
                       or (cur[3] > 320 and s[3] > 320 and d < 380)
                       or (cur[3] > 100 and s[3] > 100 and d < 250))      # ECS_H:112-135
                if hit:
                    cur[5] = s[5]                                         # read_end  := sig2 end
                    cur[3] = cur[5] - cur[4]                              # ECS_H:91-97
                    cur[7] += 1
                    continue
            else:
                if cur[3] > 150 and s[3] > 150 and d < 150:               # ECS_H:148-150
                    cur[3] = s[2] + s[3] - cur[2]                         # ECS_H:99-106
                    cur[5] = cur[4] + 1
                    cur[7] += 1
                    continue
            res.append(tuple(cur))
            cur = list(s) + [k, 1]
        res.append(tuple(cur))
        return res
    # MODE_CUTE
    first = lst[i0]
    if tp == INS:
        cur = list(first) + [i0, 1]
        last = first[2]
        for k in range(i0 + 1, i1):
            s = lst[k]
            if s[2] - last <= 100:                                        # SE:396-399
                cur[3] += s[3]
                cur[7] += 1
                last = s[2]
            else:
                res.append(tuple(cur))
                cur = list(s) + [k, 1]
                last = s[2]
        res.append(tuple(cur))
    else:
        cur = list(first) + [i0, 1]
        end = first[2] + first[3]                                         # SE:417
        for k in range(i0 + 1, i1):
            s = lst[k]
            if s[2] - end <= 0:                                           # SE:420
                cur[3] += s[3]
                cur[7] += 1
                end = s[2] + s[3]
            else:
                res.append(tuple(cur))
                cur = list(s) + [k, 1]
                end = s[2]                                                # SE:429-430 quirk
        res.append(tuple(cur))
    return res


# --------------------------------------------------------------------------- split pairs
SPLIT_HIFI, SPLIT_ONT, SPLIT_CLR, SPLIT_READS = 0, 1, 2, 3


def _pytrunc_half(x: int) -> int:
    """int(x/2) of Python: float division then truncation toward zero."""
    return int(x / 2)


def split_pair(variant, pos1, end1, rl1, lastop1, mapq1, rev1,
               pos2, firstop2, mapq2, rev2, min_mapq, max_svlen=50000):
    """One consecutive same-name record pair -> at most one signature.

    variant: SPLIT_HIFI ECS_H:307-371 | SPLIT_ONT ECS_O:307-382 (r=.5, lower .8)
           | SPLIT_CLR ECS_C:328-402 (r=.3) | SPLIT_READS ERS:147-197.
    lastop1 / firstop2 are packed ops (len<<4|op).  Returns None or
    (type, pos, len, q0, q1).
    """
    if rev1 != rev2 or mapq1 < min_mapq or mapq2 < min_mapq:
        return None
    if (lastop1 & 15) not in (OP_S, OP_H) or (firstop2 & 15) not in (OP_S, OP_H):
        return None
    Ref1e, Ref2s = end1, pos2
    Read1e = rl1 - (lastop1 >> 4)
    Read2s = firstop2 >> 4
    D = (Ref2s - Ref1e) - (Read2s - Read1e)
    if abs(D) > max_svlen:
        return None
    if variant == SPLIT_READS:
        O = Ref1e - Ref2s
        if O < 30 and D >= 30:
            return (DEL, Ref1e, D, Read1e, Read2s)
        if O < 30 and D <= -30:
            return (INS, _pytrunc_half(Ref1e + Ref2s), abs(D), Read1e, Read2s)
        return None
    if variant == SPLIT_HIFI:
        if D >= 30:
            O = Read1e - Read2s
            if abs(O) <= 3000:
                h = _pytrunc_half(O)
                return (DEL, Ref1e - h, D, Read1e - h, Read1e - h + 1)
            return None
        if D <= -30:
            O = Ref1e - Ref2s
            if O < 3000:
                svlen = abs(Read2s - Read1e + O)
                p = _pytrunc_half(Ref1e + Ref2s) if abs(O) > 400 else Ref2s
                return (INS, p, svlen, Read1e - O, Read2s)
        return None
    r = 0.5 if variant == SPLIT_ONT else 0.3
    lo = 0.8 if variant == SPLIT_ONT else r
    if D >= 30:
        O = Read1e - Read2s
        if -(D * r) <= O <= (D * r):
            return (DEL, Ref1e, D, Read1e, Read2s)
        return None
    O = Ref1e - Ref2s
    if (D * lo <= O <= abs(D) * r) and D <= -30:
        svlen = abs(Read2s - Read1e + O)
        p = _pytrunc_half(Ref1e + Ref2s) if abs(O) > 400 else Ref2s
        return (INS, p, svlen, Read1e - O, Read2s)
    return None


def split_pairs(variant, pos, ref_end, rl, first_op, last_op, mapq, is_rev, name_id,
                select, min_mapq):
    """Driver: ECS_H:421-457 / ERS:199-237.

    `select[r]` is the reference's membership filter (contigs: hp-in-name and
    mapq>=min; reads: everything).  Names with >=2 selected records, in first
    appearance order; consecutive selected records of a name in input (BAM) order.
    Returns list of (rec1, rec2, type, pos, len, q0, q1) in that order.
    """
    order: Dict[int, List[int]] = {}
    for r in range(len(pos)):
        if select[r]:
            order.setdefault(int(name_id[r]), []).append(r)
    out = []
    for nm, idxs in order.items():
        if len(idxs) < 2:
            continue
        for a, b in zip(idxs[:-1], idxs[1:]):
            s = split_pair(variant, int(pos[a]), int(ref_end[a]), int(rl[a]), int(last_op[a]),
                           int(mapq[a]), bool(is_rev[a]), int(pos[b]), int(first_op[b]),
                           int(mapq[b]), bool(is_rev[b]), min_mapq)
            if s is not None:
                out.append((a, b) + s)
    return out


# --------------------------------------------------------------------------- clustering
def stable_order_by_pos(positions):
    """sort_sig (ECS_H:170-179) with the tie order PINNED to stable."""
    return np.argsort(np.asarray(positions, dtype=np.int64), kind="stable")


def cluster_leader(pos, length, is_del, max_shift=100):
    """cluster_del / cluster_ins (ECS_H:196-288) on a pos-sorted list.

    Greedy leader clustering: first unassigned i leads; every still-unassigned j with
    |pos_i-pos_j|<=max_shift, 2*min(len)>=max(len) and, for DEL, 2*overlap>=minlen
    joins i.  Per cluster keep the first member with the strictly greatest len.
    Returns kept indices in leader order.
    """
    n = len(pos)
    cl = [-1] * n
    keep = []
    for i in range(n):
        if cl[i] != -1:
            continue
        cl[i] = i
        best = i
        for j in range(n):
            if cl[j] != -1:
                continue
            if abs(int(pos[i]) - int(pos[j])) > max_shift:
                continue
            l1, l2 = int(length[i]), int(length[j])
            if 2 * min(l1, l2) < max(l1, l2):
                continue
            if is_del:
                ovl = min(int(pos[i]) + l1, int(pos[j]) + l2) - max(int(pos[i]), int(pos[j]))
                if 2 * ovl < min(l1, l2):
                    continue
            cl[j] = i
            if l2 > int(length[best]):
                best = j
        keep.append(best)
    return keep


def pair_haps(pos1, len1, typ1, pos2, len2, typ2, max_compare=1000, max_shift=200):
    """pair_sig (ECS_H:548-569): greedy hp1 -> hp2 matching on pos-sorted lists.

    Returns (partner1, partner2): index of the partner or -1.
    """
    n1, n2 = len(pos1), len(pos2)
    p1 = [-1] * n1
    p2 = [-1] * n2
    for i in range(n1):
        for j in range(n2):
            if int(pos2[j]) - int(pos1[i]) > max_compare:
                break
            if typ1[i] != typ2[j] or p2[j] != -1:
                continue
            if abs(int(pos1[i]) - int(pos2[j])) > max_shift:
                continue
            l1, l2 = int(len1[i]), int(len2[j])
            if 2 * min(l1, l2) < max(l1, l2):
                continue
            if typ1[i] == DEL:
                ovl = min(int(pos1[i]) + l1, int(pos2[j]) + l2) - max(int(pos1[i]), int(pos2[j]))
                if 2 * ovl < min(l1, l2):
                    continue
            p1[i] = j
            p2[j] = i
            break
    return p1, p2


# --------------------------------------------------------------------------- window joins
def fp_support(call_pos, call_len, sig_pos, sig_len, max_dist=1000, max_comp_svlen=250,
               max_shift=500):
    """eval_sig + compare_sigs (FPF:87-123).  sig list in file (pos-sorted) order.

    Calls longer than max_comp_svlen get support 60 without looking.  min_size_sim is
    the driver's 0.5 (FPF:11) -> 2*min >= max; 0/0 counts as similarity 0 (FPF:94-97).
    The scan `continue`s while shift < -max_dist and `break`s at the first
    shift > max_dist, so an unsorted tail is never seen -- restated literally.
    """
    out = []
    for cp, cl in zip(call_pos, call_len):
        cp, cl = int(cp), int(cl)
        if cl > max_comp_svlen:
            out.append(60)
            continue
        sup = 0
        for sp, sl in zip(sig_pos, sig_len):
            shift = int(sp) - cp
            if shift < -max_dist:
                continue
            if shift > max_dist:
                break
            mx = max(cl, int(sl))
            sim_ok = mx != 0 and 2 * min(cl, int(sl)) >= mx
            if abs(shift) <= max_shift and sim_ok:
                sup += 1
        out.append(sup)
    return out


def ins_call_cov(call_pos, sig_pos, sig_len, flanking=1000):
    """calc_ins_call_cov (CSS:81-125): per call position, sum of INS signature lengths
    with |sig_pos - call_pos| <= flanking.  Returned per call (dict lookup by pos)."""
    out = []
    for cp in call_pos:
        cp = int(cp)
        out.append(float(sum(int(l) for p, l in zip(sig_pos, sig_len)
                             if cp - flanking <= int(p) <= cp + flanking)))
    return out


def del_call_cov(call_start, call_end, sig_start, sig_len, flanking=1000):
    """calc_del_call_cov (CSS:138-253) for calls that arrive sorted by start
    (stable ties), where the sorted/unsorted index mixing (CSS:157-160 vs 217-219)
    is the identity.  The four sweeps collect, per call window
    [start-flanking, end+flanking], the SET of sigs whose start or end lies in the
    window or which contain the window's start or end == closed-interval overlap.
    cov = sum of (-len) over that set; calls with an empty set are absent from the
    reference's dict -> 0 (CSS:331-334).  Sig interval = [pos, pos+len] (CSS:71-76).
    Duplicate (start,end) keys overwrite (CSS:253) but carry identical values here.
    """
    out = []
    for cs, ce in zip(call_start, call_end):
        lb, rb = int(cs) - flanking, int(ce) + flanking
        tot = 0
        for s, l in zip(sig_start, sig_len):
            s, l = int(s), int(l)
            e = s + l
            if s <= rb and e >= lb:
                tot += -l
        out.append(tot)
    return out


def group_sigs(chrom_id, pos, length):
    """load_sig (GTD:65-86) / load_sig_file (GTI:67-101): unique (chrom,pos,len) in
    first-occurrence order with read counts."""
    dc: Dict[Tuple[int, int, int], int] = {}
    for c, p, l in zip(chrom_id, pos, length):
        k = (int(c), int(p), int(l))
        dc[k] = dc.get(k, 0) + 1
    keys = list(dc.keys())
    return ([k[0] for k in keys], [k[1] for k in keys], [k[2] for k in keys],
            [dc[k] for k in keys])


def sig_support(cand_chrom, cand_pos, cand_svlen, sig_chrom, sig_pos, sig_len, sig_cnt,
                min_size_sim=0.6, max_shift_ratio=2.3, last_in=0, return_last=False):
    """match_varlist_siglist (GTD:92-137) / extract_sig_support (GTI:107-166), literally:
    forward scan and backward scan both start AT last_match_idx (double count), breaks
    are taken only on same-chromosome rows, last_match_idx <- min matched index.
    Returns (n_support list, match_id list = last_match_idx seen by each candidate)
    [+ the final last_match_idx with return_last].  last_in = -1 is the shard-local "the
    carried index lies in another chromosome's block" (volcanosv_b200/shard.py): the whole
    window is visited once and nothing can be double counted.
    """
    last = last_in
    sup, mids = [], []
    n = len(sig_pos)
    for cc, cp, sv in zip(cand_chrom, cand_pos, cand_svlen):
        cc, cp, sv = int(cc), int(cp), int(sv)
        max_size = sv / min_size_sim
        min_size = sv * min_size_sim
        shift = max(sv * max_shift_ratio, 500)
        lo, hi = cp - shift, cp + shift
        matched = []
        s = 0
        mids.append(last)
        for i in range(max(last, 0), n):
            if int(sig_chrom[i]) == cc:
                if lo <= int(sig_pos[i]) <= hi:
                    matched.append(i)
                    if min_size <= int(sig_len[i]) <= max_size:
                        s += int(sig_cnt[i])
                elif int(sig_pos[i]) > hi:
                    break
        for i in (range(min(last, n - 1), -1, -1) if last >= 0 else ()):
            if int(sig_chrom[i]) == cc:
                if lo <= int(sig_pos[i]) <= hi:
                    matched.append(i)
                    if min_size <= int(sig_len[i]) <= max_size:
                        s += int(sig_cnt[i])
                elif int(sig_pos[i]) < lo:
                    break
        if matched:
            last = min(matched)
        sup.append(s)
    if return_last:
        return sup, mids, last
    return sup, mids


def span_count(read_start, read_end, S, E):
    """count_reads_span_region (GTD:140-146) / check_full_cover_reads (GTI:188-196):
    number of records with reference_start < S and reference_end > E."""
    rs = np.asarray(read_start, dtype=np.int64)
    re = np.asarray(read_end, dtype=np.int64)
    return int(np.count_nonzero((rs < S) & (re > E)))


def depth_del(read_start, read_end, pos, svlen):
    """check_full_cover_reads (GTD:149-170).  svlen = |SVLEN|.  Twice the depth is
    returned as an integer (depth2) so the >1000 bp mean of two windows stays exact."""
    if svlen <= 1000:
        return 2 * span_count(read_start, read_end, pos, pos + svlen)
    ls = pos - 150
    rs_ = pos + svlen + 150
    return (span_count(read_start, read_end, ls, ls + 100)
            + span_count(read_start, read_end, rs_, rs_ + 100))


def depth_ins(read_start, read_end, pos, flanking=100):
    """GTI:188-196 with flanking 100 (GTI:327)."""
    return 2 * span_count(read_start, read_end, pos - flanking, pos + flanking)


def genotype(vtype, svlen, call_gt, n_support, depth2, thr):
    """ratio + correct_gt_eval (GTD:206-217,291-318; GTI:228,258-283).

    call_gt: 0 = '0/1', 1 = '1/1', anything else passes through.  depth2 = 2*depth.
    thr = (t_large_11, t_small_11, t_large_01, t_small_01), NaN = leave GT.
    DEL: depth 0 -> ratio 1.  INS: n/0 -> inf, 0/0 -> nan (pandas division).
    Returns (ratio, new_gt).
    """
    depth = depth2 / 2.0
    if vtype == DEL:
        ratio = 1.0 if depth2 == 0 else n_support / depth
    else:
        if depth2 == 0:
            ratio = float("nan") if n_support == 0 else float("inf")
        else:
            ratio = n_support / depth
    new_gt = call_gt
    if call_gt in (0, 1):
        large = svlen > 1000
        t = thr[(0 if large else 1) + (0 if call_gt == 1 else 2)]
        if not math.isnan(t):
            if ratio > t:
                new_gt = 1
            elif ratio <= t:
                new_gt = 0
    return ratio, new_gt


# --------------------------------------------------------------------------- SA split analysis
def sa_segments(primary, sa_entries, total_L, min_mapq):
    """organize_split_signal (SE:345-371).  primary = None or (q_s,q_e,r_s,r_e,chr,strand);
    sa_entries = [(chr, pos_1based, strand(0 '+',1 '-'), firstS, lastS, bias, mapq)] with
    firstS/lastS = leading/trailing *soft* clip only and bias = sum M,D,=,X
    (acquire_clip_pos SE:328-343).  SA pos is NOT converted to 0-based (SE:354).
    """
    seg = []
    if primary is not None:
        seg.append(tuple(primary))
        min_mapq = 0
    for (c, p, st, fs, ls, bias, mq) in sa_entries:
        if mq >= min_mapq:
            if st == 0:
                seg.append((fs, total_L - ls, p, p + bias, c, 0))
            else:
                seg.append((ls, total_L - fs, p, p + bias, c, 1))
    return seg


def _mirror(e, L):
    return (L - e[1], L - e[0]) + tuple(e[2:])


def sa_split_analysis(seg, L, sv_size=30, max_size=100000, max_parts=7):
    """analysis_split_read (SE:193-326), INS/DEL rows only (TRA rows are dropped by
    `grep -w` at SE:637-638 but the different-chromosome trigger is kept).

    Returns list of (type, pos, len, q0, q1, rev): INS sequence = query[q0:q1] of the
    query reversed `rev` times (mod 2) with Python slice semantics; pos already
    truncated the way '%d' prints the reference's float (SE:562-565).
    """
    out = []
    if not (len(seg) <= max_parts or max_parts == -1):
        return out
    sp = sorted(seg, key=lambda x: x[0])
    rev = 0
    trig = 0

    def ins_del(e1, e2, guard=True, need_gap_check=True):
        if e1[3] - e2[2] < sv_size:
            il = e2[0] + e1[3] - e2[2] - e1[1]
            if il >= sv_size:
                if e2[2] - e1[3] <= 100 and (il <= max_size or max_size == -1):
                    if guard:
                        h = int((e2[2] - e1[3]) / 2)
                        out.append((INS, int((e2[2] + e1[3]) / 2), il, e1[1] + h, e2[0] - h, rev))
            dl = e2[2] - e2[0] + e1[1] - e1[3]
            if dl >= sv_size:
                if e2[0] - e1[1] <= 100 and (dl <= max_size or max_size == -1):
                    if guard:
                        out.append((DEL, e1[3], dl, 0, 0, rev))

    if len(sp) == 2:
        e1, e2 = sp
        if e1[4] == e2[4]:
            if e1[5] == e2[5]:
                if e1[5] == 1:
                    e1, e2 = _mirror(sp[1], L), _mirror(sp[0], L)
                    rev ^= 1
                ins_del(e1, e2)
        else:
            trig = 1
    else:
        for a in range(len(sp) - 2):
            e1, e2, e3 = sp[a], sp[a + 1], sp[a + 2]
            if e1[4] == e2[4]:
                if e2[4] == e3[4]:
                    if e1[5] == e3[5] and e1[5] == e2[5]:
                        if e1[5] == 1:
                            e1, e2, e3 = _mirror(sp[a + 2], L), _mirror(sp[a + 1], L), _mirror(sp[a], L)
                            rev ^= 1
                        ins_del(e1, e2, guard=(e3[2] >= e2[3]))
                        if len(sp) - 3 == a:
                            ins_del(e2, e3)
            else:
                trig = 1
    if len(sp) >= 3 and trig == 1:
        if sp[0][4] == sp[-1][4] and sp[0][5] == sp[-1][5]:
            if sp[0][5] == 0:
                e1, e2 = sp[0], sp[-1]
            else:
                e1, e2 = _mirror(sp[-1], L), _mirror(sp[0], L)
                rev ^= 1
            dis_ref = e2[2] - e1[3]
            dis_read = e2[0] - e1[1]
            if dis_ref < 100 and dis_read - dis_ref >= sv_size and (dis_read - dis_ref <= max_size or max_size == -1):
                h = int(dis_ref / 2)
                out.append((INS, min(e2[2], e1[3]), dis_read - dis_ref, e1[1] + h, e2[0] - h, rev))
    return out


# --------------------------------------------------------------------------- edit distance
def edit_distance(a, b):
    """Global unit-cost edit distance: what remove_redundancy.py takes from
    edlib.align(a, b)["editDistance"] (RR:75-81; third-party edlib 1.3.9, mode NW).  The value
    is algorithm-independent; this is the textbook row DP (numpy prefix-min for the in-row
    dependency)."""
    import numpy as np
    n, m = len(a), len(b)
    if n == 0 or m == 0:
        return n + m
    bb = np.frombuffer(b.encode() if isinstance(b, str) else bytes(b), dtype=np.uint8)
    aa = a.encode() if isinstance(a, str) else bytes(a)
    prev = np.arange(m + 1, dtype=np.int64)
    idx = np.arange(m + 1, dtype=np.int64)
    for i in range(1, n + 1):
        cur = np.empty(m + 1, dtype=np.int64)
        cur[0] = i
        cur[1:] = np.minimum(prev[:-1] + (bb != aa[i - 1]), prev[1:] + 1)
        prev = np.minimum.accumulate(cur - idx) + idx
    return int(prev[m])
