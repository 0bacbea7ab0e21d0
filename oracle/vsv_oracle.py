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
