"""ctypes binding of libvolcano_b200.so (include/volcanosv_b200.h).

There is no CPU fallback: if the shared library is missing or no CUDA device is
present, creating a Context raises.  Host-IO methods take/return numpy arrays; the
`*_dev` variants take raw device pointers (ints) for resident pipelines.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvolcano_b200.so")

MODE_CONTIG, MODE_READS, MODE_CUTE = 0, 1, 2
DEL, INS = 0, 1
SPLIT_HIFI, SPLIT_ONT, SPLIT_CLR, SPLIT_READS = 0, 1, 2, 3
E_CUDA, E_ARG, E_CAPACITY, E_NOMEM = -1, -2, -3, -4

SIG_DTYPE = np.dtype([("rec", "<i4"), ("type", "<i4"), ("pos", "<i4"), ("len", "<i4"),
                      ("q0", "<i4"), ("q1", "<i4"), ("first", "<i4"), ("count", "<i4")])
assert SIG_DTYPE.itemsize == 32


class VsvError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("volcanosv_b200 error %d: %s" % (code, msg))
        self.code = code


_lib = None
_vp, _i64, _i32, _int = C.c_void_p, C.c_int64, C.c_int32, C.c_int

# name -> argtypes (restype is always int unless listed in _RESTYPE)
SIGNATURES = {
    "vsv_ctx_create": [_int, C.POINTER(_vp)],
    "vsv_ctx_destroy": [_vp],
    "vsv_last_error": [_vp],
    "vsv_ctx_set_device_io": [_vp, _int],
    "vsv_sync": [_vp],
    "vsv_alloc_pinned": [C.POINTER(_vp), C.c_size_t],
    "vsv_free_pinned": [_vp],
    "vsv_get_timing": [_vp, C.POINTER(C.c_float), C.POINTER(_i64)],
    "vsv_version": [],
    "vsv_timer_start": [_vp],
    "vsv_timer_stop": [_vp, C.POINTER(C.c_float)],
    "vsv_get_main_kernel_ms": [_vp, C.POINTER(C.c_float)],
    "vsv_scan_cigar": [_vp, _int, _int, _int, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _i64, C.POINTER(_i64), _vp, _vp, _vp],
    "vsv_expand_ops16": [_vp, _vp, _i64, _vp, _vp, _i64, _vp],
    "vsv_expand_ops8": [_vp, _vp, _i64, _vp, _vp, _i64, _vp],
    "vsv_expand_ops12": [_vp, _vp, _i64, _vp, _vp, _i64, _vp],
    "vsv_ops12_bytes": [_i64],
    "vsv_pack_ops12": [_vp, _i64, _int, _vp, _vp, _vp, _i64, C.POINTER(_i64)],
    "vsv_scan_reads_fused": [_vp, _int, _int, _int, _int, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _i64, C.POINTER(_i64),
                             _vp, _i64, C.POINTER(_i64), _vp, _vp, _vp],
    "vsv_merge_in_record": [_vp, _int, _vp, _i64, _vp, C.POINTER(_i64), _vp, C.POINTER(_i64)],
    "vsv_split_pairs": [_vp, _int, _int, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _i64, C.POINTER(_i64)],
    "vsv_sa_analysis": [_vp, _int, _int, _int, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _vp, _i64, C.POINTER(_i64)],
    "vsv_argsort_stable": [_vp, _vp, _i64, _vp],
    "vsv_sort_sigs": [_vp, _vp, _i64, _int, _vp],
    "vsv_sig_columns": [_vp, _vp, _i64, _vp, _vp, _vp],
    "vsv_add_i32": [_vp, _vp, _vp, _i64, _vp],
    "vsv_ge_u8": [_vp, _vp, _i32, _i64, _vp],
    "vsv_select_i32": [_vp, _vp, _i32, _vp, _vp, _i64, _vp],
    "vsv_cluster": [_vp, _int, _vp, _vp, _i64, _vp, C.POINTER(_i64)],
    "vsv_pair_haps": [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _i64, _vp, _vp],
    "vsv_fp_support": [_vp, _vp, _vp, _i64, _vp, _vp, _i64, _vp],
    "vsv_ins_cov": [_vp, _vp, _i64, _vp, _vp, _i64, _int, _vp],
    "vsv_del_cov": [_vp, _vp, _vp, _i64, _vp, _vp, _i64, _int, _vp],
    "vsv_group_sigs": [_vp, _vp, _vp, _vp, _i64, _int, _vp, _vp, _vp, _vp, C.POINTER(_i64)],
    "vsv_sig_support": [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _i64, _i32, _vp, _vp, C.POINTER(_i32)],
    "vsv_span_depth": [_vp, _vp, _vp, _i64, _vp, _vp, _i64, _vp],
    "vsv_edit_distance_batch": [_vp, _vp, _vp, _i64, _vp, _vp, _i64, _vp],
    "vsv_bam_open": [C.c_char_p, _int, C.POINTER(_vp)],
    "vsv_bam_open_ref": [C.c_char_p, C.c_char_p, _int, C.POINTER(_vp)],
    "vsv_bam_close": [_vp],
    "vsv_bam_last_error": [],
    "vsv_bam_n_ref": [_vp],
    "vsv_bam_ref": [_vp, _int, C.POINTER(C.c_char_p), C.POINTER(_i64), C.POINTER(_i64)],
    "vsv_bam_chrom_sizes": [_vp, _int, _vp],
    "vsv_bam_record_seq": [_vp, _int, _i64, _vp, _i64, C.POINTER(_i64)],
    "vsv_bam_chrom_fill": [_vp, _int, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, C.POINTER(_i64), _vp, _vp, _vp,
                           _vp, _vp, _vp],
    "vsv_genotype": [_vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp],
}
_RESTYPE = {"vsv_ctx_destroy": None, "vsv_last_error": C.c_char_p, "vsv_version": C.c_char_p, "vsv_bam_close": None,
            "vsv_bam_last_error": C.c_char_p, "vsv_ops12_bytes": _i64}


def load_library(path: Optional[str] = None):
    """dlopen the in-tree CUDA library.  Raises (never falls back) if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise VsvError(E_CUDA, "CUDA extension %s not built; run `python -m volcanosv_b200.build` "
                               "(there is no CPU fallback)" % path)
    lib = C.CDLL(path)
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = _RESTYPE.get(name, _int)
    _lib = lib
    return lib


def _arr(a, dtype):
    if a is None:
        return None
    return np.ascontiguousarray(a, dtype=dtype)


def _p(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    return C.c_void_p(a.ctypes.data)


class Context:
    """One vsv_ctx (one GPU, one stream).  Not thread-safe."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = _vp()
        rc = self.lib.vsv_ctx_create(device, C.byref(h))
        if rc != 0:
            raise VsvError(rc, "vsv_ctx_create(device=%d) failed: no usable CUDA device "
                               "(this library has no CPU path)" % device)
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.vsv_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise VsvError(rc, (self.lib.vsv_last_error(self.h) or b"").decode())

    def set_device_io(self, on: bool):
        self._chk(self.lib.vsv_ctx_set_device_io(self.h, 1 if on else 0))

    def sync(self):
        self._chk(self.lib.vsv_sync(self.h))

    def timer_start(self):
        self._chk(self.lib.vsv_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float()
        self._chk(self.lib.vsv_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def main_kernel_ms(self):
        ms = C.c_float()
        self._chk(self.lib.vsv_get_main_kernel_ms(self.h, C.byref(ms)))
        return ms.value

    def timing(self):
        ms, n = C.c_float(), _i64()
        self._chk(self.lib.vsv_get_timing(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    # ---- stage 1 ---------------------------------------------------------------
    def scan_cigar(self, mode, ops, op_off, pos, mapq, enable=None, min_len=30, min_mapq=0,
                   cap=None, totals=True):
        """-> (sigs[SIG_DTYPE], ref_total, qry_total, aux_total)."""
        ops = _arr(ops, np.uint32)
        op_off = _arr(op_off, np.int64)
        pos = _arr(pos, np.int32)
        mapq = _arr(mapq, np.uint8)
        enable = _arr(enable, np.uint8)
        n = len(pos)
        if ops.ctypes.data % 16:
            ops = _aligned_copy(ops)
        if cap is None:
            cap = max(1024, len(ops) // 16)
        rt = np.zeros(n, np.int32) if totals else None
        qt = np.zeros(n, np.int32) if totals else None
        xt = np.zeros(n, np.int32) if totals else None
        while True:
            out = _aligned_empty(cap, SIG_DTYPE)
            n_out = _i64()
            rc = self.lib.vsv_scan_cigar(self.h, mode, min_len, min_mapq, _p(ops), _p(op_off), _p(pos), _p(mapq),
                                         _p(enable), n, _p(out), cap, C.byref(n_out), _p(rt), _p(qt), _p(xt))
            if rc == E_CAPACITY:
                cap = int(n_out.value)
                continue
            self._chk(rc)
            return out[: n_out.value], rt, qt, xt

    def expand_ops16(self, ops16, exc_idx, exc_val):
        """-> the BAM-native u32 ops (host-IO mode: round trip through the device)."""
        o16 = _arr(ops16, np.uint16)
        if o16.ctypes.data % 16:
            o16 = _aligned_copy(o16)
        ei, ev = _arr(exc_idx, np.int64), _arr(exc_val, np.uint32)
        out = _aligned_empty(len(o16), np.uint32)
        self._chk(self.lib.vsv_expand_ops16(self.h, _p(o16), len(o16), _p(ei), _p(ev), len(ei), _p(out)))
        return out

    def expand_ops8(self, ops8, exc_idx, exc_val):
        o8 = _arr(ops8, np.uint8)
        if o8.ctypes.data % 16:
            o8 = _aligned_copy(o8)
        ei, ev = _arr(exc_idx, np.int64), _arr(exc_val, np.uint32)
        out = _aligned_empty(len(o8), np.uint32)
        self._chk(self.lib.vsv_expand_ops8(self.h, _p(o8), len(o8), _p(ei), _p(ev), len(ei), _p(out)))
        return out

    def expand_ops12(self, ops12, n_ops, exc_pair, exc_val):
        o12 = _arr(ops12, np.uint8)
        if len(o12) != ops12_bytes(n_ops):
            raise ValueError("expand_ops12: %d bytes for %d ops, expected %d" % (len(o12), n_ops, ops12_bytes(n_ops)))
        if o12.ctypes.data % 16:
            o12 = _aligned_copy(o12)
        ei, ev = _arr(exc_pair, np.int64), _arr(exc_val, np.uint32)
        out = _aligned_empty(int(n_ops), np.uint32)
        self._chk(self.lib.vsv_expand_ops12(self.h, _p(o12), int(n_ops), _p(ei), _p(ev), len(ei), _p(out)))
        return out

    def scan_reads_fused(self, ops, op_off, pos, mapq, enable_cute=None, min_len_reads=30, min_mapq_reads=50,
                         min_len_cute=10, min_mapq_cute=20, cap=None, totals=True):
        """ERS + SE scanners in one pass -> (reads_sigs, cute_sigs, ref_total, qry_total, aux_total)."""
        ops = _arr(ops, np.uint32)
        op_off = _arr(op_off, np.int64)
        pos = _arr(pos, np.int32)
        mapq = _arr(mapq, np.uint8)
        enable_cute = _arr(enable_cute, np.uint8)
        n = len(pos)
        if ops.ctypes.data % 16:
            ops = _aligned_copy(ops)
        cap_r = cap_c = cap if cap is not None else max(1024, len(ops) // 16)
        rt = np.zeros(n, np.int32) if totals else None
        qt = np.zeros(n, np.int32) if totals else None
        xt = np.zeros(n, np.int32) if totals else None
        while True:
            out_r, out_c = _aligned_empty(cap_r, SIG_DTYPE), _aligned_empty(cap_c, SIG_DTYPE)
            n_r, n_c = _i64(), _i64()
            rc = self.lib.vsv_scan_reads_fused(self.h, min_len_reads, min_mapq_reads, min_len_cute, min_mapq_cute,
                                               _p(ops), _p(op_off), _p(pos), _p(mapq), _p(enable_cute), n, _p(out_r),
                                               cap_r, C.byref(n_r), _p(out_c), cap_c, C.byref(n_c), _p(rt), _p(qt), _p(xt))
            if rc == E_CAPACITY:
                cap_r, cap_c = max(cap_r, int(n_r.value)), max(cap_c, int(n_c.value))
                continue
            self._chk(rc)
            return out_r[: n_r.value], out_c[: n_c.value], rt, qt, xt

    def merge_in_record(self, mode, sigs):
        sigs = np.ascontiguousarray(sigs, dtype=SIG_DTYPE)
        n = len(sigs)
        od, oi = np.empty(n, SIG_DTYPE), np.empty(n, SIG_DTYPE)
        nd, ni = _i64(), _i64()
        self._chk(self.lib.vsv_merge_in_record(self.h, mode, _p(sigs), n, _p(od), C.byref(nd), _p(oi), C.byref(ni)))
        return od[: nd.value], oi[: ni.value]

    def split_pairs(self, variant, min_mapq, pos, ref_end, readlen, first_op, last_op, mapq, is_rev, name_id,
                    select=None):
        n = len(pos)
        a = [_arr(pos, np.int32), _arr(ref_end, np.int32), _arr(readlen, np.int32), _arr(first_op, np.uint32),
             _arr(last_op, np.uint32), _arr(mapq, np.uint8), _arr(is_rev, np.uint8), _arr(name_id, np.int32),
             _arr(select, np.uint8)]
        out = np.empty(max(n, 1), SIG_DTYPE)
        n_out = _i64()
        self._chk(self.lib.vsv_split_pairs(self.h, variant, min_mapq, *[_p(x) for x in a], n, _p(out), len(out),
                                           C.byref(n_out)))
        return out[: n_out.value]

    # ---- stage 2 ---------------------------------------------------------------
    def sa_analysis(self, seg_off, q_s, q_e, r_s, r_e, chrom, strand, read_len, sv_size=30, max_size=100000,
                    max_parts=7):
        """analysis_split_read (SE:193-326): rows (rec=read, type, pos, len, q0, q1, first=rev)."""
        seg_off = _arr(seg_off, np.int64)
        n_read, n_seg = len(seg_off) - 1, int(seg_off[-1])
        cols = [_arr(x, np.int32) for x in (q_s, q_e, r_s, r_e, chrom)]
        strand, read_len = _arr(strand, np.uint8), _arr(read_len, np.int32)
        cap = 2 * n_seg + 2
        out = _aligned_empty(cap, SIG_DTYPE)
        n = _i64()
        self._chk(self.lib.vsv_sa_analysis(self.h, sv_size, max_size, max_parts, _p(seg_off), *[_p(c) for c in cols],
                                           _p(strand), _p(read_len), n_read, n_seg, _p(out), cap, C.byref(n)))
        return out[: n.value]

    def argsort_stable(self, key):
        key = _arr(key, np.int32)
        order = np.empty(len(key), np.int32)
        self._chk(self.lib.vsv_argsort_stable(self.h, _p(key), len(key), _p(order)))
        return order

    def sort_sigs(self, sigs, tie_mode=0):
        sigs = np.ascontiguousarray(sigs, dtype=SIG_DTYPE)
        out = np.empty(len(sigs), SIG_DTYPE)
        self._chk(self.lib.vsv_sort_sigs(self.h, _p(sigs), len(sigs), tie_mode, _p(out)))
        return out

    def cluster(self, is_del, pos, length):
        pos, length = _arr(pos, np.int32), _arr(length, np.int32)
        keep = np.empty(len(pos), np.int32)
        nk = _i64()
        self._chk(self.lib.vsv_cluster(self.h, 1 if is_del else 0, _p(pos), _p(length), len(pos), _p(keep), C.byref(nk)))
        return keep[: nk.value]

    def pair_haps(self, pos1, len1, typ1, pos2, len2, typ2):
        a = [_arr(x, np.int32) for x in (pos1, len1, typ1)]
        b = [_arr(x, np.int32) for x in (pos2, len2, typ2)]
        p1, p2 = np.empty(len(a[0]), np.int32), np.empty(len(b[0]), np.int32)
        self._chk(self.lib.vsv_pair_haps(self.h, _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]), _p(b[0]), _p(b[1]), _p(b[2]),
                                         len(b[0]), _p(p1), _p(p2)))
        return p1, p2

    def fp_support(self, call_pos, call_len, sig_pos, sig_len):
        cp, cl, sp, sl = (_arr(x, np.int32) for x in (call_pos, call_len, sig_pos, sig_len))
        out = np.empty(len(cp), np.int32)
        self._chk(self.lib.vsv_fp_support(self.h, _p(cp), _p(cl), len(cp), _p(sp), _p(sl), len(sp), _p(out)))
        return out

    def ins_cov(self, call_pos, sig_pos, sig_len, flanking=1000):
        cp, sp, sl = (_arr(x, np.int32) for x in (call_pos, sig_pos, sig_len))
        out = np.empty(len(cp), np.int64)
        self._chk(self.lib.vsv_ins_cov(self.h, _p(cp), len(cp), _p(sp), _p(sl), len(sp), flanking, _p(out)))
        return out

    def del_cov(self, call_start, call_end, sig_pos, sig_len, flanking=1000):
        cs, ce, sp, sl = (_arr(x, np.int32) for x in (call_start, call_end, sig_pos, sig_len))
        out = np.empty(len(cs), np.int64)
        self._chk(self.lib.vsv_del_cov(self.h, _p(cs), _p(ce), len(cs), _p(sp), _p(sl), len(sp), flanking, _p(out)))
        return out

    # ---- stage 3 ---------------------------------------------------------------
    def group_sigs(self, chrom, pos, length, min_len=0):
        c, p, l = (_arr(x, np.int32) for x in (chrom, pos, length))
        n = len(p)
        o = [np.empty(n, np.int32) for _ in range(4)]
        nu = _i64()
        self._chk(self.lib.vsv_group_sigs(self.h, _p(c), _p(p), _p(l), n, min_len, _p(o[0]), _p(o[1]), _p(o[2]), _p(o[3]),
                                          C.byref(nu)))
        return tuple(x[: nu.value] for x in o)

    def sig_support(self, cand_chrom, cand_pos, cand_svlen, sig_chrom, sig_pos, sig_len, sig_cnt, last_in=0):
        cc, cp, cl = (_arr(x, np.int32) for x in (cand_chrom, cand_pos, cand_svlen))
        sc, sp, sl, sn = (_arr(x, np.int32) for x in (sig_chrom, sig_pos, sig_len, sig_cnt))
        sup, mid = np.empty(len(cc), np.int32), np.empty(len(cc), np.int32)
        last = _i32()
        self._chk(self.lib.vsv_sig_support(self.h, _p(cc), _p(cp), _p(cl), len(cc), _p(sc), _p(sp), _p(sl), _p(sn),
                                           len(sc), last_in, _p(sup), _p(mid), C.byref(last)))
        return sup, mid, last.value

    def span_depth(self, read_start, read_end, S, E):
        rs, re, s, e = (_arr(x, np.int32) for x in (read_start, read_end, S, E))
        out = np.empty(len(s), np.int32)
        self._chk(self.lib.vsv_span_depth(self.h, _p(rs), _p(re), len(rs), _p(s), _p(e), len(s), _p(out)))
        return out

    def genotype(self, read_start, read_end, cand_pos, cand_svlen, cand_type, call_gt, n_support, thr_del, thr_ins):
        rs, re = _arr(read_start, np.int32), _arr(read_end, np.int32)
        cp, cl, ct, cg, ns = (_arr(x, np.int32) for x in (cand_pos, cand_svlen, cand_type, call_gt, n_support))
        td, ti = _arr(thr_del, np.float64), _arr(thr_ins, np.float64)
        n = len(cp)
        d2, ratio, ng = np.empty(n, np.int32), np.empty(n, np.float64), np.empty(n, np.int32)
        self._chk(self.lib.vsv_genotype(self.h, _p(rs), _p(re), len(rs), _p(cp), _p(cl), _p(ct), _p(cg), _p(ns), n,
                                        _p(td), _p(ti), _p(d2), _p(ratio), _p(ng)))
        return d2, ratio, ng


    def edit_distance_batch(self, strings, ia, ib):
        """Global edit distance of strings[ia[p]] vs strings[ib[p]] (edlib NW distance, RR:75-81)."""
        bs = [x.encode() if isinstance(x, str) else bytes(x) for x in strings]
        off = np.zeros(len(bs) + 1, np.int64)
        np.cumsum([len(b) for b in bs], out=off[1:])
        chars = np.frombuffer(b"".join(bs) + b"\0", dtype=np.uint8)
        ia, ib = _arr(ia, np.int32), _arr(ib, np.int32)
        out = np.zeros(len(ia), np.int32)
        self._chk(self.lib.vsv_edit_distance_batch(self.h, _p(chars), _p(off), len(bs), _p(ia), _p(ib), len(ia), _p(out)))
        return out


class BamFile:
    """Native BAM reader (include/volcanosv_b200.h: vsv_bam_*): the host-side replacement of
    `pysam.AlignmentFile(path).fetch(chr)` for whole-chromosome scans.  Host code only (no GPU)."""

    def __init__(self, path: str, n_threads: int = 0, ref: Optional[str] = None):
        """ref=None: every mapped record; ref="chr10": that reference only (seeks through the .bai when there
        is one); ref="": header only."""
        self.lib = load_library()
        self.h = _vp()
        if ref is None:
            rc = self.lib.vsv_bam_open(os.fsencode(path), n_threads, C.byref(self.h))
        else:
            rc = self.lib.vsv_bam_open_ref(os.fsencode(path), ref.encode(), n_threads, C.byref(self.h))
        if rc != 0:
            raise VsvError(rc, (self.lib.vsv_bam_last_error() or b"").decode())
        self.refs = []
        for i in range(self.lib.vsv_bam_n_ref(self.h)):
            name, ln, n = C.c_char_p(), _i64(), _i64()
            self.lib.vsv_bam_ref(self.h, i, C.byref(name), C.byref(ln), C.byref(n))
            self.refs.append((name.value.decode(), int(ln.value), int(n.value)))

    def close(self):
        if self.h:
            self.lib.vsv_bam_close(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def _chk(self, rc):
        if rc != 0:
            raise VsvError(rc, (self.lib.vsv_bam_last_error() or b"").decode())

    def chrom(self, name_or_tid, compact=False, want_seq: bool = True):
        """Columns of one reference: dict(pos, mapq, flag, qlen, op_off, ops | ops16 | ops8 (+ exc_idx, exc_val),
        name_off, names, seq_off, seq, sa_off, sa).  compact: False (u32 ops), True / 16 (ops16), 8 (ops8),
        12 (ops12: the u32 words of the feeder packed by vsv_pack_ops12; exc_idx holds PAIR indices, n_ops the op count)."""
        tid = name_or_tid if isinstance(name_or_tid, int) else [r[0] for r in self.refs].index(name_or_tid)
        sizes = np.zeros(5, np.int64)
        self._chk(self.lib.vsv_bam_chrom_sizes(self.h, tid, _p(sizes)))
        n, n_ops, nb, sb, ab = (int(x) for x in sizes)
        d = dict(pos=np.empty(n, np.int32), mapq=np.empty(n, np.uint8), flag=np.empty(n, np.uint16), qlen=np.empty(n, np.int32),
                 op_off=np.empty(n + 1, np.int64), name_off=np.empty(n + 1, np.int64), names=np.empty(nb, np.uint8),
                 seq_off=np.empty(n + 1, np.int64), seq=np.empty(sb if want_seq else 0, np.uint8),
                 sa_off=np.empty(n + 1, np.int64), sa=np.empty(ab, np.uint8))
        n_exc = _i64(0)
        if compact == 12:
            ops = _aligned_empty(n_ops, np.uint32)
            self._chk(self._fill(tid, d, ops, None, 0, n_exc, want_seq))
            d["ops12"], d["exc_idx"], d["exc_val"] = pack_ops12_native(ops)
            d["n_ops"] = n_ops
        elif compact:
            key = "ops8" if compact == 8 else "ops16"
            d[key] = _aligned_empty(n_ops, np.uint8 if compact == 8 else np.uint16)
            cap = max(1024, n_ops // (32 if compact == 8 else 256))
            while True:
                d["exc_idx"], d["exc_val"] = np.empty(cap, np.int64), np.empty(cap, np.uint32)
                rc = self._fill(tid, d, None, d[key], cap, n_exc, want_seq)
                if rc != E_CAPACITY:
                    break
                cap = int(n_exc.value)
            self._chk(rc)
            d["exc_idx"], d["exc_val"] = d["exc_idx"][: n_exc.value].copy(), d["exc_val"][: n_exc.value].copy()
        else:
            d["ops"] = _aligned_empty(n_ops, np.uint32)
            self._chk(self._fill(tid, d, d["ops"], None, 0, n_exc, want_seq))
        return d

    def record_seq(self, tid: int, i: int, length: int) -> str:
        """SEQ of record i of reference tid (length = its qlen)."""
        buf = C.create_string_buffer(max(int(length), 1))
        n = _i64()
        self._chk(self.lib.vsv_bam_record_seq(self.h, tid, int(i), buf, int(length), C.byref(n)))
        return buf.raw[: n.value].decode()

    def _fill(self, tid, d, ops, packed, cap, n_exc, want_seq):
        o16 = packed if packed is not None and packed.dtype == np.uint16 else None
        o8 = packed if packed is not None and packed.dtype == np.uint8 else None
        return self.lib.vsv_bam_chrom_fill(self.h, tid, _p(d["pos"]), _p(d["mapq"]), _p(d["flag"]), _p(d["qlen"]), _p(d["op_off"]),
                                           _p(ops), _p(o16), _p(o8), _p(d.get("exc_idx")), _p(d.get("exc_val")), cap, C.byref(n_exc),
                                           _p(d["name_off"]), _p(d["names"]), _p(d["seq_off"]),
                                           _p(d["seq"]) if want_seq else None, _p(d["sa_off"]), _p(d["sa"]))


def pack_ops16(ops, chunk: int = 1 << 24):
    """Host packer of the compact staging format of vsv_expand_ops16: (ops16, exc_idx, exc_val)."""
    ops = np.ascontiguousarray(ops, dtype=np.uint32)
    o16 = _aligned_empty(len(ops), np.uint16)
    idx = []
    for a in range(0, len(ops), chunk):
        w = ops[a: a + chunk]
        np.copyto(o16[a: a + chunk], w, casting="unsafe")
        e = np.flatnonzero(w >= np.uint32(0xFFF0)).astype(np.int64)
        o16[a + e] = (np.uint16(0xFFF0) | (w[e] & np.uint32(15)).astype(np.uint16))
        idx.append(e + a)
    exc_idx = np.concatenate(idx) if idx else np.zeros(0, np.int64)
    exc_val = ops[exc_idx].copy()
    return o16, exc_idx, exc_val


OPS8_M, OPS8_I, OPS8_D = 160, 47, 47      # include/volcanosv_b200.h: VSV_OPS8_*


def pack_ops8(ops, chunk: int = 1 << 24):
    """Host packer of vsv_expand_ops8: (ops8, exc_idx, exc_val).  One byte per short M/I/D run
    (M 1..160 -> 0..159, I 1..47 -> 160..206, D 1..47 -> 207..253), 0xFF + exception otherwise.
    Works through the array in chunks so the temporaries stay small next to the input."""
    ops = np.ascontiguousarray(ops, dtype=np.uint32)
    cnt = np.array([OPS8_M, OPS8_I, OPS8_D] + [0] * 13, np.uint32)
    base = np.array([0, OPS8_M, OPS8_M + OPS8_I] + [0] * 13, np.uint32)
    o8 = _aligned_empty(len(ops), np.uint8)
    idx = []
    for a in range(0, len(ops), chunk):
        w = ops[a: a + chunk]
        code, ln = w & np.uint32(15), w >> np.uint32(4)
        big = (ln < 1) | (ln > cnt[code])
        np.copyto(o8[a: a + chunk], base[code] + ln - np.uint32(1), casting="unsafe")
        e = np.flatnonzero(big).astype(np.int64)
        o8[a + e] = 0xFF
        idx.append(e + a)
    exc_idx = np.concatenate(idx) if idx else np.zeros(0, np.int64)
    exc_val = ops[exc_idx].copy()
    return o8, exc_idx, exc_val


OPS12_M, OPS12_X = 128, 8                 # include/volcanosv_b200.h: VSV_OPS12_*


def ops12_bytes(n_ops: int) -> int:
    """vsv_ops12_bytes: 12 bits per pair of ops, padded to a whole group of 8 tokens (12 bytes)."""
    n_ops = int(n_ops)
    return 0 if n_ops <= 0 else -(-(-(-n_ops // 2)) // 8) * 12


def pack_ops12(ops, chunk: int = 1 << 24):
    """Host packer of vsv_expand_ops12: (ops12, exc_pair, exc_val).  Token k of the pair (ops[2k], ops[2k+1]) =
    {M of 1..128, I/D of 1..8} in either order: (mlen-1) | (xlen-1) << 7 | is_D << 10 | indel_first << 11, two
    tokens per three bytes, little-endian; any other pair (and the lone last op of an odd array) goes to the
    exception list with both full words.  `chunk` (even) pairs at a time keep the temporaries small."""
    ops = np.ascontiguousarray(ops, dtype=np.uint32)
    n = len(ops)
    n_pair = (n + 1) // 2
    o12 = _aligned_empty(ops12_bytes(n), np.uint8)
    o12[:] = 0
    chunk = max(8, chunk - chunk % 8)
    idx = []
    for a in range(0, n_pair, chunk):
        b = min(n_pair, a + chunk)
        first = ops[2 * a: 2 * b: 2]
        second = np.zeros(b - a, np.uint32)
        tail = ops[2 * a + 1: 2 * b: 2]
        second[: len(tail)] = tail
        c1, l1, c2, l2 = first & np.uint32(15), first >> np.uint32(4), second & np.uint32(15), second >> np.uint32(4)
        indel_first = c1 != 0
        cm, lm = np.where(indel_first, c2, c1), np.where(indel_first, l2, l1)
        cx, lx = np.where(indel_first, c1, c2), np.where(indel_first, l1, l2)
        ok = (cm == 0) & (lm >= 1) & (lm <= OPS12_M) & ((cx == 1) | (cx == 2)) & (lx >= 1) & (lx <= OPS12_X)
        tok = ((lm - np.uint32(1)) | ((lx - np.uint32(1)) << np.uint32(7)) | ((cx == 2).astype(np.uint32) << np.uint32(10))
               | (indel_first.astype(np.uint32) << np.uint32(11)))
        tok = np.where(ok, tok, np.uint32(0))
        if len(tok) % 2:
            tok = np.append(tok, np.uint32(0))
        t0, t1 = tok[0::2], tok[1::2]
        out = o12[3 * (a // 2): 3 * (a // 2) + 3 * len(t0)].reshape(-1, 3)
        np.copyto(out[:, 0], t0 & np.uint32(0xFF), casting="unsafe")
        np.copyto(out[:, 1], (t0 >> np.uint32(8)) | ((t1 & np.uint32(15)) << np.uint32(4)), casting="unsafe")
        np.copyto(out[:, 2], t1 >> np.uint32(4), casting="unsafe")
        idx.append(np.flatnonzero(~ok).astype(np.int64) + a)
    exc_pair = np.concatenate(idx) if idx else np.zeros(0, np.int64)
    exc_val = np.zeros(2 * len(exc_pair), np.uint32)
    exc_val[0::2] = ops[2 * exc_pair]
    odd = 2 * exc_pair + 1
    inside = odd < n
    exc_val[1::2][inside] = ops[odd[inside]]
    return o12, exc_pair, exc_val


def pack_ops12_native(ops, threads: int = 0):
    """vsv_pack_ops12: the same (ops12, exc_pair, exc_val) as pack_ops12, packed by the library's host threads
    (no GPU involved): seconds -> tenths of a second for a chromosome of long reads."""
    lib = load_library()
    ops = np.ascontiguousarray(ops, dtype=np.uint32)
    n = len(ops)
    o12 = _aligned_empty(ops12_bytes(n), np.uint8)
    cap = max(1024, n // 64)
    while True:
        ep, ev, n_exc = np.empty(cap, np.int64), np.empty(2 * cap, np.uint32), _i64(0)
        rc = lib.vsv_pack_ops12(_p(ops), n, int(threads), _p(o12), _p(ep), _p(ev), cap, C.byref(n_exc))
        if rc == E_CAPACITY:
            cap = int(n_exc.value)
            continue
        if rc != 0:
            raise VsvError(rc, (lib.vsv_bam_last_error() or b"").decode())
        k = int(n_exc.value)
        return o12, ep[:k].copy(), ev[: 2 * k].copy()


def unpack_ops12(o12, n_ops, exc_pair, exc_val):
    """Inverse of pack_ops12 on the host (tests, debugging)."""
    n_ops = int(n_ops)
    b = np.asarray(o12, dtype=np.uint8).astype(np.uint32).reshape(-1, 3)
    tok = np.empty(2 * len(b), np.uint32)
    tok[0::2] = b[:, 0] | ((b[:, 1] & 15) << 8)
    tok[1::2] = (b[:, 1] >> 4) | (b[:, 2] << 4)
    m = ((tok & 127) + 1) << 4
    x = ((((tok >> 7) & 7) + 1) << 4) | (1 + ((tok >> 10) & 1))
    swap = (tok >> 11) == 1
    out = np.empty(2 * len(tok), np.uint32)
    out[0::2] = np.where(swap, x, m)
    out[1::2] = np.where(swap, m, x)
    ep = np.asarray(exc_pair, dtype=np.int64)
    ev = np.asarray(exc_val, dtype=np.uint32)
    out[2 * ep] = ev[0::2]
    out[2 * ep + 1] = ev[1::2]
    return out[:n_ops].copy()


def unpack_ops8(o8, exc_idx, exc_val):
    """Inverse of pack_ops8 on the host (tests, debugging)."""
    b = np.asarray(o8).astype(np.uint32)
    code = (b >= OPS8_M).astype(np.uint32) + (b >= OPS8_M + OPS8_I).astype(np.uint32)
    base = np.array([0, OPS8_M, OPS8_M + OPS8_I], np.uint32)
    out = ((b - base[code] + 1) << 4) | code
    out[exc_idx] = exc_val
    return out


def _aligned_empty(n, dtype, align=64):
    dtype = np.dtype(dtype)
    raw = np.empty(n * dtype.itemsize + align, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off: off + n * dtype.itemsize].view(dtype)


def _aligned_copy(a, align=64):
    out = _aligned_empty(len(a), a.dtype, align)
    out[:] = a
    return out
