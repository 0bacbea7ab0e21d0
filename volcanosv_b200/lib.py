"""ctypes binding of libvolcano_b200.so (include/volcanosv_b200.h).

There is no CPU fallback: if the shared library is missing or no CUDA device is
present, creating a Context raises.  Host-IO methods take/return numpy arrays; the
`*_dev` variants take raw device pointers (ints) for resident pipelines.
"""
from __future__ import annotations

import ctypes as C
import os
import sys
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VSV_B200_LIB") or os.path.join(_HERE, "libvolcano_b200.so")   # override: profiling builds of the same sources

MODE_CONTIG, MODE_READS, MODE_CUTE = 0, 1, 2
DEL, INS = 0, 1
SPLIT_HIFI, SPLIT_ONT, SPLIT_CLR, SPLIT_READS = 0, 1, 2, 3
E_CUDA, E_ARG, E_CAPACITY, E_NOMEM = -1, -2, -3, -4

SIG_DTYPE = np.dtype([("rec", "<i4"), ("type", "<i4"), ("pos", "<i4"), ("len", "<i4"),
                      ("q0", "<i4"), ("q1", "<i4"), ("first", "<i4"), ("count", "<i4")])
assert SIG_DTYPE.itemsize == 32


DTYPE_CODE = {"Hifi": 0, "ONT": 1, "CLR": 2}
OVF_READS, OVF_CUTE, OVF_DEL, OVF_INS, OVF_STAGE, OVF_POS, OVF_DUPLIST = 1, 2, 4, 8, 16, 32, 64
TAB_READS, TAB_CUTE, TAB_DEL_ROWS, TAB_INS_ROWS, TAB_REF_END, TAB_READ_LEN, TAB_IN_BED, TAB_INS_DUPS, TAB_UDEL, TAB_UINS = range(10)
JOIN_FP, JOIN_COV, JOIN_SUPPORT, JOIN_GENOTYPE = 1, 2, 4, 8
ROW_SA_REV = 256


class Params(C.Structure):
    """struct vsv_params (include/volcanosv_b200.h): every constant of the reference's operators."""
    _fields_ = [("dtype", C.c_int32), ("reads_min_len", C.c_int32), ("reads_min_mapq", C.c_int32), ("cute_min_len", C.c_int32),
                ("cute_min_mapq", C.c_int32), ("cute_min_read_len", C.c_int32), ("sa_sv_size", C.c_int32),
                ("sa_max_size", C.c_int32), ("sa_max_parts", C.c_int32), ("batch_size", C.c_int32),
                ("contig_min_len", C.c_int32), ("contig_min_mapq", C.c_int32), ("cluster_max_shift", C.c_int32),
                ("pair_max_shift", C.c_int32), ("clr_max_ins_pct", C.c_double), ("clr_min_var_dist", C.c_double),
                ("fp_max_comp", C.c_int32), ("fp_max_shift", C.c_int32), ("cov_flanking", C.c_int32), ("cov_min_size", C.c_int32),
                ("cov_lb_ins", C.c_double), ("cov_rb_ins", C.c_double), ("cov_lb_del", C.c_double), ("cov_rb_del", C.c_double),
                ("gt_ins_min_len", C.c_int32), ("gt_min_size_sim", C.c_double), ("gt_max_shift_ratio", C.c_double),
                ("gt_min_shift", C.c_double), ("gt_ins_flank", C.c_int32), ("gt_del_large", C.c_int32), ("gt_del_off", C.c_int32),
                ("gt_del_win", C.c_int32), ("thr_del", C.c_double * 4), ("thr_ins", C.c_double * 4)]


class RecordsStruct(C.Structure):
    """struct vsv_records."""
    _fields_ = [("n_rec", C.c_int64), ("n_ops", C.c_int64), ("ops_format", C.c_int32), ("ops", C.c_void_p),
                ("exc_idx", C.c_void_p), ("exc_val", C.c_void_p), ("n_exc", C.c_int64), ("op_off", C.c_void_p),
                ("pos", C.c_void_p), ("mapq", C.c_void_p), ("flag", C.c_void_p), ("qlen", C.c_void_p), ("name_id", C.c_void_p),
                ("next_same_name", C.c_void_p), ("n_sa_read", C.c_int64), ("n_sa_ent", C.c_int64), ("sa_rec", C.c_void_p),
                ("sa_off", C.c_void_p), ("sa_chrom", C.c_void_p), ("sa_pos", C.c_void_p), ("sa_strand", C.c_void_p),
                ("sa_first_s", C.c_void_p), ("sa_last_s", C.c_void_p), ("sa_bias", C.c_void_p), ("sa_mapq", C.c_void_p),
                ("sa_self_chrom", C.c_int32)]


class ShardCounts(C.Structure):
    """struct vsv_shard_counts."""
    _fields_ = [(k, C.c_int64) for k in ("n_reads_cigar", "n_reads_split", "n_cute", "n_del", "n_ins", "n_sa_rows", "n_del_dup",
                                         "n_ins_maybe_dup", "n_udel", "n_uins", "overflow", "n_stage")]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


def default_params(dtype: str = "Hifi") -> Params:
    p = Params()
    rc = load_library().vsv_params_default(C.byref(p), DTYPE_CODE[dtype])
    if rc != 0:
        raise VsvError(rc, "vsv_params_default")
    return p


def next_same_name(name_id) -> np.ndarray:
    """next_same_name column of vsv_records: for every record the next record in BAM order with the same
    interned qname, -1 if there is none (the consecutive pairs extract_sig_from_split_reads walks, ERS:227-237)."""
    name_id = np.asarray(name_id)
    n = len(name_id)
    out = np.full(n, -1, np.int32)
    if n < 2:
        return out
    o = np.argsort(name_id, kind="stable")
    same = name_id[o[1:]] == name_id[o[:-1]]
    out[o[:-1][same]] = o[1:][same]
    return out


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
    "vsv_span_depth_noalloc": [_vp, _vp, _vp, _i64, _vp, _vp, _i64, _vp],
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
    "vsv_span_depth_sorted": [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _i64, _vp],
    "vsv_params_default": [_vp, _int],
    "vsv_ctx_set_params": [_vp, _vp],
    "vsv_ctx_get_params": [_vp, _vp],
    "vsv_cigar_class_sums": [_vp, _vp, _vp, _i64, _vp, _vp, _vp],
    "vsv_shard_create": [_vp, C.POINTER(_vp)],
    "vsv_shard_destroy": [_vp],
    "vsv_shard_reserve": [_vp, _i64, _i64, _i64, _i64],
    "vsv_shard_load": [_vp, _vp],
    "vsv_shard_set_bed": [_vp, _vp, _vp, _i64, _i64],
    "vsv_shard_scan": [_vp],
    "vsv_shard_counts_get": [_vp, _vp],
    "vsv_shard_drop_ins": [_vp, _vp, _i64],
    "vsv_shard_fetch": [_vp, _int, _vp, _i64, C.POINTER(_i64)],
    "vsv_shard_candidates": [_vp, _int, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "vsv_shard_scan_kernel_ms": [_vp, C.POINTER(C.c_float)],
}
_RESTYPE = {"vsv_ctx_destroy": None, "vsv_shard_destroy": None, "vsv_last_error": C.c_char_p, "vsv_version": C.c_char_p, "vsv_bam_close": None,
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
        # at interpreter shutdown the CUDA runtime may already be unloading: leave the handle to the OS
        try:
            if not sys.is_finalizing():
                self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise VsvError(rc, (self.lib.vsv_last_error(self.h) or b"").decode())

    def set_params(self, params: "Params"):
        self._chk(self.lib.vsv_ctx_set_params(self.h, C.byref(params)))

    def get_params(self) -> "Params":
        p = Params()
        self._chk(self.lib.vsv_ctx_get_params(self.h, C.byref(p)))
        return p

    def cigar_class_sums(self, ops, op_off):
        """(sum of I lengths, sum of M lengths, number of M ops) per record (ECS_C:53-70)."""
        ops, op_off = _arr(ops, np.uint32), _arr(op_off, np.int64)
        n = len(op_off) - 1
        out = [np.zeros(n, np.int64) for _ in range(3)]
        self._chk(self.lib.vsv_cigar_class_sums(self.h, _p(ops), _p(op_off), n, _p(out[0]), _p(out[1]), _p(out[2])))
        return tuple(out)

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


class Shard:
    """One vsv_shard on a HOST-IO context: numpy in, numpy out (the operators' path).  The device-resident
    driver for benchmarks / queues of shards is volcanosv_b200.pipeline.ChromShard."""

    def __init__(self, ctx: Context):
        self.ctx, self.lib = ctx, ctx.lib
        self.h = _vp()
        ctx._chk(self.lib.vsv_shard_create(ctx.h, C.byref(self.h)))
        self._keep = []
        self.counts = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.vsv_shard_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            if not sys.is_finalizing():
                self.close()
        except Exception:  # noqa: BLE001
            pass

    def reserve(self, cap_reads=0, cap_cute=0, cap_del=0, cap_ins=0):
        self.ctx._chk(self.lib.vsv_shard_reserve(self.h, int(cap_reads), int(cap_cute), int(cap_del), int(cap_ins)))

    def load(self, ops, op_off, pos, mapq, qlen, flag=None, name_id=None, next_name=None, sa=None, self_chrom=0):
        """sa: dict(rec, off, chrom, pos, strand, first_s, last_s, bias, mapq) or None."""
        ops = _arr(ops, np.uint32)
        if len(ops) and ops.ctypes.data % 16:
            ops = _aligned_copy(ops)
        cols = dict(ops=ops, op_off=_arr(op_off, np.int64), pos=_arr(pos, np.int32), mapq=_arr(mapq, np.uint8),
                    qlen=_arr(qlen, np.int32), flag=_arr(flag, np.uint16), name=_arr(name_id, np.int32),
                    nxt=_arr(next_name, np.int32))
        r = RecordsStruct()
        r.n_rec, r.n_ops, r.ops_format = len(cols["pos"]), len(ops), 32
        ptr = lambda a: None if a is None else a.ctypes.data
        r.ops, r.op_off, r.pos, r.mapq, r.flag, r.qlen = (ptr(cols[k]) for k in ("ops", "op_off", "pos", "mapq", "flag", "qlen"))
        r.name_id, r.next_same_name = ptr(cols["name"]), ptr(cols["nxt"])
        if sa is not None and len(sa["rec"]):
            sc = dict(rec=_arr(sa["rec"], np.int32), off=_arr(sa["off"], np.int64), chrom=_arr(sa["chrom"], np.int32),
                      pos=_arr(sa["pos"], np.int32), strand=_arr(sa["strand"], np.uint8), first_s=_arr(sa["first_s"], np.int32),
                      last_s=_arr(sa["last_s"], np.int32), bias=_arr(sa["bias"], np.int32), mapq=_arr(sa["mapq"], np.uint8))
            cols.update({"sa_" + k: v for k, v in sc.items()})
            r.n_sa_read, r.n_sa_ent = len(sc["rec"]), len(sc["pos"])
            r.sa_rec, r.sa_off, r.sa_chrom, r.sa_pos, r.sa_strand = (ptr(sc[k]) for k in ("rec", "off", "chrom", "pos", "strand"))
            r.sa_first_s, r.sa_last_s, r.sa_bias, r.sa_mapq = (ptr(sc[k]) for k in ("first_s", "last_s", "bias", "mapq"))
        r.sa_self_chrom = int(self_chrom)
        self._keep = [cols]
        self.n_rec = int(r.n_rec)
        self.ctx._chk(self.lib.vsv_shard_load(self.h, C.byref(r)))

    def set_bed(self, regions, chrom_len):
        """regions: list of (start - 1000, end + 1000) as load_bed widens them (SE:69), or None."""
        if regions is None:
            self.ctx._chk(self.lib.vsv_shard_set_bed(self.h, None, None, -1, int(chrom_len)))
            return
        b0 = np.asarray([x[0] for x in regions], np.int32)
        b1 = np.asarray([x[1] for x in regions], np.int32)
        self.ctx._chk(self.lib.vsv_shard_set_bed(self.h, _p(b0) if len(b0) else None, _p(b1) if len(b1) else None, len(b0), int(chrom_len)))

    def scan(self, max_tries: int = 4):
        """Runs the step; grows a table that overflowed and retries.  Returns the counts dict."""
        for _ in range(max_tries):
            self.ctx._chk(self.lib.vsv_shard_scan(self.h))
            c = ShardCounts()
            self.ctx._chk(self.lib.vsv_shard_counts_get(self.h, C.byref(c)))
            d = c.as_dict()
            ov = d["overflow"]
            if ov & OVF_POS:
                raise VsvError(E_ARG, "a position is negative or >= 2^28: the chromosome is too long for one shard")
            if ov & OVF_DUPLIST:
                raise VsvError(E_CAPACITY, "more than 65536 potential duplicate INS rows")
            if ov == 0:
                self.counts = d
                return d
            grow = lambda n: int(n * 1.25) + 1024
            self.reserve(grow(d["n_reads_cigar"] + d["n_reads_split"]) if ov & OVF_READS else 0, grow(d["n_cute"]) if ov & OVF_CUTE else 0,
                         grow(d["n_del"]) if ov & OVF_DEL else 0, grow(d["n_ins"]) if ov & OVF_INS else 0)
        raise VsvError(E_CAPACITY, "vsv_shard_scan kept overflowing")

    def drop_ins(self, uids):
        u = np.sort(np.asarray(uids, np.int32))
        self.ctx._chk(self.lib.vsv_shard_drop_ins(self.h, _p(u) if len(u) else None, len(u)))
        c = ShardCounts()
        self.ctx._chk(self.lib.vsv_shard_counts_get(self.h, C.byref(c)))
        self.counts = c.as_dict()
        return self.counts

    def fetch(self, table):
        c = self.counts
        if table in (TAB_READS, TAB_CUTE, TAB_DEL_ROWS, TAB_INS_ROWS):
            n = {TAB_READS: c["n_reads_cigar"] + c["n_reads_split"], TAB_CUTE: c["n_cute"], TAB_DEL_ROWS: c["n_del"],
                 TAB_INS_ROWS: c["n_ins"]}[table]
            out = np.empty(n, SIG_DTYPE)
        elif table in (TAB_REF_END, TAB_READ_LEN):
            n, out = self.n_rec, np.empty(self.n_rec, np.int32)
        elif table == TAB_IN_BED:
            n, out = self.n_rec, np.empty(self.n_rec, np.uint8)
        elif table == TAB_INS_DUPS:
            n = min(c["n_ins_maybe_dup"], 65536)
            out = np.empty((n, 2), np.int32)
        else:
            n = c["n_udel"] if table == TAB_UDEL else c["n_uins"]
            out = np.empty((3, max(n, 1)), np.int32)
            got = _i64()
            self.ctx._chk(self.lib.vsv_shard_fetch(self.h, table, _p(out), max(n, 1), C.byref(got)))
            return out[0, :got.value].copy(), out[1, :got.value].copy(), out[2, :got.value].copy()
        got = _i64()
        self.ctx._chk(self.lib.vsv_shard_fetch(self.h, table, _p(out) if n else None, n, C.byref(got)))
        return out[: got.value]

    def candidates(self, what, pos, svlen, typ, gt=None, last_in=(0, 0), n_support=None):
        """-> dict with the outputs of the selected stages (+ last_out for JOIN_SUPPORT)."""
        cp, cl, ct = (_arr(x, np.int32) for x in (pos, svlen, typ))
        cg = _arr(gt, np.int32)
        n = len(cp)
        li, lo = np.asarray(last_in, np.int32), np.zeros(2, np.int32)
        o = dict(fp_support=np.zeros(n, np.int32) if what & JOIN_FP else None, cov=np.zeros(n, np.int64) if what & JOIN_COV else None,
                 n_support=(np.zeros(n, np.int32) if what & JOIN_SUPPORT else _arr(n_support, np.int32)) if what & (JOIN_SUPPORT | JOIN_GENOTYPE) else None,
                 match_id=np.zeros(n, np.int32) if what & JOIN_SUPPORT else None,
                 depth2=np.zeros(n, np.int32) if what & JOIN_GENOTYPE else None, ratio=np.zeros(n, np.float64) if what & JOIN_GENOTYPE else None,
                 new_gt=np.zeros(n, np.int32) if what & JOIN_GENOTYPE else None)
        self.ctx._chk(self.lib.vsv_shard_candidates(self.h, what, _p(cp), _p(cl), _p(ct), _p(cg), n, _p(li), _p(o["fp_support"]), _p(o["cov"]),
                                                   _p(o["n_support"]), _p(o["match_id"]), _p(o["depth2"]), _p(o["ratio"]), _p(o["new_gt"]),
                                                   _p(lo) if what & JOIN_SUPPORT else None))
        o["last_out"] = (int(lo[0]), int(lo[1]))
        return o


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
