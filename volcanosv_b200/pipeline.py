"""Device-resident large-indel hot path for one chromosome shard.

`ChromShard.run()` chains the C-ABI calls of include/volcanosv_b200.h with every
intermediate staying in HBM (the ctx is in device-IO mode): read scan (ERS), cuteSV-style
scan + in-read merge (SE), `.sigs` ordering and grouping, FP-filter support, coverage
support, supporting-signature count, spanning depth and the genotype decision.
torch is used only for device / pinned allocations and the H2D / D2H copies.

Reference call chain reproduced: LI/Raw_variant_call.py:83-96 (ERS -> FPF) and
LI/filter_GT_correction.py:111-172 (SE -> CSS -> GTD/GTI).
"""
from __future__ import annotations

import contextlib
import ctypes as C
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np
import torch

from . import lib as L

NAN = float("nan")
# LI/para/GT_correction_para_<dtype>_<type>.txt : t_large_11, t_small_11, t_large_01, t_small_01
GT_THRESHOLDS = {
    "Hifi": dict(DEL=[0.6, 0.69, 0.71, 0.91], INS=[NAN, 0.59, 0.63, 0.79]),
    "ONT": dict(DEL=[0.61, 0.61, 0.68, 0.79], INS=[NAN, NAN, 0.67, 0.72]),
    "CLR": dict(DEL=[0.55, 0.59, 0.65, 0.75], INS=[NAN, NAN, 0.64, 0.75]),
}


def _dp(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


@dataclass
class ShardResult:
    n_rec: int = 0
    n_ops: int = 0
    n_read_sigs: int = 0
    n_cute_sigs: int = 0
    n_del: int = 0
    n_ins: int = 0
    n_udel: int = 0
    n_uins: int = 0
    n_cand: int = 0
    scan_ms: list = field(default_factory=list)      # dominant kernel, per scan call
    out: Dict[str, torch.Tensor] = field(default_factory=dict)


class ChromShard:
    """Buffers and call chain for one (chromosome, GPU) shard."""

    def __init__(self, device: int = 0, dtype: str = "ONT"):
        if not torch.cuda.is_available():
            raise L.VsvError(L.E_CUDA, "no CUDA device: volcanosv_b200 has no CPU path")
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(self.device)
        self.ctx = L.Context(device)
        self.ctx.set_device_io(True)
        self.lib = self.ctx.lib
        self.h = self.ctx.h
        self.dtype = dtype
        self.thr_del = (C.c_double * 4)(*GT_THRESHOLDS[dtype]["DEL"])
        self.thr_ins = (C.c_double * 4)(*GT_THRESHOLDS[dtype]["INS"])
        self.d: Dict[str, torch.Tensor] = {}
        self.sig_cap = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self._pending = None
        # independent branches of the call chain (read-signature sort; DEL and INS `.sigs` order +
        # grouping) run on side contexts = their own streams, driven by two worker threads (the C
        # calls release the GIL): their kernels and count read-backs overlap instead of queueing
        self.side = [L.Context(device), L.Context(device)]
        for c in self.side:
            c.set_device_io(True)
        self.pool = ThreadPoolExecutor(max_workers=2)

    # ---- allocation ------------------------------------------------------------------
    def _buf(self, name, n, dtype):
        t = self.d.get(name)
        if t is None or t.numel() < n or t.dtype != dtype:
            t = torch.empty(max(int(n), 1), dtype=dtype, device=self.device)
            self.d[name] = t
        return t

    def upload(self, batch, cands, n_ops: Optional[int] = None, stream: Optional["torch.cuda.Stream"] = None):
        """H2D of one shard's inputs (packed records + candidate table).  Values are numpy
        arrays or (pinned) torch CPU tensors; `ops` may be uint32 numpy or an int32 tensor.
        With `stream` the copies are only enqueued there (pinned sources) and `finish_upload()`
        must be called before `run()`: that is how `ShardStream` overlaps the upload of the
        next shard with the kernels of the current one."""
        self.h2d_bytes = 0
        self._pending = None

        def put(name, src):
            if isinstance(src, np.ndarray):
                if src.dtype == np.uint32:
                    src = src.view(np.int32)
                src = torch.from_numpy(np.ascontiguousarray(src))
            dst = self._buf(name, src.numel(), src.dtype)
            dst[: src.numel()].copy_(src, non_blocking=True)
            self.h2d_bytes += src.numel() * src.element_size()
            return src

        with torch.cuda.stream(stream if stream is not None else torch.cuda.current_stream()):
            expand = None
            if "ops12" in batch:
                # densest staging (vsv_expand_ops12): 12 bits per pair of ops + both words of the odd pairs
                src12 = batch["ops12"]
                if isinstance(src12, np.ndarray):
                    src12 = torch.from_numpy(np.ascontiguousarray(src12.view(np.uint8)))
                n_ops = int(n_ops if n_ops is not None else batch["op_off"][-1])
                if src12.numel() != L.ops12_bytes(n_ops):
                    raise ValueError("ops12: %d bytes for %d ops" % (src12.numel(), n_ops))
                o12 = self._buf("in_ops12", src12.numel(), src12.dtype)
                o12[: src12.numel()].copy_(src12, non_blocking=True)
                self.h2d_bytes += src12.numel() * src12.element_size()
                ei = put("in_exc_idx", batch["exc_idx"]) if len(batch["exc_idx"]) else None
                put("in_exc_val", batch["exc_val"]) if len(batch["exc_val"]) else None
                expand = (o12, 0 if ei is None else ei.numel(), self._buf("in_ops", n_ops, torch.int32), 12)
            elif "ops16" in batch or "ops8" in batch:
                # compact staging (include/volcanosv_b200.h: vsv_expand_ops16 / vsv_expand_ops8): 2 bytes or
                # 1 byte per op over PCIe, widened on the device into the BAM-native u32 array the scanners read
                wide = "ops16" in batch
                src16 = batch["ops16" if wide else "ops8"]
                if isinstance(src16, np.ndarray):
                    src16 = torch.from_numpy(np.ascontiguousarray(src16.view(np.int16 if wide else np.uint8)))
                n_ops = int(n_ops if n_ops is not None else src16.numel())
                o16 = self._buf("in_ops16" if wide else "in_ops8", src16.numel(), src16.dtype)
                o16[: src16.numel()].copy_(src16, non_blocking=True)
                self.h2d_bytes += src16.numel() * src16.element_size()
                ei = put("in_exc_idx", batch["exc_idx"]) if len(batch["exc_idx"]) else None
                put("in_exc_val", batch["exc_val"]) if len(batch["exc_val"]) else None
                expand = (o16, 0 if ei is None else ei.numel(), self._buf("in_ops", n_ops, torch.int32), wide)
            else:
                put("in_ops", batch["ops"])
            for k in ("op_off", "pos", "mapq", "qlen"):
                put("in_" + k, batch[k])
            for k in ("pos", "svlen", "type", "gt"):
                put("cand_" + k, cands[k])
            done = torch.cuda.Event()
            done.record()
        self.n_rec = len(batch["pos"])
        self.n_ops = int(n_ops) if n_ops is not None else int(batch["op_off"][-1])
        self.n_cand = len(cands["pos"])
        self._pending = (done, expand)
        if stream is None:
            self.finish_upload()

    def finish_upload(self):
        """Waits for the enqueued copies and widens a compact batch on the device."""
        if self._pending is None:
            return
        done, expand = self._pending
        self._pending = None
        done.synchronize()
        if expand is not None:
            o16, n_exc, ops, wide = expand
            fn = self.lib.vsv_expand_ops12 if wide == 12 else (self.lib.vsv_expand_ops16 if wide else self.lib.vsv_expand_ops8)
            self._chk(fn(self.h, _dp(o16), self.n_ops, _dp(self.d["in_exc_idx"]) if n_exc else None,
                    _dp(self.d["in_exc_val"]) if n_exc else None, n_exc, _dp(ops)))
            self.ctx.sync()

    def _chk(self, rc):
        if rc != 0:
            raise L.VsvError(rc, (self.lib.vsv_last_error(self.h) or b"").decode())

    # ---- the hot path ----------------------------------------------------------------
    def run(self) -> ShardResult:
        lib, h, d = self.lib, self.h, self.d
        n_rec, n_ops, n_cand = self.n_rec, self.n_ops, self.n_cand
        res = ShardResult(n_rec=n_rec, n_ops=n_ops, n_cand=n_cand)
        i32, u8, i64, f64 = torch.int32, torch.uint8, torch.int64, torch.float64
        ops, op_off, pos, mapq, qlen = (d["in_" + k] for k in ("ops", "op_off", "pos", "mapq", "qlen"))
        cap = max(self.sig_cap, 1 << 16, n_ops // 64)
        sig = lambda name: self._buf(name, cap * 8, i32)  # vsv_sig = 8 x int32
        rt, qt, xt = (self._buf(k, n_rec, i32) for k in ("ref_total", "qry_total", "aux_total"))
        # -- stage 1: BOTH read scanners in one pass over the ops -------------------------
        #    stream "reads": ERS:107-145 (mapq>=50, >=30 bp); stream "cute": SE:438-478
        #    (query_length>=500, mapq>=20, >=10 bp)
        enable = self._buf("enable", n_rec, u8)
        self._chk(lib.vsv_ge_u8(h, _dp(qlen), 500, n_rec, _dp(enable)))                 # SE:439
        s_reads, s_cute = sig("sig_reads"), sig("sig_cute")
        n_r_, n_c_ = C.c_int64(), C.c_int64()
        rc = lib.vsv_scan_reads_fused(h, 30, 50, 10, 20, _dp(ops), _dp(op_off), _dp(pos), _dp(mapq), _dp(enable), n_rec,
                                      _dp(s_reads), cap, C.byref(n_r_), _dp(s_cute), cap, C.byref(n_c_), _dp(rt), _dp(qt),
                                      _dp(xt))
        if rc == L.E_CAPACITY:
            self.sig_cap = int(max(n_r_.value, n_c_.value) * 1.25) + 1024
            raise _Regrow()
        self._chk(rc)
        res.scan_ms.append(self.ctx.main_kernel_ms())
        n_r, n_c = int(n_r_.value), int(n_c_.value)
        res.n_read_sigs, res.n_cute_sigs = n_r, n_c
        # read intervals for the depth join: [pos, pos + ref_total)  (reference_end)
        rend = self._buf("read_end", n_rec, i32)
        self._chk(lib.vsv_add_i32(h, _dp(pos), _dp(rt), n_rec, _dp(rend)))
        # every buffer of the branches is taken here, on this thread
        s_reads_sorted = sig("sig_reads_sorted")
        rs_pos, rs_len = self._buf("rs_pos", cap, i32), self._buf("rs_len", cap, i32)
        s_del, s_ins = sig("sig_del"), sig("sig_ins")
        bufs = {}
        for name in ("del", "ins"):
            bufs[name] = (sig("sig_%s_sorted" % name), self._buf(name + "_pos", cap, i32), self._buf(name + "_len", cap, i32),
                          tuple(self._buf("%s_u%s" % (name, k), cap, i32) for k in "cpln"))

        cpos, clen, ctyp, cgt = (d["cand_" + k] for k in ("pos", "svlen", "type", "gt"))
        fp = self._buf("fp_support", n_cand, i32)
        cend = self._buf("cand_end", n_cand, i32)
        self._chk(lib.vsv_add_i32(h, _dp(cpos), _dp(clen), n_cand, _dp(cend)))
        cov = dict(ins=self._buf("cov_ins", n_cand, i64), **{"del": self._buf("cov_del", n_cand, i64)})
        sup = dict(ins=self._buf("sup_ins", n_cand, i32), **{"del": self._buf("sup_del", n_cand, i32)})

        def chk(c, rc):
            if rc != 0:
                raise L.VsvError(rc, (lib.vsv_last_error(c.h) or b"").decode())

        def reads_branch(c):                                                              # merge_all (ERS:254-258)
            chk(c, lib.vsv_sort_sigs(c.h, _dp(s_reads), n_r, 0, _dp(s_reads_sorted)))
            chk(c, lib.vsv_sig_columns(c.h, _dp(s_reads_sorted), n_r, _dp(rs_pos), _dp(rs_len), None))
            chk(c, lib.vsv_fp_support(c.h, _dp(cpos), _dp(clen), n_cand, _dp(rs_pos), _dp(rs_len), n_r, _dp(fp)))  # FPF
            c.sync()

        def sigs_branch(c, name, src, n, min_len):
            # `.sigs` order (SE:637-638), unique signatures (GTD:65-86 / GTI:67-101)
            srt, p_, l_, (uc, up, ul, un) = bufs[name]
            chk(c, lib.vsv_sort_sigs(c.h, _dp(src), n, 1, _dp(srt)))
            chk(c, lib.vsv_sig_columns(c.h, _dp(srt), n, _dp(p_), _dp(l_), None))
            nu = C.c_int64()
            chk(c, lib.vsv_group_sigs(c.h, None, _dp(p_), _dp(l_), n, min_len, _dp(uc), _dp(up), _dp(ul), _dp(un), C.byref(nu)))
            # candidate-side joins of this table: coverage support (CSS:81 / CSS:138), supporting signatures (GTD:92 / GTI:107)
            if name == "ins":
                chk(c, lib.vsv_ins_cov(c.h, _dp(cpos), n_cand, _dp(p_), _dp(l_), n, 1000, _dp(cov[name])))
            else:
                chk(c, lib.vsv_del_cov(c.h, _dp(cpos), _dp(cend), n_cand, _dp(p_), _dp(l_), n, 1000, _dp(cov[name])))
            chk(c, lib.vsv_sig_support(c.h, None, _dp(cpos), _dp(clen), n_cand, None, _dp(up), _dp(ul), _dp(un), int(nu.value), 0,
                                       _dp(sup[name]), None, None))
            c.sync()
            return (p_, l_, n, up, ul, un, int(nu.value))

        self.ctx.sync()                                        # the scan's outputs are complete for the other streams
        f_reads = self.pool.submit(reads_branch, self.side[0])

        # -- in-read merge of the cuteSV-style signals (SE:373-435) -------------------------
        nd, ni = C.c_int64(), C.c_int64()
        self._chk(lib.vsv_merge_in_record(h, L.MODE_CUTE, _dp(s_cute), n_c, _dp(s_del), C.byref(nd), _dp(s_ins),
                                          C.byref(ni)))
        res.n_del, res.n_ins = int(nd.value), int(ni.value)
        self.ctx.sync()

        # -- stage 2: the DEL and INS tables are independent: one on a side stream, one here ---
        f_del = self.pool.submit(sigs_branch, self.side[1], "del", s_del, res.n_del, 0)
        uniq = {"ins": sigs_branch(self.ctx, "ins", s_ins, res.n_ins, 30)}
        uniq["del"] = f_del.result()
        f_reads.result()
        res.n_udel, res.n_uins = uniq["del"][6], uniq["ins"][6]

        # -- the joins ran inside the branches; combine per candidate type -----------------
        cov_i, cov_d, sup_i, sup_d = cov["ins"], cov["del"], sup["ins"], sup["del"]
        nsup = self._buf("n_support", n_cand, i32)
        self._chk(lib.vsv_select_i32(h, _dp(ctyp), L.DEL, _dp(sup_d), _dp(sup_i), n_cand, _dp(nsup)))

        # -- stage 3: spanning depth + genotype (GTD:140-170,291-318; GTI:188-196,258-283)
        d2, ratio, ngt = self._buf("depth2", n_cand, i32), self._buf("ratio", n_cand, f64), self._buf("new_gt", n_cand, i32)
        self._chk(lib.vsv_genotype(h, _dp(pos), _dp(rend), n_rec, _dp(cpos), _dp(clen), _dp(ctyp), _dp(cgt),
                                   _dp(nsup), n_cand, self.thr_del, self.thr_ins, _dp(d2), _dp(ratio), _dp(ngt)))
        res.out = dict(fp_support=fp[:n_cand], cov_ins=cov_i[:n_cand], cov_del=cov_d[:n_cand], n_support=nsup[:n_cand],
                       depth2=d2[:n_cand], ratio=ratio[:n_cand], new_gt=ngt[:n_cand],
                       sig_reads=s_reads_sorted[: n_r * 8], sig_del=self.d["sig_del_sorted"][: res.n_del * 8],
                       sig_ins=self.d["sig_ins_sorted"][: res.n_ins * 8])
        return res

    def run_safe(self) -> ShardResult:
        """run() with one retry if a signature buffer had to grow."""
        for _ in range(4):
            try:
                return self.run()
            except _Regrow:
                for k in list(self.d):
                    if k.startswith(("sig_", "rs_", "del_", "ins_")):
                        del self.d[k]
        raise L.VsvError(L.E_CAPACITY, "signature buffers kept growing")

    def sync(self):
        self.ctx.sync()
        torch.cuda.current_stream().synchronize()

    def download(self, res: ShardResult, pinned: Optional[Dict[str, torch.Tensor]] = None,
                 stream: Optional["torch.cuda.Stream"] = None) -> Dict[str, np.ndarray]:
        """D2H of the shard's results (per-candidate table + signature arrays)."""
        self.ctx.sync()
        out = {}
        self.d2h_bytes = 0
        ctxm = torch.cuda.stream(stream) if stream is not None else contextlib.nullcontext()
        with ctxm:
            self._download_into(res, pinned, out)
            torch.cuda.current_stream().synchronize()
        return {k: v.numpy() for k, v in out.items()}

    def _download_into(self, res, pinned, out):
        for k, t in res.out.items():
            if pinned is not None:
                dst = pinned.get(k)
                if dst is None or dst.numel() < t.numel():
                    dst = torch.empty(max(t.numel(), 1), dtype=t.dtype, pin_memory=True)
                    pinned[k] = dst
                dst[: t.numel()].copy_(t, non_blocking=True)
                out[k] = dst[: t.numel()]
            else:
                out[k] = t.cpu()
            self.d2h_bytes += t.numel() * t.element_size()


class _Regrow(Exception):
    pass


class ShardStream:
    """Two shard slots on one GPU: the H2D copies of shard k+1 (their own stream) run while shard
    k is expanded, scanned, joined, genotyped and read back.  A chromosome-sharded run is a
    queue of such shards per GPU (MAIN:268), so this is the steady state of the whole-genome mode."""

    def __init__(self, device: int = 0, dtype: str = "ONT"):
        self.slots = [ChromShard(device, dtype), ChromShard(device, dtype)]
        self.up = torch.cuda.Stream(device=self.slots[0].device)
        self.down = torch.cuda.Stream(device=self.slots[0].device)
        self.n_in = 0
        self.n_out = 0

    def submit(self, batch, cands, n_ops: Optional[int] = None):
        """Enqueues the upload of the next shard (pinned host buffers); returns at once."""
        if self.n_in - self.n_out >= 2:
            raise L.VsvError(L.E_ARG, "ShardStream: both slots are busy; collect() first")
        self.slots[self.n_in % 2].upload(batch, cands, n_ops=n_ops, stream=self.up)
        self.n_in += 1

    def collect(self, pinned: Optional[Dict[str, torch.Tensor]] = None):
        """Finishes the oldest submitted shard: hot path + D2H.  Returns (ShardResult, arrays)."""
        if self.n_out >= self.n_in:
            raise L.VsvError(L.E_ARG, "ShardStream: nothing submitted")
        sh = self.slots[self.n_out % 2]
        self.n_out += 1
        sh.finish_upload()
        res = sh.run_safe()
        return res, sh.download(res, pinned, stream=self.down)
