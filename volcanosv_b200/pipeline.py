"""Device-resident large-indel hot path for one chromosome shard.

`ChromShard` drives ONE `vsv_shard` (include/volcanosv_b200.h, csrc/vsv_shard.cu) on a device-IO
context: the packed records are uploaded once, `run()` enqueues the whole read-side chain of the
reference (RVC:83-96 ERS -> FPF; FGC:111-172 SE -> CSS -> GTD/GTI) on one stream without a single
host synchronisation, and the counts / per-candidate results are read once at the end.  The
operators (volcanosv_b200.operators) format the reference's files from the same engine, so the
benchmarked path and the byte-exact path are one call chain.
torch is used only for device / pinned allocations and the H2D / D2H copies.
"""
from __future__ import annotations

import contextlib
import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np
import torch

from . import lib as L

NAN = float("nan")
# LI/para/GT_correction_para_<dtype>_<type>.txt (the library holds the same numbers: vsv_params_default)
GT_THRESHOLDS = {
    "Hifi": dict(DEL=[0.6, 0.69, 0.71, 0.91], INS=[NAN, 0.59, 0.63, 0.79]),
    "ONT": dict(DEL=[0.61, 0.61, 0.68, 0.79], INS=[NAN, NAN, 0.67, 0.72]),
    "CLR": dict(DEL=[0.55, 0.59, 0.65, 0.75], INS=[NAN, NAN, 0.64, 0.75]),
}
_REC_COLS = ("op_off", "pos", "mapq", "qlen", "flag", "name_id", "next_same_name")
_SA_COLS = ("sa_rec", "sa_off", "sa_chrom", "sa_pos", "sa_strand", "sa_first_s", "sa_last_s", "sa_bias", "sa_mapq")
_COL_DTYPE = dict(op_off=np.int64, pos=np.int32, mapq=np.uint8, qlen=np.int32, flag=np.uint16, name_id=np.int32,
                  next_same_name=np.int32, sa_rec=np.int32, sa_off=np.int64, sa_chrom=np.int32, sa_pos=np.int32,
                  sa_strand=np.uint8, sa_first_s=np.int32, sa_last_s=np.int32, sa_bias=np.int32, sa_mapq=np.uint8)


def _dp(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _as_tensor(src, name=None):
    if isinstance(src, np.ndarray):
        if name in _COL_DTYPE and src.dtype != _COL_DTYPE[name]:
            src = src.astype(_COL_DTYPE[name])
        if src.dtype == np.uint32:
            src = src.view(np.int32)
        elif src.dtype == np.uint16:
            src = src.view(np.int16)
        src = torch.from_numpy(np.ascontiguousarray(src))
    elif src.dtype == torch.uint16:
        src = src.view(torch.int16)
    elif src.dtype == torch.uint32:
        src = src.view(torch.int32)
    return src


@dataclass
class ShardResult:
    n_rec: int = 0
    n_ops: int = 0
    n_cand: int = 0
    counts: Dict[str, int] = field(default_factory=dict)
    scan_ms: list = field(default_factory=list)      # dominant kernel, per scan call
    out: Dict[str, torch.Tensor] = field(default_factory=dict)

    # names the round-1 callers used
    @property
    def n_read_sigs(self):
        return self.counts.get("n_reads_cigar", 0) + self.counts.get("n_reads_split", 0)

    @property
    def n_cute_sigs(self):
        return self.counts.get("n_cute", 0)

    @property
    def n_udel(self):
        return self.counts.get("n_udel", 0)

    @property
    def n_uins(self):
        return self.counts.get("n_uins", 0)


class ChromShard:
    """Buffers and call chain for one (chromosome, GPU) shard."""

    def __init__(self, device: int = 0, dtype: str = "ONT", params: Optional[L.Params] = None):
        if not torch.cuda.is_available():
            raise L.VsvError(L.E_CUDA, "no CUDA device: volcanosv_b200 has no CPU path")
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(self.device)
        self.ctx = L.Context(device)
        self.ctx.set_params(params if params is not None else L.default_params(dtype))
        self.ctx.set_device_io(True)
        self.lib = self.ctx.lib
        self.h = self.ctx.h
        self.dtype = dtype
        self.sh = C.c_void_p()
        self._chk(self.lib.vsv_shard_create(self.h, C.byref(self.sh)))
        self.d: Dict[str, torch.Tensor] = {}
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self._pending = None
        self.n_rec = self.n_ops = self.n_cand = 0
        self.caps = None            # (reads, cute, del, ins) once a run has measured the need

    def close(self):
        if getattr(self, "sh", None):
            self.lib.vsv_shard_destroy(self.sh)
            self.sh = None

    def __del__(self):
        try:
            import sys
            if not sys.is_finalizing():
                self.close()
        except Exception:  # noqa: BLE001
            pass

    def _chk(self, rc):
        if rc != 0:
            raise L.VsvError(rc, (self.lib.vsv_last_error(self.h) or b"").decode())

    # ---- allocation ------------------------------------------------------------------
    def _buf(self, name, n, dtype):
        t = self.d.get(name)
        if t is None or t.numel() < n or t.dtype != dtype:
            t = torch.empty(max(int(n), 1), dtype=dtype, device=self.device)
            self.d[name] = t
        return t

    def upload(self, batch, cands=None, n_ops: Optional[int] = None, stream: Optional["torch.cuda.Stream"] = None):
        """H2D of one shard's inputs (packed records + optional candidate table).  Values are numpy arrays or
        (pinned) torch CPU tensors.  `ops` (BAM-native u32) or one of the staging formats `ops16` / `ops12` /
        `ops8` with `exc_idx` / `exc_val`.  With `stream` the copies are only enqueued there (pinned sources)
        and `finish_upload()` must be called before `run()`: that is how `ShardStream` overlaps the upload of
        the next shard with the kernels of the current one."""
        self.h2d_bytes = 0
        self._pending = None
        rec = L.RecordsStruct()

        def put(name, src):
            src = _as_tensor(src, name)
            dst = self._buf("in_" + name, src.numel(), src.dtype)
            dst[: src.numel()].copy_(src, non_blocking=True)
            self.h2d_bytes += src.numel() * src.element_size()
            return dst

        with torch.cuda.stream(stream if stream is not None else torch.cuda.current_stream()):
            fmt = next((f for f in (12, 16, 8) if "ops%d" % f in batch), 32)
            key = "ops" if fmt == 32 else "ops%d" % fmt
            src = batch[key]
            if isinstance(src, np.ndarray) and fmt == 12:
                src = src.view(np.uint8)
            t_ops = put(key, src)
            if n_ops is None:
                n_ops = int(batch["op_off"][-1])
            if fmt == 12 and _as_tensor(src).numel() != L.ops12_bytes(n_ops):
                raise ValueError("ops12: %d bytes for %d ops" % (_as_tensor(src).numel(), n_ops))
            rec.ops, rec.ops_format, rec.n_ops = t_ops.data_ptr(), fmt, int(n_ops)
            if fmt != 32 and len(batch["exc_idx"]):
                rec.exc_idx, rec.exc_val = put("exc_idx", batch["exc_idx"]).data_ptr(), put("exc_val", batch["exc_val"]).data_ptr()
                rec.n_exc = len(batch["exc_idx"])
            for k in _REC_COLS:
                if k in batch and batch[k] is not None:
                    setattr(rec, k, put(k, batch[k]).data_ptr())
            n_rec = len(batch["pos"])
            rec.n_rec = n_rec
            if "sa_rec" in batch and len(batch["sa_rec"]):
                for k in _SA_COLS:
                    setattr(rec, k, put(k, batch[k]).data_ptr())
                rec.n_sa_read, rec.n_sa_ent = len(batch["sa_rec"]), len(batch["sa_pos"])
                rec.sa_self_chrom = int(batch.get("sa_self_chrom", 0))
            if cands is not None:
                for k in ("pos", "svlen", "type", "gt"):
                    put("cand_" + k, cands[k])
                self.n_cand = len(cands["pos"])
            done = torch.cuda.Event()
            done.record()
        self.n_rec, self.n_ops = n_rec, int(n_ops)
        self._pending = (done, rec)
        if stream is None:
            self.finish_upload()

    def adopt(self, dev_batch, cands=None, n_ops: Optional[int] = None):
        """Uses record columns that are ALREADY on this GPU (torch tensors, e.g. slices of a larger resident batch) without
        copying them: the engine adopts the device pointers (device-IO ctx).  BAM-native u32 ops only."""
        rec = L.RecordsStruct()
        self._pending = None
        self._adopted = dict(dev_batch)          # keeps the tensors alive
        t_ops = _as_tensor(dev_batch["ops"])
        n_rec = int(dev_batch["pos"].numel())
        n_ops = int(n_ops if n_ops is not None else t_ops.numel())
        rec.ops, rec.ops_format, rec.n_ops, rec.n_rec = t_ops.data_ptr(), 32, n_ops, n_rec
        for k in _REC_COLS:
            if dev_batch.get(k) is not None:
                setattr(rec, k, _as_tensor(dev_batch[k]).data_ptr())
        if dev_batch.get("sa_rec") is not None and dev_batch["sa_rec"].numel():
            for k in _SA_COLS:
                setattr(rec, k, _as_tensor(dev_batch[k]).data_ptr())
            rec.n_sa_read, rec.n_sa_ent = int(dev_batch["sa_rec"].numel()), int(dev_batch["sa_pos"].numel())
            rec.sa_self_chrom = int(dev_batch.get("sa_self_chrom", 0))
        if cands is not None:
            for k in ("pos", "svlen", "type", "gt"):
                src = _as_tensor(cands[k])
                self._buf("in_cand_" + k, src.numel(), src.dtype)[: src.numel()].copy_(src, non_blocking=True)
            self.n_cand = len(cands["pos"])
        torch.cuda.current_stream().synchronize()
        self.n_rec, self.n_ops = n_rec, n_ops
        self.h2d_bytes = 0
        self._chk(self.lib.vsv_shard_load(self.sh, C.byref(rec)))

    def set_candidates(self, cands):
        """Replaces the resident candidate table (pos, svlen, type, gt)."""
        for k in ("pos", "svlen", "type", "gt"):
            src = _as_tensor(cands[k])
            self._buf("in_cand_" + k, src.numel(), src.dtype)[: src.numel()].copy_(src, non_blocking=True)
        self.n_cand = len(cands["pos"])
        torch.cuda.current_stream().synchronize()

    def finish_upload(self):
        """Waits for the enqueued copies and hands the device columns to the engine (widening a staging format)."""
        if self._pending is None:
            return
        done, rec = self._pending
        self._pending = None
        done.synchronize()
        self._chk(self.lib.vsv_shard_load(self.sh, C.byref(rec)))

    # ---- the hot path ----------------------------------------------------------------
    def run(self, what: int = L.JOIN_FP | L.JOIN_COV | L.JOIN_SUPPORT | L.JOIN_GENOTYPE) -> ShardResult:
        """Enqueues scan + tables + (if a candidate table is resident) the per-candidate joins; nothing is
        waited for.  `counts()` / `download()` synchronise."""
        res = ShardResult(n_rec=self.n_rec, n_ops=self.n_ops, n_cand=self.n_cand)
        self._chk(self.lib.vsv_shard_scan(self.sh))
        n = self.n_cand
        if n and what:
            i32, i64, f64 = torch.int32, torch.int64, torch.float64
            o = dict(fp_support=self._buf("fp_support", n, i32), cov=self._buf("cov", n, i64), n_support=self._buf("n_support", n, i32),
                     match_id=self._buf("match_id", n, i32), depth2=self._buf("depth2", n, i32), ratio=self._buf("ratio", n, f64),
                     new_gt=self._buf("new_gt", n, i32))
            d = self.d
            self._chk(self.lib.vsv_shard_candidates(self.sh, what, _dp(d["in_cand_pos"]), _dp(d["in_cand_svlen"]), _dp(d["in_cand_type"]),
                                                    _dp(d["in_cand_gt"]), n, None, _dp(o["fp_support"]), _dp(o["cov"]), _dp(o["n_support"]),
                                                    _dp(o["match_id"]), _dp(o["depth2"]), _dp(o["ratio"]), _dp(o["new_gt"]), None))
            res.out = {k: v[:n] for k, v in o.items()}
        return res

    def counts(self) -> Dict[str, int]:
        c = L.ShardCounts()
        self._chk(self.lib.vsv_shard_counts_get(self.sh, C.byref(c)))
        return c.as_dict()

    def run_safe(self, what: int = 15) -> ShardResult:
        """run() + counts; a table that overflowed is grown and the step repeated; capacities far above the
        measured need are tightened for the next run (the sort works on capacity-sized arrays)."""
        for _ in range(5):
            res = self.run(what)
            c = self.counts()
            res.counts = c
            res.scan_ms.append(self.ctx.main_kernel_ms())
            ov = c["overflow"]
            need = (c["n_reads_cigar"] + c["n_reads_split"], c["n_cute"], c["n_del"], c["n_ins"])
            if ov & L.OVF_POS:
                raise L.VsvError(L.E_ARG, "a position is negative or >= 2^28: the chromosome is too long for one shard")
            if ov & L.OVF_DUPLIST:
                raise L.VsvError(L.E_CAPACITY, "more than 65536 potential duplicate INS rows")
            tight = tuple(int(x * 1.25) + 4096 for x in need)
            if ov == 0:
                if self.caps is None or any(a > 2 * b for a, b in zip(self.caps, tight)):
                    self.caps = tight
                    self._chk(self.lib.vsv_shard_reserve(self.sh, *tight))
                return res
            self.caps = tight
            self._chk(self.lib.vsv_shard_reserve(self.sh, *tight))
        raise L.VsvError(L.E_CAPACITY, "signature tables kept overflowing")

    def sync(self):
        self.ctx.sync()
        torch.cuda.current_stream().synchronize()

    def fetch(self, table):
        """A table of the engine as numpy (synchronises).  See lib.TAB_*."""
        self.ctx.sync()
        sh = L.Shard.__new__(L.Shard)          # borrow the fetch logic of the host-IO wrapper (fetch always copies to the host)
        sh.ctx, sh.lib, sh.h, sh.counts, sh.n_rec = self.ctx, self.lib, self.sh, self.counts(), self.n_rec
        try:
            return L.Shard.fetch(sh, table)
        finally:
            sh.h = None

    def download(self, res: ShardResult, pinned: Optional[Dict[str, torch.Tensor]] = None,
                 stream: Optional["torch.cuda.Stream"] = None) -> Dict[str, np.ndarray]:
        """D2H of the per-candidate results (what the VCF / TSV writers print)."""
        self.ctx.sync()
        out = {}
        self.d2h_bytes = 0
        ctxm = torch.cuda.stream(stream) if stream is not None else contextlib.nullcontext()
        with ctxm:
            for k, t in res.out.items():
                if pinned is not None:
                    dst = pinned.get(k)
                    if dst is None or dst.numel() < t.numel() or dst.dtype != t.dtype:
                        dst = torch.empty(max(t.numel(), 1), dtype=t.dtype, pin_memory=True)
                        pinned[k] = dst
                    dst[: t.numel()].copy_(t, non_blocking=True)
                    out[k] = dst[: t.numel()]
                else:
                    out[k] = t.cpu()
                self.d2h_bytes += t.numel() * t.element_size()
            torch.cuda.current_stream().synchronize()
        return {k: v.numpy() for k, v in out.items()}


def run_many(shards, what: int = 15):
    """Several resident shards in flight on one GPU: every shard's step is enqueued on its own stream (each engine has its
    own context) before any count is read, so the small latency-bound kernels of one chromosome fill the gaps of another -
    the steady state of the whole-genome mode, where a GPU owns 3 - 22 chromosomes.  A shard whose tables overflowed is
    repeated on its own (`run_safe`).  Returns the ShardResults in order."""
    results = [sh.run(what) for sh in shards]
    for i, (sh, res) in enumerate(zip(shards, results)):
        res.counts = sh.counts()
        res.scan_ms.append(sh.ctx.main_kernel_ms())
        if res.counts["overflow"]:
            results[i] = sh.run_safe(what)
    return results


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
