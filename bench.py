#!/usr/bin/env python
"""Benchmark of the large-indel hot path (BASELINE.json metric: alignment records/s, with
genotyped candidates/s beside it) on synthetic chr10 ONT-shaped alignments.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the whole hot path (ERS scan -> cuteSV-style scan + merge ->
.sigs order/grouping -> FP-filter / coverage / support joins -> spanning depth ->
genotype) over one chromosome-sized shard.  N>1 (torchrun): every rank owns one shard on
its own GPU (independent shards, no collective on the data path) -> weak scaling.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "alignment records/s"
UNIT = "records/s"


def make_workload(profile: str, scale: float, rank: int, n_cand: int):
    """chr10-shaped shard: records in 8 genomic windows (bounded generator memory)."""
    from volcanosv_b200 import synth
    n_total = max(64, int(synth.n_records(profile) * scale))
    parts, chunks = [], 8
    L = synth.CHR10_LEN
    for c in range(chunks):
        n = n_total // chunks
        # CIGAR alphabet as the aligners of each data type write it: minimap2 (ONT, CLR) uses M; pbmm2 (HiFi) =/X
        b = synth.synth_records(1000 * rank + 10 + c, n, profile, chrom_len=L // chunks, eqx=(profile == "hifi"))
        b["pos"] = (b["pos"].astype(np.int64) + c * (L // chunks)).astype(np.int32)
        parts.append(b)
    batch = {}
    off, tot = [np.zeros(1, np.int64)], 0
    for b in parts:
        off.append(b["op_off"][1:] + tot)
        tot += int(b["op_off"][-1])
    batch["op_off"] = np.concatenate(off)
    for k in ("ops", "pos", "mapq", "qlen", "flag", "name_id"):
        batch[k] = np.concatenate([b[k] for b in parts])
    pos, svlen, typ, gt = synth.synth_candidates(77 + rank, n_cand)
    cands = dict(pos=pos, svlen=svlen, type=typ, gt=gt)
    return batch, cands


class ClockSampler(threading.Thread):
    """nvidia-smi style clock / throttle-reason samples during the timed region."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.hd = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.hd, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.hd, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.hd)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.02)

    def result(self):
        s = sorted(self.samples)
        return dict(sm_mhz=(s[len(s) // 2] if s else None), sm_max_mhz=self.max_mhz, reasons=sorted(self.reasons),
                    samples=len(s))


def cpu_baseline(batch, cands, thr, target_s=12.0, threads=None):
    """Oracle C port (oracle/vsv_oracle_c.c) timed on a bounded prefix of the workload."""
    from oracle import c_oracle as CO
    threads = threads or CO.num_threads()
    n = len(batch["pos"])

    def prefix(m):
        e = int(batch["op_off"][m])
        b = dict(ops=batch["ops"][:e], op_off=batch["op_off"][: m + 1], pos=batch["pos"][:m], mapq=batch["mapq"][:m],
                 qlen=batch["qlen"][:m])
        hi = int(batch["pos"][m - 1])
        keep = cands["pos"] <= hi
        c = {k: np.ascontiguousarray(v[keep]) for k, v in cands.items()}
        return b, c

    m = max(64, n // 64)
    b, c = prefix(m)
    t0 = time.perf_counter()
    CO.shard_step(b, c, thr[0], thr[1], threads=threads)
    dt = time.perf_counter() - t0
    frac = min(1.0, (m / n) * target_s / max(dt, 1e-4))
    m2 = int(min(n, max(m, n * frac)))
    if m2 > m * 1.5:
        b, c = prefix(m2)
        t0 = time.perf_counter()
        CO.shard_step(b, c, thr[0], thr[1], threads=threads)
        dt = time.perf_counter() - t0
        m = m2
    return dict(value=m / dt, unit=UNIT, cores=threads, kind="port",
                sample="first %d of %d records (%.1f%% of the shard, %d candidates), oracle C port, %.2f s" %
                       (m, n, 100.0 * m / n, len(c["pos"]), dt)), m, dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--profile", default="ont", choices=["ont", "hifi", "clr"])
    ap.add_argument("--scale", type=float, default=1.0, help="fraction of the chr10 shard (tests use < 1)")
    ap.add_argument("--cands", type=int, default=3000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--staging", default="auto", choices=["auto", "ops12", "ops8"],
                    help="host staging format of the end-to-end leg (auto = the densest that fits the data)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dtype = {"ont": "ONT", "hifi": "Hifi", "clr": "CLR"}[args.profile]
    from volcanosv_b200.pipeline import GT_THRESHOLDS
    thr = (GT_THRESHOLDS[dtype]["DEL"], GT_THRESHOLDS[dtype]["INS"])
    workload = "chr10 %s %dx single-chromosome large-indel detection, synthetic (scale %.3g)" % (
        dtype, {"ont": 50, "hifi": 30, "clr": 89}[args.profile], args.scale)

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return
        batch, cands = make_workload(args.profile, args.scale, 0, args.cands)
        from oracle import c_oracle as CO
        base, m, _ = cpu_baseline(batch, cands, thr, target_s=8.0)
        times = []
        e = int(batch["op_off"][m])
        b = dict(ops=batch["ops"][:e], op_off=batch["op_off"][: m + 1], pos=batch["pos"][:m], mapq=batch["mapq"][:m],
                 qlen=batch["qlen"][:m])
        keep = cands["pos"] <= int(batch["pos"][m - 1])
        c = {k: np.ascontiguousarray(v[keep]) for k, v in cands.items()}
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            CO.shard_step(b, c, thr[0], thr[1], threads=base["cores"])
            if i >= args.warmup:
                times.append(time.perf_counter() - t0)
        dt = sum(times) / max(1, len(times))
        v = m / dt
        base["value"] = v
        print(json.dumps(dict(impl="reference", metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=args.steps,
                              warmup=args.warmup, ms_per_step=dt * 1e3, higher_is_better=True, scaling="weak",
                              vs_baseline=None, dtype="int32", data="synthetic",
                              config=dict(workload=workload, records_per_step=m, ops_per_step=e,
                                          note="reference is pure Python and cannot travel; timed: its C port "
                                               "(oracle/vsv_oracle_c.c) on all host threads, bounded sample per step"),
                              cpu_baseline=base,
                              e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))))
        return

    # ------------------------------------------------------------------ our arm
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (volcanosv_b200 has no CPU path)")
    torch.cuda.set_device(local)
    dist = None
    host_binding = None
    if world > 1:
        import torch.distributed as dist
        from volcanosv_b200.shard import bind_host_to_gpu
        if os.environ.get("VSV_NO_BIND") != "1":
            host_binding = bind_host_to_gpu(local)   # pinned staging buffers on the GPU's own socket
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from volcanosv_b200.pipeline import ShardStream

    batch, cands = make_workload(args.profile, args.scale, rank, args.cands)
    n_rec, n_ops = len(batch["pos"]), int(batch["op_off"][-1])
    pinned = {k: torch.from_numpy(np.ascontiguousarray(v.view(np.int32) if k == "ops" else v)).pin_memory()
              for k, v in batch.items() if k in ("ops", "op_off", "pos", "mapq", "qlen")}
    pinned_c = {k: torch.from_numpy(v).pin_memory() for k, v in cands.items()}
    stream = ShardStream(local, dtype)   # two shard slots: upload of shard k+1 overlaps the work on shard k
    shard = stream.slots[0]

    # host staging format of the end-to-end path: ops as 16-bit words + a sparse exception list
    # (vsv_expand_ops16); produced by the host packer once, like the BAM decode it follows
    from volcanosv_b200.lib import pack_ops16, pack_ops8, pack_ops12_native
    # densest staging format that fits the data: ops8 (1 B/op; noisy long reads: ~1 % exceptions) when fewer than
    # 3 % of the ops need the exception list, else ops16 (2 B/op; HiFi / =,X CIGARs).  The native BAM feeder
    # (vsv_bam_chrom_fill) writes either format straight from the BAM records, so packing is not an extra pass.
    # ops12 (12 bits per M + short-indel PAIR, vsv_expand_ops12) is denser still when the CIGARs alternate as minimap2's
    # long-read CIGARs do: taken when its total (tokens + 16 bytes per odd pair) is below ops8's.
    o8, exc_idx, exc_val = pack_ops8(batch["ops"])
    pinned16 = {k: v for k, v in pinned.items() if k != "ops"}
    bytes8 = o8.nbytes + 12 * len(exc_idx)
    o12 = None
    if args.staging in ("auto", "ops12") and len(exc_idx) < 0.03 * n_ops:
        o12, exc_pair, exc_val12 = pack_ops12_native(batch["ops"])   # vsv_pack_ops12 on the host threads
        if args.staging == "auto" and o12.nbytes + 16 * len(exc_pair) >= bytes8:
            o12 = None
    if o12 is not None:
        staging = "ops12"
        pinned16["ops12"] = torch.from_numpy(o12).pin_memory()
        exc_idx, exc_val = exc_pair, exc_val12
    elif len(exc_idx) < 0.03 * n_ops:
        staging = "ops8"
        pinned16["ops8"] = torch.from_numpy(o8).pin_memory()
    else:
        staging = "ops16"
        o16, exc_idx, exc_val = pack_ops16(batch["ops"])
        pinned16["ops16"] = torch.from_numpy(o16.view(np.int16)).pin_memory()
    del o8, o12
    pinned16["exc_idx"] = torch.from_numpy(exc_idx).pin_memory()
    pinned16["exc_val"] = torch.from_numpy(exc_val.view(np.int32)).pin_memory()

    def upload(compact=False):
        shard.upload(pinned16 if compact else pinned, pinned_c, n_ops=n_ops)

    def barrier():
        shard.sync()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    upload()
    res = None
    for _ in range(max(args.warmup, 3)):
        res = shard.run_safe()
    # ---- resident timing: CUDA events on the library's stream ------------------------
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    def launches():
        return sum(c.timing()[1] for c in [shard.ctx] + shard.side)

    l0 = launches()
    shard.ctx.timer_start()
    scan_ms = []
    for _ in range(args.steps):
        res = shard.run_safe()
        scan_ms += res.scan_ms
    ms = shard.ctx.timer_stop()
    l1 = launches()
    barrier()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    ms_per_step = ms_max / args.steps
    value = world * n_rec / (ms_per_step * 1e-3)

    # ---- end to end: pinned host buffers -> H2D -> hot path -> D2H of the results ------
    # A GPU works through a queue of shards (chromosomes, MAIN:268), so the steady state is timed:
    # every timed step enqueues one shard's H2D copies, runs the hot path of the previously
    # uploaded shard and reads its results back (K uploads + K runs + K downloads in K steps).
    pin_out = {}

    def e2e_run(host):
        stream.n_in = stream.n_out = 0
        stream.submit(host, pinned_c, n_ops=n_ops)
        stream.submit(host, pinned_c, n_ops=n_ops)
        stream.collect(pin_out)                       # warm BOTH slots (buffers, scratch arenas of every context)
        stream.collect(pin_out)                       # and the pinned output buffers
        stream.submit(host, pinned_c, n_ops=n_ops)
        barrier()                                     # one shard's upload is complete here (prologue, untimed)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            stream.submit(host, pinned_c, n_ops=n_ops)
            stream.collect(pin_out)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        sh = stream.slots[0]
        bytes_io = (int(sh.h2d_bytes), int(sh.d2h_bytes))
        stream.collect(pin_out)                       # drain
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / args.steps, bytes_io

    e2e_per_step, (h2d_b, d2h_b) = e2e_run(pinned16)
    e2e_value = world * n_rec / e2e_per_step
    e2e32_per_step, (h2d32_b, _) = e2e_run(pinned)   # BAM-native u32 ops, no compact staging
    sampler.stop_flag = True                          # clocks sampled through the resident AND the end-to-end timed regions
    sampler.join()

    # ---- resident throughput with two shards in flight (a GPU holds many chromosome shards in its
    # 180 GB; while one shard's scan owns the SMs the other's small join kernels fill the gaps).
    # Reported beside `value`, which stays the one-shard-at-a-time figure the roofline refers to.
    inflight2 = None
    try:
        import threading
        pair = stream.slots                            # both already hold the shard after the e2e runs
        for sh in pair:
            sh.finish_upload()
            sh.run_safe()
            sh.sync()
        k_each = max(1, args.steps // 2)
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()

        def work(sh):
            for _ in range(k_each):
                sh.run_safe()
            sh.sync()
        th = [threading.Thread(target=work, args=(sh,)) for sh in pair]
        for t_ in th:
            t_.start()
        for t_ in th:
            t_.join()
        torch.cuda.synchronize()
        ev1.record()
        ev1.synchronize()
        t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms2 = float(t.item()) / (2 * k_each)
        inflight2 = dict(value=world * n_rec / (ms2 * 1e-3), unit=UNIT, ms_per_step=ms2, steps=2 * k_each)
    except Exception as ex:  # noqa: BLE001
        inflight2 = dict(error=str(ex))

    # ---- genotyping stress point (candidates/s), resident ------------------------------
    geno = None
    try:
        geno = genotype_point(shard, local)
    except Exception as ex:  # noqa: BLE001
        geno = dict(error=str(ex))

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (k_scan_cigar) ---------------------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    n_sig_out = res.n_read_sigs + res.n_cute_sigs            # both streams of the fused pass
    alg_bytes = 4.0 * n_ops + 24.0 * n_rec + 32.0 * n_sig_out       # SURVEY.md §8d B1, per launch
    k_ms = float(np.mean(scan_ms)) if scan_ms else float("nan")
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    # dram bytes of that kernel from the committed `ncu --set full` capture (profiles/), per op -> per launch here
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "scan_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic = (tj["dram_bytes_read"] + tj["dram_bytes_write"]) / tj["n_ops"] * n_ops
        traffic_src = tj["source"]
    roof = dict(bound="hbm", kernel="k_scan_wt<FUSED>", achieved=achieved, peak=peak, unit="GB/s", frac=achieved / peak,
                traffic=traffic, traffic_source=traffic_src, peak_source=peak_src, alg_bytes_per_launch=alg_bytes, kernel_ms=k_ms,
                launches_per_step=1, share_of_step=k_ms / ms_per_step)
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                ms_per_step=ms_per_step, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="int32",
                data="synthetic",
                config=dict(workload=workload, records_per_gpu=n_rec, ops_per_gpu=n_ops, candidates_per_gpu=len(cands["pos"]),
                            read_sigs=res.n_read_sigs, cute_sigs=res.n_cute_sigs, unique_del=res.n_udel,
                            unique_ins=res.n_uins, parallelism="shard-per-gpu x%d, no collective" % world,
                            l2="inputs (%.2f GB/step) exceed the 126 MB L2; no flush needed" % (4.0 * n_ops / 1e9)),
                clocks=sampler.result(), gpu_launches=int(l1 - l0),
                e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=h2d_b,
                         d2h_bytes_per_step=d2h_b, ms_per_step=e2e_per_step * 1e3,
                         staging="%s + %d exceptions (vsv_expand_%s), pinned host buffers" % (staging, len(exc_idx), staging),
                         pipelining="double-buffered shards: H2D of shard k+1 overlaps kernels + D2H of shard k",
                         bam_native_u32=dict(value=world * n_rec / e2e32_per_step, ms_per_step=e2e32_per_step * 1e3,
                                             h2d_bytes_per_step=h2d32_b)),
                roofline=roof, candidates_per_s=geno, resident_two_shards_in_flight=inflight2)
    if host_binding is not None:
        line["config"]["host_binding"] = host_binding         # rank 0's; every rank binds to its own GPU's CPUs
    if not args.no_cpu_baseline and world == 1:              # reported at N=1 only (the host cores are shared by the ranks)
        base, _, _ = cpu_baseline(batch, cands, thr)
        line["cpu_baseline"] = base
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def genotype_point(shard, local, n_reads=10_000_000, n_cand=1_000_000):
    """BASELINE config 5 grid point: 10^6 candidates x 10^7 reads, resident, 3 timed runs."""
    import ctypes as C
    import torch
    from volcanosv_b200 import synth
    from volcanosv_b200.pipeline import _dp
    rs, re = synth.synth_intervals(5, n_reads)
    pos, svlen, typ, gt = synth.synth_candidates(6, n_cand)
    dev = torch.device("cuda", local)
    t = {k: torch.from_numpy(v).to(dev) for k, v in dict(rs=rs, re=re, pos=pos, svlen=svlen, typ=typ, gt=gt).items()}
    nsup = torch.zeros(n_cand, dtype=torch.int32, device=dev)
    d2 = torch.empty(n_cand, dtype=torch.int32, device=dev)
    ng = torch.empty(n_cand, dtype=torch.int32, device=dev)
    ratio = torch.empty(n_cand, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    lib, h = shard.lib, shard.h
    times = []
    for i in range(5):
        shard.ctx.timer_start()
        rc = lib.vsv_genotype(h, _dp(t["rs"]), _dp(t["re"]), n_reads, _dp(t["pos"]), _dp(t["svlen"]), _dp(t["typ"]),
                              _dp(t["gt"]), _dp(nsup), n_cand, shard.thr_del, shard.thr_ins, _dp(d2), _dp(ratio), _dp(ng))
        shard._chk(rc)
        ms = shard.ctx.timer_stop()
        if i >= 2:
            times.append(ms)
    ms = float(np.mean(times))
    alg = 8.0 * n_reads + 24.0 * n_cand        # SURVEY.md §8d B3 without the signature table
    return dict(value=n_cand / (ms * 1e-3), unit="candidates/s", reads=n_reads, candidates=n_cand, ms=ms,
                alg_GBps=alg / (ms * 1e-3) / 1e9)


if __name__ == "__main__":
    main()
