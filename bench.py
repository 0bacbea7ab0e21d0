#!/usr/bin/env python
"""Benchmark of the large-indel hot path (BASELINE.json metric: alignment records/s, with
genotyped candidates/s beside it) on synthetic alignments of the shapes BASELINE.json names.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload chr10-ont|chr10-hifi|chr10-clr|wgs-hifi|wgs-ont|sweep]

chr10-*  one "step" = one pass of the whole read-side hot path over one chromosome shard: both read scanners
         in one pass over the CIGAR ops, split pairs, SA analysis, in-read merge, one sort of all tables,
         sort -u, unique signatures, then FP support / coverage / supporting signatures / spanning depth /
         genotype for the candidate table (vsv_shard_scan + vsv_shard_candidates; ONE host sync per step, at
         its end).  N > 1: every rank owns one shard (weak scaling, no collective on the data path).
wgs-*    one step = the 22 autosomes, split over the N ranks by LPT on the op counts (strong scaling).
sweep    genotyping stress grid (BASELINE config 5): candidates/s for 10^4..10^7 candidates x 10^6..10^8 reads,
         candidates split over the ranks by position.
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
HG19 = [249250621, 243199373, 198022430, 191154276, 180915260, 171115067, 159138663, 146364022, 141213431, 135534747,
        135006516, 133851895, 115169878, 107349540, 102531392, 90354753, 81195210, 78077248, 59128983, 63025520,
        48129895, 51304566]          # LI/header:5-26
COV = {"ont": 50, "hifi": 30, "clr": 89}
DTYPE = {"ont": "ONT", "hifi": "Hifi", "clr": "CLR"}


# --------------------------------------------------------------------------------------------- workloads
def make_chrom(profile: str, chrom_len: int, n_total: int, seed: int, chunks: int = 8):
    """One chromosome-shaped batch: records generated in `chunks` genomic windows (bounded generator memory),
    read names interned, next-same-name links (ERS:227-237) and SA tags for 8 % of the primary reads."""
    from volcanosv_b200 import synth
    from volcanosv_b200.lib import next_same_name
    parts = []
    n_total = max(64, n_total)
    for c in range(chunks):
        n = n_total // chunks
        # CIGAR alphabet as the aligners of each data type write it: minimap2 (ONT, CLR) uses M; pbmm2 (HiFi) =/X
        b = synth.synth_records(seed + c, n, profile, chrom_len=chrom_len // chunks, eqx=(profile == "hifi"))
        b["pos"] = (b["pos"].astype(np.int64) + c * (chrom_len // chunks)).astype(np.int32)
        b["name_id"] = (b["name_id"].astype(np.int64) + c * n).astype(np.int32)       # names are local to a window
        parts.append(b)
    batch = {}
    off, tot = [np.zeros(1, np.int64)], 0
    for b in parts:
        off.append(b["op_off"][1:] + tot)
        tot += int(b["op_off"][-1])
    batch["op_off"] = np.concatenate(off)
    for k in ("ops", "pos", "mapq", "qlen", "flag", "name_id"):
        batch[k] = np.concatenate([b[k] for b in parts])
    batch["next_same_name"] = next_same_name(batch["name_id"])
    synth.add_sa(batch, np.random.default_rng(seed + 99), frac=0.08)
    return batch


def make_cands(seed, n, chrom_len):
    from volcanosv_b200 import synth
    pos, svlen, typ, gt = synth.synth_candidates(seed, n, chrom_len=chrom_len)
    return dict(pos=pos, svlen=svlen, type=typ, gt=gt)


def prefix(batch, m):
    """The first m records of a batch as views (no copy of the ops): a smaller chromosome of the same shape."""
    e = int(batch["op_off"][m])
    out = dict(ops=batch["ops"][:e], op_off=batch["op_off"][: m + 1])
    for k in ("pos", "mapq", "qlen", "flag", "name_id"):
        out[k] = batch[k][:m]
    nx = batch["next_same_name"][:m].copy()
    nx[nx >= m] = -1
    out["next_same_name"] = nx
    k = int(np.searchsorted(batch["sa_rec"], m))
    out["sa_rec"], out["sa_off"] = batch["sa_rec"][:k], batch["sa_off"][: k + 1]
    ee = int(batch["sa_off"][k])
    for c in ("chrom", "pos", "strand", "first_s", "last_s", "bias", "mapq"):
        out["sa_" + c] = batch["sa_" + c][:ee]
    out["sa_self_chrom"] = batch.get("sa_self_chrom", 0)
    return out


def prefix_t(tm, batch, m, pin_small=True):
    """prefix() over torch tensors `tm` of the same batch (pinned host or device): views, no copy of the ops; the links
    past the cut are rewritten in a small per-chromosome copy."""
    import torch
    e = int(batch["op_off"][m])
    out = dict(ops=tm["ops"][:e], op_off=tm["op_off"][: m + 1])
    for k in ("pos", "mapq", "qlen", "flag", "name_id"):
        out[k] = tm[k][:m]
    nx = batch["next_same_name"][:m].copy()
    nx[nx >= m] = -1
    nx = torch.from_numpy(nx)
    out["next_same_name"] = nx.to(tm["ops"].device) if tm["ops"].is_cuda else (nx.pin_memory() if pin_small else nx)
    k = int(np.searchsorted(batch["sa_rec"], m))
    out["sa_rec"], out["sa_off"] = tm["sa_rec"][:k], tm["sa_off"][: k + 1]
    ee = int(batch["sa_off"][k])
    for c in ("chrom", "pos", "strand", "first_s", "last_s", "bias", "mapq"):
        out["sa_" + c] = tm["sa_" + c][:ee]
    out["sa_self_chrom"] = batch.get("sa_self_chrom", 0)
    return out


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


# --------------------------------------------------------------------------------------------- CPU arm
def cpu_step(batch, cands, thr, threads):
    """The same stage list on the host: oracle C port (scan, merge, joins, depth, genotype on `threads` threads),
    numpy for sort / group / split pairs, Python for the SA rule set."""
    from oracle import c_oracle as CO
    return CO.shard_step2(batch, cands, thr[0], thr[1], threads=threads)


def cpu_baseline(batch, cands, thr, target_s=12.0, threads=None):
    """Oracle port timed on a bounded prefix of the workload (all host threads for the threaded parts)."""
    from oracle import c_oracle as CO
    threads = threads or CO.num_threads()
    n = len(batch["pos"])

    def sample(m):
        b = prefix(batch, m)
        keep = cands["pos"] <= int(batch["pos"][m - 1])
        return b, {k: np.ascontiguousarray(v[keep]) for k, v in cands.items()}

    m = max(64, n // 64)
    b, c = sample(m)
    t0 = time.perf_counter()
    cpu_step(b, c, thr, threads)
    dt = time.perf_counter() - t0
    frac = min(1.0, (m / n) * target_s / max(dt, 1e-4))
    m2 = int(min(n, max(m, n * frac)))
    if m2 > m * 1.5:
        b, c = sample(m2)
        t0 = time.perf_counter()
        cpu_step(b, c, thr, threads)
        dt = time.perf_counter() - t0
        m = m2
    return dict(value=m / dt, unit=UNIT, cores=threads, kind="port",
                sample="first %d of %d records (%.1f%% of the shard, %d candidates), oracle port (C: scan/merge/joins/depth/"
                       "genotype on %d threads; numpy: sort/group/split pairs on 1; Python: SA rules), %.2f s" %
                       (m, n, 100.0 * m / n, len(c["pos"]), threads, dt)), m, (b, c)


def sweep_grid(scale):
    reads = [10 ** 6, 10 ** 7, 10 ** 8]
    cands = [10 ** 4, 10 ** 5, 10 ** 6, 10 ** 7]
    if scale < 1:
        reads = [max(1000, int(r * scale)) for r in reads]
        cands = [max(100, int(c * scale)) for c in cands]
    return reads, cands


# --------------------------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=["chr10-ont", "chr10-hifi", "chr10-clr", "wgs-hifi", "wgs-ont", "sweep"])
    ap.add_argument("--profile", default=None, choices=["ont", "hifi", "clr"], help="alias of --workload chr10-<profile>")
    ap.add_argument("--scale", type=float, default=1.0, help="fraction of the named workload (tests use < 1)")
    ap.add_argument("--cands", type=int, default=3000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.workload is None:
        args.workload = "chr10-%s" % (args.profile or "ont")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "sweep":
        return run_sweep(args, rank, world, local)
    kind, profile = args.workload.split("-")
    dtype = DTYPE[profile]
    from volcanosv_b200.pipeline import GT_THRESHOLDS
    thr = (GT_THRESHOLDS[dtype]["DEL"], GT_THRESHOLDS[dtype]["INS"])
    from volcanosv_b200 import synth
    wgs = kind == "wgs"
    if wgs:
        workload = "WGS %s %dx, 22 autosomes (hg19 lengths), large-indel hot path sharded by chromosome, synthetic (scale %.3g)" % (
            dtype, COV[profile], args.scale)
    else:
        workload = "chr10 %s %dx single-chromosome large-indel detection, synthetic (scale %.3g)" % (dtype, COV[profile], args.scale)

    # ---- the shards of this process ------------------------------------------------------------------
    # chr10-*: one chr10 shard per rank.  wgs-*: a chr1-sized master batch; chromosome c is its first n_c records
    # (n_c proportional to the chromosome length, so the prefix spans ~ the chromosome), i.e. 22 shards of the named
    # shape whose CIGAR content overlaps -- what a throughput measurement needs, generated in minutes.
    if wgs:
        master_len = HG19[0]
        n_master = max(64, int(synth.n_records(profile, master_len) * args.scale))
        master = make_chrom(profile, master_len, n_master, 4242)
        n_master = len(master["pos"])
        chrom_n = [min(n_master, max(32, int(n_master * L / master_len))) for L in HG19]
        chrom_ops = [int(master["op_off"][m]) for m in chrom_n]
        from volcanosv_b200.shard import assign_lpt
        plan = assign_lpt({c: float(chrom_ops[c]) for c in range(22)}, world)
        mine = sorted(plan[rank])
        loads = [sum(chrom_ops[c] for c in p) for p in plan]
        lpt = dict(max_over_mean=max(loads) / (sum(loads) / world), chroms_per_rank=[len(p) for p in plan])
        shard_cands = [make_cands(500 + c, max(16, int(args.cands * HG19[c] / HG19[9])), int(master["pos"][chrom_n[c] - 1])) for c in mine]
        shards = None          # chromosome = prefix VIEW of the master (host: pinned once; device: uploaded once)
        total_rec, total_ops = sum(chrom_n), sum(chrom_ops)
    else:
        n_total = max(64, int(synth.n_records(profile) * args.scale))
        b = make_chrom(profile, synth.CHR10_LEN, n_total, 1000 * (rank if args.impl == "ours" else 0) + 10)
        shards = [(b, make_cands(77 + (rank if args.impl == "ours" else 0), args.cands, synth.CHR10_LEN))]
        total_rec, total_ops = world * len(b["pos"]), world * int(b["op_off"][-1])
        lpt = None

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return
        if wgs:
            # the arm runs the shards of ONE process at a time; sample the largest chromosome this arm can bound
            ref_b, ref_c = prefix(master, chrom_n[20]), make_cands(520, max(16, int(args.cands * HG19[20] / HG19[9])), int(master["pos"][chrom_n[20] - 1]))
        else:
            ref_b, ref_c = shards[0]
        base, m, (sb, sc) = cpu_baseline(ref_b, ref_c, thr, target_s=8.0)
        times = []
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            cpu_step(sb, sc, thr, base["cores"])
            if i >= args.warmup:
                times.append(time.perf_counter() - t0)
        dt = sum(times) / max(1, len(times))
        v = m / dt
        base["value"] = v
        print(json.dumps(dict(impl="reference", metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=args.steps,
                              warmup=args.warmup, ms_per_step=dt * 1e3, higher_is_better=True, scaling="strong" if wgs else "weak",
                              vs_baseline=None, dtype="int32", data="synthetic",
                              config=dict(workload=workload, records_per_step=m, ops_per_step=int(sb["op_off"][-1]),
                                          note="the reference is pure Python and cannot travel; timed: its port (oracle/) on all host "
                                               "threads over a bounded sample of the same workload, same stage list as the GPU arm; one box "
                                               "whatever N is (box-vs-box ratio)"),
                              cpu_baseline=base,
                              e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))))
        return

    # ------------------------------------------------------------------ our arm
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (volcanosv_b200 has no CPU path)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from volcanosv_b200.pipeline import ShardStream
    stream = ShardStream(local, dtype)   # two shard slots: upload of shard k+1 overlaps the work on shard k
    shard = stream.slots[0]

    def pin(b):
        keys = ("ops", "op_off", "pos", "mapq", "qlen", "flag", "name_id", "next_same_name", "sa_rec", "sa_off", "sa_chrom", "sa_pos",
                "sa_strand", "sa_first_s", "sa_last_s", "sa_bias", "sa_mapq")
        out = {}
        for k in keys:
            v = b[k]
            v = v.view(np.int32) if v.dtype == np.uint32 else (v.view(np.int16) if v.dtype == np.uint16 else v)
            out[k] = torch.from_numpy(np.ascontiguousarray(v)).pin_memory()
        out["sa_self_chrom"] = b.get("sa_self_chrom", 0)
        return out

    def barrier():
        for sl in stream.slots:
            sl.sync()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    def reduce_max(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    pin_c = lambda c: {k: torch.from_numpy(v).pin_memory() for k, v in c.items()}
    if wgs:
        pm = pin(master)                                   # ONE pinned copy of the master; every chromosome is a view of it
        host = [(prefix_t(pm, master, chrom_n[c]), pin_c(cc), chrom_ops[c]) for c, cc in zip(mine, shard_cands)]
    else:
        host = [(pin(b), pin_c(c), int(b["op_off"][-1])) for b, c in shards]
    warm = max(args.warmup, 3)

    # ---- resident timing: every shard of this rank on the device, CUDA events on the library's stream ----
    res_shards = []
    if wgs:
        from volcanosv_b200.pipeline import ChromShard
        # all chromosomes of this rank stay resident: the master is uploaded once, one engine per chromosome adopts its prefix
        dm = {k: (v.to("cuda", non_blocking=True) if hasattr(v, "to") else v) for k, v in pm.items()}
        torch.cuda.synchronize()
        for c, (hb, hc, n_ops) in zip(mine, host):
            cs = ChromShard(local, dtype)
            cs.adopt(prefix_t(dm, master, chrom_n[c]), hc, n_ops=n_ops)
            res_shards.append(cs)
    else:
        shard.upload(host[0][0], host[0][1], n_ops=host[0][2])
        res_shards = [shard]
    res = None
    for _ in range(warm):
        for cs in res_shards:
            res = cs.run_safe()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    launches = lambda: sum(cs.ctx.timing()[1] for cs in res_shards)
    l0 = launches()
    scan_ms, per_shard = [], []
    t_ev0, t_ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if len(res_shards) == 1:
        shard.ctx.timer_start()
        for _ in range(args.steps):
            res = shard.run_safe()
            scan_ms += res.scan_ms
        ms = shard.ctx.timer_stop()
        last_res = [res]
    else:
        t_ev0.record()
        last_res = []
        from volcanosv_b200.pipeline import run_many
        for _ in range(args.steps):
            # all shards of this rank in flight (one stream each); the counts of every shard are read once, after the enqueues:
            # every shard's work is complete when the last count has been read
            last_res = run_many(res_shards)
            for r in last_res:
                scan_ms += r.scan_ms
        for cs in res_shards:
            cs.sync()
        t_ev1.record()
        t_ev1.synchronize()
        ms = t_ev0.elapsed_time(t_ev1)
    l1 = launches()
    barrier()
    ms_per_step = reduce_max(ms) / args.steps
    value = total_rec / (ms_per_step * 1e-3)

    # ---- end to end: pinned host buffers (BAM-native u32 words, the boundary format of SURVEY.md 8b) -> H2D -> hot
    # path -> D2H of the per-candidate results.  A GPU works through a queue of shards (chromosomes, MAIN:268), so the
    # steady state is timed: every timed step enqueues the next shard's H2D copies, runs the hot path of the previously
    # uploaded shard and reads its results back.
    pin_out = {}

    def e2e_run(queue, steps):
        stream.n_in = stream.n_out = 0
        q = [queue[i % len(queue)] for i in range(steps * len(queue) + 3)]
        stream.submit(*q[0][:2], n_ops=q[0][2])
        stream.submit(*q[1][:2], n_ops=q[1][2])
        stream.collect(pin_out)                       # warm BOTH slots (buffers, capacities, scratch)
        stream.collect(pin_out)
        stream.submit(*q[2][:2], n_ops=q[2][2])
        barrier()                                     # one shard's upload is complete here (prologue, untimed)
        t0 = time.perf_counter()
        h2d = d2h = 0
        for i in range(steps * len(queue)):
            nxt = q[3 + i] if 3 + i < len(q) else q[i % len(queue)]
            stream.submit(*nxt[:2], n_ops=nxt[2])
            stream.collect(pin_out)
            sl = stream.slots[(stream.n_out - 1) % 2]
            h2d += int(sl.h2d_bytes)
            d2h += int(sl.d2h_bytes)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        stream.collect(pin_out)                       # drain
        return reduce_max(dt) / steps, (h2d // steps, d2h // steps)

    rot = host[2 % len(host):] + host[: 2 % len(host)] if len(host) > 1 else host
    e2e_per_step, (h2d_b, d2h_b) = e2e_run(rot, args.steps if not wgs else max(1, args.steps // 2))
    e2e_value = total_rec / e2e_per_step

    # compact staging written by the BAM feeder in its own pass (vsv_bam_chrom_fill -> ops8 / ops16) or packed from
    # the u32 words (ops12): reported BESIDE the headline with the host pack time, which is outside that timed region
    staged = None
    if not wgs:
        from volcanosv_b200.lib import pack_ops8, pack_ops12_native, pack_ops16
        ops = shards[0][0]["ops"]
        n_ops = len(ops)
        t0 = time.perf_counter()
        o12, exc_pair, exc_val12 = pack_ops12_native(ops)
        pack12_s = time.perf_counter() - t0
        o8, exc8, excv8 = pack_ops8(ops[: max(1, n_ops // 16)])
        frac8 = len(exc8) / max(1, len(o8))
        hb = {k: v for k, v in host[0][0].items() if k != "ops"}
        if frac8 < 0.03 and o12.nbytes + 16 * len(exc_pair) < 1.05 * n_ops:
            name = "ops12"
            hb.update(ops12=torch.from_numpy(o12).pin_memory(), exc_idx=torch.from_numpy(exc_pair).pin_memory(),
                      exc_val=torch.from_numpy(exc_val12.view(np.int32)).pin_memory())
            pack_s, n_exc = pack12_s, len(exc_pair)
        else:
            name = "ops16"
            t0 = time.perf_counter()
            o16, ei, ev = pack_ops16(ops)
            pack_s, n_exc = time.perf_counter() - t0, len(ei)
            hb.update(ops16=torch.from_numpy(o16.view(np.int16)).pin_memory(), exc_idx=torch.from_numpy(ei).pin_memory(),
                      exc_val=torch.from_numpy(ev.view(np.int32)).pin_memory())
        st_per_step, (st_h2d, _) = e2e_run([(hb, host[0][1], n_ops)], args.steps)
        staged = dict(format="%s + %d exceptions" % (name, n_exc), value=total_rec / st_per_step, ms_per_step=st_per_step * 1e3,
                      h2d_bytes_per_step=st_h2d, host_pack_ms=pack_s * 1e3, host_pack_threads=os.cpu_count(),
                      ms_per_step_incl_pack=(st_per_step + pack_s) * 1e3,
                      note="pack time is NOT in ms_per_step; ms_per_step_incl_pack adds it serially (no overlap)")
        del o12, o8
    to_vcf = None
    if not wgs and world == 1:
        try:
            hb = {k: (v.numpy().view(np.uint32) if k == "ops" else (v.numpy().view(np.uint16) if k == "flag" else v.numpy()))
                  for k, v in host[0][0].items() if hasattr(v, "numpy")}
            hb["sa_self_chrom"] = 0
            to_vcf = to_vcf_leg(profile, dtype, hb, max(1, min(args.steps, 3)))
        except Exception as ex:  # noqa: BLE001
            to_vcf = dict(error="%s: %s" % (type(ex).__name__, ex))
    sampler.stop_flag = True                          # clocks sampled through the resident AND the end-to-end timed regions
    sampler.join()

    # ---- genotyping stress point (candidates/s), resident ------------------------------
    geno = None
    if not wgs:
        try:
            geno = genotype_point(shard, local, cpu=(world == 1 and not args.no_cpu_baseline), thr=thr)
        except Exception as ex:  # noqa: BLE001
            geno = dict(error=str(ex))

    probe = None
    if world > 1 and not wgs:
        try:
            probe = h2d_probe(local, rank, world, dist)
        except Exception as ex:  # noqa: BLE001
            probe = dict(error="%s: %s" % (type(ex).__name__, ex))
    edit = None
    if not wgs and rank == 0:
        try:
            edit = edit_distance_point(local, cpu=(world == 1 and not args.no_cpu_baseline))
        except Exception as ex:  # noqa: BLE001
            edit = dict(error="%s: %s" % (type(ex).__name__, ex))

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (k_scan_wt<FUSED>) and of the whole step --------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    r0 = last_res[0]
    my_rec = sum(r.n_rec for r in last_res)
    my_ops = sum(r.n_ops for r in last_res)
    n_sig_out = sum(r.n_read_sigs + r.n_cute_sigs for r in last_res)
    n_rows = sum(r.counts["n_reads_cigar"] + r.counts["n_reads_split"] + r.counts["n_del"] + r.counts["n_ins"] for r in last_res)
    n_usig = sum(r.n_udel + r.n_uins for r in last_res)
    n_cand = sum(r.n_cand for r in last_res)
    b1 = 4.0 * my_ops + 24.0 * my_rec + 32.0 * n_sig_out
    b2 = 3.0 * 32.0 * n_rows + 32.0 * n_cand                                   # B2
    b3 = 8.0 * my_rec + 16.0 * n_usig + 24.0 * n_cand                          # B3
    per_launch_ms = float(np.mean(scan_ms)) if scan_ms else float("nan")
    scans_per_step = len(last_res)
    alg_launch = b1 / scans_per_step
    achieved = alg_launch / (per_launch_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "scan_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic = (tj["dram_bytes_read"] + tj["dram_bytes_write"]) / tj["n_ops"] * (my_ops / scans_per_step)
        traffic_src = tj["source"]
    roof = dict(bound="hbm", kernel="k_scan_wt<FUSED>", achieved=achieved, peak=peak, unit="GB/s", frac=achieved / peak,
                traffic=traffic, traffic_source=traffic_src, peak_source=peak_src, alg_bytes_per_launch=alg_launch,
                kernel_ms=per_launch_ms, launches_per_step=scans_per_step, share_of_step=per_launch_ms * scans_per_step / ms_per_step,
                step_alg_bytes=b1 + b2 + b3, step_frac=(b1 + b2 + b3) / (ms_per_step * 1e-3) / 1e9 / peak,
                step_note="step_frac = (B1 + B2 + B3 of SURVEY.md 8d for this rank's shards) / ms_per_step / peak")
    cfg = dict(workload=workload, records=total_rec, ops=total_ops, records_this_rank=my_rec, candidates_this_rank=n_cand,
               read_sigs=sum(r.n_read_sigs for r in last_res), cute_sigs=sum(r.n_cute_sigs for r in last_res),
               sigs_rows_del=sum(r.counts["n_del"] for r in last_res), sigs_rows_ins=sum(r.counts["n_ins"] for r in last_res),
               sa_rows=sum(r.counts["n_sa_rows"] for r in last_res), split_sigs=sum(r.counts["n_reads_split"] for r in last_res),
               unique_del=sum(r.n_udel for r in last_res), unique_ins=sum(r.n_uins for r in last_res),
               candidates="synthetic table (pos sorted, Pareto lengths)", host_syncs_per_shard=1,
               shards_in_flight=len(res_shards),
               parallelism=("22 chromosome shards over %d ranks by LPT, no collective" % world) if wgs else
               ("shard-per-gpu x%d, no collective" % world),
               l2="inputs (%.2f GB/step on this rank) exceed the 126 MB L2; no flush needed" % (4.0 * my_ops / 1e9)
               if 4.0 * my_ops / scans_per_step > 2.5e8 else
               "per-shard inputs (%.0f MB) are of the order of the 126 MB L2: consecutive shards evict each other only in the "
               "wgs workloads; chr10-hifi re-runs one 164 MB shard (larger than L2)" % (4.0 * my_ops / scans_per_step / 1e6))
    if lpt:
        cfg["lpt"] = lpt
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=warm,
                ms_per_step=ms_per_step, higher_is_better=True, scaling="strong" if wgs else "weak", vs_baseline=None, dtype="int32",
                data="synthetic", config=cfg, clocks=sampler.result(), gpu_launches=int(l1 - l0),
                e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=h2d_b, d2h_bytes_per_step=d2h_b,
                         ms_per_step=e2e_per_step * 1e3,
                         staging="BAM-native u32 CIGAR words + per-record columns in pinned host buffers (no host-side packing)",
                         pipelining="two shard slots: H2D of shard k+1 overlaps kernels + D2H of shard k",
                         feeder_staging=staged, to_final_vcf=to_vcf),
                roofline=roof, candidates_per_s=geno, edit_distance=edit)
    if probe is not None:
        line["e2e"]["h2d_probe"] = probe
    if not args.no_cpu_baseline and world == 1:              # reported at N=1 only (the host cores are shared by the ranks)
        ref_b, ref_c = (prefix(master, chrom_n[mine[-1]]), shard_cands[-1]) if wgs else shards[0]
        base, _, _ = cpu_baseline(ref_b, ref_c, thr)
        line["cpu_baseline"] = base
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def to_vcf_leg(profile, dtype, host_batch, steps, warm=1):
    """The WHOLE stage through the product driver (volcanosv_b200.operators.large_indel_chr == volcanosv-vc-large-indel.py
    -chr 10, MAIN:233-319) from packed host records to the bytes of the final phased VCF: contig scan / merge / cluster /
    hap pairing / sequences (a1-a9), read-side step on the resident shard, FP filter, redundancy (edit distance), coverage
    band filter, genotype correction, vcf-sort, phasing.  Inputs in host memory (pinned): packed read records, packed contig
    records, contig / reference sequences as lazy pseudo-random `SynthSeq`; the 10^5-10^7-row text dumps are off
    (row_dumps=False), every per-call file (VCFs, CSV, TSVs) is written."""
    import shutil
    import tempfile
    from volcanosv_b200 import operators as OP
    from volcanosv_b200 import resident as RS
    from volcanosv_b200 import synth
    cb, cnames, cseqs = synth.synth_contigs(31)
    cR = OP.Records.from_arrays(cb, names=cnames)
    L10 = synth.CHR10_LEN
    OP.register_source("mem://contigs", OP.ArraySource([("chr10", L10)], {"chr10": cR}))
    rR = OP.Records.from_arrays(host_batch)
    OP.register_source("mem://reads", OP.ArraySource([("chr10", L10)], {"chr10": rR}))
    ref = {"chr10": synth.SynthSeq(L10, 10)}
    base = "/dev/shm" if os.path.isdir("/dev/shm") else None
    times, n_calls, n_bytes, phases = [], 0, 0, {}
    for i in range(warm + steps):
        out = tempfile.mkdtemp(prefix="vsv_bench_", dir=base)
        RS.clear()                                 # the shard is rebuilt: H2D of the packed records is inside the timed region
        OP.STAGE_TIMES.clear()
        t0 = time.perf_counter()
        vcf = OP.large_indel_chr("mem://contigs", cseqs, ref, "mem://reads", out, dtype, 10, "Sample", row_dumps=False)
        data = open(vcf, "rb").read()
        dt = time.perf_counter() - t0
        if i >= warm:
            times.append(dt)
            for k, v in OP.STAGE_TIMES.items():
                phases[k] = phases.get(k, 0.0) + v * 1e3 / steps
        n_calls = sum(1 for ln in data.split(b"\n") if ln and not ln.startswith(b"#"))
        n_bytes = len(data)
        shutil.rmtree(out, ignore_errors=True)
    RS.clear()
    dt = float(np.mean(times))
    return dict(value=len(host_batch["pos"]) / dt, unit=UNIT, ms_per_step=dt * 1e3, final_vcf_calls=n_calls, final_vcf_bytes=n_bytes,
                contig_records=len(cnames), stage_ms={k: round(v, 2) for k, v in phases.items()}, note="product driver large_indel_chr(row_dumps=False): packed host records -> bytes of the "
                "phased VCF; includes the Python text handling of ~10^3 calls through 8 VCF/TSV/CSV files as the reference chains them")


def genotype_point(shard, local, n_reads=10_000_000, n_cand=1_000_000, cpu=False, thr=None):
    """BASELINE config 5 grid point: 10^6 candidates x 10^7 reads, resident (the sort of the read ends is inside)."""
    import torch
    from volcanosv_b200 import synth
    from volcanosv_b200.pipeline import _dp
    import ctypes as C
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
    p = shard.ctx.get_params()
    times = []
    for i in range(5):
        shard.ctx.timer_start()
        rc = lib.vsv_genotype(h, _dp(t["rs"]), _dp(t["re"]), n_reads, _dp(t["pos"]), _dp(t["svlen"]), _dp(t["typ"]),
                              _dp(t["gt"]), _dp(nsup), n_cand, p.thr_del, p.thr_ins, _dp(d2), _dp(ratio), _dp(ng))
        shard._chk(rc)
        ms = shard.ctx.timer_stop()
        if i >= 2:
            times.append(ms)
    ms = float(np.mean(times))
    alg = 8.0 * n_reads + 24.0 * n_cand        # SURVEY.md 8d B3 without the signature table
    out = dict(value=n_cand / (ms * 1e-3), unit="candidates/s", reads=n_reads, candidates=n_cand, ms=ms,
               alg_GBps=alg / (ms * 1e-3) / 1e9, kernel="radix sort of the read ends + k_genotype (closed-form depth)")
    if cpu:
        from oracle import c_oracle as CO
        m = 20000
        t0 = time.perf_counter()
        CO.genotype(rs, re, pos[:m], svlen[:m], typ[:m], gt[:m], np.zeros(m, np.int32), thr[0], thr[1], threads=CO.num_threads())
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = dict(value=m / dt, unit="candidates/s", cores=CO.num_threads(), kind="port",
                                   sample="first %d candidates against all %d reads, %.2f s" % (m, n_reads, dt))
    return out


def edit_pairs(seed, n_pair, lo=100, hi=100_000, frac_same=0.6):
    """Insertion-sequence pairs shaped like the redundancy stage's input (RR:92-102 keeps pairs within 500 bp whose sizes are
    within 2x): Pareto lengths in [lo, hi]; `frac_same` of the pairs are one event seen on two haplotypes (1 % divergence:
    substitutions and short indels), the rest unrelated sequences of similar size."""
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    strs, ia, ib = [], [], []
    for _ in range(n_pair):
        la = int(min(hi, lo * (1.0 - rng.random()) ** (-1.0 / 1.1)))
        a = acgt[rng.integers(0, 4, la)]
        if rng.random() < frac_same:
            b = a.copy()
            k = max(1, la // 100)
            b[rng.integers(0, la, k)] = acgt[rng.integers(0, 4, k)]
            for _ in range(max(1, la // 2000)):                # short indels
                c, ln = int(rng.integers(0, len(b))), int(rng.integers(1, 6))
                b = np.concatenate([b[:c], acgt[rng.integers(0, 4, ln)], b[c:]]) if rng.random() < 0.5 else np.concatenate([b[:c], b[c + ln:]])
        else:
            b = acgt[rng.integers(0, 4, max(1, int(la * rng.uniform(0.5, 1.0))))]
        ia.append(len(strs))
        ib.append(len(strs) + 1)
        strs += [a.tobytes(), b.tobytes()]
    return strs, ia, ib


def edit_distance_point(local, cpu=False, n_pair=1000):
    """SURVEY.md 8f-1 / RR:75-102: one redundancy-stage batch (10^3 insertion pairs of 10^2..10^5 bp) through
    vsv_edit_distance_batch with HOST buffers; the reference's own vendored edlib (oracle/_ref) beside it."""
    from volcanosv_b200 import lib as L
    ctx = L.Context(local)
    strs, ia, ib = edit_pairs(21, n_pair)
    cells = float(sum(len(strs[i]) * len(strs[j]) for i, j in zip(ia, ib)))
    ctx.edit_distance_batch(strs[:8], ia[:4], ib[:4])              # module load
    wall, kern = [], []
    for _ in range(3):
        t0 = time.perf_counter()
        d = ctx.edit_distance_batch(strs, ia, ib)
        wall.append(time.perf_counter() - t0)
        kern.append(ctx.timing()[0])
    w, k = float(np.median(wall)), float(np.median(kern))
    tot = np.array([len(strs[i]) + len(strs[j]) for i, j in zip(ia, ib)])
    out = dict(value=n_pair / w, unit="pairs/s", pairs=n_pair, ms=w * 1e3, kernel_ms=k, lengths="Pareto(1.1) in [100, 100000] bp",
               max_len=int(max(len(x) for x in strs)), similar_pairs=int(np.sum((tot - d) / tot >= 0.5)),
               full_matrix_cells=cells, cell_updates_per_s=cells / (k * 1e-3),
               note="cell_updates_per_s = full-matrix cells (len1 x len2 summed) / kernel time: the band skips most of them for the "
                    "near-identical pairs, so it is an equivalent rate, comparable with the CPU figure computed the same way",
               kernel="k_edit_banded (Myers/Hyyro 64-row blocks, Ukkonen band x8 per round, state in shared memory)")
    ctx.close()
    if cpu:
        from oracle import edlib_ref as E
        if E.available():
            m = n_pair
            t0 = time.perf_counter()
            ref = [E.distance(strs[i].decode(), strs[j].decode()) for i, j in zip(ia[:m], ib[:m])]
            dt = time.perf_counter() - t0
            c = float(sum(len(strs[i]) * len(strs[j]) for i, j in zip(ia[:m], ib[:m])))
            out["cpu_baseline"] = dict(value=m / dt, unit="pairs/s", cores=1, kind="reference", cell_updates_per_s=c / dt,
                                       sample="first %d pairs, the reference's vendored edlib (oracle/_ref/libedlib_ref.so, "
                                              "EDLIB_MODE_NW, k = -1), %.2f s" % (m, dt), parity=bool(ref == d[:m].tolist()))
    return out


def h2d_probe(local, rank, world, dist, nbytes=1 << 30, reps=4):
    """Pinned-host -> device bandwidth of this box under three schedules (N > 1 only): every rank alone, all ranks at once, and
    the ranks in two halves one after the other.  What bounds the end-to-end curve is the 'all' aggregate; if 'halves' does not
    beat it, queueing the uploads cannot help."""
    import torch
    hostbuf = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    devbuf = torch.empty(nbytes, dtype=torch.uint8, device=torch.device("cuda", local))
    devbuf.copy_(hostbuf, non_blocking=True)
    torch.cuda.synchronize()

    def timed(active):
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if active:
            for _ in range(reps):
                devbuf.copy_(hostbuf, non_blocking=True)
            torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        dist.barrier()
        return dt

    def gmax(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    gb = nbytes * reps / 1e9
    alone = [gmax(timed(rank == r)) for r in range(world)]
    t_all = gmax(timed(True))
    half = (world + 1) // 2
    t_half = gmax(timed(rank < half)) + gmax(timed(rank >= half))
    return dict(bytes_per_rank=nbytes * reps, alone_GBps_per_rank=[gb / t for t in alone], all_at_once_GBps_aggregate=world * gb / t_all,
                two_halves_GBps_aggregate=world * gb / t_half,
                note="aggregate pinned-host -> device bandwidth with all ranks copying at once vs in two halves; "
                     "e2e efficiency at N = (all_at_once aggregate) / (N x alone) because the end-to-end step is upload-bound")


def run_sweep(args, rank, world, local):
    """BASELINE config 5: candidates/s over the grid; candidates split over the ranks by position (every rank holds all
    the reads: 8 n_reads bytes), device-timed, max over ranks."""
    from volcanosv_b200 import synth
    from volcanosv_b200.pipeline import GT_THRESHOLDS
    thr = (GT_THRESHOLDS["ONT"]["DEL"], GT_THRESHOLDS["ONT"]["INS"])
    reads_grid, cand_grid = sweep_grid(args.scale)
    if args.impl == "reference":
        if rank != 0:
            return
        from oracle import c_oracle as CO
        pts = []
        tot_c, tot_t = 0, 0.0
        for nr in reads_grid:
            rs, re = synth.synth_intervals(5, nr)
            pos, svlen, typ, gt = synth.synth_candidates(6, 20000)
            t0 = time.perf_counter()
            CO.genotype(rs, re, pos, svlen, typ, gt, np.zeros(len(pos), np.int32), thr[0], thr[1], threads=CO.num_threads())
            dt = time.perf_counter() - t0
            pts.append(dict(reads=nr, candidates_sampled=len(pos), candidates_per_s=len(pos) / dt))
            tot_c += len(pos)
            tot_t += dt
        v = tot_c / tot_t
        print(json.dumps(dict(impl="reference", metric="genotyped candidates/s", value=v, unit="candidates/s", n_gpus=args.gpus,
                              steps=args.steps, warmup=args.warmup, ms_per_step=tot_t * 1e3, higher_is_better=True, scaling="strong",
                              vs_baseline=None, dtype="int32", data="synthetic",
                              config=dict(workload="genotyping stress sweep (scale %.3g)" % args.scale, points=pts,
                                          note="oracle C port on all host threads; 20000 candidates per read count: the per-candidate "
                                               "cost does not depend on the number of candidates"),
                              cpu_baseline=dict(value=v, unit="candidates/s", cores=CO.num_threads(), kind="port",
                                                sample="20000 candidates per read count"),
                              e2e=dict(value=v, unit="candidates/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))))
        return
    import torch
    import ctypes as C
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import volcanosv_b200 as V
    from volcanosv_b200 import lib as L
    from volcanosv_b200.pipeline import _dp
    ctx = V.Context(local)
    ctx.set_params(L.default_params("ONT"))
    ctx.set_device_io(True)
    dev = torch.device("cuda", local)
    p = ctx.get_params()
    sampler = ClockSampler(local)
    sampler.start()
    pts = []
    tot_c, tot_ms, launches0 = 0, 0.0, ctx.timing()[1]
    e2e_c, e2e_s, h2d, d2h = 0, 0.0, 0, 0
    for nr in reads_grid:
        rs, re = synth.synth_intervals(5, nr)
        rs_p, re_p = torch.from_numpy(rs).pin_memory(), torch.from_numpy(re).pin_memory()
        d_rs, d_re = rs_p.to(dev), re_p.to(dev)
        for nc in cand_grid:
            pos, svlen, typ, gt = synth.synth_candidates(6, nc)
            a, b = rank * nc // world, (rank + 1) * nc // world
            hc = {k: torch.from_numpy(np.ascontiguousarray(v[a:b])).pin_memory() for k, v in dict(pos=pos, svlen=svlen, typ=typ, gt=gt).items()}
            dc = {k: v.to(dev) for k, v in hc.items()}
            m = b - a
            nsup = torch.zeros(m, dtype=torch.int32, device=dev)
            d2, ng = torch.empty(m, dtype=torch.int32, device=dev), torch.empty(m, dtype=torch.int32, device=dev)
            ratio = torch.empty(m, dtype=torch.float64, device=dev)
            out_p = torch.empty(m, dtype=torch.int32).pin_memory()

            def step():
                rc = ctx.lib.vsv_genotype(ctx.h, _dp(d_rs), _dp(d_re), nr, _dp(dc["pos"]), _dp(dc["svlen"]), _dp(dc["typ"]), _dp(dc["gt"]),
                                          _dp(nsup), m, p.thr_del, p.thr_ins, _dp(d2), _dp(ratio), _dp(ng))
                ctx._chk(rc)
            for _ in range(max(args.warmup, 3)):
                step()
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            ctx.timer_start()
            for _ in range(args.steps):
                step()
            ms = ctx.timer_stop() / args.steps
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            if dist is not None:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
            # end to end: reads + candidates from pinned host memory, new genotypes back
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            d_rs.copy_(rs_p, non_blocking=True); d_re.copy_(re_p, non_blocking=True)
            for k in dc:
                dc[k].copy_(hc[k], non_blocking=True)
            torch.cuda.synchronize()
            step()
            ctx.sync()
            out_p.copy_(ng, non_blocking=True)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            if dist is not None:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
            alg = 8.0 * nr + 24.0 * m
            pts.append(dict(reads=nr, candidates=nc, ms=ms, candidates_per_s=nc / (ms * 1e-3), alg_GBps_this_rank=alg / (ms * 1e-3) / 1e9,
                            e2e_ms=dt * 1e3, e2e_candidates_per_s=nc / dt))
            tot_c += nc
            tot_ms += ms
            e2e_c += nc
            e2e_s += dt
            h2d += 8 * nr + 16 * m
            d2h += 4 * m
        del d_rs, d_re
    sampler.stop_flag = True
    sampler.join()
    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        peak = float(json.load(open(peaks_path))["hbm_gbs"]) if os.path.exists(peaks_path) else 6650.0
        big = max(pts, key=lambda x: x["reads"] * 1e3 + x["candidates"])
        line = dict(metric="genotyped candidates/s", value=tot_c / (tot_ms * 1e-3), unit="candidates/s", n_gpus=world, steps=args.steps,
                    warmup=max(args.warmup, 3), ms_per_step=tot_ms, higher_is_better=True, scaling="strong", vs_baseline=None, dtype="int32",
                    data="synthetic",
                    config=dict(workload="genotyping stress sweep: %s reads x %s candidates (scale %.3g)" % (reads_grid, cand_grid, args.scale),
                                points=pts, parallelism="candidates split by position over %d ranks, reads replicated, no collective" % world,
                                l2="read arrays of 8 n_reads bytes: 8 MB / 80 MB / 800 MB; the two smaller ones fit the 126 MB L2 (a resident "
                                   "index is the point of the closed form), candidates stream"),
                    clocks=sampler.result(), gpu_launches=int(ctx.timing()[1] - launches0),
                    e2e=dict(value=e2e_c / e2e_s, unit="candidates/s", h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h),
                    roofline=dict(bound="hbm", kernel="radix sort of read ends + k_genotype", achieved=big["alg_GBps_this_rank"], peak=peak,
                                  unit="GB/s", frac=big["alg_GBps_this_rank"] / peak, traffic=None,
                                  note="largest grid point; algorithmic bytes = 8 n_reads + 24 n_cand (SURVEY.md 8d B3)"))
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
