"""Multi-GPU partitioning of the large-indel stage: chromosome shards, no data-path collective.

The reference fans the per-chromosome chain out over processes (`joblib.Parallel` over 22
chromosomes, LI/volcanosv-vc-large-indel.py:268) and concatenates the per-chromosome bodies
in chromosome order (MAIN:271-278).  Here one process owns one GPU; chromosomes are assigned
to ranks by longest-processing-time-first on their op counts, every rank runs its shards
independently, and the results meet on the HOST (rank 0) where they are put back in the
reference's order.  The only values that cross a shard edge are scalars:
  * the genome-wide coverage median of filter_vcf_by_sig_cov_insdel.py:39,55 (<= 10^4 floats
    gathered to rank 0), and
  * `last_match_idx` of match_varlist_siglist (GTD:94): see `support_carry_in`.
torch.distributed is used for the host-side gather only (gloo for objects; bench.py uses
NCCL just for its timing barrier).
"""
from __future__ import annotations

from typing import Dict, Hashable, List, Sequence


def assign_lpt(weights: Dict[Hashable, float], n_ranks: int) -> List[List[Hashable]]:
    """Longest-processing-time-first: heaviest shard to the least loaded rank.  Deterministic
    (ties: insertion order of `weights`, then lowest rank)."""
    if n_ranks < 1:
        raise ValueError("n_ranks must be >= 1")
    order = sorted(range(len(weights)), key=lambda i: (-list(weights.values())[i], i))
    keys = list(weights.keys())
    load = [0.0] * n_ranks
    out: List[List[Hashable]] = [[] for _ in range(n_ranks)]
    for i in order:
        r = min(range(n_ranks), key=lambda j: (load[j], j))
        out[r].append(keys[i])
        load[r] += float(weights[keys[i]])
    return out


def chrom_numeric_order(names: Sequence[str]) -> List[str]:
    """chr1..chr22 in numeric order: the order MAIN:271-278 concatenates per-chromosome bodies
    (candidate order of the genotype stage)."""
    def key(n):
        s = n[3:] if n.startswith("chr") else n
        return (0, int(s)) if s.isdigit() else (1, s)
    return sorted(names, key=key)


def chrom_dictionary_order(names: Sequence[str]) -> List[str]:
    """C-locale string order: `.sigs` files (`sort -k2,2`, SE:637-638) and `vcf-sort`'s
    `sort -k1,1d` (FGC:177): chr1, chr10, ..., chr19, chr2, chr20, ..."""
    return sorted(names)


def support_carry_in(chrom: str, cand_chroms: Sequence[str], sig_chroms: Sequence[str], matched_before: bool) -> int:
    """`last_in` of vsv_sig_support for a chromosome shard (GTD:94-133 run genome-wide).

    The reference's last_match_idx starts at 0 and only ever moves to a position-matched
    signature of the chromosome being processed, so when candidates of `chrom` start it
    points into `chrom`'s own block of the unique-signature list only if it is still 0 AND
    that block starts at global index 0 (i.e. `chrom` sorts first among chromosomes that have
    signatures).  Otherwise it sits in another chromosome's block, can never qualify for the
    double count, and the shard-local carry is "none" (-1).
    `matched_before`: some earlier candidate (in candidate order) already moved the index.
    """
    if matched_before or not sig_chroms:
        return -1
    first_block = chrom_dictionary_order(sig_chroms)[0]
    return 0 if chrom == first_block else -1


def assemble_support_chain(runs, blocks):
    """Rank 0's side of the sharded supporting-signature count (GTD:92-137 run genome-wide).

    The reference's last_match_idx indexes the GENOME-WIDE unique-signature list (chromosome blocks in `.sigs` order).  Every
    shard computed its candidates with a LOCAL carry-in: variant "a" = -1 (the index points into another chromosome's block
    and can never be double counted) and, for the chromosome whose block starts at global index 0, also variant "b" = 0.
    `runs`: the chromosome runs in VCF order, each dict(chrom=..., a=(support, local_match_id, local_last_out)[, b=...]);
    `blocks`: {chrom: (first global index of its block, unique signatures in it)}.
    Returns ([(variant, support, global_match_id) per run], final global last_match_idx): walks the runs in order, keeps the global
    index, picks the variant that matches it and lifts the local indices into the global list."""
    import numpy as np
    out, global_last = [], 0
    total = sum(n for _, n in blocks.values())
    for run in runs:
        b0, nu = blocks.get(run["chrom"], (0, 0))
        li = global_last - b0 if (nu > 0 and b0 <= global_last < b0 + nu) else -1
        if li > 0 or (li == 0 and "b" not in run):
            raise ValueError("last_match_idx points inside the block of %s before its first candidate" % run["chrom"])
        tag = "b" if li == 0 else "a"
        sup, mid, last_out = run[tag]
        mid = np.asarray(mid)
        gmid = np.where(mid >= 0, b0 + mid, global_last) if total > 0 else np.zeros(len(mid), np.int64)
        out.append((tag, np.asarray(sup), gmid))
        if int(last_out) >= 0:
            global_last = b0 + int(last_out)
    return out, global_last


def gather_to_root(local: dict, group=None) -> List[dict]:
    """Host-side gather of per-rank result dicts (keyed by chromosome) to rank 0.  Returns the
    list of all ranks' dicts on rank 0, [] elsewhere.  No-op without torch.distributed."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return [local]
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    out = [None] * world if rank == 0 else None
    dist.gather_object(local, out, dst=0, group=group)
    return out if rank == 0 else []


def concat_in_order(per_rank: List[dict], order: Sequence[str]) -> list:
    """Merge the ranks' {chrom: value} dicts and list the values in `order` (missing skipped)."""
    merged = {}
    for d in per_rank:
        for k, v in d.items():
            if k in merged:
                raise ValueError("chromosome %r produced by two ranks" % (k,))
            merged[k] = v
    return [merged[c] for c in order if c in merged]


def split_records(op_off: Sequence[int], n_parts: int) -> List[tuple]:
    """Level-2 sharding of ONE chromosome (the reference cuts a chromosome into 10 Mb
    batches with read-start ownership, SE:603-612,535): contiguous record ranges [lo, hi) of a
    coordinate-sorted batch, balanced by CIGAR-op count.  The scans are per record, so the
    signatures of the parts, concatenated in part order, are exactly the signatures of the whole
    batch; sorting, grouping and the joins then run on the gathered table."""
    import numpy as np
    off = np.asarray(op_off, dtype=np.int64)
    n = len(off) - 1
    if n_parts < 1:
        raise ValueError("n_parts must be >= 1")
    cuts = [0]
    for k in range(1, n_parts):
        target = off[0] + (off[-1] - off[0]) * k // n_parts
        cuts.append(max(cuts[-1], min(n, int(np.searchsorted(off, target, side="left")))))
    cuts.append(n)
    return [(cuts[i], cuts[i + 1]) for i in range(n_parts)]


def take_records(batch: dict, lo: int, hi: int) -> dict:
    """The packed columns of records [lo, hi) with op offsets rebased to 0."""
    import numpy as np
    a, b = int(batch["op_off"][lo]), int(batch["op_off"][hi])
    out = {k: v[lo:hi] for k, v in batch.items() if k not in ("ops", "op_off")}
    out["ops"] = np.ascontiguousarray(batch["ops"][a:b])
    out["op_off"] = np.asarray(batch["op_off"][lo: hi + 1], dtype=np.int64) - a
    return out


def bind_host_to_gpu(device: int) -> dict:
    """Keep this rank's host threads (and so its first-touch pinned staging buffers) on the CPUs
    the driver reports as local to `device`.  With several ranks per node the shard uploads are
    the end-to-end bound; a staging buffer on the far socket sends every byte
    over the inter-socket link first.  Call before allocating pinned memory and before starting
    worker threads (they inherit the mask).  Best effort: returns what was done, never raises.
    """
    import os
    info = {"bound": False}
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(device)
        bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        hd = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(hd, (n_cpu + 63) // 64)
        local = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        info.update(gpu_local_cpus=len(local), allowed_cpus=len(allowed))
        try:
            info["numa_node"] = int(pynvml.nvmlDeviceGetNumaNodeId(hd))
        except Exception:
            pass
        both = local & allowed
        if len(both) >= 4 and both != allowed:
            os.sched_setaffinity(0, both)
            info["bound"] = True
        info["cpus"] = len(both) if both else len(allowed)
    except Exception as e:   # no NVML, restricted cpuset, old driver: run unbound
        info["error"] = "%s: %s" % (type(e).__name__, e)
    return info
