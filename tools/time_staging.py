"""Dev helper: per-phase timing of the staged upload paths (ops8 / ops16) through ShardStream."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from bench import make_workload  # noqa: E402
from volcanosv_b200.lib import pack_ops8, pack_ops16  # noqa: E402
from volcanosv_b200.pipeline import ShardStream  # noqa: E402

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.25
batch, cands = make_workload("ont", scale, 0, 3000)
n_ops = int(batch["op_off"][-1])
meta = {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for k, v in batch.items() if k in ("op_off", "pos", "mapq", "qlen")}
pc = {k: torch.from_numpy(v).pin_memory() for k, v in cands.items()}
ss = ShardStream(0, "ONT")
for name, packer in (("ops8", pack_ops8), ("ops16", pack_ops16)):
    o, ei, ev = packer(batch["ops"])
    host = dict(meta)
    host[name] = torch.from_numpy(o.view(np.int16) if name == "ops16" else o).pin_memory()
    host["exc_idx"] = torch.from_numpy(ei).pin_memory()
    host["exc_val"] = torch.from_numpy(ev.view(np.int32)).pin_memory()
    print(name, "n_ops", n_ops, "exceptions", len(ei), "pinned", host[name].is_pinned(), host["exc_idx"].is_pinned())
    ss.n_in = ss.n_out = 0
    ss.submit(host, pc, n_ops=n_ops)
    ss.submit(host, pc, n_ops=n_ops)
    ss.collect({})
    torch.cuda.synchronize()
    pin_out = {}
    for rep in range(4):
        t0 = time.perf_counter()
        ss.submit(host, pc, n_ops=n_ops)
        t1 = time.perf_counter()
        sh = ss.slots[ss.n_out % 2]
        sh._pending[0].synchronize()
        t2 = time.perf_counter()
        sh.finish_upload()
        t3 = time.perf_counter()
        ss.n_out += 1
        res = sh.run_safe()
        sh.sync()
        t4 = time.perf_counter()
        sh.download(res, pin_out, stream=ss.down)
        t5 = time.perf_counter()
        print("  submit %.2f wait-upload %.2f expand %.2f run %.2f download %.2f ms" %
              ((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3, (t5 - t4) * 1e3))
    ss.collect({})
