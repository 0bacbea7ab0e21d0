"""Developer helper (profiles/r02s_scan.md): event-timed k_scan_wt<FUSED> of one bench workload with the library named by
VSV_B200_LIB -- the product build, or a profiling build of the same sources with -DVSV_ABLATE_DETECT (two sums only, no
candidate / rare-code detection: an upper bound for any 'plain row' fast path) or -DVSV_ABLATE_LOOP (TMA stream + warp scans,
no per-op work: the memory-side ceiling of the kernel's structure).  Ablated builds produce wrong tables by design.
usage: VSV_B200_LIB=volcanosv_b200/libvolcano_b200_nodetect.so python tools/scan_ablation.py chr10-ont"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from volcanosv_b200 import synth  # noqa: E402
from volcanosv_b200.pipeline import ChromShard  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "chr10-ont"
profile = workload.split("-")[1]
b = bench.make_chrom(profile, synth.CHR10_LEN, synth.n_records(profile), 10)
sh = ChromShard(0, bench.DTYPE[profile])
sh.upload(b, bench.make_cands(77, 3000, synth.CHR10_LEN))
sh._chk(sh.lib.vsv_shard_reserve(sh.sh, 1 << 21, 1 << 22, 1 << 22, 1 << 22))
ms = []
for i in range(8):
    sh.run(what=0)
    sh.sync()
    if i >= 3:
        ms.append(sh.ctx.main_kernel_ms())
n_ops, n_rec = int(b["op_off"][-1]), len(b["pos"])
t = float(np.median(ms))
print(json.dumps(dict(lib=os.path.basename(os.environ.get("VSV_B200_LIB", "libvolcano_b200.so")), workload=workload, ops=n_ops,
                      scan_ms=t, ops_GBps=4.0 * n_ops / t / 1e6, frac_of_6548_8=4.0 * n_ops / t / 1e6 / 6548.8)))
