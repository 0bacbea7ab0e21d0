"""Developer helper for ncu: builds a bench workload, warms the resident step up and runs ONE step between
cudaProfilerStart/Stop (use `ncu --profile-from-start off ...`).  usage: step_profile.py [workload] [scale] [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from volcanosv_b200 import synth  # noqa: E402
from volcanosv_b200.pipeline import ChromShard  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "chr10-ont"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
profile = workload.split("-")[1]
n_total = max(64, int(synth.n_records(profile) * scale))
b = bench.make_chrom(profile, synth.CHR10_LEN, n_total, 10)
c = bench.make_cands(77, 3000, synth.CHR10_LEN)
sh = ChromShard(0, bench.DTYPE[profile])
sh.upload(b, c)
raw = bool(os.environ.get("VSV_MERGE_DBG"))      # experiments with broken outputs: no overflow handling


def step():
    if not raw:
        return sh.run_safe()
    r = sh.run()
    r.counts = sh.counts()
    return r


try:
    res = sh.run_safe()
except Exception as ex:  # noqa: BLE001
    if not raw:
        raise
    print("ignored:", ex)
    sh._chk(sh.lib.vsv_shard_reserve(sh.sh, 1 << 21, 1 << 22, 1 << 22, 1 << 22))
for _ in range(3):
    res = step()
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(steps):
    res = step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(res.counts, res.scan_ms)
