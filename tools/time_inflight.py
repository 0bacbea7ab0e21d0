"""Dev helper: resident throughput with one vs two shards in flight on one GPU."""
import sys
import threading
import time

import torch

sys.path.insert(0, ".")
from bench import make_workload  # noqa: E402
from volcanosv_b200.pipeline import ChromShard  # noqa: E402

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
batch, cands = make_workload("ont", scale, 0, 3000)
shards = [ChromShard(0, "ONT") for _ in range(3)]
for sh in shards:
    sh.upload(batch, cands)
    for _ in range(3):
        sh.run_safe()
    sh.sync()
K = 12
for nfl in (1, 2, 3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()

    def work(sh, k):
        for _ in range(k):
            sh.run_safe()
        sh.sync()
    th = [threading.Thread(target=work, args=(shards[i], K // nfl)) for i in range(nfl)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("in flight %d: %.3f ms per step" % (nfl, dt / K * 1e3))
