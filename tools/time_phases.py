"""Dev helper: wall-clock of the C-ABI calls of one resident step (each call timed to completion)."""
import sys
import time
import ctypes as C

import numpy as np
import torch

sys.path.insert(0, ".")
from bench import make_workload  # noqa: E402
from volcanosv_b200 import lib as L  # noqa: E402
from volcanosv_b200.pipeline import ChromShard  # noqa: E402

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
batch, cands = make_workload("ont", scale, 0, 3000)
sh = ChromShard(0, "ONT")
sh.upload(batch, cands)
for _ in range(3):
    sh.run_safe()
sh.sync()
lib = sh.lib
names = [n for n in L.SIGNATURES if n.startswith("vsv_") and n not in ("vsv_sync", "vsv_last_error", "vsv_get_timing", "vsv_timer_start",
                                                                      "vsv_timer_stop", "vsv_get_main_kernel_ms")]
acc = {}
orig = {}
import threading
lock = threading.Lock()
for n in names:
    fn = getattr(lib, n)
    orig[n] = fn

    def wrap(*a, _fn=fn, _n=n):
        t0 = time.perf_counter()
        rc = _fn(*a)
        lib.vsv_sync(a[0]) if _n not in ("vsv_ctx_create",) and a and isinstance(a[0], (int, C.c_void_p)) else None
        dt = time.perf_counter() - t0
        with lock:
            acc.setdefault(_n, [0, 0.0])
            acc[_n][0] += 1
            acc[_n][1] += dt
        return rc
    setattr(lib, n, wrap)
K = 10
t0 = time.perf_counter()
for _ in range(K):
    sh.run_safe()
sh.sync()
tot = (time.perf_counter() - t0) / K
print("step with per-call sync: %.3f ms" % (tot * 1e3))
for n, (c, t) in sorted(acc.items(), key=lambda x: -x[1][1]):
    print("%-28s calls/step %.1f  %.3f ms/step" % (n, c / K, t / K * 1e3))
