"""Test-only stand-in for edlib (global NW edit distance; the value is algorithm-independent)."""
import numpy as np


def align(a, b, mode="NW", task="distance", **kw):
    n, m = len(a), len(b)
    if n == 0 or m == 0:
        return {"editDistance": n + m}
    bb = np.frombuffer(b.encode(), dtype=np.uint8)
    prev = np.arange(m + 1, dtype=np.int64)
    idx = np.arange(m + 1, dtype=np.int64)
    for i in range(1, n + 1):
        sub = prev[:-1] + (bb != ord(a[i - 1]))
        cur = np.empty(m + 1, dtype=np.int64)
        cur[0] = i
        cur[1:] = np.minimum(sub, prev[1:] + 1)
        # left-to-right dependency cur[j] = min(cur[j], cur[j-1]+1): prefix-min of (cur[j]-j) + j
        cur = np.minimum.accumulate(cur - idx) + idx
        prev = cur
    return {"editDistance": int(prev[m])}
