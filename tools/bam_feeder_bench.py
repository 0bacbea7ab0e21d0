"""Dev helper: throughput of the native BAM feeder on an ONT-shaped BAM written from synthetic records."""
import os
import struct
import sys
import time
import zlib

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import bamio  # noqa: E402
from volcanosv_b200 import lib as L, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
path = "/tmp/vsv_feeder_bench.bam"
b = synth.synth_records(3, n, "ont", chrom_len=20_000_000)
rng = np.random.default_rng(1)
out = bytearray(b"BAM\1")
text = b"@HD\tVN:1.6\tSO:coordinate\n@SQ\tSN:chr10\tLN:135534747\n"
out += struct.pack("<I", len(text)) + text + struct.pack("<I", 1) + struct.pack("<I", 6) + b"chr10\0" + struct.pack("<I", 135534747)
off = b["op_off"]
for i in range(n):
    cig = b["ops"][off[i]: off[i + 1]]
    l_seq = int(b["qlen"][i])
    name = b"read%07d" % i
    if len(cig) > 65535:
        continue
    body = struct.pack("<iiBBHHHIiii", 0, int(b["pos"][i]), len(name) + 1, int(b["mapq"][i]), 4680, len(cig), int(b["flag"][i]) & 0xFFFF,
                       l_seq, -1, -1, 0)
    body += name + b"\0" + cig.tobytes() + rng.integers(0, 256, (l_seq + 1) // 2, dtype=np.uint8).tobytes() + b"\x20" * l_seq
    out += struct.pack("<I", len(body)) + body
with open(path, "wb") as fo:
    for i in range(0, len(out), 65280):
        fo.write(bamio._bgzf_block(bytes(out[i:i + 65280])))
    fo.write(bamio._bgzf_block(b""))
print("BAM: %.1f MB compressed, %.1f MB inflated, %d records, %d ops" % (os.path.getsize(path) / 1e6, len(out) / 1e6, n, int(off[-1])))
# timed in a fresh process: the numpy / BLAS worker threads of the generator above keep spinning for a while
import subprocess
code = """
import sys, time
sys.path.insert(0, ".")
from volcanosv_b200 import lib as L
path, n_ops, n_bytes = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
for nt in (1, 4, 0):
    t0 = time.perf_counter()
    bf = L.BamFile(path, n_threads=nt, ref="chr10")
    t1 = time.perf_counter()
    d = bf.chrom("chr10", compact=8, want_seq=False)
    t2 = time.perf_counter()
    print("threads %s: open (read + inflate + keep) %.3f s = %.0f MB/s inflated, fill ops8 %.3f s; decode-inclusive %.1f M ops/s" %
          (nt or "all", t1 - t0, n_bytes / 1e6 / (t1 - t0), t2 - t1, n_ops / 1e6 / (t2 - t0)))
    bf.close()
"""
subprocess.check_call([sys.executable, "-c", code, path, str(int(off[-1])), str(len(out))])
os.remove(path)
