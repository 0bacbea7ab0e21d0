"""Builds libvolcano_b200.so in-tree with nvcc for sm_100a (no JIT cache: the built
.so travels with the repo snapshot to the GPU box)."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
# profiling variants (tools/scan_ablation.sh): VSV_BUILD_TAG=_x VSV_NVCC_EXTRA="-DVSV_ABLATE_DETECT" build a separate library
TAG = os.environ.get("VSV_BUILD_TAG", "")
LIB = os.path.join(HERE, "libvolcano_b200%s.so" % TAG)
SOURCES = ["vsv_ctx.cu", "vsv_scan.cu", "vsv_merge.cu", "vsv_sa.cu", "vsv_stage2.cu", "vsv_stage3.cu", "vsv_depth.cu", "vsv_edit.cu", "vsv_bam.cu", "vsv_shard.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--extended-lambda", "-fmad=false", "-Xcompiler", "-fPIC", "-Xcompiler", "-O2"]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(HERE, "..", "include", "volcanosv_b200.h"))
    objs = []
    procs = []
    bdir = os.path.join(HERE, "build" + TAG)
    os.makedirs(bdir, exist_ok=True)
    extra = os.environ.get("VSV_NVCC_EXTRA", "").split()
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(bdir, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + hdrs):
            cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(out)
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    if force or procs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-lcudart", "-lz"]
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
