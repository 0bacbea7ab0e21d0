"""Build-container-only harness: loads the reference's own function definitions from
/root/reference (read-only, never copied) so make_golden.py can execute them on
synthetic inputs and store the OUTPUTS as fixtures.  Not imported by any test that
runs on the GPU box (the reference does not exist there).  Recipe: SURVEY.md App. D.
"""
import ast
import logging
import os
import re
from collections import Counter, defaultdict

import numpy as np
import pandas as pd

LI = "/root/reference/bin/VolcanoSV-vc/Large_INDEL/"


def available():
    return os.path.isdir(LI)


class _Tqdm:
    def __call__(self, it=None, **kw):
        return it


class FakeCigarModule:
    """Shim for the un-vendored `cigar` package (cigar==0.1.3, R/requirement.yaml:8):
    Cigar(s).items() -> iterable of (int len, op char).  Call site SE:329."""

    class Cigar:
        def __init__(self, s):
            self.s = s

        def items(self):
            return [(int(n), c) for n, c in re.findall(r"(\d+)([MIDNSHP=X])", self.s)]


class Read:
    """Duck-typed pysam.AlignedSegment (attribute list: SURVEY.md App. D)."""

    def __init__(self, qname, pos, cigar, mapq=60, is_reverse=False, chrom="chr10",
                 seq=None, flag=None, tags=()):
        self.qname = self.query_name = qname
        self.pos = self.reference_start = pos
        self.cigar = list(cigar)
        self.mapq = mapq
        self.is_reverse = is_reverse
        self.reference_name = chrom
        self.reference_end = pos + sum(l for o, l in cigar if o in (0, 2, 3, 7, 8))
        self.seq = self.query_sequence = seq
        self.query_length = len(seq) if seq else 0
        self.flag = flag if flag is not None else (16 if is_reverse else 0)
        self._tags = list(tags)

    def get_tags(self):
        return self._tags


class FakePysam:
    """fetch(chrom[,s,e]) over an in-memory coordinate-sorted list: every record on
    `chrom` overlapping [s,e)."""

    reads = []

    class AlignmentFile:
        def __init__(self, path=None):
            pass

        def fetch(self, chrom, s=None, e=None):
            for r in FakePysam.reads:
                if r.reference_name != chrom:
                    continue
                if s is not None and not (r.reference_end > s and r.reference_start < e):
                    continue
                yield r

        def close(self):
            pass


def load_defs(script, extra=None, keep_assign=False, stable_argsort=True):
    """AST-extract FunctionDef/ClassDef (and optionally top-level Assign) nodes of a
    reference script and exec them in a fresh namespace."""
    src = open(os.path.join(LI, script)).read()
    tree = ast.parse(src)
    body = []
    for n in tree.body:
        if isinstance(n, (ast.FunctionDef, ast.ClassDef)):
            body.append(n)
        elif keep_assign and isinstance(n, ast.Assign):
            body.append(n)
    npx = np
    if stable_argsort:
        class _NP:
            """numpy with argsort pinned to kind='stable' (SURVEY.md §8c pin 1)."""

            def __getattr__(self, k):
                return getattr(np, k)

            @staticmethod
            def argsort(a, *args, **kw):
                kw.setdefault("kind", "stable")
                return np.argsort(a, *args, **kw)
        npx = _NP()
    ns = {"np": npx, "pd": pd, "Counter": Counter, "defaultdict": defaultdict,
          "tqdm": _Tqdm(), "os": os, "logging": logging, "logger": logging.getLogger("ref"),
          "cigar": FakeCigarModule, "pysam": FakePysam, "re": re}
    if extra:
        ns.update(extra)
    exec(compile(ast.Module(body, []), script, "exec"), ns)
    return ns
