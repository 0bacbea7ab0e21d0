"""Test-only stand-in for pysam (absent from this image), used ONLY by
tests/golden/make_e2e_golden.py to run the reference's scripts UNMODIFIED in the build
container.  AlignmentFile reads coordinate-sorted SAM text; semantics per SURVEY.md A-18:
fetch(chr) yields every record on chr in file order, fetch(chr,s,e) those overlapping [s,e).
"""
import re

_OPS = {c: i for i, c in enumerate("MIDNSHP=X")}


class AlignedSegment:
    def __init__(self, f):
        self.qname = self.query_name = f[0]
        self.flag = int(f[1])
        self.reference_name = f[2]
        self.pos = self.reference_start = int(f[3]) - 1
        self.mapq = self.mapping_quality = int(f[4])
        self.cigarstring = f[5]
        self.cigar = self.cigartuples = [(_OPS[c], int(n)) for n, c in re.findall(r"(\d+)([MIDNSHP=X])", f[5])]
        self.reference_end = self.pos + sum(l for o, l in self.cigar if o in (0, 2, 3, 7, 8))
        self.seq = self.query_sequence = None if f[9] == "*" else f[9]
        self.query_length = 0 if self.seq is None else len(self.seq)
        self.is_reverse = bool(self.flag & 16)
        self.is_unmapped = bool(self.flag & 4)
        self._tags = []
        for t in f[11:]:
            k, ty, v = t.split(":", 2)
            self._tags.append((k, int(v) if ty == "i" else v))

    def get_tags(self):
        return list(self._tags)


class AlignmentFile:
    def __init__(self, path, mode="r", **kw):
        self.refs = []
        self.reads = []
        with open(path) as fh:
            for line in fh:
                if line.startswith("@"):
                    if line.startswith("@SQ"):
                        d = dict(x.split(":", 1) for x in line.rstrip("\n").split("\t")[1:])
                        self.refs.append((d["SN"], int(d["LN"])))
                    continue
                f = line.rstrip("\n").split("\t")
                if len(f) >= 11 and f[2] != "*":
                    self.reads.append(AlignedSegment(f))
        self.references = tuple(r[0] for r in self.refs)
        self.lengths = tuple(r[1] for r in self.refs)

    def fetch(self, contig=None, start=None, stop=None, **kw):
        if start is not None and start < 0:
            raise ValueError("start out of range (%i)" % start)
        for r in self.reads:
            if contig is not None and r.reference_name != contig:
                continue
            if start is not None and not (r.reference_end > start and r.reference_start < stop):
                continue
            yield r

    def get_index_statistics(self):
        out = []
        for name, _ in self.refs:
            n = sum(1 for r in self.reads if r.reference_name == name)
            out.append((name, n, 0, n))
        return out

    def get_reference_length(self, name):
        return dict(self.refs)[name]

    def close(self):
        pass


class FastaFile:
    """references + fetch(name) over a plain FASTA (volcanosv-vc-large-indel-otherasm.py:280-282)."""

    def __init__(self, path):
        self._seq = {}
        name = None
        with open(path) as fh:
            for line in fh:
                if line.startswith(">"):
                    name = line[1:].split()[0]
                    self._seq[name] = []
                elif name is not None:
                    self._seq[name].append(line.strip())
        self._seq = {k: "".join(v) for k, v in self._seq.items()}
        self.references = tuple(self._seq)

    def fetch(self, name):
        return self._seq[name]

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
