"""Host-side mirror of the reference's operator scripts (LI/ = bin/VolcanoSV-vc/Large_INDEL/).

Each function keeps the argv -> files contract of one reference script (SURVEY.md §8b,
Appendix C), so a single operator can be swapped and A/B-diffed against the reference's
output.  The arithmetic runs in libvolcano_b200.so (CUDA, sm_100a) through
volcanosv_b200.lib.Context; this module only decodes alignments (pysam when it is
installed, SAM text otherwise), packs struct-of-arrays buffers and formats text.  There is
no CPU fallback: creating the Context raises without a GPU.

    extract_reads_signature   LI/extract_reads_signature.py
    fp_filter                 LI/FP_filter_v1.py
    sig_extract               LI/sig_extract.py
    signature_support         LI/calculate_signature_support.py
    filter_by_sig_cov         LI/filter_vcf_by_sig_cov_insdel.py
    correct_gt                LI/correct_gt_{del,ins}_real_data.py
    filter_gt_correction      LI/filter_GT_correction.py          (driver)
    phase_vcf                 LI/volcanosv-vc-large-indel.py:202-231
"""
from __future__ import annotations

import gzip
import math
import os
import shutil
import re
from collections import Counter
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import lib as L

_OPC = {c: i for i, c in enumerate("MIDNSHP=X")}
_CIG = re.compile(r"(\d+)([MIDNSHP=X])")
NAN = float("nan")

# LI/para/GT_correction_para_<dtype>_<type>.txt: t_large_11, t_small_11, t_large_01, t_small_01
GT_THRESHOLDS = {
    "Hifi": dict(DEL=[0.6, 0.69, 0.71, 0.91], INS=[NAN, 0.59, 0.63, 0.79]),
    "ONT": dict(DEL=[0.61, 0.61, 0.68, 0.79], INS=[NAN, NAN, 0.67, 0.72]),
    "CLR": dict(DEL=[0.55, 0.59, 0.65, 0.75], INS=[NAN, NAN, 0.64, 0.75]),
}
# LI/filter_para.csv rows asm=volcano / other: (lb_ins, rb_ins, lb_del, rb_del)
FILTER_PARA = {
    ("other", "hifi"): (0.048, 2.61, 0.097, 2.59), ("other", "clr"): (0.0327, 2.476, 0.102, 2.638),
    ("other", "ont"): (0.191, 2.44, 0.123, 2.67), ("volcano", "hifi"): (0.097, 2.754, 0.2, 2.605),
    ("volcano", "clr"): (0.075, 2.383, 0.186, 3.018), ("volcano", "ont"): (0.206, 2.79, 0.242, 2.77),
}

_ctx: Optional[L.Context] = None
# The reference's intermediate files that are ROW DUMPS of the read-side tables (reads_signature/*.txt, cute_sig/*.sigs,
# INS.sigs.gte30auto: 10^5 - 10^7 text rows per chromosome).  Off: the tables stay on the device, a one-line stub marks
# the file, and everything downstream (which joins against the resident tables anyway) is unchanged.  The per-call
# files (VCFs, CSV, TSVs) are always written.
WRITE_ROW_DUMPS = True
_STUB = "# volcanosv_b200: table kept on the device (row dumps off)\n"


STAGE_TIMES: Dict[str, float] = {}          # seconds per stage of the last driver run (bench.py reports them; cleared by the caller)


class _stage:
    """with _stage("name"): ...  accumulates wall time into STAGE_TIMES (a stopwatch, no behaviour)."""

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        import time
        self.t0 = time.perf_counter()

    def __exit__(self, *exc):
        import time
        STAGE_TIMES[self.name] = STAGE_TIMES.get(self.name, 0.0) + time.perf_counter() - self.t0
        return False


def context(device: Optional[int] = None) -> L.Context:
    """The process-wide library context (one process per GPU: LOCAL_RANK picks the device)."""
    global _ctx
    if _ctx is None:
        _ctx = L.Context(int(os.environ.get("LOCAL_RANK", "0")) if device is None else device)
        import atexit
        from . import resident as RS
        atexit.register(_close_context)      # registered after resident's hook, so it runs first ... and closes the shards itself
        RS  # noqa: B018
    return _ctx


def _close_context():
    global _ctx
    from . import resident as RS
    RS._at_exit()
    if _ctx is not None:
        try:
            _ctx.close()
        except Exception:  # noqa: BLE001
            pass
        _ctx = None


# --------------------------------------------------------------------------- alignment input
class Records:
    """One chromosome's alignment records, packed (include/volcanosv_b200.h layout) plus the
    host-only fields (names, sequences, SA tags)."""

    def __init__(self):
        self.ops: List[int] = []
        self.op_off = [0]
        self.pos: List[int] = []
        self.mapq: List[int] = []
        self.flag: List[int] = []
        self.qname: List[str] = []
        self.seq: List[Optional[str]] = []
        self.sa: List[Optional[str]] = []
        self.first_op: List[int] = []
        self.last_op: List[int] = []
        self.name_key = None          # optional fixed-width byte-string array of the names (vectorised interning)

    def add(self, qname, flag, pos, mapq, cigar, seq, sa):
        packed = [(ln << 4) | op for op, ln in cigar]
        self.ops.extend(packed)
        self.op_off.append(len(self.ops))
        self.pos.append(pos)
        self.mapq.append(mapq)
        self.flag.append(flag)
        self.qname.append(qname)
        self.seq.append(seq)
        self.sa.append(sa)
        self.first_op.append(packed[0] if packed else 0)
        self.last_op.append(packed[-1] if packed else 0)

    @classmethod
    def from_arrays(cls, batch: dict, names=None, seqs=None):
        """Packed columns (ops u32, op_off i64, pos, mapq, flag, qlen [, name_id, next_same_name, sa_* columns]) -> Records.
        names / seqs: optional sequences indexed by record (default: 'r<name_id>' / no SEQ)."""
        R = cls()
        R.n = len(batch["pos"])
        R.ops, R.op_off, R.pos, R.mapq = batch["ops"], batch["op_off"], batch["pos"], batch["mapq"]
        R.flag = np.asarray(batch["flag"]).astype(np.int32)
        R.qlen = batch["qlen"]
        R.is_rev = ((R.flag & 16) != 0).astype(np.uint8)
        off = R.op_off
        has = off[1:] > off[:-1]
        last = max(len(R.ops) - 1, 0)
        R.first_op = np.where(has, R.ops[np.minimum(off[:-1], last)] if len(R.ops) else 0, 0).astype(np.uint32)
        R.last_op = np.where(has, R.ops[np.maximum(off[1:] - 1, 0)] if len(R.ops) else 0, 0).astype(np.uint32)
        nid = batch.get("name_id")
        R.name_key = None if nid is None else np.asarray(nid)
        R.qname = names if names is not None else _IdNames(nid if nid is not None else np.arange(R.n))
        R.seq = seqs if seqs is not None else _NoSeq(R.n)
        R.sa = _NoSeq(R.n)
        R.next_same_name = batch.get("next_same_name")
        if "sa_rec" in batch:
            R.sa_packed = {k[3:]: batch[k] for k in batch if k.startswith("sa_") and k != "sa_self_chrom"}
        return R

    def finish(self):
        self.n = len(self.pos)
        self.ops = np.asarray(self.ops, dtype=np.uint32)
        self.op_off = np.asarray(self.op_off, dtype=np.int64)
        self.pos = np.asarray(self.pos, dtype=np.int32)
        self.mapq = np.asarray(self.mapq, dtype=np.uint8)
        self.flag = np.asarray(self.flag, dtype=np.int32)
        self.first_op = np.asarray(self.first_op, dtype=np.uint32)
        self.last_op = np.asarray(self.last_op, dtype=np.uint32)
        self.is_rev = ((self.flag & 16) != 0).astype(np.uint8)
        self.qlen = np.asarray([0 if s is None else len(s) for s in self.seq], dtype=np.int32)
        return self


class _IdNames:
    """Read names of a packed batch that carries only interned ids: 'r<id>'."""

    def __init__(self, ids):
        self.ids = ids

    def __len__(self):
        return len(self.ids)

    def __getitem__(self, i):
        return "r%d" % int(self.ids[i])

    def __iter__(self):
        return (self[i] for i in range(len(self)))


class _NoSeq:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return None


class _LazyNames:
    """`R.qname` for BAM input: names decoded on demand from the concatenated name bytes."""

    def __init__(self, buf: bytes, off):
        self.buf, self.off = buf, off

    def __len__(self):
        return len(self.off) - 1

    def __getitem__(self, i):
        return self.buf[self.off[i]: self.off[i + 1]].decode()

    def __iter__(self):
        return (self[i] for i in range(len(self)))


def _name_key(names: np.ndarray, off: np.ndarray) -> np.ndarray:
    """Fixed-width byte-string array of the read names (vectorised interning key: np.unique on it)."""
    n = len(off) - 1
    lens = (off[1:] - off[:-1]).astype(np.int64)
    width = int(lens.max()) if n else 1
    pad = np.zeros((n, max(width, 1)), np.uint8)
    if len(names):
        row = np.repeat(np.arange(n, dtype=np.int64), lens)
        col = np.arange(len(names), dtype=np.int64) - np.repeat(off[:-1].astype(np.int64), lens)
        pad[row, col] = names
    return pad.view("S%d" % max(width, 1)).reshape(n)


class _LazySeq:
    """`R.seq[i]` for BAM input: decoded from the record on demand (None for SEQ '*')."""

    def __init__(self, bam, tid, qlen):
        self.bam, self.tid, self.qlen = bam, tid, qlen

    def __len__(self):
        return len(self.qlen)

    def __getitem__(self, i):
        n = int(self.qlen[i])
        return self.bam.record_seq(self.tid, int(i), n) if n else None


class AlignmentSource:
    """`pysam.AlignmentFile`-like view used by the operators: references() and records(chrom).
    BAM goes through the library's own feeder (vsv_bam_*: BGZF inflate + record -> packed columns),
    SAM text (plain or .gz) is parsed here, anything else (CRAM) through pysam if it is installed."""

    def __init__(self, path: str):
        self.path = path
        self._refs: List[tuple] = []
        self._by_chrom: Dict[str, Records] = {}
        kind = self._sniff(path)
        if kind == "sam":
            self._load_sam(path)
        elif kind == "bam":
            self._load_bam(path)
        else:
            self._load_pysam(path)

    @staticmethod
    def _sniff(path):
        """The container format from the first bytes, not the file name (a `samtools sort` stand-in may leave SAM text
        in a `.bam`): BGZF whose first block starts with BAM\1 -> bam; gzip or plain text starting with '@' / a
        record line -> sam; anything else (CRAM) -> pysam."""
        with open(path, "rb") as f:
            head = f.read(4)
        if head[:2] == b"\x1f\x8b":
            with gzip.open(path, "rb") as g:
                inner = g.read(4)
            return "bam" if inner == b"BAM\x01" else "sam"
        if head[:4] == b"CRAM":
            return "cram"
        return "sam"

    def _load_sam(self, path):
        with open(path, "rb") as f:
            zipped = f.read(2) == b"\x1f\x8b"
        op = gzip.open if zipped else open
        recs: Dict[str, Records] = {}
        with op(path, "rt") as fh:
            for line in fh:
                if line[0] == "@":
                    if line.startswith("@SQ"):
                        d = dict(x.split(":", 1) for x in line.rstrip("\n").split("\t")[1:])
                        self._refs.append((d["SN"], int(d["LN"])))
                    continue
                f = line.rstrip("\n").split("\t")
                if len(f) < 11 or f[2] == "*":
                    continue
                sa = None
                for t in f[11:]:
                    if t.startswith("SA:Z:"):
                        sa = t[5:]
                cigar = [(_OPC[c], int(n)) for n, c in _CIG.findall(f[5])]
                recs.setdefault(f[2], Records()).add(f[0], int(f[1]), int(f[3]) - 1, int(f[4]), cigar,
                                                     None if f[9] == "*" else f[9], sa)
        self._by_chrom = {c: r.finish() for c, r in recs.items()}

    def _load_bam(self, path):
        """BAM through the native feeder (include/volcanosv_b200.h: vsv_bam_*): only the header here; a
        chromosome's records are read when an operator asks for them (records()), so memory is bounded
        by one chromosome, as with `samfile.fetch(chr)`."""
        self._bam_path = path
        self._refs = [(n, ln) for n, ln, _ in L.BamFile(path, ref="").refs]

    def _bam_records(self, chrom):
        if chrom not in dict(self._refs):
            return Records().finish()
        bf = L.BamFile(self._bam_path, ref=chrom)
        tid = [r[0] for r in bf.refs].index(chrom)
        n_rec = bf.refs[tid][2]
        if n_rec == 0:
            return Records().finish()
        d = bf.chrom(tid, want_seq=False)
        R = Records()
        R.n = n_rec
        R.ops, R.op_off, R.pos, R.mapq = d["ops"], d["op_off"], d["pos"], d["mapq"]
        R.flag = d["flag"].astype(np.int32)
        R.qlen = d["qlen"]
        R.is_rev = ((R.flag & 16) != 0).astype(np.uint8)
        has = d["op_off"][1:] > d["op_off"][:-1]
        last = max(len(R.ops) - 1, 0)
        R.first_op = np.where(has, R.ops[np.minimum(d["op_off"][:-1], last)] if len(R.ops) else 0, 0).astype(np.uint32)
        R.last_op = np.where(has, R.ops[np.maximum(d["op_off"][1:] - 1, 0)] if len(R.ops) else 0, 0).astype(np.uint32)
        R.qname = _LazyNames(bytes(d["names"]), d["name_off"])
        R.name_key = _name_key(d["names"], d["name_off"])
        sa, so = bytes(d["sa"]), d["sa_off"]
        R.sa = [None] * n_rec
        for i in np.flatnonzero(so[1:] > so[:-1]):
            R.sa[i] = sa[so[i]: so[i + 1]].decode()
        R.seq = _LazySeq(bf, tid, R.qlen)
        return R

    def _load_pysam(self, path):
        try:
            import pysam
        except ImportError as e:  # pragma: no cover - this image has no pysam
            raise RuntimeError("CRAM input needs pysam (htslib); BAM and SAM text (.sam/.sam.gz) are read natively") from e
        af = pysam.AlignmentFile(path)
        self._refs = list(zip(af.references, af.lengths))
        recs: Dict[str, Records] = {}
        for r in af.fetch():
            sa = r.get_tag("SA") if r.has_tag("SA") else None
            recs.setdefault(r.reference_name, Records()).add(r.query_name, r.flag, r.reference_start, r.mapping_quality,
                                                             r.cigartuples, r.query_sequence, sa)
        self._by_chrom = {c: x.finish() for c, x in recs.items()}

    def references(self):
        return list(self._refs)

    def records(self, chrom: str) -> Records:
        if getattr(self, "_bam_path", None) is not None and chrom not in self._by_chrom:
            self._by_chrom = {chrom: self._bam_records(chrom)}      # one chromosome resident at a time
        return self._by_chrom.get(chrom) or Records().finish()


class ArraySource:
    """An alignment source that is ALREADY packed (include/volcanosv_b200.h layout): what a BAM decode hands over.
    `records_by_chrom` maps a reference name to a `Records` (see `Records.from_arrays`)."""

    def __init__(self, refs, records_by_chrom):
        self._refs = list(refs)
        self._by_chrom = dict(records_by_chrom)

    def references(self):
        return list(self._refs)

    def records(self, chrom: str) -> "Records":
        return self._by_chrom.get(chrom) or Records().finish()


_sources: Dict[str, AlignmentSource] = {}


def register_source(name: str, source) -> str:
    """Makes an in-memory source reachable under a pseudo path (the operators take paths, as the reference's scripts do)."""
    _sources[name] = source
    return name


def open_alignments(path: str) -> AlignmentSource:
    if path not in _sources:
        _sources[path] = AlignmentSource(path)
    return _sources[path]


def _ref_end(ctx, R: Records):
    """reference_end of every record (pos + sum of M,D,N,=,X), computed once per chromosome by the
    scan kernel's per-record totals."""
    if getattr(R, "_ref_end", None) is None:
        if R.n:
            _, rt, _, _ = ctx.scan_cigar(L.MODE_READS, R.ops, R.op_off, R.pos, R.mapq, min_len=1 << 26, min_mapq=255)
            R._ref_end = (R.pos + rt).astype(np.int32)
        else:
            R._ref_end = np.zeros(0, np.int32)
    return R._ref_end


def _write_lines(path, lines):
    with open(path, "w") as f:
        f.writelines(lines)


def _stable_by_pos(ctx, sigs):
    if len(sigs) == 0:
        return sigs
    return sigs[ctx.argsort_stable(np.ascontiguousarray(sigs["pos"]))]


# --------------------------------------------------------------------------- ERS
def _chrom_len(src, chrom):
    return dict(src.references()).get(chrom, 0)


def _read_shard(input_path: str, chrom: str):
    """The resident shard of one chromosome of a read alignment file (packed, uploaded and scanned once per process)."""
    from . import resident as RS
    src = open_alignments(input_path)
    return RS.get(context(), src, input_path, chrom, _chrom_len(src, chrom))


def _np_stable_by_pos(sigs):
    return sigs[np.argsort(sigs["pos"], kind="stable")] if len(sigs) else sigs


def extract_reads_signature(input_path: str, output_dir: str, chr_number: int, min_siglen: int = 30):
    """LI/extract_reads_signature.py -i <bam> -o <dir> -chr N: writes
    <dir>/reads_signature/chrN_{DEL_reads_cigar,INS_reads_ciga,DEL_reads_split,INS_reads_split,reads_sig}.txt
    from the resident shard's read-signature table (one scan serves ERS, SE, CSS, GTD and GTI)."""
    from . import resident as RS
    ctx = context()
    if min_siglen != ctx.get_params().reads_min_len:
        p = ctx.get_params()
        p.reads_min_len = min_siglen
        ctx.set_params(p)
        RS.clear()
    chrom = "chr%d" % chr_number
    out = os.path.join(output_dir, "reads_signature")
    os.makedirs(out, exist_ok=True)
    sh = _read_shard(input_path, chrom)
    R = sh.R
    counts = sh.scan(None if sh.bed == "unset" else (list(sh.bed) if sh.bed is not None else None))
    sig_file = "%s/%s_reads_sig.txt" % (out, chrom)
    if not WRITE_ROW_DUMPS:
        _write_lines(sig_file, [_STUB])
        RS.register_file(sig_file, chrom, sh)
        return None
    reads = sh.fetch(L.TAB_READS)
    nc = counts["n_reads_cigar"]
    sigs, sp = reads[:nc], reads[nc:]
    strand = np.where(R.is_rev != 0, "-", "+") if R.n else np.zeros(0, "U1")
    # CIGAR signatures (ERS:47-83,107-145): mapq >= 50, >= 30 bp, no per-read merge; (record, op) order, then sort_sig
    dels = _np_stable_by_pos(sigs[sigs["type"] == L.DEL])
    inss = _np_stable_by_pos(sigs[sigs["type"] == L.INS])

    def cig_rows(a, tp):
        return [[chrom, tp, int(p_), int(l_), R.qname[r_], int(q_), strand[r_], "cigar"]
                for p_, l_, r_, q_ in zip(a["pos"], a["len"], a["rec"], a["q0"])]
    del_c, ins_c = cig_rows(dels, "DEL"), cig_rows(inss, "INS")
    fmt = lambda rows: ["\t".join(str(x) for x in r) + "\n" for r in rows]
    _write_lines("%s/%s_DEL_reads_cigar.txt" % (out, chrom), fmt(del_c))
    _write_lines("%s/%s_INS_reads_ciga.txt" % (out, chrom), fmt(ins_c))     # file name typo is the reference's
    # split-alignment signatures (ERS:147-248): names in first-appearance order, consecutive records of a name (the engine
    # appends them unordered: restore the reference's order), then sort_sig
    if len(sp):
        sp = sp[np.lexsort((sp["rec"], sh.name_first[sh.name_id[sp["rec"]]]))]
    sdel = _np_stable_by_pos(sp[sp["type"] == L.DEL])
    sins = _np_stable_by_pos(sp[sp["type"] == L.INS])

    def split_rows(a, tp, label):
        return [[chrom, tp, int(p_), int(l_), R.qname[r_], int(q0), int(q1), strand[r_], label]
                for p_, l_, r_, q0, q1 in zip(a["pos"], a["len"], a["rec"], a["q0"], a["q1"])]
    del_s = split_rows(sdel, "DEL", "split-alignemnt")                       # typos are the reference's (ERS:192,195)
    ins_s = split_rows(sins, "INS", "split_alignment")
    _write_lines("%s/%s_DEL_reads_split.txt" % (out, chrom), fmt(del_s))
    _write_lines("%s/%s_INS_reads_split.txt" % (out, chrom), fmt(ins_s))
    allrows = del_c + ins_c + del_s + ins_s                                  # ERS:279-282, then sort_sig
    if allrows:
        order = np.argsort(np.asarray([r[2] for r in allrows], dtype=np.int64), kind="stable")
        allrows = [allrows[i] for i in order]
    _write_lines(sig_file, fmt(allrows))
    RS.register_file(sig_file, chrom, sh)
    return allrows


# --------------------------------------------------------------------------- FPF
def _read_vcf(path):
    header, body = [], []
    with open(path) as f:
        for line in f:
            (header if line[0] == "#" else body).append(line)
    return header, body


def fp_filter(input_path: str, signature_dir: str, output_path: str):
    """LI/FP_filter_v1.py -i <raw.vcf> -sigd <reads_signature dir> -o <filtered.vcf> (FPF:87-202)."""
    from . import resident as RS
    ctx = context()
    header, body = _read_vcf(input_path)
    by_chr: Dict[str, List[str]] = {}
    for line in body:
        by_chr.setdefault(line.split()[0], []).append(line)
    out = list(header)
    for i in range(1, 23):
        chrom = "chr%d" % i
        if chrom not in by_chr:
            continue
        lines = by_chr[chrom]
        cpos = np.asarray([int(x.split()[1]) for x in lines], dtype=np.int32)
        clen = np.asarray([abs(len(x.split()[4]) - len(x.split()[3])) for x in lines], dtype=np.int32)
        sig_file = "%s/%s_reads_sig.txt" % (signature_dir, chrom)
        owner = RS.owner_of(sig_file)
        if owner is not None and chrom in owner:          # written from a resident shard in this process: join on the device table
            sup = owner[chrom].candidates(L.JOIN_FP, cpos, clen, np.zeros(len(cpos), np.int32))["fp_support"]
        else:                                             # a pre-extracted signature directory (-rdsig): columns 3-4 of the file
            spos, slen = [], []
            with open(sig_file) as f:
                for line in f:
                    d = line.split()
                    spos.append(int(d[2]))
                    slen.append(int(d[3]))
            sup = ctx.fp_support(cpos, clen, np.asarray(spos, dtype=np.int32), np.asarray(slen, dtype=np.int32))
        out += [ln for ln, s in zip(lines, sup) if s > 0]
    _write_lines(output_path, out)


# --------------------------------------------------------------------------- SE
def _sigs_chrom(sh, chrom, clen, batches, min_mapq, bed_regions):
    """The `.sigs` / `.reads` lines of ONE chromosome from its resident shard (SE:514-572): (DEL line set, INS line set,
    {`_chrom_lo_hi.reads` file name: lines})."""
    R = sh.R
    # batches own the records that START inside them (SE:522,535); the regions of the BED file are attached to the
    # tasks they overlap and a record is tested against its own task's regions only (SE:58-80) - both on the device
    sh.scan(bed_regions)
    in_bed = sh.fetch(L.TAB_IN_BED).astype(bool)
    ref_end = sh.fetch(L.TAB_REF_END)
    batch_of = (R.pos // batches) if clen >= batches else np.zeros(R.n, np.int64)
    has_rows = np.zeros(int(batch_of.max()) + 1 if R.n else 1, bool)
    d_rows, i_rows = sh.fetch(L.TAB_DEL_ROWS), sh.fetch(L.TAB_INS_ROWS)
    pieces = sh._ins_piece_index() if len(i_rows) else None
    del_lines, ins_lines = set(), set()
    for m in i_rows:                                                                     # INS rows: SE:394-415 / SE:220-232
        r = int(m["rec"])
        ins_lines.add("INS\t%s\t%d\t%d\t%s\t%s\n" % (chrom, int(m["pos"]), int(m["len"]), R.qname[r], sh.ins_sequence(m, pieces)))
    for pos_, len_, r in zip(d_rows["pos"], d_rows["len"], d_rows["rec"]):
        del_lines.add("DEL\t%s\t%d\t%d\t%s\n" % (chrom, int(pos_), int(len_), R.qname[r]))
    has_rows[batch_of[d_rows["rec"]]] = True
    has_rows[batch_of[i_rows["rec"]]] = True
    # reads.sigs: per batch (file-name order of `cat signatures/*.reads`), records with mapq >= min (SE:536-541); a
    # batch without a candidate row writes no files (SE:548-550)
    keep = has_rows[batch_of] & in_bed & (R.mapq >= min_mapq)
    prim = np.isin(R.flag, (0, 16)).astype(np.int64)
    names = {}
    for bt in np.unique(batch_of[keep]):
        lo = int(bt) * batches if clen >= batches else 0
        hi = min(clen, lo + batches) if clen >= batches else clen
        idx = np.flatnonzero(keep & (batch_of == bt))
        names["_%s_%d_%d.reads" % (chrom, lo, hi)] = ["%s\t%d\t%d\t%d\t%s\n" % (chrom, int(R.pos[r]), int(ref_end[r]), int(prim[r]), R.qname[r])
                                                     for r in idx]
    return del_lines, ins_lines, names


_sigs_key = lambda ln: (ln.split("\t")[1], int(ln.split("\t")[2]), ln)          # sort -k2,2 -k3,3n, C locale (SE:637-638)


def sig_extract(input_path: str, reference: str, work_dir: str, include_bed: Optional[str] = None, batches: int = 10_000_000,
                min_mapq: int = 20, min_read_len: int = 500, min_siglength: int = 10, min_size: int = 30,
                max_size: int = 100000, max_split_parts: int = 7):
    """LI/sig_extract.py <bam> <ref> <work_dir> [-include_bed bed]: writes DEL.sigs, INS.sigs,
    reads.sigs (SE:514-644).  merge thresholds are the script defaults (DEL 0, INS 100).  The rows come from the
    resident shard of each chromosome (CIGAR scan + in-read merge + SA analysis on the device); the sequence column
    of the INS rows is cut from the reads here."""
    from . import resident as RS
    ctx = context()
    if isinstance(reference, str) and not os.path.isfile(reference):
        raise FileNotFoundError("[Errno 2] No such file: '%s'" % reference)
    os.makedirs(os.path.join(work_dir, "signatures"), exist_ok=True)
    p = ctx.get_params()
    want = (min_siglength, min_mapq, min_read_len, min_size, max_size, max_split_parts, batches)
    if want != (p.cute_min_len, p.cute_min_mapq, p.cute_min_read_len, p.sa_sv_size, p.sa_max_size, p.sa_max_parts, p.batch_size):
        (p.cute_min_len, p.cute_min_mapq, p.cute_min_read_len, p.sa_sv_size, p.sa_max_size, p.sa_max_parts, p.batch_size) = want
        ctx.set_params(p)
        RS.clear()
    src = open_alignments(input_path)
    bed: Optional[Dict[str, list]] = None
    if include_bed is not None:
        bed = {}
        with open(include_bed) as f:
            for line in f:
                x = line.strip().split("\t")
                bed.setdefault(x[0], []).append((int(x[1]) - 1000, int(x[2]) + 1000))          # SE:69
    del_lines, ins_lines, read_lines = set(), set(), []
    owners = {}
    for chrom, clen in src.references():
        sh = RS.get(ctx, src, input_path, chrom, clen)
        if sh.R.n == 0:
            continue
        owners[chrom] = sh
        if not WRITE_ROW_DUMPS:
            sh.scan(None if bed is None else sorted(bed.get(chrom, [])))
            continue
        d, i, names = _sigs_chrom(sh, chrom, clen, batches, min_mapq, None if bed is None else sorted(bed.get(chrom, [])))
        del_lines |= d
        ins_lines |= i
        for nm in sorted(names):
            read_lines.append((nm, names[nm]))
    if WRITE_ROW_DUMPS:
        _write_lines(os.path.join(work_dir, "DEL.sigs"), sorted(del_lines, key=_sigs_key))
        _write_lines(os.path.join(work_dir, "INS.sigs"), sorted(ins_lines, key=_sigs_key))
        out = []
        for nm, lines in sorted(read_lines, key=lambda x: x[0]):
            out += lines
        _write_lines(os.path.join(work_dir, "reads.sigs"), out)
    else:
        for fn in ("DEL.sigs", "INS.sigs", "reads.sigs"):
            _write_lines(os.path.join(work_dir, fn), [_STUB])
    for fn in ("DEL.sigs", "INS.sigs"):
        for chrom, sh in owners.items():
            RS.register_file(os.path.join(work_dir, fn), chrom, sh)


# --------------------------------------------------------------------------- CSS + coverage filter
def _load_sigs(path):
    """.sigs rows -> {chrom: (pos[], len[])} in file order (CSS:55-79, GTD:65-86)."""
    d: Dict[str, list] = {}
    with open(path) as f:
        for line in f:
            x = line.split()
            d.setdefault(x[1], ([], []))
            d[x[1]][0].append(int(x[2]))
            d[x[1]][1].append(int(x[3]))
    return {c: (np.asarray(p, dtype=np.int32), np.asarray(l, dtype=np.int32)) for c, (p, l) in d.items()}


def _fl(x: float) -> str:
    """pandas to_csv float text (repr: shortest round trip; inf / empty for nan)."""
    if math.isnan(x):
        return ""
    if math.isinf(x):
        return "inf" if x > 0 else "-inf"
    return repr(float(x))


def signature_support(vcffile: str, cutesv_dir: str, chr_num: Optional[int] = None, flanking: int = 1000, min_size: int = 30):
    """LI/calculate_signature_support.py -v <vcf> -ct <cute_sig dir> [-chr N]: writes
    <vcf-basename>_cutesv_sig_support_mins30_fl1000.csv (CSS:32-366)."""
    from . import resident as RS
    ctx = context()
    own_i, own_d = RS.owner_of(os.path.join(cutesv_dir, "INS.sigs")), RS.owner_of(os.path.join(cutesv_dir, "DEL.sigs"))
    resident = own_i is not None and own_d is not None and (flanking, min_size) == (ctx.get_params().cov_flanking, 30)
    # -presig (a pre-extracted cute_sig directory): the tables are read from the files; otherwise they are on the device already
    sig_ins = {} if resident else _load_sigs(os.path.join(cutesv_dir, "INS.sigs"))
    sig_del = {} if resident else _load_sigs(os.path.join(cutesv_dir, "DEL.sigs"))
    calls = _css_calls(vcffile, min_size)
    chroms = ["chr%d" % chr_num] if chr_num is not None else ["chr%d" % i for i in range(1, 23)]
    rows, any_float = [], False
    for chrom in chroms:
        cl = calls["INS"].get(chrom)
        if cl:
            starts = np.asarray([c[0] for c in cl], dtype=np.int32)
            if resident:
                cov = own_i[chrom].candidates(L.JOIN_COV, starts, np.asarray([c[2] for c in cl], np.int32), np.full(len(cl), L.INS, np.int32))["cov"] \
                    if chrom in own_i else np.zeros(len(cl), np.int64)
            else:
                sp, sl = sig_ins.get(chrom, (np.zeros(0, np.int32), np.zeros(0, np.int32)))
                cov = ctx.ins_cov(starts, sp, sl, flanking)
            any_float = True
            rows += [list(c) + [float(v)] for c, v in zip(cl, cov)]
    for chrom in chroms:
        cl = calls["DEL"].get(chrom)
        if cl and resident:
            cov = own_d[chrom].candidates(L.JOIN_COV, np.asarray([c[0] for c in cl], np.int32), np.asarray([-c[2] for c in cl], np.int32),
                                          np.full(len(cl), L.DEL, np.int32))["cov"] if chrom in own_d else np.zeros(len(cl), np.int64)
            rows += [list(c) + [int(v)] for c, v in zip(cl, cov)]
        elif cl:
            sp, sl = sig_del.get(chrom, (np.zeros(0, np.int32), np.zeros(0, np.int32)))
            cov = ctx.del_cov(np.asarray([c[0] for c in cl], dtype=np.int32), np.asarray([c[1] for c in cl], dtype=np.int32),
                              sp, sl, flanking)
            rows += [list(c) + [int(v)] for c, v in zip(cl, cov)]
    return _css_write(vcffile, rows, any_float, min_size, flanking)


def _css_calls(vcffile, min_size=30):
    """load_vcf (CSS:32-53): {type: {chrom: [(start, end, svlen, id, gt, type)]}} in file order, |SVLEN| >= min_size."""
    calls = {"INS": {}, "DEL": {}}
    _, body = _read_vcf(vcffile)
    for line in body:
        for tp in ("INS", "DEL"):
            if "SVTYPE=%s" % tp in line:
                d = line.split()
                svlen = int(d[7].split("SVLEN=")[1].split(";")[0])
                if abs(svlen) >= min_size:
                    start = int(d[1])
                    end = start + 1 if tp == "INS" else start - svlen
                    calls[tp].setdefault(d[0], []).append((start, end, svlen, d[2], d[-1].split(":")[0], tp))
    return calls


def _css_write(vcffile, rows, any_float, min_size=30, flanking=1000):
    """The CSV of CSS:361-366: the INS coverages are floats, so the shared column prints DEL sums as floats too."""
    base = os.path.basename(vcffile).split(".")[0]
    csv = os.path.join(os.path.dirname(vcffile), "%s_cutesv_sig_support_mins%d_fl%d.csv" % (base, min_size, flanking))
    out = ["start,end,svlen,svid,gt,svtype,cov,rel_cov\n"]
    for (start, end, svlen, svid, gt, tp, cov) in rows:
        covs = _fl(float(cov)) if any_float else str(int(cov))
        out.append("%d,%d,%d,%s,%s,%s,%s,%s\n" % (start, end, svlen, svid, gt, tp, covs, _fl(cov / svlen)))
    _write_lines(csv, out)
    return csv


def _median(v):
    v = sorted(v)
    n = len(v)
    # np.quantile(x, 0.5) (linear interpolation): lerp of the two middle order statistics
    if n % 2:
        return v[n // 2]
    a, b = v[n // 2 - 1], v[n // 2]
    return float(np.quantile(np.asarray([a, b], dtype=np.float64), 0.5))


def filter_by_sig_cov(input_path: str, dtype: str, asm: str = "volcano", vtype: str = "INSDEL"):
    """LI/filter_vcf_by_sig_cov_insdel.py -i <vcf> -d hifi|ont|clr -a volcano -v DEL: keeps calls
    whose relative signature coverage lies in the median-scaled band (fp64; genome-wide median)."""
    lb_ins_r, rb_ins_r, lb_del_r, rb_del_r = FILTER_PARA[(asm, dtype)]
    csv = input_path.replace(".vcf", "_cutesv_sig_support_mins30_fl1000.csv")
    rec = {"INS": [], "DEL": []}
    with open(csv) as f:
        next(f)
        for line in f:
            d = line.rstrip("\n").split(",")
            rec[d[5]].append((d[3], float(d[6]) / int(d[2])))
    keep = set()
    for tp, lo_r, hi_r, skip in (("INS", lb_ins_r, rb_ins_r, "DEL"), ("DEL", lb_del_r, rb_del_r, "INS")):
        if not rec[tp]:
            continue
        if vtype != skip:
            med = float(np.quantile(np.asarray([r for _, r in rec[tp]], dtype=np.float64), 0.5))
            lo, hi = med * lo_r, med * hi_r
            keep |= {i for i, r in rec[tp] if lo <= r <= hi}
        else:
            keep |= {i for i, _ in rec[tp]}
    outvcf = input_path.replace(".vcf", "_filter_%s.vcf" % vtype)
    header, body = _read_vcf(input_path)
    _write_lines(outvcf, header + [ln for ln in body if ln.split()[2] in keep])
    return outvcf


# --------------------------------------------------------------------------- GTD / GTI
def _chrom_id(name: str, table: Dict[str, int]) -> int:
    return table.setdefault(name, len(table))


def _correct_gt_from_file(ctx, bamfile, sigfile, is_del, dtype, chrom, pos, svlen, gcode):
    """-presig: the unique signatures come from a `.sigs` FILE (GTD:65-86 / GTI:67-101: INS keeps len >= 30 on integer chrN)."""
    ctab: Dict[str, int] = {}
    s_chrom, s_pos, s_len = [], [], []
    with open(sigfile) as f:
        for line in f:
            d = line.split()
            if not is_del:
                if int(d[3]) < 30 or not d[1][3:].isdigit():
                    continue
            s_chrom.append(_chrom_id(d[1], ctab))
            s_pos.append(int(d[2]))
            s_len.append(int(d[3]))
    if s_pos:
        uc, up, ul, un = ctx.group_sigs(np.asarray(s_chrom, np.int32), np.asarray(s_pos, np.int32), np.asarray(s_len, np.int32))
    else:
        uc = up = ul = un = np.zeros(0, np.int32)
    inv = {v: k for k, v in ctab.items()}
    if not is_del:
        _write_lines(sigfile + ".gte30auto", ["%d\t%d\t%d\t%d\n" % (int(inv[int(c)][3:]), p, l, k) for c, p, l, k in zip(uc, up, ul, un)])
    cc = np.asarray([_chrom_id(c, ctab) for c in chrom], dtype=np.int32)
    n = len(pos)
    if n == 0:
        z = np.zeros(0, np.int32)
        return z, z, z, np.zeros(0, np.float64), z
    if len(up):
        sup, match_id, _ = ctx.sig_support(cc, pos, svlen, uc, up, ul, un, last_in=0)
    else:
        sup, match_id = np.zeros(n, np.int32), np.zeros(n, np.int32)
    depth2, ratio, newgt = np.zeros(n, np.int32), np.zeros(n, np.float64), np.zeros(n, np.int32)
    src = open_alignments(bamfile)
    thr_d, thr_i = GT_THRESHOLDS[dtype]["DEL"], GT_THRESHOLDS[dtype]["INS"]
    carr = np.asarray(chrom)
    for c in dict.fromkeys(chrom):            # depth per chromosome (reads of that chromosome, BAM order)
        idx = np.flatnonzero(carr == c)
        R = src.records(c)
        rend = _ref_end(ctx, R)
        d2, rr, ng = ctx.genotype(R.pos, rend, pos[idx], svlen[idx], np.full(len(idx), L.DEL if is_del else L.INS, np.int32),
                                  gcode[idx], sup[idx], thr_d, thr_i)
        depth2[idx], ratio[idx], newgt[idx] = d2, rr, ng
    return sup, match_id, depth2, ratio, newgt


def _correct_gt_resident(owner, bamfile, sigfile, is_del, chrom, pos, svlen, gcode):
    """The `.sigs` file was written from resident shards in this process: supporting signatures, spanning depth and the
    genotype decision run against the device tables, one call per run of same-chromosome candidates.  The reference's
    last_match_idx (GTD:94-133) indexes the GENOME-WIDE unique list (chromosomes in `.sigs` order = C-locale string order):
    it is carried from call to call as a global index and handed to a shard only while it points into that shard's block."""
    tb = 0 if is_del else 1
    key = "n_udel" if is_del else "n_uins"
    chroms = sorted(c for c, sh in owner.items() if sh.counts and sh.counts[key] > 0 and (is_del or c[3:].isdigit()))
    block, off = {}, 0
    for c in chroms:
        block[c] = (off, owner[c].counts[key])
        off += owner[c].counts[key]
    if not is_del and WRITE_ROW_DUMPS:
        out = []
        for c in chroms:
            up, ul, un = owner[c].fetch(L.TAB_UINS)
            out += ["%d\t%d\t%d\t%d\n" % (int(c[3:]), p_, l_, k_) for p_, l_, k_ in zip(up, ul, un)]
        _write_lines(sigfile + ".gte30auto", out)
    n = len(pos)
    sup, match_id, depth2, newgt = (np.zeros(n, np.int32) for _ in range(4))
    ratio = np.zeros(n, np.float64)
    if n == 0:
        return sup, match_id, depth2, ratio, newgt
    typ = np.full(n, L.DEL if is_del else L.INS, np.int32)
    src = open_alignments(bamfile)
    global_last = 0
    carr = np.asarray(chrom)
    cut = np.flatnonzero(np.concatenate([[True], carr[1:] != carr[:-1]]))
    for a, b in zip(cut, list(cut[1:]) + [n]):
        c = chrom[a]
        sh = owner.get(c)
        if sh is None:                                     # a chromosome without a single read signature: still needs its reads
            sh = _read_shard(bamfile, c)
            sh.scan(None if sh.bed == "unset" else (list(sh.bed) if sh.bed is not None else None))
        b0, nu = block.get(c, (0, 0))
        li = global_last - b0 if (nu > 0 and b0 <= global_last < b0 + nu) else -1
        last_in = (li, -1) if is_del else (-1, li)
        o = sh.candidates(L.JOIN_SUPPORT | L.JOIN_GENOTYPE, pos[a:b], svlen[a:b], typ[a:b], gcode[a:b], last_in=last_in)
        sup[a:b], depth2[a:b], ratio[a:b] = o["n_support"], o["depth2"], o["ratio"]
        newgt[a:b] = o["new_gt"] if o["new_gt"] is not None else gcode[a:b]
        mid = np.asarray(o["match_id"])
        match_id[a:b] = np.where(mid >= 0, b0 + mid, global_last) if off > 0 else 0
        lo = o["last_out"][tb]
        if lo >= 0:
            global_last = b0 + lo
    return sup, match_id, depth2, ratio, newgt


def correct_gt(input_path: str, output_path: str, bamfile: str, sigfile: str, dtype: str, vtype: str):
    """LI/correct_gt_del_real_data.py / correct_gt_ins_real_data.py -i <vcf> -o <tsv> -bam <bam>
    -sig <DEL.sigs|INS.sigs> -d <dtype> -v DEL|INS: writes <tsv>, <tsv>.newgt and <vcf>.newgt.<vtype>
    (INS also <sigfile>.gte30auto)."""
    ctx = context()
    is_del = vtype == "DEL"
    lines, chrom, pos, svid, gt, svlen = _gt_calls(input_path, vtype)
    from . import resident as RS
    n = len(lines)
    gcode = np.asarray([0 if g == "0/1" else 1 if g == "1/1" else 2 for g in gt], dtype=np.int32)
    for i in range(n):                           # the reference's pysam.fetch raises on a negative start (GTI:190, GTD:160)
        lo = pos[i] - 100 if not is_del else (pos[i] - 150 if svlen[i] > 1000 else pos[i])
        if lo < 0:
            raise L.VsvError(L.E_ARG, "candidate %s is closer than %d bp to the chromosome start: the reference's "
                                      "fetch() fails here" % (svid[i], pos[i] - lo))
    ctx.set_params(L.default_params(dtype))
    owner = RS.owner_of(sigfile)
    if owner is not None:
        sup, match_id, depth2, ratio, newgt = _correct_gt_resident(owner, bamfile, sigfile, is_del, chrom, pos, svlen, gcode)
    else:
        sup, match_id, depth2, ratio, newgt = _correct_gt_from_file(ctx, bamfile, sigfile, is_del, dtype, chrom, pos, svlen, gcode)
    _gt_write(input_path, output_path, vtype, lines, chrom, pos, svid, gt, svlen, sup, match_id, depth2, ratio, newgt)


def _gt_calls(input_path, vtype):
    """The candidate lines of one type (GTD:250-289 / GTI:303-322): lines, chrom, pos, id, gt, svlen (DEL: |SVLEN|)."""
    _, body = _read_vcf(input_path)
    lines = [ln for ln in body if "SVTYPE=%s" % vtype in ln]
    chrom = [ln.split()[0] for ln in lines]
    pos = np.asarray([int(ln.split()[1]) for ln in lines], dtype=np.int32)
    svid = [ln.split()[2] for ln in lines]
    gt = [ln.split()[-1].split(":")[0] for ln in lines]
    sv_signed = [int(ln.split("SVLEN=")[1].split(";")[0]) for ln in lines]
    svlen = np.asarray([abs(x) for x in sv_signed] if vtype == "DEL" else sv_signed, dtype=np.int32)
    return lines, chrom, pos, svid, gt, svlen


def _gt_write(input_path, output_path, vtype, lines, chrom, pos, svid, gt, svlen, sup, match_id, depth2, ratio, newgt):
    """<tsv>, <tsv>.newgt and <vcf>.newgt.<vtype> (GTD:219-234,320-340,379-382; GTI:222-229,285-301,341-344)."""
    is_del = vtype == "DEL"
    n = len(lines)
    new_gt = [g if k == 2 else ("1/1" if k == 1 else "0/1") for g, k in zip(gt, newgt)]
    rows = []
    if is_del:
        any_half = bool(np.any(svlen > 1000))     # np.array(depth_list): float as soon as one mean-of-two-windows exists
        hdr = "svlen\tsvid\tcall_gt\tn_support\tn_cov\tn_ratio"
        all_int_ratio = n > 0 and bool(np.all(depth2 == 0))
        for i in range(n):
            cov = _fl(depth2[i] / 2.0) if any_half else str(int(depth2[i]) // 2)
            rat = "1" if all_int_ratio else _fl(float(ratio[i]))
            rows.append("%d\t%s\t%s\t%d\t%s\t%s" % (svlen[i], svid[i], gt[i], sup[i], cov, rat))
    else:
        hdr = "chrom\tpos\tsvlen\tcall_gt\tsvid\tmatch_id\tn_support\tn_cov\tn_ratio"
        for i in range(n):
            rows.append("%d\t%d\t%d\t%s\t%s\t%d\t%d\t%d\t%s" % (int(chrom[i][3:]), pos[i], svlen[i], gt[i], svid[i], match_id[i],
                                                               sup[i], int(depth2[i]) // 2, _fl(float(ratio[i]))))
    _write_lines(output_path, [hdr + "\n"] + [r + "\n" for r in rows])
    # the reference re-reads its TSV with pandas and writes it back with the new column (GTD:379-382,
    # GTI:341-344); pandas' default float parser is not round-trip exact, so the n_ratio TEXT of the
    # .newgt file can differ from the first file in the last digit: do the same round trip
    import pandas as pd
    df = pd.read_csv(output_path, sep="\t")
    df["new_gt"] = new_gt
    df.to_csv(output_path + ".newgt", sep="\t", index=False)
    by_id = dict(zip(svid, new_gt))
    out = []
    for ln in lines:
        d = ln.split()
        d[-1] = by_id[d[2]]
        out.append("\t".join(d) + "\n")
    _write_lines(input_path + ".newgt.%s" % vtype, out)



# --------------------------------------------------------------------------- FGC driver + final steps
def vcf_sort(lines: Sequence[str]) -> List[str]:
    """vcftools `vcf-sort` default: header, then body by `sort -k1,1d -k2,2n` (FGC:177; SURVEY §8c pin 4).
    -d (dictionary order) compares only blanks and alphanumerics of the key; ties -> whole line."""
    header = [ln for ln in lines if ln.startswith("#")]
    body = [ln for ln in lines if not ln.startswith("#")]
    dk = lambda s: "".join(ch for ch in s if ch.isalnum() or ch in " \t")
    body.sort(key=lambda ln: (dk(ln.split("\t")[0]), int(ln.split("\t")[1]), ln))
    return header + body


def filter_gt_correction(vcffile: str, bamfile: Optional[str], reference: Optional[str], dtype: str,
                         chr_num: Optional[int] = None, pre_cutesig: Optional[str] = None, rank: int = 0, world: int = 1,
                         barrier=None, mine: Optional[Sequence[int]] = None):
    """LI/filter_GT_correction.py -vcf <vcf> -bam <bam> -ref <ref> -dtype <Hifi|CLR|ONT> [-chr N] [-presig dir].
    With world > 1 (whole-genome mode, one process per GPU) every rank runs the BAM-scale half for the chromosomes in
    `mine` (`_fgc_sharded`); rank 0 only concatenates, takes the genome-wide median and writes the files."""
    if world > 1 and chr_num is None and pre_cutesig is None:
        return _fgc_sharded(vcffile, bamfile, reference, dtype, list(mine or []), rank, world, barrier)
    if rank != 0:
        return None
    wd = os.path.dirname(vcffile)
    if pre_cutesig is None:
        sigdir = os.path.join(wd, "cute_sig")
        if os.path.exists(sigdir):
            import shutil
            shutil.rmtree(sigdir)                                                            # FGC:64-66
        os.makedirs(sigdir)
        bed = None
        if chr_num is not None:
            if isinstance(reference, str):
                with open(reference + ".fai") as f:
                    chr_len = int(f.readlines()[chr_num - 1].split()[1])
            else:
                chr_len = len(reference["chr%d" % chr_num])
            bed = os.path.join(sigdir, "sample.bed")
            _write_lines(bed, ["chr%d\t1\t%d\n" % (chr_num, chr_len)])
        with _stage("fgc.sig_extract"):
            sig_extract(bamfile, reference, sigdir, include_bed=bed)
    else:
        sigdir = pre_cutesig
    with _stage("fgc.signature_support"):
        signature_support(vcffile, sigdir, chr_num)
    with _stage("fgc.filter_by_sig_cov"):
        filtered = filter_by_sig_cov(vcffile, dtype.lower(), "volcano", "DEL")                 # FGC:27,141-145
    gtdir = os.path.join(wd, "GT_Correction")
    os.makedirs(gtdir, exist_ok=True)
    with _stage("fgc.correct_gt"):
        correct_gt(filtered, os.path.join(gtdir, "bnd_del_real.tsv"), bamfile, os.path.join(sigdir, "DEL.sigs"), dtype, "DEL")
        correct_gt(filtered, os.path.join(gtdir, "bnd_ins_real.tsv"), bamfile, os.path.join(sigdir, "INS.sigs"), dtype, "INS")
    header, _ = _read_vcf(filtered)
    _write_lines(os.path.join(wd, "header"), header)
    merged = header + open(filtered + ".newgt.DEL").readlines() + open(filtered + ".newgt.INS").readlines()
    final = os.path.join(wd, "variants_filtered_GT_corrected.vcf")
    _write_lines(final, vcf_sort(merged))
    return final


def _fgc_sharded(vcffile, bamfile, reference, dtype, mine, rank, world, barrier):
    """filter_GT_correction.py for the whole genome, sharded like the reference shards it (sig_extract over 10 Mb tasks,
    SE:600-632; the depth join per candidate, GTD:201-204), here by chromosome over the ranks.  Every rank: `.sigs` rows,
    coverage support, supporting signatures, depth and genotype of ITS chromosomes from the shards it already holds.
    Rank 0: concatenation in the reference's orders, the genome-wide coverage median (filter_vcf_by_sig_cov_insdel.py:39,55)
    and the files.  Two values cross chromosome edges through the host: that median and last_match_idx (GTD:94)."""
    from . import resident as RS
    from . import shard as SH
    ctx = context()
    wd = os.path.dirname(vcffile)
    sigdir = os.path.join(wd, "cute_sig")
    parts = os.path.join(sigdir, "signatures")
    if rank == 0:
        if os.path.exists(sigdir):
            shutil.rmtree(sigdir)                                                            # FGC:64-66
        os.makedirs(parts)
    barrier()
    ctx.set_params(L.default_params(dtype))
    P = ctx.get_params()
    src = open_alignments(bamfile)
    refs = dict(src.references())
    autos = {"chr%d" % c for c in range(1, 23)}
    my = ["chr%d" % c for c in mine if "chr%d" % c in refs]
    if rank == 0:
        my += [c for c in refs if c not in autos]          # other references of the BAM: `.sigs` rows only
    calls = _css_calls(vcffile)
    shards = {}
    for chrom in my:
        sh = RS.get(ctx, src, bamfile, chrom, refs[chrom])
        if sh.R.n == 0:
            continue
        shards[chrom] = sh
        d, i, names = _sigs_chrom(sh, chrom, refs[chrom], P.batch_size, P.cute_min_mapq, None)
        _write_lines(os.path.join(parts, "_%s.DEL.part" % chrom), sorted(d, key=_sigs_key))
        _write_lines(os.path.join(parts, "_%s.INS.part" % chrom), sorted(i, key=_sigs_key))
        for nm, lines in names.items():
            _write_lines(os.path.join(parts, nm), lines)
        up, ul, un = sh.fetch(L.TAB_UINS)
        ci, cd = calls["INS"].get(chrom, []), calls["DEL"].get(chrom, [])
        cov_i = sh.candidates(L.JOIN_COV, np.asarray([c[0] for c in ci], np.int32), np.asarray([c[2] for c in ci], np.int32),
                              np.full(len(ci), L.INS, np.int32))["cov"]
        cov_d = sh.candidates(L.JOIN_COV, np.asarray([c[0] for c in cd], np.int32), np.asarray([-c[2] for c in cd], np.int32),
                              np.full(len(cd), L.DEL, np.int32))["cov"]
        np.savez(os.path.join(parts, "_%s.tables.npz" % chrom), n_udel=sh.counts["n_udel"], n_uins=sh.counts["n_uins"], uins=np.stack([up, ul, un]),
                 cov_ins=cov_i, cov_del=cov_d)
    barrier()
    chrom_files = sorted(fn[1:-len(".tables.npz")] for fn in os.listdir(parts) if fn.endswith(".tables.npz"))     # C-locale order
    tables = {c: np.load(os.path.join(parts, "_%s.tables.npz" % c)) for c in chrom_files}
    filtered = vcffile.replace(".vcf", "_filter_DEL.vcf")
    if rank == 0:
        for tp in ("DEL", "INS"):
            with open(os.path.join(sigdir, tp + ".sigs"), "w") as fo:
                for c in chrom_files:                      # lines sort by (chromosome, position, line): per-chromosome sorted parts in string order
                    with open(os.path.join(parts, "_%s.%s.part" % (c, tp))) as fi:
                        shutil.copyfileobj(fi, fo)
        with open(os.path.join(sigdir, "reads.sigs"), "w") as fo:
            for fn in sorted(f for f in os.listdir(parts) if f.endswith(".reads")):
                with open(os.path.join(parts, fn)) as fi:
                    shutil.copyfileobj(fi, fo)
        rows, any_float = [], False
        for tp, key in (("INS", "cov_ins"), ("DEL", "cov_del")):
            for c in ["chr%d" % k for k in range(1, 23)]:
                cl = calls[tp].get(c)
                if cl:
                    cov = tables[c][key] if c in tables else np.zeros(len(cl), np.int64)
                    any_float |= tp == "INS"
                    rows += [list(x) + [float(v) if tp == "INS" else int(v)] for x, v in zip(cl, cov)]
        _css_write(vcffile, rows, any_float)
        filter_by_sig_cov(vcffile, dtype.lower(), "volcano", "DEL")                           # FGC:27,141-145
    barrier()
    # ---- genotype correction of this rank's chromosomes ---------------------------------------------------------------
    gtdir = os.path.join(wd, "GT_Correction")
    res_path = lambda c, tp: os.path.join(parts, "_%s.gt.%s.npz" % (c, tp))
    per_type = {}
    for tp in ("DEL", "INS"):
        is_del = tp == "DEL"
        lines, chrom, pos, svid, gt, svlen = _gt_calls(filtered, tp)
        per_type[tp] = (lines, chrom, pos, svid, gt, svlen)
        n = len(lines)
        for i in range(n):
            lo = pos[i] - 100 if not is_del else (pos[i] - 150 if svlen[i] > 1000 else pos[i])
            if lo < 0:
                raise L.VsvError(L.E_ARG, "candidate %s is closer than %d bp to the chromosome start: the reference's "
                                          "fetch() fails here" % (svid[i], pos[i] - lo))
        gcode = np.asarray([0 if g == "0/1" else 1 if g == "1/1" else 2 for g in gt], dtype=np.int32)
        carr = np.asarray(chrom)
        key = "n_udel" if is_del else "n_uins"
        with_sigs = [c for c in chrom_files if int(tables[c][key]) > 0 and (is_del or c[3:].isdigit())]
        first_block = with_sigs[0] if with_sigs else None
        for c in my:
            idx = np.flatnonzero(carr == c)
            if len(idx) == 0:
                continue
            if len(idx) != idx[-1] - idx[0] + 1:
                raise L.VsvError(L.E_ARG, "candidates of %s are not contiguous in %s" % (c, filtered))
            sh = shards.get(c) or RS.get(ctx, src, bamfile, c, refs[c])
            sh.scan(None)
            typ = np.full(len(idx), L.DEL if is_del else L.INS, np.int32)
            out = {}
            for li in ((-1, 0) if c == first_block else (-1,)):   # the block at global index 0 may be entered with last_match_idx == 0
                o = sh.candidates(L.JOIN_SUPPORT | L.JOIN_GENOTYPE, pos[idx], svlen[idx], typ, gcode[idx],
                                  last_in=(li, -1) if is_del else (-1, li))
                tag = "a" if li < 0 else "b"
                out.update({"sup_" + tag: o["n_support"], "mid_" + tag: o["match_id"], "gt_" + tag: o["new_gt"] if o["new_gt"] is not None else gcode[idx],
                            "ratio_" + tag: o["ratio"], "last_" + tag: np.int64(o["last_out"][0 if is_del else 1]), "d2": o["depth2"]})
            np.savez(res_path(c, tp), **out)
    barrier()
    if rank != 0:
        return None
    os.makedirs(gtdir, exist_ok=True)
    for tp in ("DEL", "INS"):
        is_del = tp == "DEL"
        lines, chrom, pos, svid, gt, svlen = per_type[tp]
        n = len(lines)
        sup, match_id, depth2, newgt = (np.zeros(n, np.int32) for _ in range(4))
        ratio = np.zeros(n, np.float64)
        key = "n_udel" if is_del else "n_uins"
        with_sigs = [c for c in chrom_files if int(tables[c][key]) > 0 and (is_del or c[3:].isdigit())]
        block, off = {}, 0
        for c in with_sigs:
            block[c] = (off, int(tables[c][key]))
            off += int(tables[c][key])
        if not is_del:
            out = []
            for c in with_sigs:
                up, ul, un = tables[c]["uins"]
                out += ["%d\t%d\t%d\t%d\n" % (int(c[3:]), p_, l_, k_) for p_, l_, k_ in zip(up, ul, un)]
            _write_lines(os.path.join(sigdir, "INS.sigs.gte30auto"), out)
        carr = np.asarray(chrom)
        runs, idxs = [], []
        for c in dict.fromkeys(chrom):                 # chromosomes in VCF order (each one contiguous: checked by its rank)
            idx = np.flatnonzero(carr == c)
            r = np.load(res_path(c, tp))
            run = dict(chrom=c, a=(r["sup_a"], r["mid_a"], int(r["last_a"])))
            if "sup_b" in r.files:
                run["b"] = (r["sup_b"], r["mid_b"], int(r["last_b"]))
            runs.append(run)
            idxs.append((idx, r))
        chosen, _ = SH.assemble_support_chain(runs, block)
        for (idx, r), (tag, s_, m_) in zip(idxs, chosen):
            sup[idx], match_id[idx] = s_, m_
            depth2[idx], ratio[idx], newgt[idx] = r["d2"], r["ratio_" + tag], r["gt_" + tag]
        _gt_write(filtered, os.path.join(gtdir, "bnd_%s_real.tsv" % tp.lower()), tp, lines, chrom, pos, svid, gt, svlen, sup, match_id,
                  depth2, ratio, newgt)
    header, _ = _read_vcf(filtered)
    _write_lines(os.path.join(wd, "header"), header)
    merged = header + open(filtered + ".newgt.DEL").readlines() + open(filtered + ".newgt.INS").readlines()
    final = os.path.join(wd, "variants_filtered_GT_corrected.vcf")
    _write_lines(final, vcf_sort(merged))
    return final


def phase_vcf(infile: str, outfile: str, header_file: str):
    """MAIN:202-231: 0/1 -> 1|0 (first TIG_REGION contig is hp1) or 0|1, everything else -> 1|1; PS tag."""
    out = open(header_file).readlines()
    with open(infile) as f:
        for line in f:
            if line[0] == "#":
                continue
            d = line.split()
            hps = d[7].split("TIG_REGION=")[1].split(";")[0].split(",")
            ps = hps[0].split("_")[0][2:]
            gt = ("1|0" if "hp1" in hps[0] else "0|1") if d[-1] == "0/1" else "1|1"
            d[7] = d[7] + ";PS=" + ps
            d[-1] = gt
            out.append("\t".join(d) + "\n")
    _write_lines(outfile, out)


# --------------------------------------------------------------------------- ECS (contig signatures -> raw VCF)
_HG19 = [249250621, 243199373, 198022430, 191154276, 180915260, 171115067, 159138663, 146364022, 141213431, 135534747,
         135006516, 133851895, 115169878, 107349540, 102531392, 90354753, 81195210, 78077248, 59128983, 63025520,
         48129895, 51304566]
_INFO_LINES = [
    '##INFO=<ID=SVTYPE,Number=1,Type=String,Description="Type of SV:DEL=Deletion, CON=Contraction, INS=Insertion, DUP=Duplication, INV=Inversion">\n',
    '##INFO=<ID=SVLEN,Number=.,Type=Integer,Description="Difference in length between REF and ALT alleles">\n',
    '##INFO=<ID=TIG_REGION,Number=.,Type=String,Description="Contig region where variant was found (one per alt with h1 before h2 for homozygous calls)">\n',
    '##INFO=<ID=QUERY_STRAND,Number=.,Type=String,Description="Strand of variant in the contig relative to the reference (order follows TIG_REGION)">\n',
    '##INFO=<ID=SIG_SOURCE,Number=.,Type=String,Description="Source of the variant call signature (order follows TIG_REGION)">\n',
    '##INFO=<ID=TIG_MAPQ,Number=.,Type=String,Description="Mapping quality of the contigs (order follows TIG_REGION)">\n',
    '##INFO=<ID=CollapseId,Number=1,Type=Integer,Description="collapse match ID">\n',
]


def default_header() -> List[str]:
    """The per-chromosome VCF header the reference ships as LI/header (sample column HG002; RVC:42-43)."""
    h = ["##fileformat=VCFv4.2\n", '##FILTER=<ID=PASS,Description="All filters passed">\n', "##fileDate=20180605\n",
         "##reference=GRCh37\n"]
    h += ["##contig=<ID=chr%d,length=%d>\n" % (i + 1, n) for i, n in enumerate(_HG19)]
    h += _INFO_LINES
    h += ['##FORMAT=<ID=GT,Number=1,Type=String,Description="Genotype">\n',
          "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tHG002\n"]
    return h


def final_header_info(prefix: str) -> List[str]:
    """LI/header_info with "Sample" -> prefix (MAIN:125-129)."""
    lines = _INFO_LINES + ['##INFO=<ID=PS,Number=.,Type=Integer,Description="phase block name">\n',
                           '##FORMAT=<ID=GT,Number=1,Type=String,Description="Genotype">\n',
                           "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tSample\n"]
    return [ln.replace("Sample", prefix) for ln in lines]


def load_contigs(path: str) -> Dict[str, str]:
    """load_contigs (ECS_H:605-622): the record name is the whole header line after '>'."""
    if not isinstance(path, str):          # already in memory: {name: sequence-like}
        return path
    d: Dict[str, List[str]] = {}
    cur = None
    op = gzip.open if path.endswith(".gz") else open
    with op(path, "rt") as f:
        for line in f:
            if ">" in line:
                cur = line[1:-1]
            else:
                d.setdefault(cur, []).append(line[:-1])
    return {k: "".join(v) for k, v in d.items()}


_RC = {"N": "N", "A": "T", "T": "A", "G": "C", "C": "G"}


def _revcomp(seq: str) -> str:
    return "".join(_RC[c] for c in seq.upper()[::-1])        # other letters raise, as ECS_H:626-632


def _cluster_rows(ctx, rows, is_del):
    """sort_sig + cluster_del / cluster_ins (ECS_H:170-288) on 10-field signature rows."""
    if not rows:
        return []
    order = ctx.argsort_stable(np.asarray([r[2] for r in rows], dtype=np.int32))
    rows = [rows[i] for i in order]
    keep = ctx.cluster(is_del, [r[2] for r in rows], [r[3] for r in rows])
    return [rows[i] for i in keep]


def _sort_rows(ctx, rows):
    if not rows:
        return []
    order = ctx.argsort_stable(np.asarray([r[2] for r in rows], dtype=np.int32))
    return [rows[i] for i in order]


def clr_gate(ctx, ops, op_off):
    """The CLR contig gate (ECS_C:53-70,425-427): ins_pct = I / (M + I) <= 0.13 or var_dist = M / #M >= 200, from the
    per-record class sums of the device (vsv_cigar_class_sums); the two ratios are fp64 divisions as in the reference.  A
    contig without a single M or I op makes the reference divide by zero: such a record is simply not enabled here."""
    p = ctx.get_params()
    i_sum, m_sum, m_cnt = ctx.cigar_class_sums(ops, op_off)
    with np.errstate(divide="ignore", invalid="ignore"):
        ip = i_sum.astype(np.float64) / (m_sum + i_sum).astype(np.float64)
        vd = m_sum.astype(np.float64) / m_cnt.astype(np.float64)
    return (ip <= p.clr_max_ins_pct) | (vd >= p.clr_min_var_dist)


def _contig_one_hap(ctx, R: Records, chrom, hp, dtype, min_cigar_mapq, min_split_mapq, sig_dir):
    strand = ["-" if x else "+" for x in R.is_rev]
    has_hp = np.asarray([hp in nm for nm in R.qname], dtype=bool)
    enable = has_hp.copy()
    if dtype == "CLR" and R.n:                             # ECS_C:53-70,425-427: ins_pct <= 0.13 or mean M run >= 200
        enable &= clr_gate(ctx, R.ops, R.op_off)
    sigs, rt, qt, xt = ctx.scan_cigar(L.MODE_CONTIG, R.ops, R.op_off, R.pos, R.mapq, enable=enable.astype(np.uint8),
                                      min_len=30, min_mapq=min_cigar_mapq)
    d, s = ctx.merge_in_record(L.MODE_CONTIG, sigs)

    def rows_of(a, tp):
        return [[chrom, tp, int(x["pos"]), int(x["len"]), R.qname[x["rec"]], int(x["q0"]), int(x["q1"]), strand[x["rec"]],
                 "cigar", int(R.mapq[x["rec"]])] for x in a]
    del_c, ins_c = _sort_rows(ctx, rows_of(d, "DEL")), _sort_rows(ctx, rows_of(s, "INS"))
    fmt = lambda rows: ["\t".join(str(x) for x in r) + "\n" for r in rows]
    if sig_dir:
        _write_lines("%s/%s_DEL_contig_cigar_%s.txt" % (sig_dir, chrom, hp), fmt(del_c))
        _write_lines("%s/%s_INS_contig_cigar_%s.txt" % (sig_dir, chrom, hp), fmt(ins_c))
    del_cc, ins_cc = _cluster_rows(ctx, del_c, True), _cluster_rows(ctx, ins_c, False)
    # split alignments (ECS_H:421-476): names with >= 2 records passing hp + mapq
    ok = has_hp & (R.mapq >= min_split_mapq)
    cnt = Counter(nm for nm, k in zip(R.qname, ok) if k)
    first_seen: Dict[str, int] = {}
    for nm, k in zip(R.qname, ok):
        if k and nm not in first_seen:
            first_seen[nm] = len(first_seen)
    select = np.asarray([1 if (k and cnt[nm] > 1) else 0 for nm, k in zip(R.qname, ok)], dtype=np.uint8)
    name_id = np.asarray([first_seen.get(nm, 0) for nm in R.qname], dtype=np.int32)
    variant = {"Hifi": L.SPLIT_HIFI, "ONT": L.SPLIT_ONT, "CLR": L.SPLIT_CLR}[dtype]
    sp = ctx.split_pairs(variant, min_split_mapq, R.pos, (R.pos + rt).astype(np.int32), (qt + xt).astype(np.int32), R.first_op,
                         R.last_op, R.mapq, R.is_rev, name_id, select) if R.n >= 2 else np.zeros(0, L.SIG_DTYPE)

    def split_rows(tp, code):
        return [[chrom, tp, int(x["pos"]), int(x["len"]), R.qname[x["rec"]], int(x["q0"]), int(x["q1"]), strand[x["rec"]],
                 "split-alignment", "%d-%d" % (int(R.mapq[x["rec"]]), int(R.mapq[x["first"]]))] for x in sp if int(x["type"]) == code]
    del_s, ins_s = _sort_rows(ctx, split_rows("DEL", L.DEL)), _sort_rows(ctx, split_rows("INS", L.INS))
    if sig_dir:
        _write_lines("%s/%s_DEL_contig_split_%s.txt" % (sig_dir, chrom, hp), fmt(del_s))
        _write_lines("%s/%s_INS_contig_split_%s.txt" % (sig_dir, chrom, hp), fmt(ins_s))
    del_sc, ins_sc = _cluster_rows(ctx, del_s, True), _cluster_rows(ctx, ins_s, False)
    ins_final = _cluster_rows(ctx, ins_cc + ins_sc, False)                                     # merge_all, ECS_H:478-499
    del_final = _cluster_rows(ctx, del_cc + del_sc, True)
    return _sort_rows(ctx, ins_final + del_final)


def extract_contig_signature(bam_path: str, contig_path: str, ref_path: str, output_dir: str, chr_number: Optional[int] = None,
                             dtype: str = "Hifi", header_path: Optional[str] = None, min_cigar_mapq: int = 50,
                             min_split_mapq: int = 50):
    """LI/extract_contig_signature_{Hifi,ONT,CLR}.py -bam -contig -header -ref -o -chr: writes
    <o>/volcano_variant_chrN.vcf (+ the signature/ side files)."""
    ctx = context()
    sig_dir = os.path.join(output_dir, "signature")
    os.makedirs(sig_dir, exist_ok=True)
    dc_contig = load_contigs(contig_path)
    dc_ref = load_contigs(ref_path)
    header = open(header_path).readlines() if header_path else default_header()
    src = open_alignments(bam_path)
    for i in ([chr_number] if chr_number is not None else range(1, 23)):
        chrom = "chr%d" % i
        ref_seq = dc_ref[chrom]
        R = src.records(chrom)
        hp1 = _contig_one_hap(ctx, R, chrom, "hp1", dtype, min_cigar_mapq, min_split_mapq, sig_dir)
        hp2 = _contig_one_hap(ctx, R, chrom, "hp2", dtype, min_cigar_mapq, min_split_mapq, sig_dir)
        tcode = lambda rows: [L.DEL if r[1] == "DEL" else L.INS for r in rows]
        if hp1 and hp2:
            p1, p2 = ctx.pair_haps([r[2] for r in hp1], [r[3] for r in hp1], tcode(hp1), [r[2] for r in hp2], [r[3] for r in hp2],
                                   tcode(hp2))
        else:
            p1, p2 = [-1] * len(hp1), [-1] * len(hp2)
        info = lambda r: "%s:%d-%d" % (r[4], r[5], r[6])
        paired = []
        for k, s1 in enumerate(hp1):                                                         # ECS_H:571-592
            if p1[k] < 0:
                paired.append(s1 + ["0/1", info(s1), s1[7], s1[8], str(s1[9])])
            else:
                s2 = hp2[int(p1[k])]
                extra = ["1/1", info(s1) + "," + info(s2), s1[7] + "," + s2[7], s1[8] + "," + s2[8], str(s1[9]) + "," + str(s2[9])]
                paired.append((s1 if s1[3] > s2[3] else s2) + extra)
        for k, s2 in enumerate(hp2):
            if p2[k] < 0:
                paired.append(s2 + ["0/1", info(s2), s2[7], s2[8], str(s2[9])])
        paired = _sort_rows(ctx, paired)
        out, n_ins, n_del = list(header), 0, 0
        for sig in paired:                                                                   # add_seq_to_sig + write_vcf
            if sig[4] not in dc_contig:
                continue
            if sig[1] == "INS":
                sc = dc_contig[sig[4]]
                seq = _revcomp(sc[-sig[6]:-sig[5]]) if sig[7] == "-" else sc[sig[5]:sig[6]]
            else:
                seq = ref_seq[sig[2]:sig[2] + sig[3]]
            anchor = ref_seq[sig[2] - 1]
            if sig[1] == "DEL":
                ref_a, alt_a = anchor + seq, anchor
                n_del += 1
                k = n_del
            else:
                ref_a, alt_a = anchor, anchor + seq
                n_ins += 1
                k = n_ins
            svinfo = "SVLEN=%d;SVTYPE=%s;TIG_REGION=%s;QUERY_STRAND=%s;SIG_SOURCE=%s;TIG_MAPQ=%s" % (
                len(alt_a) - len(ref_a), sig[1], sig[11], sig[12], sig[13], sig[14])
            out.append("%s\t%d\tvolcano.%s.%s.%d\t%s\t%s\t%d\tPASS\t%s\tGT\t%s\n" % (chrom, sig[2], chrom, sig[1], k, ref_a.upper(),
                                                                                   alt_a.upper(), 20, svinfo, sig[10]))
        _write_lines(os.path.join(output_dir, "volcano_variant_chr%d.vcf" % i), out)


# --------------------------------------------------------------------------- RR (remove_redundancy)
def _sort_calls(ctx, calls):
    """sort_sig (RR:26-48): chr1..chr22, stable by POS inside each."""
    out = []
    for i in range(1, 23):
        c = [x for x in calls if x[0] == "chr%d" % i]
        if c:
            order = ctx.argsort_stable(np.asarray([x[1] for x in c], dtype=np.int32))
            out += [c[k] for k in order]
    return out


def _components(links):
    """Connected components in networkx order (RR:156-160): nodes in edge-insertion order, a
    component is numbered when its first node is reached."""
    adj: Dict[str, List[str]] = {}
    for a, b in links:
        adj.setdefault(a, []).append(b)
        adj.setdefault(b, []).append(a)
    seen, comps = set(), []
    for n in adj:
        if n in seen:
            continue
        comp, stack = [], [n]
        seen.add(n)
        while stack:
            v = stack.pop()
            comp.append(v)
            for w in adj[v]:
                if w not in seen:
                    seen.add(w)
                    stack.append(w)
        comps.append(comp)
    return comps


def remove_redundancy(input_path: str, output_dir: str, dist_thresh: int = 500, dist_thresh_del: int = 3000,
                      overlap_thresh: float = 0.0, size_sim_thresh: float = 0.5, size_sim_thresh_del: float = 0.1,
                      seq_sim_thresh: float = 0.5):
    """LI/remove_redundancy.py -i <vcf> -o <dir>: writes volcano_variant_no_redundancy.vcf and
    volcano_variant_redundancy.vcf (RR:51-291).  Insertion pairs are compared by global edit distance
    on the GPU (vsv_edit_distance_batch replaces edlib, RR:75-81).  Tie pin: among the
    longest members of a component the one earliest in VCF order is kept."""
    ctx = context()
    os.makedirs(output_dir, exist_ok=True)
    header, dels, inss, dc = [], [], [], {}
    order_of: Dict[str, int] = {}
    with open(input_path) as f:
        for line in f:
            if line[0] == "#":
                header.append(line)
                continue
            d = line.split()
            d[1] = int(d[1])
            d[3], d[4] = d[3].upper(), d[4].upper()
            order_of.setdefault(d[2], len(order_of))
            dc[d[2]] = d
            if "SVTYPE=DEL" in line:
                dels.append(d)
            elif "SVTYPE=INS" in line:
                inss.append(d)
    header = header[:-2] + ['##INFO=<ID=CollapseId,Number=1,Type=Integer,Description="collapse match ID">\n'] + header[-2:]
    dels, inss = _sort_calls(ctx, dels), _sort_calls(ctx, inss)
    svlen = lambda x: abs(len(x[3]) - len(x[4]))

    def windows(calls, thresh):
        """(i, j) with i != j, same chromosome, |POS_j - POS_i| <= thresh, in the reference's loop order."""
        for i in range(1, 23):
            c = [x for x in calls if x[0] == "chr%d" % i]
            lo = 0
            for a in range(len(c)):
                while c[lo][1] < c[a][1] - thresh:
                    lo += 1
                b = lo
                while b < len(c) and c[b][1] <= c[a][1] + thresh:
                    if b != a:
                        yield c[a], c[b]
                    b += 1
    links_del = []
    for s1, s2 in windows(dels, dist_thresh_del):
        l1, l2 = svlen(s1), svlen(s2)
        if min(l1, l2) / max(l1, l2) >= size_sim_thresh_del:
            ov = (min(s1[1] + l1, s2[1] + l2) - max(s1[1], s2[1])) / max(l1, l2)
            if ov >= overlap_thresh:
                links_del.append((s1[2], s2[2]))
    cand = []
    for s1, s2 in windows(inss, dist_thresh):
        l1, l2 = svlen(s1), svlen(s2)
        if min(l1, l2) / max(l1, l2) >= size_sim_thresh:
            cand.append((s1, s2))
    links_ins = []
    if cand:
        ids = {}
        strs = []
        for s1, s2 in cand:
            for s in (s1, s2):
                if s[2] not in ids:
                    ids[s[2]] = len(strs)
                    strs.append(s[4])
        uniq = {}
        for s1, s2 in cand:                                  # each unordered pair once
            k = (min(ids[s1[2]], ids[s2[2]]), max(ids[s1[2]], ids[s2[2]]))
            uniq.setdefault(k, len(uniq))
        keys = list(uniq)
        dist = ctx.edit_distance_batch(strs, [k[0] for k in keys], [k[1] for k in keys])
        for s1, s2 in cand:
            k = (min(ids[s1[2]], ids[s2[2]]), max(ids[s1[2]], ids[s2[2]]))
            tot = len(s1[4]) + len(s2[4])
            if (tot - int(dist[uniq[k]])) / tot >= seq_sim_thresh:
                links_ins.append((s1[2], s2[2]))

    def pick(comps):
        retain, remove = {}, {}
        for i, comp in enumerate(comps):
            best = max(comp, key=lambda x: (svlen(dc[x]), -order_of[x]))
            retain[best] = i
            for x in comp:
                if x != best:
                    remove[x] = i
        return retain, remove
    ret_d, rem_d = pick(_components(links_del))
    ret_i, rem_i = pick(_components(links_ins))
    keep, drop = [], []
    for idx, sig in dc.items():
        if idx in ret_d:
            sig[7] += ";CollapseId=DEL%d" % ret_d[idx]
            keep.append(sig)
        elif idx in ret_i:
            sig[7] += ";CollapseId=INS%d" % ret_i[idx]
            keep.append(sig)
        elif idx in rem_d:
            sig[7] += ";CollapseId=DEL%d" % rem_d[idx]
            drop.append(sig)
        elif idx in rem_i:
            sig[7] += ";CollapseId=INS%d" % rem_i[idx]
            drop.append(sig)
        else:
            keep.append(sig)
    fmt = lambda calls: ["\t".join(str(x) for x in c) + "\n" for c in _sort_calls(ctx, calls)]
    _write_lines(os.path.join(output_dir, "volcano_variant_redundancy.vcf"), header + fmt(drop))
    _write_lines(os.path.join(output_dir, "volcano_variant_no_redundancy.vcf"), header + fmt(keep))


# --------------------------------------------------------------------------- drivers (RVC, MAIN)
def align_contigs(contig_path: str, reference_path: str, output_dir: str, n_thread: int = 10, mem_per_thread: str = "1G",
                  preset: str = "asm5", bam_path: Optional[str] = None) -> str:
    """The alignment step of LI/Raw_variant_call.py (RVC:45-58; otherasm:232-245 with preset asm20), verbatim: the contig BAM is
    made by the external tools the reference uses.  Alignment itself is out of scope: without minimap2 and
    samtools on PATH this raises and names -cbam, which hands over alignments made elsewhere."""
    import subprocess
    missing = [t for t in ("minimap2", "samtools") if shutil.which(t) is None]
    if missing:
        raise RuntimeError("%s not found on PATH: volcanosv-vc-large-indel.py aligns the contigs with `minimap2 -a -x %s --cs -r2k | "
                           "samtools sort` (Raw_variant_call.py:49-58).  Install them, or pass the sorted contig alignments with -cbam."
                           % (" and ".join(missing), preset))
    os.makedirs(output_dir, exist_ok=True)
    if bam_path is None:
        bam_path = output_dir + "/" + contig_path.split("/")[-1].split(".")[0] + ".sorted.bam"
    cmd = "minimap2 -a -x %s --cs -r2k -t %d %s %s | samtools sort -@ %d -m %s > %s" % (preset, n_thread, reference_path, contig_path,
                                                                                       n_thread, mem_per_thread, bam_path)
    if subprocess.Popen(cmd, shell=True).wait() != 0 or not os.path.exists(bam_path) or os.path.getsize(bam_path) == 0:
        raise RuntimeError("contig alignment failed: %s" % cmd)
    subprocess.Popen("samtools index " + bam_path, shell=True).wait()
    return bam_path


def raw_variant_call(contig_bam: Optional[str], contig_path: str, reference_path: str, output_dir: str, dtype: str, chr_num: int,
                     rbam_file: Optional[str] = None, signature_dir: Optional[str] = None, header_file: Optional[str] = None,
                     n_thread: int = 10, mem_per_thread: str = "1G"):
    """LI/Raw_variant_call.py (RVC:45-105).  contig_bam=None runs the reference's own alignment command (`align_contigs`);
    a path skips it (-cbam: the sorted contig alignments made elsewhere)."""
    os.makedirs(output_dir, exist_ok=True)
    if contig_bam is None:
        contig_bam = align_contigs(contig_path, reference_path, output_dir, n_thread, mem_per_thread)
    with _stage("contig_signature"):
        extract_contig_signature(contig_bam, contig_path, reference_path, output_dir, chr_num, dtype, header_file)
    header = open(header_file).readlines() if header_file else default_header()
    body = []
    import glob
    for p in sorted(glob.glob(os.path.join(output_dir, "volcano_variant_chr*vcf"))):
        body += [ln for ln in open(p) if "#" not in ln]                                     # RVC:77 `grep -v "#"`
    raw = os.path.join(output_dir, "volcano_raw_variant.vcf")
    _write_lines(raw, header + body)
    if signature_dir is None:
        with _stage("reads_signature"):
            extract_reads_signature(rbam_file, output_dir, chr_num)
        signature_dir = os.path.join(output_dir, "reads_signature")
    filtered = os.path.join(output_dir, "volcano_variant_filtered.vcf")
    with _stage("fp_filter"):
        fp_filter(raw, signature_dir, filtered)
    with _stage("remove_redundancy"):
        remove_redundancy(filtered, os.path.join(output_dir, "final_vcf"))
    return os.path.join(output_dir, "final_vcf", "volcano_variant_no_redundancy.vcf")


def _write_fai(reference: str) -> List[tuple]:
    """`samtools faidx <ref>` (MAIN:78-96): writes <ref>.fai, returns [(name, length)] in file order.  As faidx: LF or CRLF
    line ends, a record without sequence lines has length 0, and a sequence line longer than the first one of its record
    (or a short line that is not the last) is an error."""
    out = []
    with open(reference, "rb") as fa:
        data = fa.read().split(b"\n")
    rows = []
    pos, name, ln, lb, lw, start, short_seen = 0, None, 0, 0, 0, 0, False
    for row in data:
        raw = len(row) + 1                     # bytes of the line incl. its LF
        body = row[:-1] if row.endswith(b"\r") else row
        if body.startswith(b">"):
            if name is not None:
                rows.append((name, ln, start, lb, lw))
            name, ln, lb, lw, start, short_seen = body[1:].split()[0].decode(), 0, 0, 0, pos + raw, False
        elif body and name is not None:
            if lb == 0:
                lb, lw = len(body), raw
            elif len(body) > lb or short_seen:
                raise ValueError("%s: different line length in sequence '%s' (samtools faidx refuses such a FASTA)" % (reference, name))
            if len(body) < lb:
                short_seen = True
            ln += len(body)
        pos += raw
    if name is not None:
        rows.append((name, ln, start, lb, lw))
    with open(reference + ".fai", "w") as f:
        for r in rows:
            f.write("%s\t%d\t%d\t%d\t%d\n" % r)
            out.append((r[0], r[1]))
    return out


def _write_chrom_ref(ref: Dict[str, str], ref_dir: str, chr_num: int) -> str:
    os.makedirs(ref_dir, exist_ok=True)
    one = os.path.join(ref_dir, "chr%d.fa" % chr_num)
    seq = ref["chr%d" % chr_num]
    _write_lines(one, [">chr%d\n" % chr_num] + [seq[i:i + 60] + "\n" for i in range(0, len(seq), 60)])
    return one


def large_indel_wgs(contig_bam_pattern: Optional[str], input_dir: str, reference: str, rbam_file: str, output_dir: str, dtype: str,
                    prefix: str = "Sample", read_signature_dir: Optional[str] = None, pre_cutesig: Optional[str] = None,
                    chroms: Sequence[int] = tuple(range(1, 23)), rank: int = 0, world: int = 1, barrier=None,
                    contig_fasta_pattern: Optional[str] = None, n_thread_align: int = 10):
    """LI/volcanosv-vc-large-indel.py WITHOUT -chr (MAIN:233-319): chr1..chr22 through the raw-call
    chain, bodies concatenated in numeric chromosome order into raw_variants_wgs.vcf, ONE
    genome-wide filter + genotype correction, phasing.

    `contig_bam_pattern` holds `{chr}` (the sorted contig alignments of each chromosome,
    RVC:49-58 output).  With world > 1 every rank (one per GPU) takes the chromosomes that
    `shard.assign_lpt` gives it (weight = size of the contig alignments + the chromosome length as a
    proxy for the reads); files meet in `output_dir`, `barrier()` (e.g. torch.distributed.barrier)
    separates the per-chromosome phase from the genome-wide one, which rank 0 runs.  Returns the
    final VCF path on rank 0, None elsewhere."""
    from . import shard as SH
    os.makedirs(output_dir, exist_ok=True)
    ref = load_contigs(reference)
    header_file = os.path.join(output_dir, "VCF_header")
    ref_dir = os.path.join(output_dir, "ref_by_chr")
    if rank == 0:
        contigs = _write_fai(reference)
        _write_lines(header_file, ["##fileformat=VCFv4.2\n"] + ["##contig=<ID=%s,length=%d>\n" % c for c in contigs] +
                     final_header_info(prefix))
    weights = {}
    fasta_of = lambda c: contig_fasta_pattern.format(chr=c) if contig_fasta_pattern else \
        "%s/chr%d/assembly/final_contigs/%s_final_contigs.fa" % (input_dir, c, prefix)
    for c in chroms:
        p = contig_bam_pattern.format(chr=c) if contig_bam_pattern else fasta_of(c)
        weights[c] = float(os.path.getsize(p) if os.path.exists(p) else 0) + float(len(ref.get("chr%d" % c, "")))
    mine = SH.assign_lpt(weights, world)[rank]
    for c in sorted(mine):
        one = _write_chrom_ref(ref, ref_dir, c)
        raw_variant_call(contig_bam_pattern.format(chr=c) if contig_bam_pattern else None, fasta_of(c), one,
                         os.path.join(output_dir, "chr%d" % c), dtype, c, rbam_file, read_signature_dir, n_thread=n_thread_align)
    if barrier is not None:
        barrier()
    raw = os.path.join(output_dir, "raw_variants_wgs.vcf")
    if rank == 0:
        header, body = [], []
        for c in chroms:                                                                   # MAIN:271-278
            header, b = _read_vcf(os.path.join(output_dir, "chr%d" % c, "final_vcf", "volcano_variant_no_redundancy.vcf"))
            body += b
        _write_lines(raw, header + body)
    if barrier is not None:
        barrier()
    # the genome-wide half (FGC:111-178): every rank works on the chromosomes whose shards it holds
    final = filter_gt_correction(raw, rbam_file, reference, dtype, None, pre_cutesig, rank=rank, world=world, barrier=barrier, mine=sorted(mine))
    if rank != 0:
        return None
    outfile = os.path.join(output_dir, "%s_volcanosv_large_indel.vcf" % prefix)
    phase_vcf(final, outfile, header_file)
    return outfile


def large_indel_otherasm(input_path: str, contig_bam: Optional[str], reference: str, rbam_file: str, output_dir: str, dtype: str,
                         prefix: str = "Sample", read_signature_dir: Optional[str] = None, pre_cutesig: Optional[str] = None,
                         rank: int = 0, world: int = 1, barrier=None, n_thread: int = 11, mem_per_thread: str = "768M"):
    """LI/volcanosv-vc-large-indel-otherasm.py (any haplotype-resolved assembly): ONE whole-genome contig
    FASTA (`input_path`, 'hp1'/'hp2' in the contig names) and ONE sorted contig alignment file
    (`contig_bam`; the reference makes it with minimap2 -x asm20, otherasm:232-245, out of scope).
    Contigs are split by the chromosome they align to (otherasm:249-293), then the whole-genome
    chain of `large_indel_wgs` runs with Raw_variant_call-otherasm.py's contract (-cbam)."""
    folder = os.path.join(output_dir, "contig_by_chr")
    if contig_bam is None:                                 # otherasm:305,328: <o>/<prefix>.sorted.bam with preset asm20
        contig_bam = output_dir + "/" + prefix + ".sorted.bam"
        if rank == 0:
            align_contigs(input_path, reference, output_dir, n_thread, mem_per_thread, "asm20", contig_bam)
        if barrier is not None:
            barrier()
    if rank == 0:
        os.makedirs(folder, exist_ok=True)
        src = open_alignments(contig_bam)
        seqs = load_contigs(input_path)
        for c in range(1, 23):
            names = list(dict.fromkeys(src.records("chr%d" % c).qname))
            if names:
                _write_lines(os.path.join(folder, "chr%d.fasta" % c), [">%s\n%s\n" % (n, seqs[n]) for n in names if n in seqs])
    if barrier is not None:
        barrier()
    return large_indel_wgs(contig_bam, "", reference, rbam_file, output_dir, dtype, prefix, read_signature_dir, pre_cutesig,
                           rank=rank, world=world, barrier=barrier, contig_fasta_pattern=os.path.join(folder, "chr{chr}.fasta"))


def large_indel_otherasm_2hap(hp1_fa: str, hp2_fa: str, contig_bam: Optional[str], reference: str, rbam_file: str, output_dir: str,
                              dtype: str, prefix: str = "Sample", read_signature_dir: Optional[str] = None,
                              pre_cutesig: Optional[str] = None, rank: int = 0, world: int = 1, barrier=None, n_thread: int = 11,
                              mem_per_thread: str = "768M"):
    """LI/volcanosv-vc-large-indel-otherasm-2hap.py: the two haplotypes arrive as two FASTA files.  The
    reference aligns each, merges the two BAMs into one sorted file (2hap:250-258, minimap2 / samtools: out
    of scope, `contig_bam` is that file) and concatenates the FASTAs to `contig_by_chr/merged.fa`
    (2hap:274-279) before the same per-chromosome split as the one-file entry (2hap:281-311)."""
    folder = os.path.join(output_dir, "contig_by_chr")
    merged = os.path.join(folder, "merged.fa")
    if contig_bam is None:                                 # 2hap:250-258: one alignment per haplotype, merged and sorted
        import subprocess
        contig_bam = output_dir + "/" + prefix + ".sorted.bam"
        if rank == 0:
            hp = [align_contigs(fa, reference, output_dir, n_thread, mem_per_thread, "asm20", output_dir + "/hp%d.bam" % (k + 1))
                  for k, fa in enumerate((hp1_fa, hp2_fa))]
            cmd = "samtools merge -u - %s %s | samtools sort -o %s;samtools index %s" % (hp[0], hp[1], contig_bam, contig_bam)
            subprocess.Popen(cmd, shell=True).wait()
            if not os.path.exists(contig_bam):
                raise RuntimeError("merging the haplotype alignments failed: %s" % cmd)
        if barrier is not None:
            barrier()
    if rank == 0:
        os.makedirs(folder, exist_ok=True)
        with open(merged, "wb") as fo:                                                     # cat hp1 hp2 > merged.fa
            for src in (hp1_fa, hp2_fa):
                with open(src, "rb") as fi:
                    shutil.copyfileobj(fi, fo)
    if barrier is not None:
        barrier()
    return large_indel_otherasm(merged, contig_bam, reference, rbam_file, output_dir, dtype, prefix, read_signature_dir,
                                pre_cutesig, rank=rank, world=world, barrier=barrier)


def large_indel_chr(contig_bam: Optional[str], contig_path: str, reference: str, rbam_file: str, output_dir: str, dtype: str,
                    chr_num: int, prefix: str = "Sample", read_signature_dir: Optional[str] = None,
                    pre_cutesig: Optional[str] = None, n_thread_align: int = 10, row_dumps: bool = True):
    """LI/volcanosv-vc-large-indel.py ... -chr N (MAIN:233-319) with the contig alignments supplied:
    header, per-chromosome chain, filter + genotype correction, phasing.
    `contig_path` / `reference` may be in-memory {name: sequence} mappings and the alignment "paths" names registered with
    `register_source` (packed records straight from a decoder); row_dumps=False keeps the read-side tables on the device
    (see WRITE_ROW_DUMPS): same final VCF, none of the 10^5-10^7-row text files in between."""
    global WRITE_ROW_DUMPS
    was = WRITE_ROW_DUMPS
    WRITE_ROW_DUMPS = bool(row_dumps)
    try:
        return _large_indel_chr(contig_bam, contig_path, reference, rbam_file, output_dir, dtype, chr_num, prefix, read_signature_dir,
                                pre_cutesig, n_thread_align)
    finally:
        WRITE_ROW_DUMPS = was


def _large_indel_chr(contig_bam, contig_path, reference, rbam_file, output_dir, dtype, chr_num, prefix, read_signature_dir, pre_cutesig,
                     n_thread_align):
    out_chr = os.path.join(output_dir, "chr%d" % chr_num)
    os.makedirs(out_chr, exist_ok=True)
    ref = load_contigs(reference)
    vcf_header = ["##fileformat=VCFv4.2\n"] + ["##contig=<ID=%s,length=%d>\n" % (k, len(v)) for k, v in ref.items()
                                             if k == "chr%d" % chr_num]
    header_file = os.path.join(out_chr, "VCF_header")
    _write_lines(header_file, vcf_header + final_header_info(prefix))
    if isinstance(reference, str):
        _write_fai(reference)                                                              # samtools faidx (MAIN:88-96)
        one = _write_chrom_ref(ref, os.path.join(out_chr, "ref_by_chr"), chr_num)
    else:
        one = ref                                                                          # in memory: nothing to split or index
    raw = raw_variant_call(contig_bam, contig_path, one, out_chr, dtype, chr_num, rbam_file, read_signature_dir, n_thread=n_thread_align)
    with _stage("filter_gt_correction"):
        final = filter_gt_correction(raw, rbam_file, reference, dtype, chr_num, pre_cutesig)
    outfile = os.path.join(out_chr, "%s_volcanosv_large_indel_chr%d.vcf" % (prefix, chr_num))
    with _stage("phase_vcf"):
        phase_vcf(final, outfile, header_file)
    return outfile
