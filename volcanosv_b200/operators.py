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


def context(device: Optional[int] = None) -> L.Context:
    """The process-wide library context (one process per GPU: LOCAL_RANK picks the device)."""
    global _ctx
    if _ctx is None:
        _ctx = L.Context(int(os.environ.get("LOCAL_RANK", "0")) if device is None else device)
    return _ctx


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
        if path.endswith((".sam", ".sam.gz")):
            self._load_sam(path)
        elif path.endswith(".bam"):
            self._load_bam(path)
        else:
            self._load_pysam(path)

    def _load_sam(self, path):
        op = gzip.open if path.endswith(".gz") else open
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
        names, no = bytes(d["names"]), d["name_off"]
        R.qname = [names[no[i]: no[i + 1]].decode() for i in range(n_rec)]
        sa, so = bytes(d["sa"]), d["sa_off"]
        R.sa = [sa[so[i]: so[i + 1]].decode() if so[i + 1] > so[i] else None for i in range(n_rec)]
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


_sources: Dict[str, AlignmentSource] = {}


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
def extract_reads_signature(input_path: str, output_dir: str, chr_number: int, min_siglen: int = 30):
    """LI/extract_reads_signature.py -i <bam> -o <dir> -chr N: writes
    <dir>/reads_signature/chrN_{DEL_reads_cigar,INS_reads_ciga,DEL_reads_split,INS_reads_split,reads_sig}.txt"""
    ctx = context()
    chrom = "chr%d" % chr_number
    out = os.path.join(output_dir, "reads_signature")
    os.makedirs(out, exist_ok=True)
    R = open_alignments(input_path).records(chrom)
    strand = ["-" if x else "+" for x in R.is_rev]
    # CIGAR signatures (ERS:47-83,107-145): mapq >= 50, >= 30 bp, no per-read merge
    sigs, rt, qt, xt = ctx.scan_cigar(L.MODE_READS, R.ops, R.op_off, R.pos, R.mapq, min_len=30, min_mapq=50)
    dels = _stable_by_pos(ctx, sigs[sigs["type"] == L.DEL])
    inss = _stable_by_pos(ctx, sigs[sigs["type"] == L.INS])

    def cig_rows(a, tp):
        return [[chrom, tp, int(s["pos"]), int(s["len"]), R.qname[s["rec"]], int(s["q0"]), strand[s["rec"]], "cigar"] for s in a]
    del_c, ins_c = cig_rows(dels, "DEL"), cig_rows(inss, "INS")
    fmt = lambda rows: ["\t".join(str(x) for x in r) + "\n" for r in rows]
    _write_lines("%s/%s_DEL_reads_cigar.txt" % (out, chrom), fmt(del_c))
    _write_lines("%s/%s_INS_reads_ciga.txt" % (out, chrom), fmt(ins_c))     # file name typo is the reference's
    # split-alignment signatures (ERS:147-248): every name seen more than once, no mapq filter
    cnt = Counter(R.qname)
    first_seen: Dict[str, int] = {}
    for nm in R.qname:
        if nm not in first_seen:
            first_seen[nm] = len(first_seen)
    name_id = np.asarray([first_seen[nm] for nm in R.qname], dtype=np.int32)
    select = np.asarray([1 if cnt[nm] > 1 else 0 for nm in R.qname], dtype=np.uint8)
    ref_end = (R.pos + rt).astype(np.int32)
    readlen = (qt + xt).astype(np.int32)                                     # ops {M,=,X,I,S} + H (ERS:151)
    sp = ctx.split_pairs(L.SPLIT_READS, 0, R.pos, ref_end, readlen, R.first_op, R.last_op, R.mapq, R.is_rev, name_id,
                         select) if R.n >= 2 else np.zeros(0, L.SIG_DTYPE)
    sdel = _stable_by_pos(ctx, sp[sp["type"] == L.DEL])
    sins = _stable_by_pos(ctx, sp[sp["type"] == L.INS])

    def split_rows(a, tp, label):
        return [[chrom, tp, int(s["pos"]), int(s["len"]), R.qname[s["rec"]], int(s["q0"]), int(s["q1"]), strand[s["rec"]],
                 label] for s in a]
    del_s = split_rows(sdel, "DEL", "split-alignemnt")                       # typos are the reference's (ERS:192,195)
    ins_s = split_rows(sins, "INS", "split_alignment")
    _write_lines("%s/%s_DEL_reads_split.txt" % (out, chrom), fmt(del_s))
    _write_lines("%s/%s_INS_reads_split.txt" % (out, chrom), fmt(ins_s))
    allrows = del_c + ins_c + del_s + ins_s                                  # ERS:279-282, then sort_sig
    if allrows:
        order = ctx.argsort_stable(np.asarray([r[2] for r in allrows], dtype=np.int32))
        allrows = [allrows[i] for i in order]
    _write_lines("%s/%s_reads_sig.txt" % (out, chrom), fmt(allrows))
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
        spos, slen = [], []
        with open("%s/%s_reads_sig.txt" % (signature_dir, chrom)) as f:
            for line in f:
                d = line.split()
                spos.append(int(d[2]))
                slen.append(int(d[3]))
        sup = ctx.fp_support(cpos, clen, np.asarray(spos, dtype=np.int32), np.asarray(slen, dtype=np.int32))
        out += [ln for ln, s in zip(lines, sup) if s > 0]
    _write_lines(output_path, out)


# --------------------------------------------------------------------------- SE
def _parse_sa(entry):
    """acquire_clip_pos + the SA fields organize_split_signal uses (SE:328-358)."""
    chrom, pos, strand, cg, mq, _nm = entry.split(",")
    items = [(int(n), ch) for n, ch in _CIG.findall(cg)]
    fs = items[0][0] if items[0][1] == "S" else 0
    ls = items[-1][0] if items[-1][1] == "S" else 0
    bias = sum(n for n, ch in items if ch in "MD=X")
    return chrom, int(pos), 0 if strand == "+" else 1, fs, ls, bias, int(mq)


def sig_extract(input_path: str, reference: str, work_dir: str, include_bed: Optional[str] = None, batches: int = 10_000_000,
                min_mapq: int = 20, min_read_len: int = 500, min_siglength: int = 10, min_size: int = 30,
                max_size: int = 100000, max_split_parts: int = 7):
    """LI/sig_extract.py <bam> <ref> <work_dir> [-include_bed bed]: writes DEL.sigs, INS.sigs,
    reads.sigs (SE:514-644).  merge thresholds are the script defaults (DEL 0, INS 100)."""
    ctx = context()
    if not os.path.isfile(reference):
        raise FileNotFoundError("[Errno 2] No such file: '%s'" % reference)
    os.makedirs(os.path.join(work_dir, "signatures"), exist_ok=True)
    src = open_alignments(input_path)
    bed: Optional[Dict[str, list]] = None
    if include_bed is not None:
        bed = {}
        with open(include_bed) as f:
            for line in f:
                s = line.strip().split("\t")
                bed.setdefault(s[0], []).append((int(s[1]) - 1000, int(s[2]) + 1000))          # SE:69
    del_lines, ins_lines, read_lines = set(), set(), []
    chrom_ids: Dict[str, int] = {}
    for chrom, clen in src.references():
        R = src.records(chrom)
        if R.n == 0:
            continue
        # batches own the records that START inside them (SE:522,535); a batch with no candidate
        # writes no files (SE:548-550), which only matters for reads.sigs
        edges = list(range(0, clen, batches)) if clen >= batches else [0]
        rt0 = None
        enable = (R.qlen >= min_read_len).astype(np.uint8)                                    # SE:439
        sigs, rt, qt, xt = ctx.scan_cigar(L.MODE_CUTE, R.ops, R.op_off, R.pos, R.mapq, enable=enable, min_len=min_siglength,
                                          min_mapq=min_mapq)
        ref_end = R.pos + rt + xt                                                            # CUTE totals: aux = N
        in_bed = np.ones(R.n, dtype=bool)
        if bed is not None:
            in_bed[:] = False
            for (b0, b1) in bed.get(chrom, []):
                in_bed |= ~((ref_end <= b0) | (R.pos >= b1))                                   # SE:528-533
        d, s = ctx.merge_in_record(L.MODE_CUTE, sigs)
        rows_by_batch: Dict[int, int] = {}
        inss = sigs[sigs["type"] == L.INS]
        for m in s:                                                                          # INS rows: SE:394-415
            r = int(m["rec"])
            if not in_bed[r]:
                continue
            seq = R.seq[r]
            piece = "".join(seq[int(inss[k]["q0"]):int(inss[k]["q1"])] for k in range(int(m["first"]), int(m["first"]) + int(m["count"])))
            ins_lines.add("INS\t%s\t%d\t%d\t%s\t%s\n" % (chrom, int(m["pos"]), int(m["len"]), R.qname[r], piece))
            rows_by_batch[int(R.pos[r]) // batches if clen >= batches else 0] = 1
        for m in d:
            r = int(m["rec"])
            if not in_bed[r]:
                continue
            del_lines.add("DEL\t%s\t%d\t%d\t%s\n" % (chrom, int(m["pos"]), int(m["len"]), R.qname[r]))
            rows_by_batch[int(R.pos[r]) // batches if clen >= batches else 0] = 1
        # SA-tag split analysis (SE:495-511,345-371): primary records (flag 0 / 16) carrying an SA tag
        seg_off, cols, rlen, owners = [0], [[] for _ in range(6)], [], []
        for r in range(R.n):
            if R.sa[r] is None or int(R.flag[r]) not in (0, 16) or not enable[r] or not in_bed[r]:
                continue
            segs = []
            mq_min = min_mapq
            qlen = int(R.qlen[r])
            if int(R.mapq[r]) >= min_mapq:
                f, l = int(R.first_op[r]), int(R.last_op[r])
                left = (f >> 4) if (f & 15) in (4, 5) else 0
                right = (l >> 4) if (l & 15) in (4, 5) else 0
                if int(R.flag[r]) == 0:
                    segs.append((left, qlen - right, int(R.pos[r]), int(ref_end[r]), chrom, 0))
                else:
                    segs.append((right, qlen - left, int(R.pos[r]), int(ref_end[r]), chrom, 1))
                mq_min = 0                                                                  # SE:348-350
            for e in R.sa[r].split(";")[:-1]:
                c, p, st, fs, ls, bias, mq = _parse_sa(e)
                if mq >= mq_min:
                    segs.append((fs, qlen - ls, p, p + bias, c, 0) if st == 0 else (ls, qlen - fs, p, p + bias, c, 1))
            for sg in segs:
                for k in range(4):
                    cols[k].append(sg[k])
                cols[4].append(chrom_ids.setdefault(sg[4], len(chrom_ids)))
                cols[5].append(sg[5])
            seg_off.append(seg_off[-1] + len(segs))
            rlen.append(qlen)
            owners.append(r)
        if owners:
            sa_rows = ctx.sa_analysis(seg_off, *cols, rlen, sv_size=min_size, max_size=max_size, max_parts=max_split_parts)
            for x in sa_rows:
                r = owners[int(x["rec"])]
                if int(x["type"]) == L.INS:
                    q = R.seq[r][::-1] if int(x["first"]) else R.seq[r]
                    ins_lines.add("INS\t%s\t%d\t%d\t%s\t%s\n" % (chrom, int(x["pos"]), int(x["len"]), R.qname[r],
                                                                q[int(x["q0"]):int(x["q1"])]))
                else:
                    del_lines.add("DEL\t%s\t%d\t%d\t%s\n" % (chrom, int(x["pos"]), int(x["len"]), R.qname[r]))
                rows_by_batch[int(R.pos[r]) // batches if clen >= batches else 0] = 1
        # reads.sigs: per batch (file-name order of `cat signatures/*.reads`), records with mapq >= min (SE:536-541)
        per_batch: Dict[int, List[str]] = {}
        for r in range(R.n):
            b = int(R.pos[r]) // batches if clen >= batches else 0
            if b in rows_by_batch and in_bed[r] and int(R.mapq[r]) >= min_mapq:
                per_batch.setdefault(b, []).append("%s\t%d\t%d\t%d\t%s\n" % (chrom, int(R.pos[r]), int(ref_end[r]),
                                                                            1 if int(R.flag[r]) in (0, 16) else 0, R.qname[r]))
        names = {}
        for b in per_batch:
            lo = b * batches if clen >= batches else 0
            hi = min(clen, lo + batches) if clen >= batches else clen
            names["_%s_%d_%d.reads" % (chrom, lo, hi)] = per_batch[b]
        for nm in sorted(names):
            read_lines.append((nm, names[nm]))
    key = lambda ln: (ln.split("\t")[1], int(ln.split("\t")[2]), ln)          # sort -k2,2 -k3,3n, C locale (SE:637-638)
    _write_lines(os.path.join(work_dir, "DEL.sigs"), sorted(del_lines, key=key))
    _write_lines(os.path.join(work_dir, "INS.sigs"), sorted(ins_lines, key=key))
    out = []
    for nm, lines in sorted(read_lines, key=lambda x: x[0]):
        out += lines
    _write_lines(os.path.join(work_dir, "reads.sigs"), out)


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
    ctx = context()
    sig_ins = _load_sigs(os.path.join(cutesv_dir, "INS.sigs"))
    sig_del = _load_sigs(os.path.join(cutesv_dir, "DEL.sigs"))
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
    chroms = ["chr%d" % chr_num] if chr_num is not None else ["chr%d" % i for i in range(1, 23)]
    rows, any_float = [], False
    for chrom in chroms:
        cl = calls["INS"].get(chrom)
        if cl:
            sp, sl = sig_ins.get(chrom, (np.zeros(0, np.int32), np.zeros(0, np.int32)))
            cov = ctx.ins_cov(np.asarray([c[0] for c in cl], dtype=np.int32), sp, sl, flanking)
            any_float = True
            rows += [list(c) + [float(v)] for c, v in zip(cl, cov)]
    for chrom in chroms:
        cl = calls["DEL"].get(chrom)
        if cl:
            sp, sl = sig_del.get(chrom, (np.zeros(0, np.int32), np.zeros(0, np.int32)))
            cov = ctx.del_cov(np.asarray([c[0] for c in cl], dtype=np.int32), np.asarray([c[1] for c in cl], dtype=np.int32),
                              sp, sl, flanking)
            rows += [list(c) + [int(v)] for c, v in zip(cl, cov)]
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


def correct_gt(input_path: str, output_path: str, bamfile: str, sigfile: str, dtype: str, vtype: str):
    """LI/correct_gt_del_real_data.py / correct_gt_ins_real_data.py -i <vcf> -o <tsv> -bam <bam>
    -sig <DEL.sigs|INS.sigs> -d <dtype> -v DEL|INS: writes <tsv>, <tsv>.newgt and <vcf>.newgt.<vtype>
    (INS also <sigfile>.gte30auto)."""
    ctx = context()
    is_del = vtype == "DEL"
    _, body = _read_vcf(input_path)
    lines = [ln for ln in body if "SVTYPE=%s" % vtype in ln]
    chrom = [ln.split()[0] for ln in lines]
    pos = np.asarray([int(ln.split()[1]) for ln in lines], dtype=np.int32)
    svid = [ln.split()[2] for ln in lines]
    gt = [ln.split()[-1].split(":")[0] for ln in lines]
    sv_signed = [int(ln.split("SVLEN=")[1].split(";")[0]) for ln in lines]
    svlen = np.asarray([abs(x) for x in sv_signed] if is_del else sv_signed, dtype=np.int32)
    # unique signatures in file order (GTD:65-86 / GTI:67-101: INS keeps len >= 30 on integer chrN)
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
    n_sig = len(s_pos)
    if n_sig:
        uc, up, ul, un = ctx.group_sigs(np.asarray(s_chrom, np.int32), np.asarray(s_pos, np.int32), np.asarray(s_len, np.int32))
    else:
        uc = up = ul = un = np.zeros(0, np.int32)
    inv = {v: k for k, v in ctab.items()}
    if not is_del:
        _write_lines(sigfile + ".gte30auto", ["%d\t%d\t%d\t%d\n" % (int(inv[int(c)][3:]), p, l, n) for c, p, l, n in zip(uc, up, ul, un)])
    cc = np.asarray([_chrom_id(c, ctab) for c in chrom], dtype=np.int32)
    n = len(lines)
    if n:
        if len(up):
            sup, match_id, _ = ctx.sig_support(cc, pos, svlen, uc, up, ul, un, last_in=0)
        else:
            sup, match_id = np.zeros(n, np.int32), np.zeros(n, np.int32)
        for i in range(n):                       # the reference's pysam.fetch raises on a negative start (GTI:190, GTD:160)
            lo = pos[i] - 100 if not is_del else (pos[i] - 150 if svlen[i] > 1000 else pos[i])
            if lo < 0:
                raise L.VsvError(L.E_ARG, "candidate %s is closer than %d bp to the chromosome start: the reference's "
                                          "fetch() fails here" % (svid[i], pos[i] - lo))
        depth2 = np.zeros(n, np.int32)
        ratio = np.zeros(n, np.float64)
        newgt = np.zeros(n, np.int32)
        src = open_alignments(bamfile)
        gcode = np.asarray([0 if g == "0/1" else 1 if g == "1/1" else 2 for g in gt], dtype=np.int32)
        thr_d, thr_i = GT_THRESHOLDS[dtype]["DEL"], GT_THRESHOLDS[dtype]["INS"]
        for c in dict.fromkeys(chrom):            # depth per chromosome (reads of that chromosome, BAM order)
            idx = np.asarray([i for i in range(n) if chrom[i] == c], dtype=np.int64)
            R = src.records(c)
            rend = _ref_end(ctx, R)
            d2, rr, ng = ctx.genotype(R.pos, rend, pos[idx], svlen[idx], np.full(len(idx), L.DEL if is_del else L.INS, np.int32),
                                      gcode[idx], sup[idx], thr_d, thr_i)
            depth2[idx], ratio[idx], newgt[idx] = d2, rr, ng
    else:
        sup = match_id = depth2 = newgt = np.zeros(0, np.int32)
        ratio = np.zeros(0, np.float64)
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
                         chr_num: Optional[int] = None, pre_cutesig: Optional[str] = None):
    """LI/filter_GT_correction.py -vcf <vcf> -bam <bam> -ref <ref> -dtype <Hifi|CLR|ONT> [-chr N] [-presig dir]."""
    wd = os.path.dirname(vcffile)
    if pre_cutesig is None:
        sigdir = os.path.join(wd, "cute_sig")
        if os.path.exists(sigdir):
            import shutil
            shutil.rmtree(sigdir)                                                            # FGC:64-66
        os.makedirs(sigdir)
        bed = None
        if chr_num is not None:
            with open(reference + ".fai") as f:
                chr_len = int(f.readlines()[chr_num - 1].split()[1])
            bed = os.path.join(sigdir, "sample.bed")
            _write_lines(bed, ["chr%d\t1\t%d\n" % (chr_num, chr_len)])
        sig_extract(bamfile, reference, sigdir, include_bed=bed)
    else:
        sigdir = pre_cutesig
    signature_support(vcffile, sigdir, chr_num)
    filtered = filter_by_sig_cov(vcffile, dtype.lower(), "volcano", "DEL")                     # FGC:27,141-145
    gtdir = os.path.join(wd, "GT_Correction")
    os.makedirs(gtdir, exist_ok=True)
    correct_gt(filtered, os.path.join(gtdir, "bnd_del_real.tsv"), bamfile, os.path.join(sigdir, "DEL.sigs"), dtype, "DEL")
    correct_gt(filtered, os.path.join(gtdir, "bnd_ins_real.tsv"), bamfile, os.path.join(sigdir, "INS.sigs"), dtype, "INS")
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


def _contig_one_hap(ctx, R: Records, chrom, hp, dtype, min_cigar_mapq, min_split_mapq, sig_dir):
    strand = ["-" if x else "+" for x in R.is_rev]
    has_hp = np.asarray([hp in nm for nm in R.qname], dtype=bool)
    enable = has_hp.copy()
    if dtype == "CLR":                                     # ECS_C:53-70,425-427: ins_pct <= 0.13 or mean M run >= 200
        codes, lens = (R.ops & 15).astype(np.int64), (R.ops >> 4).astype(np.int64)
        for r in range(R.n):
            a, b = int(R.op_off[r]), int(R.op_off[r + 1])
            c, l = codes[a:b], lens[a:b]
            m_sum, m_cnt, i_sum = int(l[c == 0].sum()), int((c == 0).sum()), int(l[c == 1].sum())
            ip = i_sum / (m_sum + i_sum)
            vd = m_sum / m_cnt
            enable[r] = enable[r] and (ip <= 0.13 or vd >= 200)
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
def raw_variant_call(contig_bam: str, contig_path: str, reference_path: str, output_dir: str, dtype: str, chr_num: int,
                     rbam_file: Optional[str] = None, signature_dir: Optional[str] = None, header_file: Optional[str] = None):
    """LI/Raw_variant_call.py after its alignment step (RVC:65-105; minimap2/samtools are out of scope:
    the caller hands over the sorted contig alignments)."""
    os.makedirs(output_dir, exist_ok=True)
    extract_contig_signature(contig_bam, contig_path, reference_path, output_dir, chr_num, dtype, header_file)
    header = open(header_file).readlines() if header_file else default_header()
    body = []
    import glob
    for p in sorted(glob.glob(os.path.join(output_dir, "volcano_variant_chr*vcf"))):
        body += [ln for ln in open(p) if "#" not in ln]                                     # RVC:77 `grep -v "#"`
    raw = os.path.join(output_dir, "volcano_raw_variant.vcf")
    _write_lines(raw, header + body)
    if signature_dir is None:
        extract_reads_signature(rbam_file, output_dir, chr_num)
        signature_dir = os.path.join(output_dir, "reads_signature")
    filtered = os.path.join(output_dir, "volcano_variant_filtered.vcf")
    fp_filter(raw, signature_dir, filtered)
    remove_redundancy(filtered, os.path.join(output_dir, "final_vcf"))
    return os.path.join(output_dir, "final_vcf", "volcano_variant_no_redundancy.vcf")


def _write_fai(reference: str) -> List[tuple]:
    """`samtools faidx <ref>` (MAIN:78-96): writes <ref>.fai, returns [(name, length)] in file order."""
    out = []
    with open(reference, "rb") as fa:
        data = fa.read().split(b"\n")
    rows = []
    pos, name, ln, lb, start = 0, None, 0, None, 0
    for row in data:
        if row.startswith(b">"):
            if name is not None:
                rows.append((name, ln, start, lb, lb + 1))
            name, ln, lb, start = row[1:].split()[0].decode(), 0, None, pos + len(row) + 1
        elif row:
            lb = lb or len(row)
            ln += len(row)
        pos += len(row) + 1
    if name is not None:
        rows.append((name, ln, start, lb, lb + 1))
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


def large_indel_wgs(contig_bam_pattern: str, input_dir: str, reference: str, rbam_file: str, output_dir: str, dtype: str,
                    prefix: str = "Sample", read_signature_dir: Optional[str] = None, pre_cutesig: Optional[str] = None,
                    chroms: Sequence[int] = tuple(range(1, 23)), rank: int = 0, world: int = 1, barrier=None,
                    contig_fasta_pattern: Optional[str] = None):
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
    for c in chroms:
        p = contig_bam_pattern.format(chr=c)
        weights[c] = float(os.path.getsize(p) if os.path.exists(p) else 0) + float(len(ref.get("chr%d" % c, "")))
    mine = SH.assign_lpt(weights, world)[rank]
    for c in sorted(mine):
        one = _write_chrom_ref(ref, ref_dir, c)
        contig_path = contig_fasta_pattern.format(chr=c) if contig_fasta_pattern else \
            "%s/chr%d/assembly/final_contigs/%s_final_contigs.fa" % (input_dir, c, prefix)
        raw_variant_call(contig_bam_pattern.format(chr=c), contig_path, one, os.path.join(output_dir, "chr%d" % c), dtype, c,
                         rbam_file, read_signature_dir)
    if barrier is not None:
        barrier()
    if rank != 0:
        return None
    header, body = [], []
    for c in chroms:                                                                       # MAIN:271-278
        header, b = _read_vcf(os.path.join(output_dir, "chr%d" % c, "final_vcf", "volcano_variant_no_redundancy.vcf"))
        body += b
    raw = os.path.join(output_dir, "raw_variants_wgs.vcf")
    _write_lines(raw, header + body)
    final = filter_gt_correction(raw, rbam_file, reference, dtype, None, pre_cutesig)
    outfile = os.path.join(output_dir, "%s_volcanosv_large_indel.vcf" % prefix)
    phase_vcf(final, outfile, header_file)
    return outfile


def large_indel_otherasm(input_path: str, contig_bam: str, reference: str, rbam_file: str, output_dir: str, dtype: str,
                         prefix: str = "Sample", read_signature_dir: Optional[str] = None, pre_cutesig: Optional[str] = None,
                         rank: int = 0, world: int = 1, barrier=None):
    """LI/volcanosv-vc-large-indel-otherasm.py (any haplotype-resolved assembly): ONE whole-genome contig
    FASTA (`input_path`, 'hp1'/'hp2' in the contig names) and ONE sorted contig alignment file
    (`contig_bam`; the reference makes it with minimap2 -x asm20, otherasm:232-245, out of scope).
    Contigs are split by the chromosome they align to (otherasm:249-293), then the whole-genome
    chain of `large_indel_wgs` runs with Raw_variant_call-otherasm.py's contract (-cbam)."""
    folder = os.path.join(output_dir, "contig_by_chr")
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


def large_indel_otherasm_2hap(hp1_fa: str, hp2_fa: str, contig_bam: str, reference: str, rbam_file: str, output_dir: str,
                              dtype: str, prefix: str = "Sample", read_signature_dir: Optional[str] = None,
                              pre_cutesig: Optional[str] = None, rank: int = 0, world: int = 1, barrier=None):
    """LI/volcanosv-vc-large-indel-otherasm-2hap.py: the two haplotypes arrive as two FASTA files.  The
    reference aligns each, merges the two BAMs into one sorted file (2hap:250-258, minimap2 / samtools: out
    of scope, `contig_bam` is that file) and concatenates the FASTAs to `contig_by_chr/merged.fa`
    (2hap:274-279) before the same per-chromosome split as the one-file entry (2hap:281-311)."""
    folder = os.path.join(output_dir, "contig_by_chr")
    merged = os.path.join(folder, "merged.fa")
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


def large_indel_chr(contig_bam: str, contig_path: str, reference: str, rbam_file: str, output_dir: str, dtype: str,
                    chr_num: int, prefix: str = "Sample", read_signature_dir: Optional[str] = None,
                    pre_cutesig: Optional[str] = None):
    """LI/volcanosv-vc-large-indel.py ... -chr N (MAIN:233-319) with the contig alignments supplied:
    header, per-chromosome chain, filter + genotype correction, phasing."""
    out_chr = os.path.join(output_dir, "chr%d" % chr_num)
    os.makedirs(out_chr, exist_ok=True)
    ref = load_contigs(reference)
    vcf_header = ["##fileformat=VCFv4.2\n"] + ["##contig=<ID=%s,length=%d>\n" % (k, len(v)) for k, v in ref.items()
                                             if k == "chr%d" % chr_num]
    header_file = os.path.join(out_chr, "VCF_header")
    _write_lines(header_file, vcf_header + final_header_info(prefix))
    _write_fai(reference)                                                                  # samtools faidx (MAIN:88-96)
    one = _write_chrom_ref(ref, os.path.join(out_chr, "ref_by_chr"), chr_num)
    raw = raw_variant_call(contig_bam, contig_path, one, out_chr, dtype, chr_num, rbam_file, read_signature_dir)
    final = filter_gt_correction(raw, rbam_file, reference, dtype, chr_num, pre_cutesig)
    outfile = os.path.join(out_chr, "%s_volcanosv_large_indel_chr%d.vcf" % (prefix, chr_num))
    phase_vcf(final, outfile, header_file)
    return outfile
