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


def context(device: int = 0) -> L.Context:
    global _ctx
    if _ctx is None:
        _ctx = L.Context(device)
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


class AlignmentSource:
    """`pysam.AlignmentFile`-like view used by the operators: references() and records(chrom).
    BAM/CRAM go through pysam (htslib stays the decoder, as in the reference); SAM text
    (plain or .gz) is parsed here so the operators also run where pysam is absent."""

    def __init__(self, path: str):
        self.path = path
        self._refs: List[tuple] = []
        self._by_chrom: Dict[str, Records] = {}
        if path.endswith((".sam", ".sam.gz")):
            self._load_sam(path)
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

    def _load_pysam(self, path):
        try:
            import pysam
        except ImportError as e:  # pragma: no cover - this image has no pysam
            raise RuntimeError("BAM input needs pysam (htslib); SAM text (.sam/.sam.gz) is read natively") from e
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
    _write_lines(output_path + ".newgt", [hdr + "\tnew_gt\n"] + [r + "\t" + g + "\n" for r, g in zip(rows, new_gt)])
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
