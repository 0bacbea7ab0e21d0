"""Native BAM feeder (vsv_bam_*, host code: runs without a GPU) against the SAM-text loader on the
same records.  BAM files are written by tests/bamio.py from the committed SAM fixtures."""
import gzip
import os
import sys

import numpy as np
import pytest


sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bamio  # noqa: E402
from volcanosv_b200 import lib as L

HERE = os.path.dirname(os.path.abspath(__file__))
E2E = os.path.join(HERE, "golden", "e2e", "input")


def sam_records(path):
    refs, recs = [], []
    with gzip.open(path, "rt") as fh:
        for line in fh:
            if line[0] == "@":
                if line.startswith("@SQ"):
                    d = dict(x.split(":", 1) for x in line.rstrip("\n").split("\t")[1:])
                    refs.append((d["SN"], int(d["LN"])))
                continue
            recs.append(line.rstrip("\n").split("\t"))
    return refs, recs


def check(bam, sam, compact):
    refs, recs = sam_records(sam)
    bf = L.BamFile(bam, n_threads=3)
    assert [(n, ln) for n, ln, _ in bf.refs] == refs
    for tid, (chrom, _, n_rec) in enumerate(bf.refs):
        mine = [f for f in recs if f[2] == chrom]
        assert n_rec == len(mine)
        d = bf.chrom(chrom, compact=compact)
        ops = d["ops"] if not compact else None
        if compact == 8:
            assert np.all(np.diff(d["exc_idx"]) > 0) and np.all(d["ops8"][d["exc_idx"]] == 0xFF)
            ops = L.unpack_ops8(d["ops8"], d["exc_idx"], d["exc_val"])
        elif compact == 12:
            assert np.all(np.diff(d["exc_idx"]) > 0) and len(d["ops12"]) == L.ops12_bytes(d["n_ops"])
            ops = L.unpack_ops12(d["ops12"], d["n_ops"], d["exc_idx"], d["exc_val"])
        elif compact:
            ops = d["ops16"].astype(np.uint32)
            assert np.all(np.diff(d["exc_idx"]) > 0)
            assert np.all((d["ops16"][d["exc_idx"]] >> 4) == 0xFFF)
            ops[d["exc_idx"]] = d["exc_val"]
        names = bytes(d["names"])
        seq = bytes(d["seq"])
        sa = bytes(d["sa"])
        for i, f in enumerate(mine):
            cig = [(int(n) << 4) | bamio._OPC[c] for n, c in bamio.re.findall(r"(\d+)([MIDNSHP=X])", f[5])]
            assert d["pos"][i] == int(f[3]) - 1 and d["mapq"][i] == int(f[4]) and d["flag"][i] == int(f[1])
            assert list(ops[d["op_off"][i]: d["op_off"][i + 1]]) == cig
            assert names[d["name_off"][i]: d["name_off"][i + 1]].decode() == f[0]
            want_seq = "" if f[9] == "*" else f[9]
            assert d["qlen"][i] == len(want_seq)
            assert seq[d["seq_off"][i]: d["seq_off"][i + 1]].decode() == want_seq
            want_sa = "".join(t[5:] for t in f[11:] if t.startswith("SA:Z:"))
            assert sa[d["sa_off"][i]: d["sa_off"][i + 1]].decode() == want_sa
    bf.close()


@pytest.mark.parametrize("compact", [False, True, 8, 12])
def test_reads_and_contigs_round_trip(tmp_path, compact):
    for name in ("reads", "contigs"):
        bam = str(tmp_path / (name + ".bam"))
        bamio.sam_to_bam(os.path.join(E2E, name + ".sam.gz"), bam, block=20000 if name == "reads" else 65000)
        check(bam, os.path.join(E2E, name + ".sam.gz"), compact)


def test_long_cigar_unmapped_and_long_ops(tmp_path):
    """CG:B,I long CIGAR (> 65535 ops), an unmapped record (skipped), ops of 4095+ bases (exception list)."""
    rng = np.random.default_rng(5)
    n = 70000
    cig = "".join("%d%s" % (int(l), "MID"[i % 3] if i % 3 else "M") for i, l in enumerate(rng.integers(1, 9, n)))
    cig = re_fix(cig)
    qlen = sum(int(a) for a, c in bamio.re.findall(r"(\d+)([MIDNSHP=X])", cig) if c in "MIS=X")
    sam = tmp_path / "x.sam"
    with open(sam, "w") as f:
        f.write("@HD\tVN:1.6\n@SQ\tSN:chrA\tLN:900000\n@SQ\tSN:chrB\tLN:5000000\n")
        f.write("long\t0\tchrA\t11\t60\t%s\t*\t0\t0\t%s\t*\tSA:Z:chrB,5,+,10S20M,60,1;\n" % (cig, "ACGT" * (qlen // 4) + "A" * (qlen % 4)))
        f.write("nomap\t4\t*\t0\t0\t*\t*\t0\t0\tACGTN\t*\n")
        f.write("big\t16\tchrB\t101\t7\t5S4095M70000D4094I100000N3M\t*\t0\t0\t*\t*\n")
    with open(sam, "rb") as fi, gzip.open(str(sam) + ".gz", "wb") as fo:
        fo.write(fi.read())
    bam = str(tmp_path / "x.bam")
    bamio.sam_to_bam(str(sam) + ".gz", bam)
    check(bam, str(sam) + ".gz", False)
    check(bam, str(sam) + ".gz", True)
    check(bam, str(sam) + ".gz", 8)
    d = L.BamFile(bam).chrom("chrB", compact=True)
    assert len(d["exc_idx"]) == 3 and list(d["exc_val"] >> 4) == [4095, 70000, 100000]


def re_fix(cig):
    """merge adjacent equal ops (a valid CIGAR never repeats an op)"""
    out = []
    for n, c in bamio.re.findall(r"(\d+)([MIDNSHP=X])", cig):
        if out and out[-1][1] == c:
            out[-1][0] += int(n)
        else:
            out.append([int(n), c])
    return "".join("%d%s" % (n, c) for n, c in out)


def test_errors(tmp_path):
    with pytest.raises(L.VsvError):
        L.BamFile(str(tmp_path / "missing.bam"))
    p = tmp_path / "bad.bam"
    p.write_bytes(b"not a bam file at all, just bytes")
    with pytest.raises(L.VsvError):
        L.BamFile(str(p))
    with gzip.open(str(tmp_path / "plain.gz"), "wb") as f:     # gzip but not BGZF
        f.write(b"BAM\1")
    with pytest.raises(L.VsvError):
        L.BamFile(str(tmp_path / "plain.gz"))


def test_alignment_source_bam_equals_sam(tmp_path):
    """operators.AlignmentSource: a .bam through the native feeder == the same records from SAM text."""
    from volcanosv_b200 import operators as OP
    for nm in ("reads", "contigs"):
        bam = str(tmp_path / (nm + ".bam"))
        bamio.sam_to_bam(os.path.join(E2E, nm + ".sam.gz"), bam)
        a = OP.AlignmentSource(os.path.join(E2E, nm + ".sam.gz"))
        b = OP.AlignmentSource(bam)
        assert a.references() == b.references()
        A, B = a.records("chr1"), b.records("chr1")
        for k in ("ops", "op_off", "pos", "mapq", "flag", "qlen", "is_rev", "first_op", "last_op"):
            assert np.array_equal(getattr(A, k), getattr(B, k)), k
        assert A.n == B.n and list(A.qname) == list(B.qname) and A.sa == B.sa
        assert all(A.seq[i] == B.seq[i] for i in range(0, A.n, 5))
        assert b.records("chrNone").n == 0


def _columns(bf, chrom):
    d = bf.chrom(chrom)
    return {k: (bytes(v) if v.dtype == np.uint8 and k in ("names", "seq", "sa") else v.tolist()) for k, v in d.items()}


@pytest.mark.parametrize("window", [None, "700", "70000"])
def test_open_one_reference_streaming(tmp_path, monkeypatch, window):
    """vsv_bam_open_ref: one reference of a 22-reference BAM == the same reference of the full load; with the
    .bai (seek), without it (scan + stop after the reference), and with read windows far smaller than a
    BGZF block / a record (VSV_BAM_WINDOW), so records and blocks straddle window edges."""
    if window:
        monkeypatch.setenv("VSV_BAM_WINDOW", window)
    wgs = os.path.join(HERE, "golden", "e2e_wgs", "input")
    bam = str(tmp_path / "reads.bam")
    bamio.sam_to_bam(os.path.join(wgs, "reads.sam.gz"), bam, block=30000, bai=True)
    full = L.BamFile(bam, n_threads=2)
    assert len(full.refs) == 22
    for chrom in ("chr1", "chr10", "chr22"):
        one = L.BamFile(bam, n_threads=2, ref=chrom)
        assert [(n, ln) for n, ln, _ in one.refs] == [(n, ln) for n, ln, _ in full.refs]
        assert [n for _, _, n in one.refs if n] == [dict((a, c) for a, _, c in full.refs)[chrom]]
        assert _columns(one, chrom) == _columns(full, chrom)
    os.remove(bam + ".bai")
    one = L.BamFile(bam, ref="chr7")
    assert _columns(one, "chr7") == _columns(full, "chr7")
    hdr = L.BamFile(bam, ref="")
    assert len(hdr.refs) == 22 and all(n == 0 for _, _, n in hdr.refs)
    with pytest.raises(L.VsvError):
        L.BamFile(bam, ref="chrUn")


def _raw_bam(records, refs=(("chrA", 100000),), l_text_override=None, n_ref_override=None):
    """A one-block BAM from hand-made record bodies (malformed on purpose)."""
    import struct
    out = bytearray(b"BAM\1")
    out += struct.pack("<I", 0 if l_text_override is None else l_text_override)
    out += struct.pack("<I", len(refs) if n_ref_override is None else n_ref_override)
    for n, ln in refs:
        out += struct.pack("<I", len(n) + 1) + n.encode() + b"\0" + struct.pack("<I", ln)
    for body in records:
        out += struct.pack("<I", len(body)) + body
    return bamio._bgzf_block(bytes(out)) + bamio._bgzf_block(b"")


def _cg_record(n_claimed, n_present, l_seq=8):
    """<l_seq>S<n>N placeholder CIGAR + a CG:B,I tag that CLAIMS n_claimed ops but carries n_present."""
    import struct
    name = b"r\0"
    cigar = [(l_seq << 4) | 4, (50 << 4) | 3]
    body = struct.pack("<iiBBHHHIiii", 0, 10, len(name), 60, 4681, len(cigar), 0, l_seq, -1, -1, 0)
    body += name + struct.pack("<2I", *cigar) + bytes((l_seq + 1) // 2) + b"\xff" * l_seq
    body += b"CGBI" + struct.pack("<I", n_claimed) + struct.pack("<%dI" % n_present, *([(5 << 4) | 0] * n_present))
    return body


def test_malformed_records_are_rejected_not_overread(tmp_path):
    """a CG:B,I count that runs past the record, a scalar tag cut short, implausible header sizes -> an
    error (or a clean load), never a read past the record."""
    good = tmp_path / "good.bam"
    good.write_bytes(_raw_bam([_cg_record(3, 3)]))
    bf = L.BamFile(str(good))
    d = bf.chrom("chrA")
    assert d["op_off"].tolist() == [0, 3] and d["ops"].tolist() == [(5 << 4)] * 3
    for name, blob in (("cg_count_past_record", _raw_bam([_cg_record(1 << 20, 3)])),
                       ("cg_count_huge", _raw_bam([_cg_record(0xFFFFFFF0, 1)])),
                       ("scalar_tag_cut", _raw_bam([_cg_record(3, 3)[:-10] + b"XYi\x01"])),
                       ("l_text_huge", _raw_bam([_cg_record(3, 3)], l_text_override=0x7FFFFFF0)),
                       ("n_ref_huge", _raw_bam([_cg_record(3, 3)], n_ref_override=0x7FFFFFF0))):
        p = tmp_path / (name + ".bam")
        p.write_bytes(blob)
        with pytest.raises(L.VsvError):
            f = L.BamFile(str(p))
            f.chrom("chrA")


def test_stale_index_falls_back_to_the_linear_scan(tmp_path):
    """A .bai of another file (different reference count) or one whose offsets point at records of another reference
    must not make the feeder return part of a chromosome."""
    wgs = os.path.join(HERE, "golden", "e2e_wgs", "input")
    bam = str(tmp_path / "reads.bam")
    bamio.sam_to_bam(os.path.join(wgs, "reads.sam.gz"), bam, block=30000, bai=True)
    full = L.BamFile(bam, n_threads=2)
    want = _columns(full, "chr7")
    good = open(bam + ".bai", "rb").read()
    # (a) index with another reference count
    import struct
    open(bam + ".bai", "wb").write(b"BAI\1" + struct.pack("<i", 3) + good[8:])
    assert _columns(L.BamFile(bam, ref="chr7"), "chr7") == want
    # (b) same count, but chr7's chunk start replaced by chr2's: the record there is not chr7's
    other = str(tmp_path / "other.bam")
    bamio.sam_to_bam(os.path.join(wgs, "reads.sam.gz"), other, block=30000, bai=True)
    idx = bytearray(open(other + ".bai", "rb").read())
    # per reference: n_bin(4) [bin(4) n_chunk(4) chunks(16 each)]* n_intv(4); the writer emits 2 bins: (0, 1 chunk) and the pseudo-bin (2 chunks)
    q, starts = 8, []
    for r in range(22):
        n_bin = struct.unpack_from("<i", idx, q)[0]
        q += 4
        for _ in range(n_bin):
            b_, n_chunk = struct.unpack_from("<Ii", idx, q)
            q += 8
            if b_ != 37450:
                starts.append(q)
            q += 16 * n_chunk
        q += 4 + 8 * struct.unpack_from("<i", idx, q)[0]
    idx[starts[6]: starts[6] + 8] = idx[starts[1]: starts[1] + 8]
    open(bam + ".bai", "wb").write(bytes(idx))
    assert _columns(L.BamFile(bam, ref="chr7"), "chr7") == want


def test_htsbox_written_bam(tmp_path):
    """BAM + BAI written by the reference's vendored htsbox (htslib lite-r346), not by tests/bamio.py: the native feeder
    must give the same columns as the SAM-text loader - whole file, and one reference at a time through htslib's own index
    (real bins, chunks and linear index; multi-member BGZF as htslib cuts it)."""
    from volcanosv_b200 import operators as OP
    hb = os.path.join(HERE, "golden", "htsbox")
    wgs_sam = os.path.join(HERE, "golden", "e2e_wgs", "input", "reads.sam.gz")
    a = OP.AlignmentSource(wgs_sam)
    b = OP.AlignmentSource(os.path.join(hb, "reads_wgs.bam"))
    assert a.references() == b.references() and len(a.references()) == 22
    for chrom, _ in a.references():                       # b.records() opens one reference through the .bai each time
        A, B = a.records(chrom), b.records(chrom)
        assert A.n == B.n and A.n > 0
        for k in ("ops", "op_off", "pos", "mapq", "flag", "qlen", "is_rev", "first_op", "last_op"):
            assert np.array_equal(getattr(A, k), getattr(B, k)), (chrom, k)
        assert list(A.qname) == list(B.qname) and A.sa == B.sa
        assert all(A.seq[i] == B.seq[i] for i in range(0, A.n, 7))
    full = L.BamFile(os.path.join(hb, "reads_wgs.bam"), n_threads=2)
    one = L.BamFile(os.path.join(hb, "reads_wgs.bam"), n_threads=2, ref="chr13")
    assert _columns(one, "chr13") == _columns(full, "chr13")
    # without the index: linear scan, stops after the reference
    shutil_bam = str(tmp_path / "noidx.bam")
    import shutil
    shutil.copy(os.path.join(hb, "reads_wgs.bam"), shutil_bam)
    assert _columns(L.BamFile(shutil_bam, ref="chr13"), "chr13") == _columns(full, "chr13")
    # contigs with real sequences (4-bit SEQ decode incl. odd lengths)
    ca = OP.AlignmentSource(os.path.join(E2E, "contigs.sam.gz")).records("chr1")
    cb = OP.AlignmentSource(os.path.join(hb, "contigs_chr1.bam")).records("chr1")
    assert ca.n == cb.n and np.array_equal(ca.ops, cb.ops) and list(ca.qname) == list(cb.qname)
    assert all(ca.seq[i] == cb.seq[i] for i in range(ca.n))
