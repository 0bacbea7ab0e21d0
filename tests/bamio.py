"""Test-side BAM writer (SAM/BAM specification v1.6): turns SAM text into a BGZF-compressed BAM so
the native reader (volcanosv_b200.lib.BamFile) can be checked against the SAM-text loader.  Not part
of the product."""
import gzip
import re
import struct
import zlib

_OPC = {c: i for i, c in enumerate("MIDNSHP=X")}
_NT = {c: i for i, c in enumerate("=ACMGRSVTWYHKDBN")}


def _bgzf_block(data: bytes) -> bytes:
    co = zlib.compressobj(6, zlib.DEFLATED, -15)
    comp = co.compress(data) + co.flush()
    bsize = len(comp) + 25
    return (struct.pack("<BBBBIBBHBBHH", 31, 139, 8, 4, 0, 0, 255, 6, 66, 67, 2, bsize) + comp +
            struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data)))


def _reg2bin(beg, end):
    end -= 1
    for sh, off in ((14, 4681), (17, 585), (20, 73), (23, 9), (26, 1)):
        if beg >> sh == end >> sh:
            return off + (beg >> sh)
    return 0


def _aux(tag: str) -> bytes:
    k, ty, v = tag.split(":", 2)
    if ty == "i":
        return k.encode() + b"i" + struct.pack("<i", int(v))
    if ty == "f":
        return k.encode() + b"f" + struct.pack("<f", float(v))
    if ty == "A":
        return k.encode() + b"A" + v.encode()
    if ty == "B":
        sub, *vals = v.split(",")
        fmt = {"c": "b", "C": "B", "s": "h", "S": "H", "i": "i", "I": "I", "f": "f"}[sub]
        return k.encode() + b"B" + sub.encode() + struct.pack("<I", len(vals)) + b"".join(
            struct.pack("<" + fmt, float(x) if sub == "f" else int(x)) for x in vals)
    return k.encode() + b"Z" + v.encode() + b"\0"


def sam_to_bam(sam_path: str, bam_path: str, block: int = 60000, extra_tags=("NM:i:3", "ms:B:s,1,-2,3"), bai: bool = False):
    """Every mapped/unmapped record of the SAM file, in file order; `extra_tags` are added in front of
    the record's own tags so the reader has to walk past other aux types.  bai=True also writes
    <bam_path>.bai (per reference: its records as one chunk in bin 0 plus the metadata pseudo-bin; no
    linear index entries) - enough for readers that seek to the start of a reference."""
    op = gzip.open if sam_path.endswith(".gz") else open
    refs, text, recs = [], [], []
    with op(sam_path, "rt") as fh:
        for line in fh:
            if line[0] == "@":
                text.append(line)
                if line.startswith("@SQ"):
                    d = dict(x.split(":", 1) for x in line.rstrip("\n").split("\t")[1:])
                    refs.append((d["SN"], int(d["LN"])))
                continue
            recs.append(line.rstrip("\n").split("\t"))
    rid = {n: i for i, (n, _) in enumerate(refs)}
    out = bytearray(b"BAM\1")
    t = "".join(text).encode()
    out += struct.pack("<I", len(t)) + t + struct.pack("<I", len(refs))
    for n, ln in refs:
        out += struct.pack("<I", len(n) + 1) + n.encode() + b"\0" + struct.pack("<I", ln)
    rec_off = []                                   # (refID, offset of the record in the uncompressed stream)
    for f in recs:
        name, flag, rname, pos, mapq, cig, seq = f[0], int(f[1]), f[2], int(f[3]) - 1, int(f[4]), f[5], f[9]
        rec_off.append((rid.get(rname, -1), len(out)))
        cigar = [(int(n) << 4) | _OPC[c] for n, c in re.findall(r"(\d+)([MIDNSHP=X])", cig)] if cig != "*" else []
        ref_len = sum(w >> 4 for w in cigar if (w & 15) in (0, 2, 3, 7, 8))
        l_seq = 0 if seq == "*" else len(seq)
        aux = b"".join(_aux(x) for x in list(extra_tags) + f[11:])
        if len(cigar) > 65535:
            aux += b"CGBI" + struct.pack("<I", len(cigar)) + struct.pack("<%dI" % len(cigar), *cigar)
            cigar = [(l_seq << 4) | 4, (ref_len << 4) | 3]
        sq = bytearray((l_seq + 1) // 2)
        for i, c in enumerate(seq if l_seq else ""):
            sq[i >> 1] |= _NT.get(c, 15) << (4 if i % 2 == 0 else 0)
        body = struct.pack("<iiBBHHHIiii", rid.get(rname, -1), pos, len(name) + 1, mapq, _reg2bin(pos, pos + max(ref_len, 1)),
                           len(cigar), flag, l_seq, -1, -1, 0)
        body += name.encode() + b"\0" + struct.pack("<%dI" % len(cigar), *cigar) + bytes(sq) + b"\xff" * l_seq + aux
        out += struct.pack("<I", len(body)) + body
    coff = []                                      # compressed offset of every block
    with open(bam_path, "wb") as fo:
        for i in range(0, len(out), block):
            coff.append(fo.tell())
            fo.write(_bgzf_block(bytes(out[i:i + block])))
        coff.append(fo.tell())
        fo.write(_bgzf_block(b""))
    if bai:
        def voff(u):
            return (coff[u // block] << 16) | (u % block)
        with open(bam_path + ".bai", "wb") as fo:
            fo.write(b"BAI\1" + struct.pack("<i", len(refs)))
            for r in range(len(refs)):
                mine = [u for t, u in rec_off if t == r]
                if not mine:
                    fo.write(struct.pack("<ii", 0, 0))
                    continue
                nxt = [u for t, u in rec_off if u > mine[-1]]
                end = voff(nxt[0]) if nxt else voff(len(out) - 1) + 1
                fo.write(struct.pack("<i", 2))
                fo.write(struct.pack("<IiQQ", 0, 1, voff(mine[0]), end))
                fo.write(struct.pack("<IiQQQQ", 37450, 2, 1, 2, len(mine), 0))   # pseudo-bin: offsets there are not chunk starts
                fo.write(struct.pack("<i", 0))
            fo.write(struct.pack("<Q", sum(1 for t, _ in rec_off if t < 0)))
