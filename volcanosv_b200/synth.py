"""Seeded synthetic alignment records of the shapes BASELINE.md / SURVEY.md §8d name.

There is no network, so the zenodo chr10 BAMs cannot be fetched; every test and the
benchmark run on these generators (numpy PCG64, fixed seeds).  Records come out in
the packed struct-of-arrays layout of include/volcanosv_b200.h.
"""
from __future__ import annotations

import numpy as np

CHR10_LEN = 135_534_747          # hg19 chr10, LI/header:14

# profile -> (mean read length, ops per record, large-SV op rate, 10..29bp op rate)
PROFILES = {
    "hifi": dict(read_len=15_000, ops=150, sv_rate=0.004, mid_rate=0.004, cov=30),
    "ont": dict(read_len=20_000, ops=2_400, sv_rate=0.0010, mid_rate=0.006, cov=50),
    "clr": dict(read_len=12_000, ops=3_000, sv_rate=0.0010, mid_rate=0.008, cov=89),
    "contig": dict(read_len=300_000, ops=3_000, sv_rate=0.002, mid_rate=0.004, cov=2),
}


def n_records(profile: str, chrom_len: int = CHR10_LEN) -> int:
    p = PROFILES[profile]
    return int(chrom_len * p["cov"] / p["read_len"])


def synth_records(seed: int, n_rec: int, profile: str = "hifi", chrom_len: int = CHR10_LEN,
                  eqx: bool = False, clip_rate: float = 0.3, n_names: int | None = None):
    """Vectorised generator.  Each record: [clip] M (I|D M)* [clip]; match lengths are
    drawn so the record spans ~read_len; I/D lengths are geometric (1..9 bp) with
    `mid_rate` 10..29 bp and `sv_rate` >=30 bp (Pareto tail, <=50 kb).  Records are
    sorted by position (BAM coordinate order).  eqx=True writes '='/'X' instead of 'M'
    for half of the match ops (read BAMs only; the contig scanner rejects them).
    Returns dict of numpy arrays: ops u32, op_off i64, pos i32, mapq u8, flag u16,
    name_id i32, qlen i32, hp u8.
    """
    p = PROFILES[profile]
    rng = np.random.Generator(np.random.PCG64(seed))
    # odd inner op count: M (X M)*
    k = np.maximum(1, rng.poisson(p["ops"] / 2, n_rec)).astype(np.int64)
    inner = 2 * k + 1
    lead = rng.random(n_rec) < clip_rate
    tail = rng.random(n_rec) < clip_rate
    nops = inner + lead + tail
    op_off = np.zeros(n_rec + 1, dtype=np.int64)
    np.cumsum(nops, out=op_off[1:])
    total = int(op_off[-1])
    rec_of = np.repeat(np.arange(n_rec, dtype=np.int64), nops)
    idx = np.arange(total, dtype=np.int64) - op_off[rec_of]
    j = idx - lead[rec_of]                      # index inside the inner run
    is_lead = (idx == 0) & lead[rec_of]
    is_tail = (idx == nops[rec_of] - 1) & tail[rec_of]
    is_match = (~is_lead) & (~is_tail) & ((j & 1) == 0)
    is_var = (~is_lead) & (~is_tail) & ((j & 1) == 1)

    code = np.zeros(total, dtype=np.uint32)
    length = np.zeros(total, dtype=np.int64)
    # matches
    mlen = p["read_len"] / np.maximum(1, k[rec_of] + 1)
    length[is_match] = np.maximum(1, rng.exponential(1.0, int(is_match.sum())) * mlen[is_match]).astype(np.int64)
    if eqx:
        r = rng.random(total)
        code[is_match & (r < 0.35)] = 7
        code[is_match & (r >= 0.35) & (r < 0.5)] = 8
    # variants
    nv = int(is_var.sum())
    vcode = np.where(rng.random(nv) < 0.55, 1, 2).astype(np.uint32)
    u = rng.random(nv)
    vlen = rng.geometric(0.5, nv).astype(np.int64)
    vlen = np.minimum(vlen, 9)
    mid = u < p["mid_rate"]
    vlen[mid] = rng.integers(10, 30, int(mid.sum()))
    big = (u >= p["mid_rate"]) & (u < p["mid_rate"] + p["sv_rate"])
    vlen[big] = np.minimum(50_000, (30 * (1 + rng.pareto(1.2, int(big.sum())))).astype(np.int64))
    code[is_var] = vcode
    length[is_var] = vlen
    # clips
    hard = rng.random(total) < 0.3
    code[is_lead | is_tail] = np.where(hard[is_lead | is_tail], 5, 4).astype(np.uint32)
    length[is_lead | is_tail] = rng.integers(10, 5000, int((is_lead | is_tail).sum()))

    ops = ((length.astype(np.uint64) << 4) | code).astype(np.uint32)
    pos = np.sort(rng.integers(1000, max(2000, chrom_len - 2 * p["read_len"]), n_rec)).astype(np.int32)
    mapq = np.where(rng.random(n_rec) < 0.9, 60, rng.integers(0, 60, n_rec)).astype(np.uint8)
    rev = rng.random(n_rec) < 0.5
    flag = np.where(rev, 16, 0).astype(np.uint16)
    supp = rng.random(n_rec) < 0.04
    flag[supp] |= 2048
    if n_names is None:
        n_names = max(1, int(n_rec * 0.95))
    name_id = rng.integers(0, n_names, n_rec).astype(np.int32)
    # query length (SEQ length) = every query-consuming op but hard clips
    qmask = np.isin(code, [0, 1, 4, 7, 8])
    ql = np.zeros(n_rec, dtype=np.int64)
    np.add.at(ql, rec_of[qmask], length[qmask])
    hp = rng.integers(1, 3, n_rec).astype(np.uint8)
    return dict(ops=ops, op_off=op_off, pos=pos, mapq=mapq, flag=flag, name_id=name_id,
                qlen=ql.astype(np.int32), hp=hp)


def synth_intervals(seed: int, n_reads: int, chrom_len: int = CHR10_LEN, median_len: int = 20_000,
                    sigma: float = 0.6, max_len: int = 500_000):
    """Stress-sweep read intervals: start U[0,L), ONT-like log-normal
    length, sorted by start.  Returns (start i32, end i32)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    start = np.sort(rng.integers(0, chrom_len, n_reads)).astype(np.int32)
    ln = np.minimum(max_len, np.maximum(200, rng.lognormal(np.log(median_len), sigma, n_reads))).astype(np.int64)
    end = np.minimum(np.int64(2**31 - 1), start.astype(np.int64) + ln).astype(np.int32)
    return start, end


def synth_candidates(seed: int, n_cand: int, chrom_len: int = CHR10_LEN):
    """Sorted candidate calls: pos, svlen (30*Pareto(1.2), <=50 kb), type (0 DEL,1 INS),
    call GT (0 '0/1', 1 '1/1')."""
    rng = np.random.Generator(np.random.PCG64(seed))
    pos = np.sort(rng.integers(1000, chrom_len - 60_000, n_cand)).astype(np.int32)
    svlen = np.minimum(50_000, (30 * (1 + rng.pareto(1.2, n_cand)))).astype(np.int32)
    typ = (rng.random(n_cand) < 0.55).astype(np.int32)
    gt = (rng.random(n_cand) < 0.6).astype(np.int32)
    return pos, svlen, typ, gt


def synth_sig_table(seed: int, cand_pos, cand_len, per_cand: int = 10):
    """Unique-signature table around the candidates (pos-sorted; pos,len,count)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n = len(cand_pos) * per_cand
    src = rng.integers(0, len(cand_pos), n)
    pos = cand_pos[src].astype(np.int64) + rng.integers(-800, 800, n)
    ln = np.maximum(10, (cand_len[src] * rng.uniform(0.4, 1.8, n))).astype(np.int32)
    cnt = rng.integers(1, 12, n).astype(np.int32)
    o = np.argsort(pos, kind="stable")
    return np.maximum(0, pos[o]).astype(np.int32), ln[o], cnt[o]


def add_sa(b, rng, frac=0.08, self_chrom=0):
    """8 % of the reads are split with SA tags.  Plausible SA tags for a share of the primary records: a second segment behind the primary one with a
    reference gap (deletion), a query gap (insertion), a third segment, another chromosome, low mapq ..."""
    n = len(b["pos"])
    off, ops = b["op_off"], b["ops"]
    codes, lens = ops & 15, (ops >> 4).astype(np.int64)
    ref_adv = np.isin(codes, [0, 2, 3, 7, 8])
    csum = np.concatenate([[0], np.cumsum(np.where(ref_adv, lens, 0))])
    ref_end = b["pos"] + (csum[off[1:]] - csum[off[:-1]])
    pick = np.flatnonzero((rng.random(n) < frac) & np.isin(b["flag"], [0, 16]))
    cols = {k: [] for k in ("chrom", "pos", "strand", "first_s", "last_s", "bias", "mapq")}
    sa_off = [0]
    for r in pick:
        a, e = int(off[r]), int(off[r + 1])
        l = int(ops[e - 1])
        right = (l >> 4) if (l & 15) in (4, 5) else 0
        qlen = int(b["qlen"][r])
        q_end = qlen - right
        st = 1 if b["flag"][r] == 16 else 0
        k = int(rng.integers(1, 4)) if rng.random() < 0.9 else int(rng.integers(7, 10))
        p = int(ref_end[r])
        for j in range(k):
            kind = rng.random()
            gap_ref = int(rng.integers(30, 4000)) if kind < 0.45 else int(rng.integers(-20, 60))
            gap_q = int(rng.integers(30, 3000)) if 0.45 <= kind < 0.8 else int(rng.integers(0, 80))
            bias = int(rng.integers(500, 5000))
            fs = min(max(q_end + gap_q, 0), max(qlen - 1, 0))
            cols["chrom"].append(self_chrom if rng.random() < 0.9 else self_chrom + 1)
            cols["pos"].append(max(1, p + gap_ref + 1))
            cols["strand"].append(st if rng.random() < 0.9 else 1 - st)
            if st == 0:
                cols["first_s"].append(fs); cols["last_s"].append(max(0, qlen - fs - bias))
            else:   # '-' entries swap the roles of the clips (SE:356-357)
                cols["last_s"].append(fs); cols["first_s"].append(max(0, qlen - fs - bias))
            cols["bias"].append(bias)
            cols["mapq"].append(int(rng.choice([60, 60, 60, 30, 10, 0])))
            p += gap_ref + bias
            q_end = fs + bias
        sa_off.append(len(cols["pos"]))
    b["sa_rec"] = pick.astype(np.int32)
    b["sa_off"] = np.asarray(sa_off, np.int64)
    for k, v in cols.items():
        b["sa_" + k] = np.asarray(v, np.uint8 if k in ("strand", "mapq") else np.int32)
    b["sa_self_chrom"] = self_chrom
    return b


def add_duplicates(b, rng, n_dup=40):
    """Copies of records under the same read name right behind the original: every `.sigs` row of the copy is
    a `sort -u` duplicate (DEL) or a potential one (INS)."""
    n = len(b["pos"])
    src = np.sort(rng.choice(n, n_dup, replace=False))
    order = np.sort(np.concatenate([np.arange(n), src]), kind="stable")
    nops = np.diff(b["op_off"])
    out = {k: (v[order] if isinstance(v, np.ndarray) and len(v) == n else v) for k, v in b.items() if k not in ("ops", "op_off")}
    new_nops = nops[order]
    out["op_off"] = np.concatenate([[0], np.cumsum(new_nops)]).astype(np.int64)
    idx = np.concatenate([np.arange(b["op_off"][r], b["op_off"][r + 1]) for r in order]) if len(order) else np.zeros(0, np.int64)
    out["ops"] = np.ascontiguousarray(b["ops"][idx])
    return out


class SynthSeq:
    """A deterministic pseudo-random DNA sequence of a given length that is never materialised: base i is a hash of
    (seed, i) (splitmix64 finaliser, vectorised), so a 135 Mb reference or 300 kb contigs cost nothing to hold and a
    slice costs a few numpy operations.  Supports len(), integer and slice indexing (Python semantics); returns str."""
    _BASES = np.frombuffer(b"ACGT", dtype=np.uint8)

    def __init__(self, length: int, seed: int):
        self.n = int(length)
        self.seed = np.uint64((int(seed) * 0x9E3779B97F4A7C15 + 0xD1B54A32D192ED03) & 0xFFFFFFFFFFFFFFFF)

    def __len__(self):
        return self.n

    def _bases(self, a, b):
        with np.errstate(over="ignore"):
            x = np.arange(a, b, dtype=np.uint64) + self.seed
            x ^= x >> np.uint64(30)
            x *= np.uint64(0xBF58476D1CE4E5B9)
            x ^= x >> np.uint64(27)
            x *= np.uint64(0x94D049BB133111EB)
            x ^= x >> np.uint64(31)
        return self._BASES[(x & np.uint64(3)).astype(np.intp)]

    def __getitem__(self, k):
        if isinstance(k, slice):
            a, b, step = k.indices(self.n)
            if step != 1:
                raise IndexError("SynthSeq: step slices are not supported")
            return "" if b <= a else self._bases(a, b).tobytes().decode()
        k = int(k)
        if k < 0:
            k += self.n
        if not 0 <= k < self.n:
            raise IndexError("SynthSeq index out of range")
        return chr(self._bases(k, k + 1)[0])


def synth_contigs(seed: int, chrom_len: int = CHR10_LEN, contig_len: int = 300_000, chrom: str = "chr10"):
    """Contig-to-reference alignments of two haplotypes tiling one chromosome: packed records with the
    M/I/D/S/H alphabet the contig scanner accepts, names `PS<s>_<e>_hp<h>_<k>` (the hp tag selects the haplotype,
    ECS_H:392), and the contig sequences as `SynthSeq` (length = read length incl. hard clips).  Returns (batch, names,
    {name: SynthSeq})."""
    n = max(4, int(2 * chrom_len / contig_len))
    b = synth_records(seed, n, "contig", chrom_len=chrom_len, eqx=False, clip_rate=0.05)
    codes, lens = b["ops"] & 15, (b["ops"] >> 4).astype(np.int64)
    off = b["op_off"]
    csum = np.concatenate([[0], np.cumsum(np.where(np.isin(codes, [0, 1, 4, 5]), lens, 0))])
    readlen = csum[off[1:]] - csum[off[:-1]]
    names, seqs = [], {}
    for k in range(n):
        s = int(b["pos"][k])
        nm = "PS%d_%d_hp%d_%d" % (s + 1, s + contig_len, int(b["hp"][k]), k)
        names.append(nm)
        seqs[nm] = SynthSeq(int(readlen[k]), seed * 1_000_003 + k)
    b["flag"] = (b["flag"] & 16).astype(np.uint16)            # primary alignments only: one record per contig name
    b["name_id"] = np.arange(n, dtype=np.int32)
    return b, names, seqs
