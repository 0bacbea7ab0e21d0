"""Parity of the CUDA path (through the C-ABI, ctypes) against the oracle and the
reference-generated golden vectors.  Integer work: bit-exact."""
import numpy as np
import pytest

import oracle as O
import volcanosv_b200 as V
from volcanosv_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = V.Context(0)
    yield c
    c.close()


def sig_tuples(a, fields=("rec", "type", "pos", "len", "q0", "q1")):
    return [tuple(int(x[f]) for f in fields) for x in a]


# ------------------------------------------------------------------ stage 1
@pytest.mark.parametrize("name,mode,minlen", [("contig_scan", V.MODE_CONTIG, 30), ("reads_scan", V.MODE_READS, 30)])
def test_scan_golden(ctx, golden, name, mode, minlen):
    g = golden(name)
    ops, off = O.pack_cigar([c["cigar"] for c in g])
    pos = np.array([c["pos"] for c in g], np.int32)
    mapq = np.full(len(g), 60, np.uint8)
    sigs, rt, qt, xt = ctx.scan_cigar(mode, ops, off, pos, mapq, min_len=minlen)
    for r, c in enumerate(g):
        assert int(pos[r]) + int(rt[r]) == c["oref"] and int(qt[r]) == c["oq"]
    if mode == V.MODE_CONTIG:
        d, s = ctx.merge_in_record(mode, sigs)
        for r, c in enumerate(g):
            assert [[int(x["pos"]), int(x["len"]), int(x["q0"]), int(x["q1"])] for x in d[d["rec"] == r]] == c["dels"]
            assert [[int(x["pos"]), int(x["len"]), int(x["q0"]), int(x["q1"])] for x in s[s["rec"] == r]] == c["inss"]
    else:
        for r, c in enumerate(g):
            m = sigs[sigs["rec"] == r]
            assert [[int(x["pos"]), int(x["len"]), int(x["q0"])] for x in m[m["type"] == V.DEL]] == c["dels"]
            assert [[int(x["pos"]), int(x["len"]), int(x["q0"])] for x in m[m["type"] == V.INS]] == c["inss"]


def test_cute_scan_merge_golden(ctx, golden):
    g = [c for c in golden("parse_read") if len(c["seq"]) >= 500]
    ops, off = O.pack_cigar([c["cigar"] for c in g])
    pos = np.array([c["pos"] for c in g], np.int32)
    mapq = np.array([c["mapq"] for c in g], np.uint8)
    sigs, rt, qt, xt = ctx.scan_cigar(V.MODE_CUTE, ops, off, pos, mapq, min_len=10, min_mapq=20)
    d, s = ctx.merge_in_record(V.MODE_CUTE, sigs)
    inss = sigs[sigs["type"] == V.INS]
    for r, c in enumerate(g):
        rows = []
        for m in s[s["rec"] == r]:
            piece = "".join(c["seq"][int(inss[k]["q0"]):int(inss[k]["q1"])] for k in range(m["first"], m["first"] + m["count"]))
            rows.append(["INS", int(m["pos"]), int(m["len"]), piece])
        for m in d[d["rec"] == r]:
            rows.append(["DEL", int(m["pos"]), int(m["len"])])
        # golden rows = cigar-derived rows first, then SA-derived rows (checked in the oracle test)
        assert rows == c["rows"][: len(rows)]


@pytest.mark.parametrize("profile,mode,minlen,minmq,eqx", [
    ("hifi", V.MODE_READS, 30, 50, True), ("ont", V.MODE_CUTE, 10, 20, True), ("contig", V.MODE_CONTIG, 30, 50, False),
    ("hifi", V.MODE_CUTE, 10, 20, False)])
def test_scan_vs_oracle_random(ctx, profile, mode, minlen, minmq, eqx):
    n = {"hifi": 300, "ont": 40, "contig": 30}[profile]
    b = synth.synth_records(7, n, profile, eqx=eqx)
    en = (np.arange(n) % 7 != 3).astype(np.uint8)
    sigs, rt, qt, xt = ctx.scan_cigar(mode, b["ops"], b["op_off"], b["pos"], b["mapq"], enable=en, min_len=minlen,
                                      min_mapq=minmq)
    es, ert, eqt, ext = O.scan_cigar(b["ops"], b["op_off"], b["pos"], b["mapq"], mode, minlen, minmq, enable=en)
    assert sig_tuples(sigs) == es
    assert len(es) > 0
    assert np.array_equal(rt, ert) and np.array_equal(qt, eqt) and np.array_equal(xt, ext)
    d, s = ctx.merge_in_record(mode, sigs)
    ed, ei = O.merge_in_record(es, mode)
    f8 = ("rec", "type", "pos", "len", "q0", "q1", "first", "count")
    assert sig_tuples(d, f8) == ed and sig_tuples(s, f8) == ei


def test_scan_edge_cases(ctx):
    # empty records, a record of one op, a record spanning many tiles, capacity growth
    cig = [[], [(0, 5)], [(2, 40)], [], [(5, 7), (0, 3), (1, 31), (0, 2)]]
    big = []
    for i in range(9000):
        big += [(0, 10), (1, 30 + (i % 5)), (0, 7), (2, 10 + (i % 40))]
    cig.append(big)
    cig.append([(4, 3), (0, 100), (2, 30), (0, 1)])
    cig.append([])
    ops, off = O.pack_cigar(cig)
    pos = np.arange(len(cig), dtype=np.int32) * 1000
    mapq = np.full(len(cig), 60, np.uint8)
    for mode in (V.MODE_CONTIG, V.MODE_READS, V.MODE_CUTE):
        sigs, rt, qt, xt = ctx.scan_cigar(mode, ops, off, pos, mapq, min_len=30, cap=16)
        es, ert, eqt, ext = O.scan_cigar(ops, off, pos, mapq, mode, 30, 0)
        assert sig_tuples(sigs) == es and len(es) > 9000
        assert np.array_equal(rt, ert) and np.array_equal(qt, eqt) and np.array_equal(xt, ext)
    # nothing at all
    sigs, rt, qt, xt = ctx.scan_cigar(V.MODE_READS, np.zeros(0, np.uint32), np.zeros(1, np.int64), np.zeros(0, np.int32),
                                      np.zeros(0, np.uint8))
    assert len(sigs) == 0
    # records but zero ops
    sigs, rt, qt, xt = ctx.scan_cigar(V.MODE_READS, np.zeros(0, np.uint32), np.zeros(4, np.int64), np.zeros(3, np.int32),
                                      np.zeros(3, np.uint8))
    assert len(sigs) == 0 and rt.tolist() == [0, 0, 0]


def test_scan_full_size_properties(ctx):
    """chr10 HiFi-shaped batch (BASELINE config 1 scale / 10): size-independent checks."""
    n = 27000
    b = synth.synth_records(11, n, "hifi", eqx=True)
    sigs, rt, qt, xt = ctx.scan_cigar(V.MODE_READS, b["ops"], b["op_off"], b["pos"], b["mapq"], min_len=30, min_mapq=50)
    code = b["ops"] & 15
    ln = (b["ops"] >> 4).astype(np.int64)
    rec_of = np.repeat(np.arange(n), np.diff(b["op_off"]))
    # totals == segmented sums computed with numpy
    ref = np.zeros(n, np.int64)
    np.add.at(ref, rec_of[np.isin(code, [0, 2, 3, 7, 8])], ln[np.isin(code, [0, 2, 3, 7, 8])])
    assert np.array_equal(ref, rt)
    qry = np.zeros(n, np.int64)
    np.add.at(qry, rec_of[np.isin(code, [0, 1, 4, 7, 8])], ln[np.isin(code, [0, 1, 4, 7, 8])])
    assert np.array_equal(qry, qt)
    # emitted set == ops with len>=30 and code in {I,D} of records with mapq>=50, in order
    sel = np.isin(code, [1, 2]) & (ln >= 30) & (b["mapq"][rec_of] >= 50)
    assert len(sigs) == int(sel.sum())
    assert np.array_equal(sigs["rec"], rec_of[sel]) and np.array_equal(sigs["len"], ln[sel])
    assert np.array_equal(sigs["type"], (code[sel] == 1).astype(np.int32))
    # idempotence
    sigs2, *_ = ctx.scan_cigar(V.MODE_READS, b["ops"], b["op_off"], b["pos"], b["mapq"], min_len=30, min_mapq=50)
    assert np.array_equal(sigs.view(np.uint8), sigs2.view(np.uint8))


def _rand_ops_batch(seed, n_rec, mean_ops, codes, big_rate=0.02):
    """Random records over an arbitrary op-code alphabet (incl. N, P, H, reserved codes)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    nops = rng.poisson(mean_ops, n_rec).astype(np.int64)
    off = np.zeros(n_rec + 1, np.int64)
    np.cumsum(nops, out=off[1:])
    tot = int(off[-1])
    code = rng.choice(np.array(codes, np.uint32), tot)
    ln = rng.integers(0, 25, tot).astype(np.uint32)
    big = rng.random(tot) < big_rate
    ln[big] = rng.integers(25, 4000, int(big.sum()))
    ops = (ln << 4) | code
    pos = np.sort(rng.integers(0, 10_000_000, n_rec)).astype(np.int32)
    mapq = rng.integers(0, 61, n_rec).astype(np.uint8)
    return ops.astype(np.uint32), off, pos, mapq


@pytest.mark.parametrize("n_rec,mean_ops,codes", [
    (60, 900, list(range(9))),                 # every BAM op incl. N, P, mid-record H
    (20000, 2, [0, 1, 2, 4, 5, 3]),            # > 128 records per tile: global record-table path
    (300, 120, list(range(16))),               # reserved codes 9..15
])
def test_scan_rare_ops_and_dense_records(ctx, n_rec, mean_ops, codes):
    ops, off, pos, mapq = _rand_ops_batch(n_rec, n_rec, mean_ops, codes)
    en = (np.arange(n_rec) % 5 != 1).astype(np.uint8)
    for mode, minlen, minmq in ((V.MODE_CONTIG, 30, 50), (V.MODE_READS, 30, 50), (V.MODE_CUTE, 10, 20)):
        sigs, rt, qt, xt = ctx.scan_cigar(mode, ops, off, pos, mapq, enable=en, min_len=minlen, min_mapq=minmq)
        es, ert, eqt, ext = O.scan_cigar(ops, off, pos, mapq, mode, minlen, minmq, enable=en)
        assert sig_tuples(sigs) == es and len(es) > 0
        assert np.array_equal(rt, ert) and np.array_equal(qt, eqt) and np.array_equal(xt, ext)
        assert (sigs["count"] == 1).all()


@pytest.mark.parametrize("case", ["ont", "hifi", "rare", "dense"])
def test_scan_reads_fused_equals_two_scans(ctx, case):
    """vsv_scan_reads_fused == (READS scan, CUTE scan), bit for bit, and == the oracle."""
    if case in ("ont", "hifi"):
        n = {"ont": 60, "hifi": 600}[case]
        b = synth.synth_records(21, n, case, eqx=True)
        ops, off, pos, mapq, qlen = b["ops"], b["op_off"], b["pos"], b["mapq"], b["qlen"]
    else:
        n = 20000 if case == "dense" else 80
        ops, off, pos, mapq = _rand_ops_batch(5, n, 2 if case == "dense" else 700, list(range(9)))
        qlen = np.random.Generator(np.random.PCG64(1)).integers(0, 1000, n).astype(np.int32)
    en = (qlen >= 500).astype(np.uint8)
    r, c, rt, qt, xt = ctx.scan_reads_fused(ops, off, pos, mapq, enable_cute=en, cap=64)
    er, ert, eqt, ext = O.scan_cigar(ops, off, pos, mapq, V.MODE_READS, 30, 50)
    ec, *_ = O.scan_cigar(ops, off, pos, mapq, V.MODE_CUTE, 10, 20, enable=en)
    assert sig_tuples(r) == er and sig_tuples(c) == ec
    assert len(ec) > 0
    assert np.array_equal(rt, ert) and np.array_equal(qt, eqt) and np.array_equal(xt, ext)
    r2, *_ = ctx.scan_cigar(V.MODE_READS, ops, off, pos, mapq, min_len=30, min_mapq=50)
    c2, *_ = ctx.scan_cigar(V.MODE_CUTE, ops, off, pos, mapq, enable=en, min_len=10, min_mapq=20, totals=False)
    assert np.array_equal(r.view(np.uint8), r2.view(np.uint8)) and np.array_equal(c.view(np.uint8), c2.view(np.uint8))


def test_scan_staging_retry_and_candidate_overflow(ctx):
    """(a) most records fail the mapq gate, so the raw candidate count is far above the
    output count: with cap == n_out the internal staging must grow and retry; (b) a tile
    whose per-warp candidate lists overflow (every op is a candidate) takes the rescan path."""
    ops, off, pos, mapq = _rand_ops_batch(9, 400, 600, [0, 1, 2, 4], big_rate=0.3)
    mapq[:] = 0
    mapq[::40] = 60
    es, *_ = O.scan_cigar(ops, off, pos, mapq, V.MODE_READS, 30, 50)
    assert 0 < len(es) < 0.1 * int(((ops >> 4) >= 30).sum())
    sigs, *_ = ctx.scan_cigar(V.MODE_READS, ops, off, pos, mapq, min_len=30, min_mapq=50, cap=len(es))
    assert sig_tuples(sigs) == es
    en = np.ones(400, np.uint8)
    r, c, *_ = ctx.scan_reads_fused(ops, off, pos, mapq, enable_cute=en, cap=len(es))
    ec, *_ = O.scan_cigar(ops, off, pos, mapq, V.MODE_CUTE, 10, 20, enable=en)
    assert sig_tuples(r) == es and sig_tuples(c) == ec
    # (b) dense candidates: alternating 40I / 35D only
    cig = [[(1, 40), (2, 35)] * 3000, [(0, 5)], [(2, 31), (1, 32)] * 2500]
    ops, off = O.pack_cigar(cig)
    pos = np.array([10, 20, 30], np.int32)
    mapq = np.full(3, 60, np.uint8)
    for mode in (V.MODE_CONTIG, V.MODE_READS, V.MODE_CUTE):
        sigs, rt, qt, xt = ctx.scan_cigar(mode, ops, off, pos, mapq, min_len=30)
        es, ert, eqt, ext = O.scan_cigar(ops, off, pos, mapq, mode, 30, 0)
        assert sig_tuples(sigs) == es and len(es) == 11000
        assert np.array_equal(rt, ert) and np.array_equal(qt, eqt)


def test_expand_ops16(ctx):
    """vsv_expand_ops16 restores the BAM-native ops (incl. exceptions and the ragged tail)."""
    from volcanosv_b200.lib import pack_ops16
    rng = np.random.default_rng(11)
    for n in (0, 1, 7, 8, 9, 4099, 100003):
        ln = rng.integers(0, 6000, n).astype(np.uint32)
        if n:
            ln[rng.integers(0, n, max(1, n // 50))] = rng.integers(4095, 1 << 27, max(1, n // 50))
        ops = (ln << 4) | rng.integers(0, 9, n).astype(np.uint32)
        o16, ei, ev = pack_ops16(ops)
        assert np.array_equal(ctx.expand_ops16(o16, ei, ev), ops)


def test_fused_scan_ont_scale_positions(ctx):
    """ONT-shaped shard at 1/16 of the BASELINE chr10 size (~5x10^7 ops, far more tiles than resident
    warps, records spanning several tiles): EVERY signature's position and query offset against numpy
    prefix sums over the whole op array, for both streams of the fused pass."""
    n = synth.n_records("ont") // 16
    b = synth.synth_records(77, n, "ont", eqx=False)
    ops, off = b["ops"], b["op_off"]
    n_ops = int(off[-1])
    code = (ops & 15).astype(np.uint8)
    ln = (ops >> 4).astype(np.int64)
    rec_of = np.repeat(np.arange(n, dtype=np.int64), np.diff(off))
    enable = (b["qlen"] >= 500).astype(np.uint8)
    ra, rb, rt, qt, xt = ctx.scan_reads_fused(ops, off, b["pos"], b["mapq"], enable_cute=enable)

    def prefix(member):
        x = np.where(np.isin(code, member), ln, 0)
        c = np.cumsum(x)
        excl = c - x                                   # exclusive prefix over the whole array
        base = np.concatenate([[0], c])[off[:-1]]      # value at each record start
        return excl - base[rec_of]

    hc = np.where((np.diff(off) > 0) & (code[np.minimum(off[:-1], n_ops - 1)] == 5), ln[np.minimum(off[:-1], n_ops - 1)], 0)
    isid = np.isin(code, [1, 2])
    # stream A = extract_reads_signature (ERS:47-83): mapq >= 50, >= 30 bp; ref = M,D,N,=,X; query = M,I,S,=,X (+ leading H)
    sel = isid & (ln >= 30) & (b["mapq"][rec_of] >= 50)
    assert len(ra) == int(sel.sum())
    assert np.array_equal(ra["rec"], rec_of[sel]) and np.array_equal(ra["len"], ln[sel])
    assert np.array_equal(ra["pos"], (b["pos"][rec_of] + prefix([0, 2, 3, 7, 8]))[sel])
    assert np.array_equal(ra["q0"], (prefix([0, 1, 4, 7, 8]) + hc[rec_of])[sel])
    # stream B = parse_read's CIGAR walk (SE:438-478): mapq >= 20, query_length >= 500, >= 10 bp; N moves no reference counter
    sel = isid & (ln >= 10) & (b["mapq"][rec_of] >= 20) & (enable[rec_of] != 0)
    assert len(rb) == int(sel.sum())
    assert np.array_equal(rb["rec"], rec_of[sel]) and np.array_equal(rb["len"], ln[sel])
    assert np.array_equal(rb["pos"], (b["pos"][rec_of] + prefix([0, 2, 7, 8]))[sel])
    # per-record totals
    tot = np.zeros(n, np.int64)
    np.add.at(tot, rec_of, np.where(np.isin(code, [0, 2, 3, 7, 8]), ln, 0))
    assert np.array_equal(tot, rt)


def test_expand_ops8(ctx):
    """vsv_expand_ops8 == the BAM-native words, incl. exceptions and a ragged tail."""
    from volcanosv_b200.lib import pack_ops8
    rng = np.random.default_rng(21)
    for n in (0, 1, 15, 16, 17, 100003):
        code = rng.choice([0, 1, 2, 3, 4, 5, 8, 12], n).astype(np.uint32) if n else np.zeros(0, np.uint32)
        ln = rng.integers(0, 200, n).astype(np.uint32)
        if n > 50:
            ln[rng.integers(0, n, max(1, n // 50))] = rng.integers(4095, 1 << 27, max(1, n // 50))
        ops = (ln << 4) | code
        o8, ei, ev = pack_ops8(ops)
        assert np.array_equal(ctx.expand_ops8(o8, ei, ev), ops)


def test_expand_ops12(ctx):
    """vsv_expand_ops12 == the BAM-native words: alternating M / short-indel pairs in both phases, odd pairs
    through the exception list, ragged tails and odd lengths."""
    from volcanosv_b200.lib import pack_ops12, unpack_ops12
    rng = np.random.default_rng(22)
    for n in (0, 1, 2, 15, 16, 17, 31, 32, 33, 100003, 1 << 20):
        code = rng.choice([0, 1, 2, 3, 4, 5, 8, 12], n).astype(np.uint32) if n else np.zeros(0, np.uint32)
        ln = rng.integers(0, 200, n).astype(np.uint32)
        ops = (ln << 4) | code
        if n > 50:      # mostly minimap2-like: M, indel, M, indel ... with a phase flip in the middle
            m = (rng.integers(1, 140, n).astype(np.uint32) << 4)
            x = (rng.integers(1, 11, n).astype(np.uint32) << 4) | rng.integers(1, 3, n).astype(np.uint32)
            alt = np.where(np.arange(n) % 2 == 0, m, x)
            alt[n // 2:] = np.where(np.arange(n - n // 2) % 2 == (n // 2 + 1) % 2, m[n // 2:], x[n // 2:])
            keep = rng.random(n) < 0.97
            ops = np.where(keep, alt, ops)
        o12, ep, ev = pack_ops12(ops)
        if n > 50:
            assert 0 < len(ep) < 0.35 * n
        assert np.array_equal(unpack_ops12(o12, n, ep, ev), ops)
        assert np.array_equal(ctx.expand_ops12(o12, n, ep, ev), ops)


def test_record_range_parts_concatenate_to_the_whole(ctx):
    """Level-2 sharding (shard.split_records): scanning contiguous record ranges separately and
    concatenating in part order == scanning the chromosome batch at once (both fused streams)."""
    from volcanosv_b200 import shard as SH
    b = synth.synth_records(91, 4000, "ont", eqx=True)
    enable = (b["qlen"] >= 500).astype(np.uint8)
    ra, rb, rt, qt, xt = ctx.scan_reads_fused(b["ops"], b["op_off"], b["pos"], b["mapq"], enable_cute=enable)
    pa, pb, prt = [], [], []
    for lo, hi in SH.split_records(b["op_off"], 5):
        t = SH.take_records(b, lo, hi)
        a_, b_, rt_, _, _ = ctx.scan_reads_fused(t["ops"], t["op_off"], t["pos"], t["mapq"], enable_cute=enable[lo:hi])
        for s_ in (a_, b_):
            s_["rec"] += lo
        pa.append(a_); pb.append(b_); prt.append(rt_)
    assert np.array_equal(np.concatenate(pa).view(np.uint8), ra.view(np.uint8))
    assert np.array_equal(np.concatenate(pb).view(np.uint8), rb.view(np.uint8))
    assert np.array_equal(np.concatenate(prt), rt)


def _segmented_alphabet_batch(seed):
    """Consecutive stretches of different op alphabets, each tens of tiles long: what a warp of the streaming kernel meets when
    it walks many tiles (byte path -> v6 loop for 32 tiles -> byte path again)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    segs = [("m", 300_000), ("eqx", 150_000), ("m", 700_000), ("clips", 120_000), ("rare", 90_000), ("m", 600_000),
            ("reserved", 40_000), ("long_id", 150_000), ("m", 250_000)]
    parts = []
    for kind, n in segs:
        code = rng.choice(np.array([0, 0, 0, 1, 2], np.uint32), n)
        ln = rng.integers(0, 40, n).astype(np.uint32)
        if kind == "eqx":
            code = rng.choice(np.array([7, 7, 8, 1, 2], np.uint32), n)
        elif kind == "clips":                       # lengths >= 256 on M / S in ~1 op of 200 (a few lanes per tile)
            big = rng.random(n) < 0.005
            ln[big] = rng.integers(256, 70000, int(big.sum()))
            code[big] = rng.choice(np.array([0, 4], np.uint32), int(big.sum()))
        elif kind == "rare":                        # N / H / P everywhere: more than six exact lanes per tile
            code = rng.choice(np.array([0, 1, 2, 3, 5, 6], np.uint32), n)
        elif kind == "reserved":
            code = rng.choice(np.arange(16, dtype=np.uint32), n)
        elif kind == "long_id":                     # I / D of 128..255 bases (the byte path's -128 threshold) and of >= 256
            big = rng.random(n) < 0.003
            ln[big] = rng.choice(np.array([127, 128, 129, 255, 256, 257, 4000], np.uint32), int(big.sum()))
            code[big] = rng.choice(np.array([1, 2], np.uint32), int(big.sum()))
        parts.append((ln << 4) | code)
    ops = np.concatenate(parts).astype(np.uint32)
    n_rec = 700
    cuts = np.sort(rng.choice(np.arange(1, len(ops)), n_rec - 1, replace=False))
    off = np.concatenate([[0], cuts, [len(ops)]]).astype(np.int64)
    pos = np.sort(rng.integers(0, 100_000_000, n_rec)).astype(np.int32)
    mapq = rng.integers(0, 61, n_rec).astype(np.uint8)
    return ops, off, pos, mapq


@pytest.mark.parametrize("grid", ["1", "3", ""])
def test_scan_byte_path_switching(ctx, grid, monkeypatch):
    """The streaming kernel's byte path (dp4a sums, valid for codes <= 7 and lengths < 256) and its per-warp fall-back to the
    general loop, on an input whose alphabet changes every few dozen tiles; VSV_SCAN_GRID = 1 / 3 makes every warp walk
    hundreds of consecutive tiles.  Signatures and per-record totals against the oracle, every mode."""
    ops, off, pos, mapq = _segmented_alphabet_batch(17)
    if grid:
        monkeypatch.setenv("VSV_SCAN_GRID", grid)
    en = (np.arange(len(pos)) % 7 != 3).astype(np.uint8)
    r, c, rt, qt, xt = ctx.scan_reads_fused(ops, off, pos, mapq, enable_cute=en)
    er, ert, eqt, ext = O.scan_cigar(ops, off, pos, mapq, V.MODE_READS, 30, 50)
    ec, *_ = O.scan_cigar(ops, off, pos, mapq, V.MODE_CUTE, 10, 20, enable=en)
    assert sig_tuples(r) == er and sig_tuples(c) == ec and len(er) > 1000 and len(ec) > 1000
    assert np.array_equal(rt, ert) and np.array_equal(qt, eqt) and np.array_equal(xt, ext)
    for mode, minlen, minmq in ((V.MODE_READS, 30, 50), (V.MODE_CUTE, 10, 20)):
        sigs, rt, qt, xt = ctx.scan_cigar(mode, ops, off, pos, mapq, enable=en, min_len=minlen, min_mapq=minmq)
        es, ert, eqt, ext = O.scan_cigar(ops, off, pos, mapq, mode, minlen, minmq, enable=en)
        assert sig_tuples(sigs) == es
        assert np.array_equal(rt, ert) and np.array_equal(qt, eqt) and np.array_equal(xt, ext)
