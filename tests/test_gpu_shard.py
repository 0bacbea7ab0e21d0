"""Resident shard engine (csrc/vsv_shard.cu: vsv_shard_*) against the oracle's restatement of the same step
(oracle.c_oracle.shard_step2: RVC:83-96 + FGC:111-172 with the reference's product semantics), bit-exact on
every table and every per-candidate output, through the C-ABI in both IO modes."""
import numpy as np
import pytest

import volcanosv_b200 as V
from oracle import c_oracle as CO
from volcanosv_b200 import lib as L
from volcanosv_b200 import synth
from volcanosv_b200.pipeline import ChromShard, GT_THRESHOLDS

pytestmark = pytest.mark.gpu

ALL = L.JOIN_FP | L.JOIN_COV | L.JOIN_SUPPORT | L.JOIN_GENOTYPE
SIG_FIELDS = ("rec", "type", "pos", "len", "q0", "q1")


def make(profile, n, seed, L_chrom=4_000_000, sa=True, dups=True, n_cand=4000):
    rng = np.random.default_rng(seed)
    # both alphabets the aligners write: 'M' (the streaming kernel's byte path) for three of the seeds, '='/'X' for the rest
    b = synth.synth_records(seed, n, profile, chrom_len=L_chrom, eqx=seed not in (31, 9, 13))
    if dups:
        b = synth.add_duplicates(b, rng)
    b["next_same_name"] = L.next_same_name(b["name_id"])
    if sa:
        synth.add_sa(b, rng)
    pos, svlen, typ, gt = synth.synth_candidates(seed + 1, n_cand, chrom_len=L_chrom)
    return b, dict(pos=pos, svlen=svlen, type=typ, gt=gt)


def canon(rows, with_count=True):
    f = SIG_FIELDS + (("count",) if with_count else ())
    a = np.stack([rows[k].astype(np.int64) for k in f], 1)
    return a[np.lexsort(a.T[::-1])]


def check_tables(fetch, counts, exp):
    reads = fetch(L.TAB_READS)
    nc = counts["n_reads_cigar"]
    assert nc == len(exp["sig_reads_cigar"]) and counts["n_reads_split"] == len(exp["sig_reads_split"])
    for f in SIG_FIELDS:
        assert np.array_equal(reads[f][:nc], exp["sig_reads_cigar"][f]), ("reads cigar", f)
    assert np.array_equal(canon(reads[nc:], False), canon(exp["sig_reads_split"], False))
    assert np.array_equal(np.sort(reads["first"][nc:]), np.sort(exp["sig_reads_split"]["first"]))
    cute = fetch(L.TAB_CUTE)
    assert len(cute) == len(exp["sig_cute"])
    for f in SIG_FIELDS:
        assert np.array_equal(cute[f], exp["sig_cute"][f]), ("cute", f)
    assert np.array_equal(fetch(L.TAB_REF_END), exp["ref_end"])
    assert np.array_equal(fetch(L.TAB_READ_LEN), exp["readlen"])
    assert np.array_equal(fetch(L.TAB_IN_BED), exp["in_bed"])
    for tab, key in ((L.TAB_DEL_ROWS, "rows_del"), (L.TAB_INS_ROWS, "rows_ins")):
        rows = fetch(tab)
        assert np.array_equal(canon(rows), canon(exp[key])), key
        merged = rows[rows["count"] > 0]
        # `first` of a merged row = index of its first raw signature
        assert np.array_equal(cute["rec"][merged["first"]], merged["rec"]) and np.array_equal(cute["pos"][merged["first"]], merged["pos"])
        uid = rows["first"]
        assert len(np.unique(uid)) == len(uid), "row uids must be unique"
    assert counts["n_sa_rows"] == exp["n_sa_rows"]
    assert counts["n_del_dup"] == exp["n_del_dup"] and counts["n_ins_maybe_dup"] == exp["n_ins_maybe_dup"]
    for tab, key in ((L.TAB_UDEL, "udel"), (L.TAB_UINS, "uins")):
        got = fetch(tab)
        for a, e, nm in zip(got, exp[key], ("pos", "len", "cnt")):
            assert np.array_equal(a, e), (key, nm)


def check_cands(got, exp):
    for k in ("fp_support", "cov", "n_support", "depth2", "new_gt"):
        assert np.array_equal(np.asarray(got[k]), exp[k]), k
    r = np.asarray(got["ratio"])
    assert np.array_equal(np.isnan(r), np.isnan(exp["ratio"]))
    m = ~np.isnan(exp["ratio"])
    assert np.array_equal(r[m], exp["ratio"][m])


@pytest.mark.parametrize("profile,dtype,n", [("hifi", "Hifi", 20000), ("ont", "ONT", 1500), ("clr", "CLR", 800)])
def test_resident_step_matches_oracle(profile, dtype, n):
    b, cands = make(profile, n, 31)
    sh = ChromShard(0, dtype)
    sh.upload(b, cands)
    res = sh.run_safe()
    got = sh.download(res)
    thr = GT_THRESHOLDS[dtype]
    exp = CO.shard_step2(b, cands, thr["DEL"], thr["INS"], threads=2)
    assert res.counts["overflow"] == 0
    check_tables(sh.fetch, res.counts, exp)
    check_cands(got, exp)
    assert exp["n_sa_rows"] > 0 and exp["n_del_dup"] > 0 and exp["n_ins_maybe_dup"] > 0
    assert profile != "hifi" or len(exp["sig_reads_split"]) > 0
    assert (exp["n_support"] > 0).sum() > 50 and (exp["depth2"] > 0).sum() > 100 and (exp["fp_support"] > 0).sum() > 50
    # idempotent: a second run on the same resident inputs gives the same results
    res2 = sh.run_safe()
    got2 = sh.download(res2)
    assert res2.counts == res.counts
    for k in got:
        assert np.array_equal(got[k], got2[k], equal_nan=True) if got[k].dtype.kind == "f" else np.array_equal(got[k], got2[k]), k


def test_host_io_shard_with_bed_and_staged_calls():
    """lib.Shard (host pointers): -include_bed gate with the reference's per-task region lists (SE:58-80), and
    the joins called one stage at a time as the operators do (FP on the raw calls, coverage, then support +
    genotype on a subset with a carried last_match_idx)."""
    b, cands = make("hifi", 12000, 77, L_chrom=30_000_000, n_cand=3000)
    chrom_len = 30_000_000
    bed = [(2_000_000 - 1000, 9_990_000 + 1000), (10_020_000 - 1000, 10_500_000 + 1000), (21_000_000 - 1000, 24_000_000 + 1000)]
    ctx = V.Context(0)
    ctx.set_params(L.default_params("Hifi"))
    sh = L.Shard(ctx)
    sh.load(b["ops"], b["op_off"], b["pos"], b["mapq"], b["qlen"], b["flag"], b["name_id"], b["next_same_name"],
            sa={k[3:]: b[k] for k in b if k.startswith("sa_") and k != "sa_self_chrom"}, self_chrom=0)
    sh.set_bed(bed, chrom_len)
    counts = sh.scan()
    thr = GT_THRESHOLDS["Hifi"]
    exp = CO.shard_step2(b, cands, thr["DEL"], thr["INS"], threads=2, bed=bed, chrom_len=chrom_len)
    assert 0 < exp["in_bed"].sum() < len(exp["in_bed"])
    check_tables(sh.fetch, counts, exp)
    fp = sh.candidates(L.JOIN_FP, cands["pos"], cands["svlen"], cands["type"])
    assert np.array_equal(fp["fp_support"], exp["fp_support"])
    cv = sh.candidates(L.JOIN_COV, cands["pos"], cands["svlen"], cands["type"])
    assert np.array_equal(cv["cov"], exp["cov"])
    full = sh.candidates(L.JOIN_SUPPORT | L.JOIN_GENOTYPE, cands["pos"], cands["svlen"], cands["type"], cands["gt"])
    check_cands(dict(full, fp_support=fp["fp_support"], cov=cv["cov"]), exp)
    # two halves with the carry == one call (the whole-genome mode hands last_match_idx from shard to shard)
    h = len(cands["pos"]) // 2
    first = sh.candidates(L.JOIN_SUPPORT, cands["pos"][:h], cands["svlen"][:h], cands["type"][:h])
    second = sh.candidates(L.JOIN_SUPPORT, cands["pos"][h:], cands["svlen"][h:], cands["type"][h:], last_in=first["last_out"])
    assert np.array_equal(np.concatenate([first["n_support"], second["n_support"]]), full["n_support"])
    assert np.array_equal(np.concatenate([first["match_id"], second["match_id"]]), full["match_id"])
    assert second["last_out"] == full["last_out"]
    # genotype alone with n_support as an input
    g = sh.candidates(L.JOIN_GENOTYPE, cands["pos"], cands["svlen"], cands["type"], cands["gt"], n_support=full["n_support"])
    assert np.array_equal(g["new_gt"], full["new_gt"]) and np.array_equal(g["depth2"], full["depth2"])
    sh.close()
    ctx.close()


def test_confirmed_ins_duplicates_are_dropped():
    b, cands = make("hifi", 8000, 5, sa=False)
    ctx = V.Context(0)
    sh = L.Shard(ctx)
    sh.load(b["ops"], b["op_off"], b["pos"], b["mapq"], b["qlen"], b["flag"], b["name_id"], b["next_same_name"])
    c0 = sh.scan()
    dups = sh.fetch(L.TAB_INS_DUPS)
    assert len(dups) == c0["n_ins_maybe_dup"] > 0
    rows = sh.fetch(L.TAB_INS_ROWS)
    by_uid = {int(u): i for i, u in enumerate(rows["first"])}
    for later, earlier in dups:   # the pair really is equal in (pos, len, read)
        a, e = rows[by_uid[int(later)]], rows[by_uid[int(earlier)]]
        assert a["pos"] == e["pos"] and a["len"] == e["len"] and b["name_id"][a["rec"]] == b["name_id"][e["rec"]]
    before = sh.candidates(L.JOIN_COV | L.JOIN_SUPPORT, cands["pos"], cands["svlen"], cands["type"])
    u0 = sh.fetch(L.TAB_UINS)
    c1 = sh.drop_ins(dups[:, 0])
    assert c1["n_ins_maybe_dup"] == 0 and c1["n_uins"] == c0["n_uins"] and c1["n_udel"] == c0["n_udel"]
    u1 = sh.fetch(L.TAB_UINS)
    assert np.array_equal(u0[0], u1[0]) and np.array_equal(u0[1], u1[1])
    big = rows["len"][[by_uid[int(x)] for x in dups[:, 0]]] >= 30
    assert int(u0[2].sum() - u1[2].sum()) == int(big.sum())          # one read fewer per dropped row (>= 30 bp rows are grouped)
    after = sh.candidates(L.JOIN_COV | L.JOIN_SUPPORT, cands["pos"], cands["svlen"], cands["type"])
    ins = cands["type"] == 1
    assert np.all(after["cov"][ins] <= before["cov"][ins]) and np.any(after["cov"][ins] < before["cov"][ins])
    assert np.array_equal(after["cov"][~ins], before["cov"][~ins])
    # the oracle agrees once the same rows are removed from its input
    sh.close()
    ctx.close()


def test_overflow_regrows_and_positions_out_of_range_fail():
    b, cands = make("ont", 600, 9, sa=False, dups=False, n_cand=500)
    ctx = V.Context(0)
    ctx.set_params(L.default_params("ONT"))
    sh = L.Shard(ctx)
    sh.load(b["ops"], b["op_off"], b["pos"], b["mapq"], b["qlen"], b["flag"], b["name_id"], b["next_same_name"])
    sh.reserve(64, 64, 64, 64)              # far too small: every table overflows once
    counts = sh.scan()
    thr = GT_THRESHOLDS["ONT"]
    exp = CO.shard_step2(b, cands, thr["DEL"], thr["INS"], threads=2)
    check_tables(sh.fetch, counts, exp)
    out = sh.candidates(ALL, cands["pos"], cands["svlen"], cands["type"], cands["gt"])
    check_cands(out, exp)
    far = b["pos"].copy()
    far += (1 << 28)
    sh.load(b["ops"], b["op_off"], far, b["mapq"], b["qlen"], b["flag"])
    with pytest.raises(V.VsvError):
        sh.scan()
    sh.close()
    ctx.close()


def test_many_candidates_chain_across_tiles():
    """The last_match_idx carry is a look-back scan over 1024-candidate tiles: 300 k candidates on a dense table."""
    b, _ = make("hifi", 30000, 13, L_chrom=3_000_000, sa=False, dups=False, n_cand=10)
    pos, svlen, typ, gt = synth.synth_candidates(14, 300_000, chrom_len=3_000_000)
    cands = dict(pos=pos, svlen=svlen, type=typ, gt=gt)
    sh = ChromShard(0, "Hifi")
    sh.upload(b, cands)
    got = sh.download(sh.run_safe())
    thr = GT_THRESHOLDS["Hifi"]
    exp = CO.shard_step2(b, cands, thr["DEL"], thr["INS"], threads=2)
    check_cands(got, exp)


def test_staging_formats_give_identical_results():
    from volcanosv_b200.lib import pack_ops8, pack_ops12, pack_ops16
    b, cands = make("ont", 400, 41, sa=False, dups=False, n_cand=500)
    b["ops"][::501] = (np.uint32(70000) << 4) | (b["ops"][::501] & 15)           # long ops -> exception lists
    sh = ChromShard(0, "ONT")
    sh.upload(b, cands)
    a = sh.download(sh.run_safe())
    ca = sh.counts()
    for name, packer in (("ops16", pack_ops16), ("ops8", pack_ops8), ("ops12", pack_ops12)):
        packed, ei, ev = packer(b["ops"])
        b2 = {k: v for k, v in b.items() if k != "ops"}
        b2.update({name: packed, "exc_idx": ei, "exc_val": ev})
        sh.upload(b2, cands, n_ops=len(b["ops"]))
        c = sh.download(sh.run_safe())
        assert sh.counts() == ca, name
        for k in a:
            assert np.array_equal(a[k], c[k], equal_nan=True) if a[k].dtype.kind == "f" else np.array_equal(a[k], c[k]), (name, k)


def test_closed_form_depth_edge_cases():
    """Zero-length records (end == start), windows outside the reads, E < S never occurs but E == S does."""
    rng = np.random.default_rng(3)
    n = 50_000
    rs = np.sort(rng.integers(0, 100_000, n)).astype(np.int32)
    ln = rng.integers(0, 3000, n)
    ln[rng.random(n) < 0.05] = 0
    ln[rng.random(n) < 0.01] = 90_000
    re = (rs + ln).astype(np.int32)
    S = rng.integers(-500, 101_000, 20_000).astype(np.int32)
    E = (S + rng.integers(0, 1500, len(S))).astype(np.int32)
    ctx = V.Context(0)
    got = ctx.span_depth(rs, re, S, E)
    exp = CO.span_depth(rs, re, S, E, threads=2)
    assert np.array_equal(got, exp)
    brute = np.array([int(((rs < s) & (re > e)).sum()) for s, e in zip(S[:300], E[:300])])
    assert np.array_equal(got[:300], brute)
    ctx.close()
