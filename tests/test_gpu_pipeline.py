"""Device-resident shard pipeline (volcanosv_b200.pipeline.ChromShard) against the
oracle's restatement of the same call chain, bit-exact on every output."""
import numpy as np
import pytest

import volcanosv_b200 as V
from oracle import c_oracle as CO
from volcanosv_b200 import synth
from volcanosv_b200.pipeline import ChromShard, GT_THRESHOLDS

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("profile,dtype,n", [("hifi", "Hifi", 20000), ("ont", "ONT", 1500)])
def test_shard_matches_oracle(profile, dtype, n):
    L = 4_000_000
    b = synth.synth_records(31, n, profile, chrom_len=L, eqx=True)
    pos, svlen, typ, gt = synth.synth_candidates(32, 4000, chrom_len=L)
    cands = dict(pos=pos, svlen=svlen, type=typ, gt=gt)
    sh = ChromShard(0, dtype)
    sh.upload(b, cands)
    res = sh.run_safe()
    got = sh.download(res)
    exp = CO.shard_step(b, cands, GT_THRESHOLDS[dtype]["DEL"], GT_THRESHOLDS[dtype]["INS"], threads=2)
    for k in ("fp_support", "cov_ins", "cov_del", "n_support", "depth2", "new_gt"):
        assert np.array_equal(got[k], exp[k]), k
    assert np.array_equal(np.isnan(got["ratio"]), np.isnan(exp["ratio"]))
    m = ~np.isnan(exp["ratio"])
    assert np.array_equal(got["ratio"][m], exp["ratio"][m])
    for k in ("sig_reads", "sig_del", "sig_ins"):
        g = got[k].view(V.SIG_DTYPE)
        e = exp[k]
        assert len(g) == len(e) and len(e) > 0, k
        for f in ("rec", "type", "pos", "len", "q0", "q1", "count"):
            assert np.array_equal(g[f], e[f]), (k, f)
    assert (exp["n_support"] > 0).sum() > 100 and (exp["depth2"] > 0).sum() > 100
    # idempotent: a second run on the same resident inputs gives the same bytes
    got2 = sh.download(sh.run_safe())
    for k in got:
        assert np.array_equal(got[k], got2[k], equal_nan=True) if got[k].dtype.kind == "f" else np.array_equal(got[k], got2[k])


def test_compact_staging_gives_identical_results():
    """upload() through the ops16 staging format == upload() of the BAM-native ops."""
    from volcanosv_b200.lib import pack_ops16
    L = 3_000_000
    b = synth.synth_records(41, 3000, "hifi", chrom_len=L, eqx=True)
    b["ops"][::501] = (np.uint32(70000) << 4) | (b["ops"][::501] & 15)           # long ops -> exception list
    pos, svlen, typ, gt = synth.synth_candidates(42, 500, chrom_len=L)
    cands = dict(pos=pos, svlen=svlen, type=typ, gt=gt)
    sh = ChromShard(0, "Hifi")
    sh.upload(b, cands)
    a = sh.download(sh.run_safe())
    o16, ei, ev = pack_ops16(b["ops"])
    assert len(ei) > 0
    b2 = {k: v for k, v in b.items() if k != "ops"}
    b2.update(ops16=o16, exc_idx=ei, exc_val=ev)
    sh.upload(b2, cands, n_ops=len(o16))
    c = sh.download(sh.run_safe())
    assert sh.h2d_bytes < 0.62 * (4 * len(o16) + 17 * len(b["pos"]) + 16 * 500)
    for k in a:
        assert np.array_equal(a[k], c[k], equal_nan=True) if a[k].dtype.kind == "f" else np.array_equal(a[k], c[k]), k
    from volcanosv_b200.lib import pack_ops8
    o8, ei, ev = pack_ops8(b["ops"])
    b3 = {k: v for k, v in b.items() if k != "ops"}
    b3.update(ops8=o8, exc_idx=ei, exc_val=ev)
    sh.upload(b3, cands, n_ops=len(o8))
    e = sh.download(sh.run_safe())
    for k in a:
        assert np.array_equal(a[k], e[k], equal_nan=True) if a[k].dtype.kind == "f" else np.array_equal(a[k], e[k]), k
    from volcanosv_b200.lib import pack_ops12
    o12, ep, ev = pack_ops12(b["ops"])
    b4 = {k: v for k, v in b.items() if k != "ops"}
    b4.update(ops12=o12, exc_idx=ep, exc_val=ev)
    sh.upload(b4, cands, n_ops=len(b["ops"]))
    f = sh.download(sh.run_safe())
    for k in a:
        assert np.array_equal(a[k], f[k], equal_nan=True) if a[k].dtype.kind == "f" else np.array_equal(a[k], f[k]), k


def test_shard_stream_overlapped_uploads_match_single_shard():
    """ShardStream (upload of shard k+1 in flight while shard k runs) == one ChromShard per shard."""
    import torch
    from volcanosv_b200.lib import pack_ops16
    from volcanosv_b200.pipeline import ShardStream
    L = 2_000_000
    shards = []
    for seed in (51, 52, 53):
        b = synth.synth_records(seed, 2000, "ont", chrom_len=L)
        pos, svlen, typ, gt = synth.synth_candidates(seed + 100, 300, chrom_len=L)
        shards.append((b, dict(pos=pos, svlen=svlen, type=typ, gt=gt)))
    want = []
    one = ChromShard(0, "ONT")
    for b, c in shards:
        one.upload(b, c)
        want.append(one.download(one.run_safe()))

    def pin(b):
        o16, ei, ev = pack_ops16(b["ops"])
        h = {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for k, v in b.items() if k != "ops"}
        h.update(ops16=torch.from_numpy(o16.view(np.int16)).pin_memory(), exc_idx=torch.from_numpy(ei).pin_memory(),
                 exc_val=torch.from_numpy(ev.view(np.int32)).pin_memory())
        return h, len(o16)

    ss = ShardStream(0, "ONT")
    with pytest.raises(V.VsvError):
        ss.collect()
    host = [pin(b) for b, _ in shards]
    got = []
    ss.submit(host[0][0], shards[0][1], n_ops=host[0][1])
    for i in range(len(shards)):
        if i + 1 < len(shards):
            ss.submit(host[i + 1][0], shards[i + 1][1], n_ops=host[i + 1][1])
        _, arrays = ss.collect()
        got.append({k: v.copy() for k, v in arrays.items()})
    for a, c in zip(want, got):
        for k in a:
            assert np.array_equal(a[k], c[k], equal_nan=True) if a[k].dtype.kind == "f" else np.array_equal(a[k], c[k]), k
