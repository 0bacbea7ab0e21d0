"""ShardStream (two shard slots per GPU, uploads overlapped with the running shard) on the resident engine.
The step itself is pinned against the oracle in tests/test_gpu_shard.py."""
import numpy as np
import pytest

import volcanosv_b200 as V
from volcanosv_b200 import synth
from volcanosv_b200.pipeline import ChromShard

pytestmark = pytest.mark.gpu


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
        h = {k: torch.from_numpy(np.ascontiguousarray(v.view(np.int16) if v.dtype == np.uint16 else v)).pin_memory()
             for k, v in b.items() if k != "ops"}
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
