"""vsv_span_depth_noalloc (csrc/vsv_depth.cu): the span-depth entry whose temporaries live in a fixed file-scope device
array, for callers that cannot allocate.  No counterpart in the reference;
checked here against the existing entry point vsv_span_depth -- integers, so bit-exact -- through the C-ABI, on two calls in
a row that share the array.  Collected last on purpose (zz)."""
import ctypes as C

import numpy as np
import pytest

from volcanosv_b200 import lib as L

pytestmark = pytest.mark.gpu

MAX_READS = 1 << 21          # VSV_NOALLOC_MAX_READS (include/volcanosv_b200.h)


def _case(seed, n_reads, n_win, chrom_len=50_000_000):
    rng = np.random.default_rng(seed)
    start = np.sort(rng.integers(0, chrom_len, n_reads)).astype(np.int32)
    ln = np.minimum(500_000, np.maximum(200, rng.lognormal(np.log(20_000), 0.6, n_reads))).astype(np.int64)
    end = (start.astype(np.int64) + ln).astype(np.int32)
    S = rng.integers(0, chrom_len, n_win).astype(np.int32)
    E = (S.astype(np.int64) + rng.choice(np.array([0, 1, 100, 200, 5000, 200_000]), n_win)).astype(np.int32)
    if n_win > 8:
        S[3] = np.iinfo(np.int32).min          # VSV_WIN_OFF: depth 0
        E[5] = S[5] - 10                       # empty window
    return start, end, S, E


def _expected(start, end, S, E):
    ctx = L.Context(0)                         # host-IO context: the existing entry point, numpy in / numpy out
    try:
        out = np.zeros(len(S), np.int32)
        ctx._chk(ctx.lib.vsv_span_depth(ctx.h, L._p(start), L._p(end), len(start), L._p(S), L._p(E), len(S), L._p(out)))
        return out
    finally:
        ctx.close()


def _dp(t):
    return C.c_void_p(t.data_ptr())


def test_span_depth_noalloc_matches_span_depth_twice_in_a_row():
    import torch
    dev = torch.device("cuda", 0)
    ctx = L.Context(0)
    ctx.set_device_io(True)
    try:
        # second call: fewer reads than the first, so stale ends of call 1 stay in the array behind the ones it sorts
        for seed, n_reads, n_win in ((1, 300_000, 20_000), (2, 70_001, 5_003), (3, 0, 64)):
            start, end, S, E = _case(seed, n_reads, n_win)
            exp = _expected(start, end, S, E)
            t = [torch.from_numpy(a).to(dev) for a in (start, end, S, E)]
            out = torch.full((n_win,), -7, dtype=torch.int32, device=dev)
            free = []
            for rep in range(2):               # the same call again: kernels are loaded by now, free device memory must not move
                out.fill_(-7)
                rc = ctx.lib.vsv_span_depth_noalloc(ctx.h, _dp(t[0]), _dp(t[1]), n_reads, _dp(t[2]), _dp(t[3]), n_win, _dp(out))
                ctx._chk(rc)
                ctx.sync()
                torch.cuda.synchronize()
                free.append(torch.cuda.mem_get_info(0)[0])
                got = out.cpu().numpy()
                assert np.array_equal(got, exp), (seed, rep, int((got != exp).sum()))
            assert free[0] == free[1]
            if n_reads:
                assert exp.max() > 10
            # plain numpy statement of count_reads_span_region (GTD:140-146) on a sample of the windows
            for w in range(0, n_win, max(1, n_win // 50)):
                if S[w] != np.iinfo(np.int32).min:
                    assert got[w] == int(np.sum((start < S[w]) & (end > E[w])))
    finally:
        ctx.close()


def test_span_depth_noalloc_rejects_what_it_cannot_hold():
    import torch
    dev = torch.device("cuda", 0)
    small = torch.zeros(16, dtype=torch.int32, device=dev)
    ctx = L.Context(0)
    try:
        # host-IO context: refused (staging host buffers would allocate)
        rc = ctx.lib.vsv_span_depth_noalloc(ctx.h, _dp(small), _dp(small), 4, _dp(small), _dp(small), 4, _dp(small))
        assert rc == L.E_ARG and b"device pointers only" in ctx.lib.vsv_last_error(ctx.h)
        ctx.set_device_io(True)
        # more reads than the fixed scratch holds: refused before any memory is touched
        rc = ctx.lib.vsv_span_depth_noalloc(ctx.h, _dp(small), _dp(small), MAX_READS + 1, _dp(small), _dp(small), 4, _dp(small))
        assert rc == L.E_CAPACITY and b"VSV_NOALLOC_MAX_READS" in ctx.lib.vsv_last_error(ctx.h)
        rc = ctx.lib.vsv_span_depth_noalloc(ctx.h, None, _dp(small), 4, _dp(small), _dp(small), 4, _dp(small))
        assert rc == L.E_ARG
        assert ctx.lib.vsv_span_depth_noalloc(ctx.h, None, None, 0, None, None, 0, None) == 0
    finally:
        ctx.close()


def test_span_depth_noalloc_at_the_limit():
    """n_reads == VSV_NOALLOC_MAX_READS: the sort's work space must fit the fixed array."""
    import torch
    dev = torch.device("cuda", 0)
    start, end, S, E = _case(9, MAX_READS, 4096, chrom_len=240_000_000)
    exp = _expected(start, end, S, E)
    ctx = L.Context(0)
    ctx.set_device_io(True)
    try:
        t = [torch.from_numpy(a).to(dev) for a in (start, end, S, E)]
        out = torch.empty(len(S), dtype=torch.int32, device=dev)
        ctx._chk(ctx.lib.vsv_span_depth_noalloc(ctx.h, _dp(t[0]), _dp(t[1]), MAX_READS, _dp(t[2]), _dp(t[3]), len(S), _dp(out)))
        ctx.sync()
        assert np.array_equal(out.cpu().numpy(), exp)
    finally:
        ctx.close()
