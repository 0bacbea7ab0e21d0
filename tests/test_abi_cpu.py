"""CPU-side checks of the C-ABI boundary: the library loads and exports every symbol
include/volcanosv_b200.h declares; without a GPU, context creation fails loudly
(no CPU fallback)."""
import ctypes
import os
import re

import pytest

import volcanosv_b200 as V
from volcanosv_b200 import lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "volcanosv_b200.h")).read()
    return sorted(set(re.findall(r"\b(vsv_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    assert os.path.exists(V.LIB_PATH), "run python -m volcanosv_b200.build"
    lib = ctypes.CDLL(V.LIB_PATH)
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(L.SIGNATURES) == names, "ctypes table and header disagree"
    L.load_library()


def test_version_string():
    lib = L.load_library()
    assert b"volcanosv_b200" in lib.vsv_version()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(V.VsvError):
        V.Context(0)


def test_sig_struct_layout():
    assert V.SIG_DTYPE.itemsize == 32
    assert V.SIG_DTYPE.names == ("rec", "type", "pos", "len", "q0", "q1", "first", "count")


def test_pack_ops16_layout():
    """Host packer of the compact staging format (vsv_expand_ops16): short ops verbatim, long ops as
    a marker + exception entry; unpacking on the host restores the BAM-native words."""
    import numpy as np
    from volcanosv_b200.lib import pack_ops16
    rng = np.random.default_rng(3)
    ln = rng.integers(0, 5000, 10000).astype(np.uint32)
    ln[::97] = rng.integers(4095, 1 << 27, len(ln[::97]))
    ops = (ln << 4) | rng.integers(0, 9, 10000).astype(np.uint32)
    o16, ei, ev = pack_ops16(ops)
    assert o16.dtype == np.uint16 and o16.ctypes.data % 16 == 0
    back = o16.astype(np.uint32)
    assert np.all((back[ei] >> 4) == 0xFFF) and np.all((back[ei] & 15) == (ops[ei] & 15))
    back[ei] = ev
    assert np.array_equal(back, ops) and np.all(np.diff(ei) > 0) and len(ei) == int((ln >= 4095).sum())


def test_cli_mirrors_every_reference_operator():
    """One command per LI script on the hot path."""
    from volcanosv_b200 import cli
    for name in ("extract_contig_signature_Hifi.py", "extract_contig_signature_ONT.py", "extract_contig_signature_CLR.py",
                 "extract_reads_signature.py", "FP_filter_v1.py", "remove_redundancy.py", "sig_extract.py",
                 "calculate_signature_support.py", "filter_vcf_by_sig_cov_insdel.py", "correct_gt_del_real_data.py",
                 "correct_gt_ins_real_data.py", "filter_GT_correction.py", "volcanosv-vc-large-indel.py",
                 "Raw_variant_call-otherasm.py", "volcanosv-vc-large-indel-otherasm.py",
                 "volcanosv-vc-large-indel-otherasm-2hap.py"):
        assert name in cli.COMMANDS
    import pytest
    with pytest.raises(SystemExit):
        cli.main(["no_such_script.py"])


def test_pack_ops8_layout():
    """ops8: one byte per short M/I/D run, 0xFF + exception otherwise; inverse on the host."""
    import numpy as np
    from volcanosv_b200.lib import pack_ops8, unpack_ops8
    rng = np.random.default_rng(9)
    n = 5000
    code = rng.choice([0, 1, 2, 4, 5, 7, 8], n, p=[.5, .2, .2, .04, .02, .02, .02]).astype(np.uint32)
    ln = rng.integers(0, 200, n).astype(np.uint32)
    ln[::97] = rng.integers(4095, 1 << 27, len(ln[::97]))
    ops = (ln << 4) | code
    o8, ei, ev = pack_ops8(ops)
    assert o8.dtype == np.uint8 and o8.ctypes.data % 16 == 0
    cap = np.array([160, 47, 47] + [0] * 13)[code]
    big = (ln < 1) | (ln > cap)
    assert np.array_equal(ei, np.flatnonzero(big)) and np.array_equal(ev, ops[big])
    assert np.all(o8[big] == 0xFF) and not np.any(o8[~big] == 0xFF)
    assert o8[(code == 0) & (ln == 1)].tolist() == [0] * int(((code == 0) & (ln == 1)).sum())
    assert np.all(o8[(code == 2) & (ln == 47)] == 253)
    assert np.array_equal(unpack_ops8(o8, ei, ev), ops)


def test_pack_ops12_layout():
    """ops12: 12 bits per {M 1..128, I/D 1..8} pair in either order, odd pairs in the exception list with both
    words; token k at bits [12k, 12k+12) of the byte stream; inverse on the host."""
    import numpy as np
    from volcanosv_b200.lib import pack_ops12, unpack_ops12, ops12_bytes, load_library
    ops = np.array([(5 << 4) | 0, (3 << 4) | 2,          # M5 D3  -> token 4 | 2<<7 | 1<<10
                    (8 << 4) | 1, (128 << 4) | 0,        # I8 M128 -> 127 | 7<<7 | 1<<11
                    (129 << 4) | 0, (1 << 4) | 1,        # M too long -> exception
                    (4 << 4) | 4], np.uint32)            # lone soft clip at an odd tail -> exception
    o12, ep, ev = pack_ops12(ops)
    assert len(o12) == 12 == ops12_bytes(len(ops)) and list(ep) == [2, 3]
    assert list(ev) == [(129 << 4) | 0, (1 << 4) | 1, (4 << 4) | 4, 0]
    t0, t1 = 4 | (2 << 7) | (1 << 10), 127 | (7 << 7) | (1 << 11)
    assert list(o12[:3]) == [t0 & 0xFF, (t0 >> 8) | ((t1 & 15) << 4), t1 >> 4]
    assert np.array_equal(unpack_ops12(o12, len(ops), ep, ev), ops)
    rng = np.random.default_rng(5)
    for n in (0, 1, 16, 17, 4097):
        m = rng.integers(1, 200, n).astype(np.uint32) << 4
        x = (rng.integers(1, 12, n).astype(np.uint32) << 4) | rng.integers(1, 3, n).astype(np.uint32)
        w = np.where((np.arange(n) + (np.arange(n) > n // 3)) % 2 == 0, m, x).astype(np.uint32)
        for chunk in (8, 1 << 24):
            o12, ep, ev = pack_ops12(w, chunk=chunk)
            assert len(o12) == ops12_bytes(n) and np.all(np.diff(ep) > 0)
            assert np.array_equal(unpack_ops12(o12, n, ep, ev), w)
    lib = load_library()
    for n in (0, 1, 2, 16, 17, 12345):
        assert lib.vsv_ops12_bytes(n) == ops12_bytes(n)


def test_native_ops12_packer_matches_the_numpy_statement():
    """vsv_pack_ops12 (host threads, no GPU) writes the same bytes and the same exception list as lib.pack_ops12,
    whatever the thread count; capacity query pattern."""
    import ctypes as C
    import numpy as np
    from volcanosv_b200 import lib as L
    from volcanosv_b200 import synth
    rng = np.random.default_rng(12)
    cases = [np.zeros(0, np.uint32), np.array([(7 << 4) | 4], np.uint32),
             synth.synth_records(77, 300, "ont")["ops"], synth.synth_records(78, 300, "hifi", eqx=True)["ops"][:100001]]
    n = 70001
    m = rng.integers(1, 140, n).astype(np.uint32) << 4
    x = (rng.integers(1, 11, n).astype(np.uint32) << 4) | rng.integers(1, 3, n).astype(np.uint32)
    cases.append(np.where((np.arange(n) + (np.arange(n) > n // 2)) % 2 == 0, m, x).astype(np.uint32))
    for ops in cases:
        want = L.pack_ops12(ops)
        for threads in (1, 3, 0):
            got = L.pack_ops12_native(ops, threads=threads)
            for a, b in zip(got, want):
                assert a.dtype == b.dtype and np.array_equal(a, b)
    lib = L.load_library()
    ops = cases[-1]
    o12 = np.empty(L.ops12_bytes(len(ops)), np.uint8)
    n_exc = C.c_int64(0)
    rc = lib.vsv_pack_ops12(L._p(ops), len(ops), 2, L._p(o12), None, None, 0, C.byref(n_exc))
    assert rc == L.E_CAPACITY and n_exc.value == len(L.pack_ops12(ops)[1]) and np.array_equal(o12, L.pack_ops12(ops)[0])
    assert lib.vsv_pack_ops12(None, 5, 1, None, None, None, 0, C.byref(n_exc)) == L.E_ARG


def test_ops12_round_trip_property():
    """Any u32 ops array survives pack -> unpack in both packers (hypothesis: arbitrary words, skewed to the
    M / short-indel alphabet so that tokens, both orders and exceptions all occur)."""
    import numpy as np
    from hypothesis import given, settings, strategies as st
    from volcanosv_b200 import lib as L
    word = st.one_of(
        st.builds(lambda ln: (ln << 4), st.integers(0, 140)),
        st.builds(lambda ln, c: (ln << 4) | c, st.integers(0, 10), st.sampled_from([1, 2])),
        st.integers(0, 2 ** 32 - 1))

    @settings(max_examples=150, deadline=None)
    @given(st.lists(word, max_size=70))
    def run(words):
        ops = np.array(words, dtype=np.uint32)
        a = L.pack_ops12(ops, chunk=8)
        b = L.pack_ops12_native(ops, threads=2)
        assert all(np.array_equal(x, y) for x, y in zip(a, b))
        assert np.array_equal(L.unpack_ops12(a[0], len(ops), a[1], a[2]), ops)

    run()
