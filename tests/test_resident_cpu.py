"""Host-side packing of the resident shard engine (volcanosv_b200/resident.py, lib.next_same_name): name interning,
same-name links (ERS:227-237), SA-tag pre-parse (SE:328-358) -- against plain Python restatements.  No GPU."""
import re
from collections import defaultdict

import numpy as np

from volcanosv_b200 import lib as L
from volcanosv_b200 import resident as RS


def test_intern_names_ids_and_first_appearance():
    rng = np.random.default_rng(3)
    names = ["read_%d" % rng.integers(0, 40) for _ in range(300)] + ["x", "", "x", "read_0/1"]
    ids, first = RS.intern_names(names)
    assert len(ids) == len(names)
    seen = {}
    for i, (nm, k) in enumerate(zip(names, ids)):
        assert seen.setdefault(nm, int(k)) == int(k)          # one id per name
        assert int(first[k]) <= i and names[int(first[k])] == nm
    assert len(set(seen.values())) == len(seen)               # different names, different ids
    for nm, k in seen.items():
        assert int(first[k]) == names.index(nm)               # first appearance in record order
    ids2, first2 = RS.intern_names(np.asarray(names))
    assert np.array_equal(ids, ids2) and np.array_equal(first, first2)
    e_ids, e_first = RS.intern_names([])
    assert len(e_ids) == 0 and len(e_first) == 0


def test_next_same_name_links_consecutive_records_of_a_name():
    rng = np.random.default_rng(5)
    for n in (0, 1, 2, 57, 1000):
        name_id = rng.integers(0, max(1, n // 3), n).astype(np.int32)
        nxt = L.next_same_name(name_id)
        by = defaultdict(list)
        for i, k in enumerate(name_id):
            by[int(k)].append(i)
        exp = np.full(n, -1, np.int32)
        for idx in by.values():                               # ERS:227-237 walks consecutive pairs of one name in BAM order
            for a, b in zip(idx, idx[1:]):
                exp[a] = b
        assert np.array_equal(nxt, exp)


def _parse_sa_plain(entry):
    chrom, pos, strand, cg, mq, _nm = entry.split(",")
    items = [(int(n), ch) for n, ch in re.findall(r"(\d+)([MIDNSHP=X])", cg)]
    fs = items[0][0] if items[0][1] == "S" else 0
    ls = items[-1][0] if items[-1][1] == "S" else 0
    return chrom, int(pos), 0 if strand == "+" else 1, fs, ls, sum(n for n, ch in items if ch in "MD=X"), int(mq)


def test_parse_sa_and_pack_sa_columns():
    entries = ["chr1,1000,+,100S50M2D30M20S,60,3", "chr2,5,-,80M5I15M,0,1", "chr1,77,+,10H40=2X8M,255,0", "chrUn_x,9,-,30S70M,300,9"]
    for e in entries:
        assert RS.parse_sa(e) == _parse_sa_plain(e)
    # hard clips are not soft clips (acquire_clip_pos looks at 'S' only, SE:328-343)
    assert RS.parse_sa(entries[2])[3:6] == (0, 0, 50)

    class R:
        pass
    R.n = 5
    R.sa = [None, entries[0] + ";" + entries[1] + ";", None, entries[3] + ";", entries[2] + ";"]
    cols = RS.pack_sa(R, "chr1")
    assert cols["rec"].tolist() == [1, 3, 4] and cols["off"].tolist() == [0, 2, 3, 4]
    assert cols["chrom"].tolist() == [0, 1, 2, 0]             # the shard's chromosome is id 0, others in order of appearance
    assert cols["pos"].tolist() == [1000, 5, 9, 77] and cols["strand"].tolist() == [0, 1, 1, 0]
    assert cols["first_s"].tolist() == [100, 0, 30, 0] and cols["last_s"].tolist() == [20, 0, 0, 0]
    assert cols["bias"].tolist() == [82, 95, 70, 50]
    assert cols["mapq"].tolist() == [60, 0, 255, 255]          # mapq column is u8: values above 255 saturate
    R.sa = [None] * 5
    assert RS.pack_sa(R, "chr1") is None
