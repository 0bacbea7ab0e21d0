"""The a13 edit-distance oracle (oracle.vsv_oracle.edit_distance, the DP the GPU kernel is checked against) pinned to the
REFERENCE'S OWN vendored edlib (oracle/_ref/libedlib_ref.so, compiled by oracle/Makefile from
/root/reference/.../submodules/edlib/edlib.cpp): `edlib.align(a, b)["editDistance"]` of remove_redundancy.py:75-81."""
import random

import pytest

from oracle import edlib_ref as E
from oracle import vsv_oracle as O

pytestmark = pytest.mark.skipif(not E.available(), reason="oracle/_ref/libedlib_ref.so not built (needs /root/reference at build time)")


def test_oracle_dp_equals_reference_edlib():
    rng = random.Random(20261105)
    cases = [("", ""), ("A", ""), ("", "ACGT"), ("ACGT", "ACGT"), ("AAAA", "TTTT"), ("ACGTACGT", "ACGACGT")]
    for _ in range(300):
        a = "".join(rng.choice("ACGT") for _ in range(rng.randint(0, 400)))
        if rng.random() < 0.5:     # a noisy copy: the case remove_redundancy cares about
            b = "".join(c if rng.random() > 0.1 else rng.choice("ACGT") for c in a)
            b = b[: rng.randint(0, len(b))] + "".join(rng.choice("ACGT") for _ in range(rng.randint(0, 30))) + b[rng.randint(0, len(b)):]
        else:
            b = "".join(rng.choice("ACGTN") for _ in range(rng.randint(0, 400)))
        cases.append((a, b))
    for a, b in cases:
        assert O.edit_distance(a, b) == E.distance(a, b), (len(a), len(b))


def test_similarity_threshold_of_remove_redundancy_on_reference_edlib():
    """(len1 + len2 - ed) / (len1 + len2) >= 0.5 (RR:75-81): the planted near-duplicate of the e2e fixture passes, unrelated sequences fail."""
    rng = random.Random(7)
    s1 = "".join(rng.choice("ACGT") for _ in range(300))
    s2 = "".join(c if rng.random() > 0.08 else rng.choice("ACGT") for c in s1) + "".join(rng.choice("ACGT") for _ in range(11))
    s3 = "".join(rng.choice("ACGT") for _ in range(300))
    sim = lambda a, b: (len(a) + len(b) - E.distance(a, b)) / (len(a) + len(b))
    assert sim(s1, s2) >= 0.5
    assert sim(s1, s3) < 0.9
