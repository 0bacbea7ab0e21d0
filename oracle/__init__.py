"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the VolcanoSV-vc large-indel hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import it, and
there only as the checker / the timed CPU baseline, never as the thing shipped.

Parity status: the reference ships NO tests or fixtures for this path,
so the oracle is pinned against outputs of the reference's own functions executed in
the build container (``tests/golden/make_golden.py`` -> ``tests/golden/*.json``).
"""
from .vsv_oracle import *  # noqa: F401,F403
