"""Pins applied while the reference's scripts run as the e2e oracle:
np.argsort -> kind='stable' (the reference's default is unstable, so tie order would depend on
the numpy build)."""
import numpy as _np

_orig = _np.argsort


def _stable_argsort(a, *args, **kw):
    if len(args) < 2:
        kw.setdefault("kind", "stable")
    return _orig(a, *args, **kw)


_np.argsort = _stable_argsort
