"""Test-only stub (the reference imports matplotlib.pyplot at module top and never plots)."""
