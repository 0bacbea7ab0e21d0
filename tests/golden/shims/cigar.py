"""Test-only stand-in for the un-vendored `cigar` package (cigar==0.1.3): call site SE:329."""
import re


class Cigar:
    def __init__(self, s):
        self.s = s

    def items(self):
        return [(int(n), c) for n, c in re.findall(r"(\d+)([MIDNSHP=X])", self.s)]
