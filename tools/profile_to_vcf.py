"""Developer helper: cProfile of the to-final-VCF leg of bench.py at a small read scale (the host-side cost is the contig side + text)."""
import cProfile
import os
import pstats
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from volcanosv_b200 import synth  # noqa: E402

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.03
b = bench.make_chrom("ont", synth.CHR10_LEN, int(synth.n_records("ont") * scale), 10)
b["sa_self_chrom"] = 0
bench.to_vcf_leg("ont", "ONT", b, 1, warm=1)
pr = cProfile.Profile()
pr.enable()
out = bench.to_vcf_leg("ont", "ONT", b, 2, warm=0)
pr.disable()
print(out)
pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
