"""Developer helper: prints the per-launch times of an `ncu --metrics gpu__time_duration.sum --csv` log."""
import csv
import sys

for f in sys.argv[1:]:
    lines = [l for l in open(f) if not l.startswith("==")]
    seq = []
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1000 if u == "ns" else (v * 1000 if u == "ms" else v)
        seq.append((row["Kernel Name"][:58], v))
    print(f, "launches", len(seq), "total us %.1f" % sum(v for _, v in seq))
    for n, v in seq:
        print("  %-58s %8.1f" % (n, v))
