"""Opcode mix of the hot region (instructions executed >= a threshold) from `ncu --page source --csv`."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
thresh, unit = int(sys.argv[2]), float(sys.argv[3])
hdr = next(r for r in rows if "# Samples" in r)
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows if len(r) == len(hdr) and r is not hdr and r[ix["# Samples"]].isdigit()]
mix, tot = collections.Counter(), 0
for r in data:
    n = int(r[ix["Instructions Executed"]] or 0)
    if n < thresh:
        continue
    op = re.sub(r"^@!?U?P\d+\s+", "", r[ix["Source"]].strip()).split()[0].split(".")[0]
    mix[op] += n
    tot += n
print("hot-region warp instructions", tot, "-> per unit", round(tot / unit, 1))
print("  ".join(f"{k}:{v / unit:.1f}" for k, v in mix.most_common(24)))
