"""Hottest SASS instructions (by stall samples) of one kernel from `ncu --page source --csv` output."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr = next(r for r in rows if "# Samples" in r)
ix = {h: i for i, h in enumerate(hdr)}


def num(r, k):
    try:
        return int(r[ix[k]])
    except (ValueError, IndexError):
        return 0


data = [r for r in rows if len(r) == len(hdr) and r is not hdr and r[ix["# Samples"]].isdigit()]
tot = sum(num(r, "# Samples") for r in data)
print("samples", tot, "warp instructions", sum(num(r, "Instructions Executed") for r in data), "sass lines", len(data))
top = sorted(enumerate(data), key=lambda ir: -num(ir[1], "# Samples"))[:n]
for i, r in sorted(top):
    print(f"{i:5d} {num(r, '# Samples'):6d} {num(r, 'Instructions Executed'):9d}  {r[ix['Source']].strip()[:100]}")
