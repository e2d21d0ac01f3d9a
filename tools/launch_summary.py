"""ncu launch list (gpu__time_duration.sum CSV of `bench.py`) -> per-kernel share table (markdown)."""
import collections
import csv
import gzip
import sys

src, steps = sys.argv[1], int(sys.argv[2])
op = gzip.open if src.endswith(".gz") else open
with op(src, "rt") as f:
    rows = list(csv.reader(l for l in f if not l.startswith("==")))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
agg = collections.defaultdict(lambda: [0, 0.0])
for row in rows[1:]:
    if len(row) < len(hdr) or row[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    v = float(row[ix["Metric Value"]].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}[row[ix["Metric Unit"]]]
    name = row[ix["Kernel Name"]].replace("void ", "").replace("bqa::<unnamed>::", "")
    name = name.split("(")[0]
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print(f"{sum(v[0] for v in agg.values())} launches, {tot / 1e3:.2f} ms of kernel time over {steps} optimizer steps "
      f"({tot / 1e3 / steps:.3f} ms/step; ncu per-launch times are serialised and cold-cache)\n")
print("| kernel | launches/step | avg us | us/step | share |")
print("|---|---:|---:|---:|---:|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if v[1] / tot < 0.0005:
        continue
    print(f"| `{k[:70]}` | {v[0] / steps:.1f} | {v[1] / v[0]:.1f} | {v[1] / steps:.0f} | {v[1] / tot * 100:.1f}% |")
