"""SASS opcode histogram per kernel of libbqa.so: the mnemonics that prove the Blackwell-native paths
(UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG / UTMAREDG = TMA load / store / reduce,
UTCBAR = tcgen05.commit, HMMA = legacy mma.sync, which must be absent).
    python tools/sass_ops.py > profiles/r02_sass_ops.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ml-recipe-distributed-pytorch_b200", "libbqa.so")
OPS = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTCBAR", "UTCATOMSWS", "SYNCS", "MUFU.EX2",
       "HMMA", "HGMMA"]
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
demangle = {}
cur, counts, totals = None, collections.defaultdict(collections.Counter), collections.Counter()
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        totals[cur] += 1
        base = op.split(".")[0]
        for key in OPS:
            if op == key or op.startswith(key + ".") or (key == base):
                counts[cur][key] += 1
names = subprocess.run(["c++filt"] + list(totals), capture_output=True, text=True).stdout.splitlines()
print(f"{'kernel':70s} {'instr':>6s} " + " ".join(f"{o:>12s}" for o in OPS))
rows = []
for mangled, nm in zip(totals, names):
    nm = nm.replace("void ", "").replace("bqa::(anonymous namespace)::", "").replace("(anonymous namespace)::", "")
    nm = re.sub(r"\(.*", "", nm)
    rows.append((nm, mangled))
for nm, mangled in sorted(rows):
    c = counts[mangled]
    if not any(c.values()) and totals[mangled] < 400:
        continue
    print(f"{nm[:70]:70s} {totals[mangled]:6d} " + " ".join(f"{c.get(o, 0):12d}" for o in OPS))
tot = collections.Counter()
for c in counts.values():
    tot.update(c)
print(f"{'TOTAL':70s} {sum(totals.values()):6d} " + " ".join(f"{tot.get(o, 0):12d}" for o in OPS))
