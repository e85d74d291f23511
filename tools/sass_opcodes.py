#!/usr/bin/env python
"""Per-kernel SASS opcode counts of the in-tree libbax.so (cuobjdump -sass): the instructions that prove which hardware
path a kernel takes — DMMA (FP64 tensor pipe), UBLKCP (cp.async.bulk, the TMA engine's non-tensor copies), SYNCS
(mbarrier), SHFL, MUFU, DFMA/DADD/DMUL, LDS/STS, LDG/STG — plus the absence of UTCMMA / LDTM / UTMALDG (tcgen05 has no
FP64 kind; the bulk copies here are 1-D contiguous runs).    python tools/sass_opcodes.py > profiles/r02/sass_opcodes.txt"""
import collections, os, re, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(root, "extrinsic-camera-calibration-from-a-moving-person_b200", "libbax.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout.split("\n")
cols = ["DMMA", "UBLKCP", "SYNCS", "UTCMMA", "LDTM", "UTMALDG", "DFMA", "DADD", "DMUL", "MUFU", "SHFL", "LDS", "STS", "LDG", "STG", "BAR", "ATOM"]
name, per = None, collections.OrderedDict()
for l in sass:
    m = re.search(r"Function : (\S+)", l)
    if m:
        name = subprocess.run(["cu++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = name.replace("void ", "")
        name = name[:name.index(">(") + 1] if ">(" in name else name.split("(")[0]
        per[name] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P[T\d]+\s+)?([A-Z0-9_]+)", l)
    if m and name:
        op = m.group(1)
        per[name]["total"] += 1
        for c in cols:
            if op.startswith(c):
                per[name][c] += 1
print(f"{'kernel':46s} {'total':>7s} " + " ".join(f"{c:>7s}" for c in cols))
for k, v in sorted(per.items(), key=lambda kv: -kv[1]["total"]):
    print(f"{k[:46]:46s} {v['total']:7d} " + " ".join(f"{v[c]:7d}" for c in cols))
tot = collections.Counter()
for v in per.values():
    tot.update(v)
print(f"{'ALL KERNELS':46s} {tot['total']:7d} " + " ".join(f"{tot[c]:7d}" for c in cols))
