#!/usr/bin/env python
"""Join the SASS-level source page of an ncu report (ncu -i X.ncu-rep --page source --csv) to source lines through the
line table of the in-tree libbax.so (cuobjdump -xelf + nvdisasm -g): executed warp-instructions and stall samples per
source line of one kernel.
    python tools/ncu_source_lines.py source.csv <mangled-kernel-prefix> [cubin-name=rsx]"""
import collections, csv, os, re, subprocess, sys, tempfile
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(root, "extrinsic-camera-calibration-from-a-moving-person_b200", "libbax.so")
src_csv, prefix = sys.argv[1], sys.argv[2]
cub = sys.argv[3] if len(sys.argv) > 3 else "rsx"
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub + ".sm_100a.cubin")], capture_output=True, text=True).stdout.split("\n")
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]; ia, ie, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
data = rows[2:]; base = int(data[0][ia], 16)
ex, sm = {}, {}
for r in data:
    try:
        off = int(r[ia], 16) - base; ex[off] = int(r[ie]); sm[off] = int(r[isamp])
    except Exception:
        pass
start = [i for i, l in enumerate(sass) if l.startswith(".text." + prefix)][0]
end = [i for i, l in enumerate(sass) if l.startswith(".text.") and i > start][0]
cur = None; agg = collections.Counter(); samp = collections.Counter(); stat = collections.Counter()
for l in sass[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+\S", l)
    if m and cur:
        off = int(m.group(1), 16)
        if off in ex:
            agg[cur] += ex[off]; samp[cur] += sm[off]; stat[cur] += 1
tot, tots = sum(agg.values()), sum(samp.values())
print(f"kernel {prefix}: {tot} executed warp-instructions, {tots} samples, {sum(stat.values())} SASS instructions matched")
import signal
signal.signal(signal.SIGPIPE, signal.SIG_DFL)
for (f, ln), c in agg.most_common(40):
    print(f"{f}:{ln:<5d} executed {100 * c / tot:5.1f} %  samples {100 * samp[(f, ln)] / max(tots, 1):5.1f} %  static {stat[(f, ln)]}")
