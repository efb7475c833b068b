#!/usr/bin/env python
"""Attribute an ncu report's per-SASS-instruction counters to CUDA source lines.
usage: ncu_lines.py report.ncu-rep object.o kernel-substring [top]
The SASS page of the report and `nvdisasm -g` of the same object list the kernel's instructions in the same order."""
import collections, csv, os, re, subprocess, sys, tempfile

rep, obj, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = rows[1]
srows = rows[2:]
ci = {h: i for i, h in enumerate(hdr)}
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
# the kernel's section
kernel_full = rows[0][1]
want_f = "ILb1E" if "(bool)1" in kernel_full else "ILb0E"
start = [i for i, l in enumerate(dis) if l.startswith(".text.") and kname in l and (want_f in l or "ILb" not in l)][0]
lines = []
cur = ("?", 0)
for l in dis[start + 1:]:
    if l.startswith("//-------") or l.startswith(".text."):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append(cur)
assert len(lines) == len(srows), (len(lines), len(srows))
agg = collections.defaultdict(lambda: collections.Counter())
for (f, ln), r in zip(lines, srows):
    a = agg[(f, ln)]
    a["inst"] += int(r[ci["Instructions Executed"]])
    a["samples"] += int(r[ci["# Samples"]])
    a["wf"] += int(r[ci["L1 Wavefronts Shared"]] or 0)
    a["wf_ideal"] += int(r[ci["L1 Wavefronts Shared Ideal"]] or 0)
tot_i = sum(a["inst"] for a in agg.values()); tot_s = sum(a["samples"] for a in agg.values()); tot_w = sum(a["wf"] for a in agg.values())
print(f"total warp-instructions {tot_i}, samples {tot_s}, shared wavefronts {tot_w}")
text = {}
for (f, ln), a in sorted(agg.items(), key=lambda x: -x[1][os.environ.get("NCU_SORT", "inst")])[:top]:
    if f not in text:
        try: text[f] = open(os.path.join(os.path.dirname(os.path.abspath(obj)), f)).read().splitlines()
        except OSError: text[f] = []
    s = text[f][ln - 1].strip()[:90] if 0 < ln <= len(text[f]) else ""
    print(f"{f}:{ln:4d} inst {100 * a['inst'] / tot_i:5.1f}% samples {100 * a['samples'] / tot_s:5.1f}% wf {a['wf']:>10d} (ideal {a['wf_ideal']:>10d}) | {s}")
