#!/usr/bin/env python
"""Per-source-line view of an ncu report without the GUI: joins `ncu --page source --csv` (SASS
rows with executed-instruction and stall-sample counts) with `nvdisasm -g` line info of the same
cubin, by instruction order inside the kernel.

    python tools/ncu_lines.py <report.ncu-rep> <kernel-regex> [lib.so] [min_pct]
"""
import csv
import re
import subprocess
import sys
import tempfile
from collections import defaultdict
from pathlib import Path

rep, kern = sys.argv[1], sys.argv[2]
lib = str(Path(sys.argv[3]).resolve()) if len(sys.argv) > 3 else str(Path(__file__).resolve().parents[1] / "ugu_b200/lib/libugu_b200_raster.so")
min_pct = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0

raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
# first kernel instance only
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]
body = []
for r in rows[h + 1:]:
    if not r or r[0] in ("Kernel Name", "Address"):
        break
    body.append(r)
ci = {n: hdr.index(n) for n in ("Source", "# Samples", "Instructions Executed", "Thread Instructions Executed")}

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, capture_output=True)
dis = []
for cubin in sorted(Path(tmp).glob("*.cubin")):  # one per .cu of the library: keep the one that holds the kernel
    d = subprocess.run(["nvdisasm", "-g", "-c", str(cubin)], capture_output=True, text=True).stdout.splitlines()
    if any(l.startswith(".text.") and re.search(kern, l) for l in d):
        dis = d
        break
# locate the function
def parse(start):
    out, cur = [], 0
    for l in dis[start + 1:]:
        if l.startswith("\t.section") or l.startswith(".text."):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = int(m.group(2))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4}\*/", l):
            out.append((cur, l.split("*/", 1)[1].strip().rstrip(";")))
    return out


# several instantiations may match the regex: take the one whose instruction count fits the profile
cands = [i for i, l in enumerate(dis) if l.startswith(".text.") and re.search(kern, l)]
start = min(cands, key=lambda i: abs(len(parse(i)) - len(body)))
lines = parse(start)
if len(lines) != len(body):
    print(f"warning: {len(lines)} disassembled vs {len(body)} profiled instructions", file=sys.stderr)
src = Path(re.search(r'//## File "([^"]+)"', "\n".join(dis[start:start + 50])).group(1)).read_text().splitlines()
agg = defaultdict(lambda: [0, 0, 0])
for (ln, _), r in zip(lines, body):
    a = agg[ln]
    a[0] += int(r[ci["Instructions Executed"]] or 0)
    a[1] += int(r[ci["# Samples"]] or 0)
    a[2] += int(r[ci["Thread Instructions Executed"]] or 0)
ti = sum(a[0] for a in agg.values()) or 1
ts = sum(a[1] for a in agg.values()) or 1
print(f"kernel {kern}: {ti} warp instructions, {ts} stall samples")
for ln in sorted(agg):
    a = agg[ln]
    if a[0] / ti * 100 >= min_pct or a[1] / ts * 100 >= min_pct:
        text = src[ln - 1].strip() if 0 < ln <= len(src) else "?"
        print(f"{ln:5d} inst {a[0] / ti * 100:5.1f}%  samples {a[1] / ts * 100:5.1f}%  lanes {a[2] / max(a[0], 1):4.1f} | {text[:100]}")
