#!/usr/bin/env python
"""Registers / spills / shared memory of every kernel in ugu_b200/lib/build.log (ptxas -v output)."""
import re
import subprocess
import sys
from pathlib import Path

log = Path(sys.argv[1] if len(sys.argv) > 1 else Path(__file__).resolve().parents[1] / "ugu_b200/lib/build.log").read_text()
names = re.findall(r"Compiling entry function '(\S+)'", log)
dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
blocks = log.split("Compiling entry function '")[1:]
for name, blk in zip(dem, blocks):
    regs = re.search(r"Used (\d+) registers", blk)
    spill = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", blk)
    smem = re.search(r"(\d+) bytes smem", blk)
    short = re.sub(r"\(.*", "", name).replace("ugu_b200::", "").replace("void ", "")
    filt = sys.argv[2] if len(sys.argv) > 2 else ""
    if filt and filt not in short:
        continue
    print(f"{short:60s} regs {regs.group(1):>3s}  spill st/ld {spill.group(1):>3s}/{spill.group(2):>3s}  smem {smem.group(1) if smem else 0}")
