"""Compact table of registers / stack / spills per kernel from an `nvcc -Xptxas=-v` log (build.py --force prints one).
usage: python lattice-gaussian-mcmc_b200/build.py --force > /tmp/build.log 2>&1; python tools/spill_report.py /tmp/build.log [filter]"""
import re
import subprocess
import sys

log = open(sys.argv[1]).read().splitlines()
flt = sys.argv[2] if len(sys.argv) > 2 else ""
name = None
for i, line in enumerate(log):
    m = re.search(r"Compiling entry function '(\S+)'", line)
    if m:
        name = m.group(1)
        continue
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
    if m and name and "Function properties for " + name in log[i - 1]:
        regs = re.search(r"Used (\d+) registers", log[i + 1])
        dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        dem = re.sub(r"\(anonymous namespace\)::|lgs::", "", dem).split("(")[0]
        if flt in dem:
            print(f"{dem:70s} regs {regs.group(1) if regs else '?':>3s}  stack {m.group(1):>4s}  spill st {m.group(2):>4s} ld {m.group(3):>4s}")
