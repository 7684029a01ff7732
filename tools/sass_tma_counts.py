#!/usr/bin/env python
"""SASS evidence of the TMA path: per kernel of libfdtd3d_b200.so the number of UTMALDG (TMA tensor loads), SYNCS
(mbarrier operations), LDG / STG.128 instructions, from `cuobjdump -sass` (needs no GPU).
Usage: python tools/sass_tma_counts.py > profiles/sass_tma.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fdtd3d_b200", "libfdtd3d_b200.so")
out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
names = subprocess.run(["/usr/local/cuda/bin/cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), stdout=subprocess.PIPE,
                       text=True).stdout.splitlines()
counts = []
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = collections.Counter()
        counts.append(cur)
        continue
    if cur is None:
        continue
    for key, pat in (("UTMALDG.2D", r"\bUTMALDG\.2D"), ("UTMALDG.3D", r"\bUTMALDG\.3D"), ("SYNCS", r"\bSYNCS\."), ("LDG", r"\bLDG\."),
                     ("STG.128", r"\bSTG\.E\.128"), ("LDS", r"\bLDS"), ("DFMA", r"\bDFMA"), ("FFMA", r"\bFFMA")):
        if re.search(pat, line):
            cur[key] += 1
print("# cuobjdump -sass fdtd3d_b200/libfdtd3d_b200.so (sm_100a): instruction counts per kernel (tools/sass_tma_counts.py)")
print("%-12s %-12s %-7s %-6s %-8s %-6s %-6s %-6s  kernel" % ("UTMALDG.3D", "UTMALDG.2D", "SYNCS", "LDG", "STG.128", "LDS", "DFMA", "FFMA"))
tot = collections.Counter()
for name, c in sorted(zip(names, counts)):
    short = re.sub(r"fdtd::\(anonymous namespace\)::", "", name)
    short = re.sub(r"\(fdtd::StepParams<.*", "", short)
    if not (c["UTMALDG.3D"] or c["UTMALDG.2D"] or "phase" in short or "fused" in short):
        continue
    tot.update(c)
    print("%-12d %-12d %-7d %-6d %-8d %-6d %-6d %-6d  %s" % (c["UTMALDG.3D"], c["UTMALDG.2D"], c["SYNCS"], c["LDG"], c["STG.128"], c["LDS"],
                                                              c["DFMA"], c["FFMA"], short[:150]))
print("# total over the kernels listed: UTMALDG.3D %d, SYNCS %d" % (tot["UTMALDG.3D"], tot["SYNCS"]))
