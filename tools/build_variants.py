#!/usr/bin/env python
"""Builds tuning variants of the library into fdtd3d_b200/variants/lib_<name>.so (see tools/bench_variants.py).

usage: tools/build_variants.py name=DEF1=V1,DEF2=V2 [name=...]      e.g.  b2s2=FDTD_TMA_MINB=2,FDTD_TMA_STAGES=2
Prints registers / spills of the double real UPML+TF/SF kernels of each variant."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fdtd3d_b200.build as B  # noqa: E402

vdir = os.path.join(ROOT, "fdtd3d_b200", "variants")
os.makedirs(vdir, exist_ok=True)
for spec in sys.argv[1:]:
    name, defs = spec.split("=", 1)
    bdir = os.path.join(vdir, "build_" + name)
    B.build(force=True, defines=defs.split(","), lib=os.path.join(vdir, "lib_%s.so" % name), build_dir=bdir)
    log = open(os.path.join(bdir, "kernels3d_tma.cu.o.log")).read()
    blocks = re.findall(r"Function properties for (\S+)\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\nptxas info\s+: Used (\d+) registers", log)
    names = subprocess.run(["c++filt"] + [b[0] for b in blocks], stdout=subprocess.PIPE, text=True).stdout.split("\n")
    for b, d in zip(blocks, names):
        m = re.search(r"(\w+)<(.*?)>\(", d)
        if m and (m.group(2).startswith("double, 1, true, true, false") or m.group(2).startswith("float, 1, true, false, false")):
            print("%-10s %-16s %-36s spill %4s/%4s regs %s" % (name, m.group(1), m.group(2), b[2], b[3], b[4]), flush=True)
