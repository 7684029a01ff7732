#!/usr/bin/env python
"""Registers / spills of the 3D half-step kernels from a forced verbose build (ptxas -v).  usage: tools/ptxas_regs.py [filter]"""
import contextlib
import io
import os
import re
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    flt = sys.argv[1] if len(sys.argv) > 1 else "phase_3d_tma"
    out = subprocess.run([sys.executable, "-c", "import fdtd3d_b200.build as B; B.build(force=True, verbose=True)"],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))).stdout
    blocks = re.findall(r"Function properties for (\S+)\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\nptxas info\s+: Used (\d+) registers", out)
    names = subprocess.run(["c++filt"] + [b[0] for b in blocks], stdout=subprocess.PIPE, text=True).stdout.split("\n")
    for b, d in zip(blocks, names):
        if flt in d:
            m = re.search(r"(\w+)<(.*?)>\(", d)
            print("%-18s %-40s stack %4s spill %4s/%4s regs %s" % (m.group(1) if m else d[:18], m.group(2) if m else "", b[1], b[2], b[3], b[4]))
    if "error" in out:
        print(out[-3000:])


if __name__ == "__main__":
    main()
