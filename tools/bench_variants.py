#!/usr/bin/env python
"""Runs bench.py (short, device-resident numbers only) once per library build found in
fdtd3d_b200/variants/ plus the default build, and prints one summary line each.  Tuning aid."""
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
libs = [None] + sorted(glob.glob(os.path.join(ROOT, "fdtd3d_b200", "variants", "lib_*.so")))
extra = sys.argv[1:]
for lib in libs:
    env = dict(os.environ)
    if lib:
        env["FDTD3D_B200_LIB"] = lib
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "6", "--warmup", "2", "--no-e2e",
                        "--no-cpu-baseline", "--no-parity", "--quick"] + extra, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    name = os.path.basename(lib) if lib else "default"
    try:
        j = json.loads(r.stdout.strip().splitlines()[-1])
        rf = j["roofline"]
        print("%-22s %8.0f Mcell/s  E %.3f ms  H %.3f ms  one-pass %s ms  frac %.3f  of needed %.3f" % (name, j["value"], rf["e_phase_ms"], rf["h_phase_ms"],
              "%.3f" % rf["one_pass_ms"] if rf.get("one_pass_ms") else "-", rf["frac"], rf.get("whole_step_frac_of_needed_bytes") or 0), flush=True)
    except Exception:
        print(name, "FAILED", r.stdout[-400:], flush=True)
