#!/usr/bin/env python
"""Diagnostic: error growth of the float kernels over many steps (fast and exact mode) against the float oracle and
against the double oracle, family-wise relative errors.  Usage: python tools/float_drift.py [steps] [n]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import oracle_lib as O  # noqa: E402
from gpu_util import family_rel_errs, run_gpu  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
args = dict(size=(n, n, n), pml=8, tfsf=14, use_pml=True, use_tfsf=True, theta=65, phi=25, psi=40)
cfg_f = O.make_config(value_type=4, exact_mode=0, **args)
cfg_d = O.make_config(value_type=8, exact_mode=0, **args)
keys = O.field_names(cfg_f) + ["EInc", "HInc"]
ref_f = O.run(cfg_f, steps, O.reference_setup(cfg_f))
ref_d = O.run(cfg_d, steps, O.reference_setup(cfg_d))
ref_d32 = {k: v.astype(np.float32) for k, v in ref_d.items()}


def fmt(e):
    return " ".join("%s=%.1e" % (k, e[k]) for k in keys)


print("float oracle vs double oracle:", fmt(family_rel_errs(ref_f, ref_d32, keys)))
for exact in (1, 0):
    cfg = O.make_config(value_type=4, exact_mode=exact, **args)
    out = run_gpu(cfg, steps, O.reference_setup(cfg))
    print("gpu float exact=%d vs float oracle :" % exact, fmt(family_rel_errs(out, ref_f, keys)))
    print("gpu float exact=%d vs double oracle:" % exact, fmt(family_rel_errs(out, ref_d32, keys)))
