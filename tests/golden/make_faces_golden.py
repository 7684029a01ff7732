#!/usr/bin/env python
"""TF/SF face membership of every cell, from the reference's own floating-point predicates
(YeeGridLayout::doNeedTFSFUpdate{Ex..Hz}Border, linked by oracle/build_ref.sh into oracle/_ref/ref_faces), for the
SURVEY §8c case (N = 32, L = 8: 24 boxes of 272 cells) and a ragged one.  One byte per cell and component, bit d =
LayoutDirection d.  Build container only.  Usage: python tests/golden/make_faces_golden.py"""
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CASES = {"n32_l8": ((32, 32, 32), (8, 8, 8)), "ragged": ((26, 33, 30), (5, 9, 7)), "thin": ((12, 9, 7), (2, 3, 1))}

out = {}
for name, (n, l) in CASES.items():
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "f.bin")
        subprocess.run([os.path.join(ROOT, "oracle", "_ref", "ref_faces")] + [str(v) for v in n + l] + [path], check=True)
        out[name] = np.fromfile(path, dtype=np.uint8).reshape((6,) + n)
        out[name + "_cfg"] = np.array(n + l)
    print(name, [int((out[name] & (1 << d) != 0).sum()) for d in range(6)])
np.savez_compressed(os.path.join(HERE, "tfsf_faces.faces.npz"), **out)
