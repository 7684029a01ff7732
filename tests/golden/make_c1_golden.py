#!/usr/bin/env python
"""BASELINE.json configs[0] exactly as written (the README launch example): 3D 80^3, PML 10, TF/SF 20,
dx = 0.0005, wavelength = 0.02, 100 steps, double -- real and complex -- run ONCE by the unmodified reference
solver (oracle/_ref/ref_3d_d, ref_3d_dc; about 4 minutes each on one core) and condensed into small fixtures:

  xor[field]    uint64[80]   XOR of the raw bits of every i-plane (bit-exact check of all 512 000 cells per field)
  sumsq[field]  float64      sum |f|^2 over the grid
  plane_<i>[field]           the full i-planes PLANES (tolerance checks of the fast modes need values, not checksums)
  EInc, HInc                 the whole incident line

A full dump would be 50-100 MB; this is ~3 MB (real) / ~6 MB (complex).  The script also runs the Tier-2 oracle on the
same configuration and refuses to write the fixture unless every field is bit-equal to the reference dump, so the
GPU tests may use the oracle as a full-field stand-in for the reference at this size.

Usage: python tests/golden/make_c1_golden.py [c1_d] [c1_dc]      (build container only: needs oracle/_ref)
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from refio import read_records  # noqa: E402

N, PML, TFSF, STEPS = 80, 10, 20, 100
DX, WAVELENGTH = "0.0005", "0.02"
PLANES = (10, 20, 40, 59, 70)
CASES = {"c1_d": "3d_d", "c1_dc": "3d_dc"}
FIELDS = ["Ex", "Ey", "Ez", "Hx", "Hy", "Hz", "Dx", "Dy", "Dz", "Bx", "By", "Bz"]


def condense(rec):
    """The compact form of a dict of full fields (also used by the tests on the GPU's output)."""
    out = {}
    for name in FIELDS:
        a = np.ascontiguousarray(rec[name])
        bits = a.view(np.uint64).reshape(N, -1)
        out["xor_" + name] = np.bitwise_xor.reduce(bits, axis=1)
        out["sumsq_" + name] = np.array(float((np.abs(a.astype(np.complex128)) ** 2).sum()))
        planes = a.reshape(N, -1)
        for i in PLANES:
            out["plane_%d_%s" % (i, name)] = planes[i].copy()
    out["EInc"] = rec["EInc"].copy()
    out["HInc"] = rec["HInc"].copy()
    return out


def main():
    import oracle_lib as O
    for name in sys.argv[1:] or list(CASES):
        variant = CASES[name]
        cplx = variant.endswith("c")
        exe = os.path.join(ROOT, "oracle", "_ref", "ref_" + variant)
        with tempfile.TemporaryDirectory() as tmp:
            out = os.path.join(tmp, "dump.bin")
            args = [exe] + [str(v) for v in (N, N, N, PML, TFSF, STEPS, 1, 1)] + [DX, WAVELENGTH, "90", "0", "90", out]
            subprocess.run(args, check=True, stdout=subprocess.DEVNULL, cwd=tmp)
            rec = read_records(out)
        sc = rec.pop("scalars")
        cfg = O.make_config(size=(N, N, N), pml=PML, tfsf=TFSF, use_pml=True, use_tfsf=True, value_type=8,
                            complex_values=cplx, dx=float(DX), wavelength=float(WAVELENGTH), exact_mode=1)
        assert cfg.dt == float(sc[0]) and cfg.rel_phase_velocity == float(sc[3])
        ref2 = O.run(cfg, STEPS, O.reference_setup(cfg))
        for key in FIELDS + ["EInc", "HInc"]:
            want = rec[key] if cplx else rec[key].real
            assert np.array_equal(ref2[key], want), "Tier-2 oracle differs from the reference dump in " + key
        full = {k: (rec[k] if cplx else np.ascontiguousarray(rec[k].real)) for k in FIELDS + ["EInc", "HInc"]}
        small = condense(full)
        meta = dict(variant=variant, size=[N, N, N], pml=PML, tfsf=TFSF, steps=STEPS, dx=float(DX), wavelength=float(WAVELENGTH),
                    theta=90, phi=0, psi=90, dt=float(sc[0]), rel_phase_velocity=float(sc[3]), planes=list(PLANES),
                    oracle_bit_equal=True)
        path = os.path.join(HERE, name + ".c1.npz")
        np.savez_compressed(path, meta=np.array(repr(meta)), **small)
        print("%-8s %8.1f KiB  max|Ez| = %.3g" % (name, os.path.getsize(path) / 1024, float(np.abs(full["Ez"]).max())))


if __name__ == "__main__":
    main()
