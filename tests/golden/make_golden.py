#!/usr/bin/env python
"""Generates the golden fixtures under tests/golden/ from the UNMODIFIED reference solver.

Runs the Tier-1 driver binaries oracle/_ref/ref_* (built from /root/reference by
oracle/build_ref.sh; only possible in the build container, where /root/reference exists) on a list
of small cases and stores every grid's `cur` layer, in the reference's linear order, in one
compressed .npz per case, together with the scalars the reference derived (dt, relPhaseVelocity).
The fixtures travel with the repo; the reference does not.

Usage: python tests/golden/make_golden.py [case-name ...]
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from refio import read_records  # noqa: E402

# name: (variant, nx, ny, nz, pml, tfsf, steps, use_pml, use_tfsf, theta, phi, psi)
CASES = {
    # 3D double real (reference needs shim 3, see oracle/build_ref.sh)
    "3d_d_pml_tfsf":        ("3d_d", 32, 32, 32, 6, 10, 30, 1, 1, 90, 0, 90),
    "3d_d_pml_tfsf_angled": ("3d_d", 32, 32, 32, 6, 10, 30, 1, 1, 60, 30, 45),
    "3d_d_pml_tfsf_ragged": ("3d_d", 26, 33, 30, 5, 9, 24, 1, 1, 90, 90, 0),
    "3d_d_pml_src":         ("3d_d", 32, 32, 32, 6, 10, 30, 1, 0, 90, 0, 90),
    "3d_d_tfsf":            ("3d_d", 32, 32, 32, 6, 10, 30, 0, 1, 90, 0, 90),
    "3d_d_vacuum_src":      ("3d_d", 32, 32, 32, 6, 10, 30, 0, 0, 90, 0, 90),
    # 3D double complex (reference builds unmodified)
    "3d_dc_pml_tfsf":       ("3d_dc", 24, 24, 24, 5, 8, 24, 1, 1, 90, 0, 90),
    "3d_dc_vacuum_src":     ("3d_dc", 24, 24, 24, 5, 8, 24, 0, 0, 90, 0, 90),
    # 3D float real
    "3d_f_pml_tfsf":        ("3d_f", 32, 32, 32, 6, 10, 30, 1, 1, 90, 0, 90),
    "3d_f_vacuum_src":      ("3d_f", 32, 32, 32, 6, 10, 30, 0, 0, 90, 0, 90),
    # 2D TMz (reference needs shim 4); the point source sits at Ez(70, Ny/2)
    "tmz_d_pml_src":        ("tmz_d", 96, 80, 1, 8, 14, 60, 1, 0, 90, 0, 90),
    "tmz_d_pml_tfsf":       ("tmz_d", 96, 80, 1, 8, 14, 60, 1, 1, 90, 0, 90),
    "tmz_d_tfsf":           ("tmz_d", 96, 80, 1, 8, 14, 60, 0, 1, 90, 0, 90),
    "tmz_d_vacuum_src":     ("tmz_d", 96, 80, 1, 8, 14, 60, 0, 0, 90, 0, 90),
    "tmz_dc_pml_tfsf":      ("tmz_dc", 96, 80, 1, 8, 14, 60, 1, 1, 90, 0, 90),
    "tmz_f_pml_src":        ("tmz_f", 96, 80, 1, 8, 14, 60, 1, 0, 90, 0, 90),
    "tmz_f_pml_tfsf":       ("tmz_f", 96, 80, 1, 8, 14, 60, 1, 1, 90, 0, 90),
    # TMz at phi = 45 degrees, the one oblique angle SchemeTMz.cpp:40-49 accepts (relPhi != 1), ragged sizes; complex point
    # source; float and complex at oblique 3D angles
    "tmz_d_pml_tfsf_45":     ("tmz_d", 90, 77, 1, 7, 13, 50, 1, 1, 90, 45, 90),
    "tmz_f_pml_tfsf_45":     ("tmz_f", 90, 77, 1, 7, 13, 50, 1, 1, 90, 45, 90),
    "tmz_dc_pml_src":        ("tmz_dc", 96, 80, 1, 8, 14, 40, 1, 0, 90, 0, 90),
    "3d_f_pml_tfsf_angled":  ("3d_f", 30, 27, 25, 5, 8, 24, 1, 1, 55, 25, 40),
    "3d_dc_pml_tfsf_angled": ("3d_dc", 26, 23, 22, 4, 7, 20, 1, 1, 65, 35, 50),
    # large enough for a sigma-free core box in the kernels' tile granularity: fast mode is then compared
    # with the reference itself where it skips D/B, not only with the oracle
    "3d_d_pml_tfsf_core":   ("3d_d", 48, 36, 28, 5, 9, 30, 1, 1, 60, 30, 45),
    "3d_d_pml_src_core":    ("3d_d", 48, 36, 28, 5, 9, 30, 1, 0, 90, 0, 90),
    "3d_f_pml_src_core":    ("3d_f", 72, 40, 20, 4, 8, 30, 1, 0, 90, 0, 90),
    # --save-res: the reference subtracts the incident field inside the TF/SF box before saving (Scheme3D.cpp:2767-2921)
    "3d_d_pml_tfsf_saveres": ("3d_d", 30, 28, 26, 5, 8, 40, 1, 1, 70, 40, 55, "saveres"),
    # the reference's own DATDumper writes Ez and Hy of this run: the fixture keeps the files' bytes (dat_Ez, dat_Hy)
    "3d_d_pml_tfsf_dat":    ("3d_d", 24, 22, 20, 4, 7, 20, 1, 1, 65, 35, 50, "dat"),
}
DX, WAVELENGTH = "0.0005", "0.02"


def main():
    names = sys.argv[1:] or list(CASES)
    for name in names:
        variant, nx, ny, nz, pml, tfsf, steps, use_pml, use_tfsf, th, ph, ps = CASES[name][:12]
        save_res = len(CASES[name]) > 12 and CASES[name][12] == "saveres"
        dump_dat = len(CASES[name]) > 12 and CASES[name][12] == "dat"
        exe = os.path.join(ROOT, "oracle", "_ref", "ref_" + variant)
        with tempfile.TemporaryDirectory() as tmp:
            out = os.path.join(tmp, "dump.bin")
            args = [exe] + [str(v) for v in (nx, ny, nz, pml, tfsf, steps, use_pml, use_tfsf)] + [DX, WAVELENGTH, str(th), str(ph), str(ps), out]
            if save_res or dump_dat:
                args.append("1" if save_res else "0")
            if dump_dat:
                args.append("1")
            subprocess.run(args, check=True, stdout=subprocess.DEVNULL, cwd=tmp)
            rec = read_records(out)
            if dump_dat:
                for comp in ("Ez", "Hy"):
                    rec["dat_" + comp] = np.fromfile(os.path.join(tmp, "current[%d]_rank-0_%s.dat" % (steps, comp)), dtype=np.uint8)
        meta = dict(variant=variant, size=[nx, ny, nz], pml=pml, tfsf=tfsf, steps=steps, use_pml=use_pml,
                    use_tfsf=use_tfsf, theta=th, phi=ph, psi=ps, dx=float(DX), wavelength=float(WAVELENGTH),
                    save_res=int(save_res))
        sc = rec.pop("scalars")
        meta.update(dt=float(sc[0]), rel_phase_velocity=float(sc[3]), source_frequency=float(sc[2]),
                    angle_teta=float(sc[4]), angle_phi=float(sc[5]), angle_psi=float(sc[6]))
        # the 3-D sigma grids only vary along their own axis: keep the 1-D profile and check that here
        m = (nx + 1, ny + 1, nz + 1) if variant.startswith("3d") else (nx + 1, ny + 1)
        for ax, key in enumerate(["SigmaX", "SigmaY", "SigmaZ"]):
            g = rec.pop(key).real.reshape(m)
            if ax >= len(m):
                assert not g.any()
                continue
            prof = np.moveaxis(g, ax, 0).reshape(m[ax], -1)
            assert (prof == prof[:, :1]).all(), key
            rec[key] = prof[:, 0].copy()
        rec["Eps"] = rec["Eps"].real.copy()
        rec["Mu"] = rec["Mu"].real.copy()
        assert (rec["Mu"] == 1).all()
        del rec["Mu"]
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, meta=np.array(repr(meta)), **rec)
        print("%-24s %8.1f KiB" % (name, os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
