"""Loads the golden fixtures of tests/golden/ (dumps of the unmodified reference solver)."""
import ast
import glob
import os

import numpy as np

import oracle_lib as O

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
VARIANT = {  # variant -> (dims, value_type, complex)
    "3d_d": (3, 8, False), "3d_dc": (3, 8, True), "3d_f": (3, 4, False),
    "tmz_d": (2, 8, False), "tmz_dc": (2, 8, True), "tmz_f": (2, 4, False),
}


def case_names():
    # *.c1.npz are the condensed BASELINE configs[0] fixtures (make_c1_golden.py), loaded by load_c1
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")) if not p.endswith((".c1.npz", ".faces.npz")))


def load_c1(name):
    """(meta, arrays) of a condensed 80^3 x 100 steps fixture: 'c1_d' (real double) or 'c1_dc' (complex double)."""
    z = np.load(os.path.join(GOLDEN_DIR, name + ".c1.npz"))
    meta = ast.literal_eval(str(z["meta"]))
    return meta, {k: z[k] for k in z.files if k != "meta"}


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = ast.literal_eval(str(z["meta"]))
    data = {k: z[k] for k in z.files if k != "meta"}
    return meta, data


def config_from_meta(meta, exact_mode=1):
    dims, vt, cplx = VARIANT[meta["variant"]]
    size = meta["size"][:dims]
    cfg = O.make_config(dims=dims, size=size, pml=meta["pml"], tfsf=meta["tfsf"], use_pml=bool(meta["use_pml"]),
                        use_tfsf=bool(meta["use_tfsf"]), value_type=vt, complex_values=cplx, dx=meta["dx"],
                        wavelength=meta["wavelength"], theta=meta["theta"], phi=meta["phi"], psi=meta["psi"],
                        exact_mode=exact_mode)
    return cfg


def inputs_from_golden(cfg, data):
    """eps / sigma exactly as the reference produced them (mu == 1 everywhere, checked at dump time)."""
    inp = {"Eps": data["Eps"], "Mu": np.ones_like(data["Eps"])}
    for a in range(cfg.dims):
        inp["Sigma" + "XYZ"[a]] = data["Sigma" + "XYZ"[a]]
    return inp
