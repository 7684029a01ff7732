"""Helpers shared by the GPU parity tests: run the CUDA path through the C ABI on given inputs."""
import numpy as np

import fdtd3d_b200 as F
import oracle_lib as O


def run_gpu(cfg, nsteps, inputs=None, fields=None, order=F.ORDER_REFERENCE, chunks=None, fills=None):
    """Mirror of oracle_lib.run(): same inputs, same output dict, computed by libfdtd3d_b200.
    `fills`: {field: value} set with fdtd_fill instead of an upload (e.g. a homogeneous Mu)."""
    inputs = inputs or {}
    fields = fields or {}
    with F.Solver(cfg) as s:
        for key, value in (fills or {}).items():
            s.fill(key, value)
        for a in range(cfg.dims):
            key = "Sigma" + "XYZ"[a]
            if key in inputs:
                s.set_sigma(a, inputs[key])
        for key in ("Eps", "Mu"):
            if key in inputs:
                s.upload(key, inputs[key])
        for name in O.field_names(cfg):
            if name in fields:
                s.upload(name, fields[name])
        if cfg.use_tfsf:
            for name in ("EInc", "HInc"):
                if name in fields:
                    s.upload(name, fields[name])
        if chunks:
            for n in chunks:
                s.run(n)
        else:
            s.run(nsteps)
        out = {name: s.download(name) for name in O.field_names(cfg)}
        if cfg.use_tfsf:
            out["EInc"] = s.download("EInc")
            out["HInc"] = s.download("HInc")
        out["_launches"] = s.launch_count
    return out


def rel_err(a, b):
    """max |a-b| / max |b|."""
    scale = np.abs(b).max()
    if scale == 0:
        return float(np.abs(a).max())
    return float(np.abs(a - b).max() / scale)


def family(name):
    """Fields that share a physical scale: E, H, D, B components; the incident lines on their own."""
    return name if name in ("EInc", "HInc") else name[0]


def family_rel_errs(out, ref, keys):
    """{field: max |out-ref| / max |ref over the field's family|}.

    The tolerance of BASELINE.json is a *relative* one; components that are identically zero in exact
    arithmetic (e.g. Ex, Ey, Hx, Hz for the default plane wave: |.| ~ 1e-16 of Ez) hold nothing but
    rounding noise in the reference itself, so the error of every component is measured against the
    largest magnitude of its field family ("compare with an absolute floor")."""
    scale = {}
    for k in keys:
        scale[family(k)] = max(scale.get(family(k), 0.0), float(np.abs(ref[k]).max()))
    errs = {}
    for k in keys:
        sc = scale[family(k)]
        d = float(np.abs(out[k] - ref[k]).max())
        errs[k] = d / sc if sc > 0 else d
    return errs


def field_rel_errs(out, ref, keys, floor=1e-9):
    """{field: max |out-ref| / max(max |ref of that field|, floor * max |ref over the field's family|)}.

    The per-field figure SURVEY.md §8c asks for: every component against its OWN largest magnitude, with an absolute
    floor so that a component that is zero in exact arithmetic (nothing but rounding noise, ~1e-16 of its family) does
    not turn noise into a relative error of order one.  family_rel_errs is the bar of the tests; this one is reported
    beside it and asserted for every component that carries signal (significant(): own maximum >= 1e-3 of the family's
    in double, >= 0.05 in float)."""
    fam = {}
    for k in keys:
        fam[family(k)] = max(fam.get(family(k), 0.0), float(np.abs(ref[k]).max()))
    errs = {}
    for k in keys:
        own = float(np.abs(ref[k]).max())
        sc = max(own, floor * fam[family(k)])
        d = float(np.abs(out[k] - ref[k]).max())
        errs[k] = d / sc if sc > 0 else d
    return errs


def significant(ref, keys, frac=None, value_type=8):
    """The components whose own maximum is at least `frac` of their family's.

    Default frac = 8 ulp / tolerance of the value type: 1e-3 in double (8 * 1.1e-16 / 1e-12), 0.05 in float
    (8 * 6e-8 / 1e-5).  Every update adds terms of the FAMILY's magnitude (the curl of the other field), so two
    implementations that round those terms differently differ by a few ulps of the family's scale in every component;
    a component smaller than that fraction of its family (e.g. Ex of a plane wave whose polarisation projects 1 % onto
    x) cannot agree to the relative tolerance of its OWN maximum in that arithmetic, whatever the implementation.
    Such components stay under the family-relative bar of family_rel_errs."""
    if frac is None:
        frac = 1e-3 if value_type == 8 else 0.05
    fam = {}
    for k in keys:
        fam[family(k)] = max(fam.get(family(k), 0.0), float(np.abs(ref[k]).max()))
    return [k for k in keys if fam[family(k)] > 0 and float(np.abs(ref[k]).max()) >= frac * fam[family(k)]]


def random_inputs(cfg, seed):
    """Random eps/mu in [1,4], sigma profiles from the oracle, random small initial fields."""
    rng = np.random.default_rng(seed)
    rd, fd = O.real_dtype(cfg), O.field_dtype(cfg)
    mz = cfg.size[2] + 1 if cfg.dims == 3 else 1
    nm = (cfg.size[0] + 1) * (cfg.size[1] + 1) * mz
    inputs = {"Eps": rng.uniform(1.0, 4.0, nm).astype(rd), "Mu": rng.uniform(1.0, 2.0, nm).astype(rd)}
    for a in range(cfg.dims):
        inputs["Sigma" + "XYZ"[a]] = O.sigma_profile(cfg, a) if cfg.use_pml else np.zeros(cfg.size[a] + 1, rd)
    n = O.n_cells(cfg)
    fields = {}
    for name in O.field_names(cfg):
        scale = 1.0 if name[0] in "ED" else 2.6e-3
        if name[0] == "D":
            scale = 8.85e-12
        if name[0] == "B":
            scale = 3.3e-9
        v = rng.standard_normal(n)
        if cfg.complex_values:
            v = v + 1j * rng.standard_normal(n)
        fields[name] = (v * scale).astype(fd)
    return inputs, fields
