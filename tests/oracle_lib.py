"""ctypes wrapper of the Tier-2 oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Also hosts the small helpers that turn "CLI-like" parameters into an fdtd_config the way the
reference's main() + initScheme() do, using the oracle's restatement of those functions.
"""
import ctypes as C
import os
import subprocess
import numpy as np

from fdtd3d_b200 import _capi as capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_lib = None


def lib():
    global _lib
    if _lib is None:
        so = os.path.join(ORACLE_DIR, "liboracle.so")
        src = [os.path.join(ORACLE_DIR, f) for f in ("oracle_capi.cpp", "fdtd_oracle.hpp")]
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
        _lib = C.CDLL(so)
        _lib.oracle_angle_from_degrees.restype = C.c_double
        _lib.oracle_angle_from_degrees.argtypes = [C.c_int, C.c_double]
        _lib.oracle_source_frequency.restype = C.c_double
        _lib.oracle_source_frequency.argtypes = [C.c_int, C.c_double]
        _lib.oracle_init_scheme.argtypes = [C.POINTER(capi.Config)]
        _lib.oracle_init_sigma.argtypes = [C.POINTER(capi.Config), C.c_int, C.c_void_p]
        _lib.oracle_init_eps_sphere.argtypes = [C.POINTER(capi.Config), C.c_void_p] + [C.c_double] * 5
        _lib.oracle_run.argtypes = [C.POINTER(capi.Config), C.POINTER(capi.HostArrays), C.c_int64, C.c_int]
        _lib.oracle_run_scattered.argtypes = [C.POINTER(capi.Config), C.POINTER(capi.HostArrays), C.c_int64, C.c_int]
    return _lib


def real_dtype(cfg):
    return np.float32 if cfg.value_type == 4 else np.float64


def field_dtype(cfg):
    if cfg.complex_values:
        return np.complex64 if cfg.value_type == 4 else np.complex128
    return real_dtype(cfg)


def make_config(dims=3, size=(48, 48, 48), pml=10, tfsf=20, use_pml=True, use_tfsf=True, value_type=8,
                complex_values=False, dx=0.0005, wavelength=0.02, theta=90.0, phi=0.0, psi=90.0,
                exact_mode=1, source_pos=None):
    """fdtd_config for the given CLI-like parameters, dt / relPhaseVelocity from the oracle's initScheme."""
    L = lib()
    cfg = capi.Config()
    cfg.abi_version = capi.ABI_VERSION
    cfg.dims = dims
    cfg.value_type = value_type
    cfg.complex_values = int(complex_values)
    cfg.use_pml = int(use_pml)
    cfg.use_tfsf = int(use_tfsf)
    cfg.exact_mode = exact_mode
    size = tuple(size) + (1,) * (3 - len(size))
    if dims == 2:
        size = (size[0], size[1], 1)
    for a in range(3):
        cfg.size[a] = size[a]
        cfg.pml_size[a] = pml if (dims == 3 or a < 2) else 0
        cfg.tfsf_size[a] = tfsf if (dims == 3 or a < 2) else 0
    if source_pos is None:
        # Scheme3D.cpp:2011-2013 / SchemeTMz.cpp:1347
        source_pos = (size[0] // 2, size[1] // 2, size[2] // 2) if dims == 3 else (70, size[1] // 2, 0)
    for a in range(3):
        cfg.source_pos[a] = source_pos[a]
    fp = np.float32 if value_type == 4 else np.float64
    cfg.dx = float(fp(dx))
    cfg.source_frequency = L.oracle_source_frequency(value_type, float(fp(wavelength)))
    cfg.angle_teta = L.oracle_angle_from_degrees(value_type, theta)
    cfg.angle_phi = L.oracle_angle_from_degrees(value_type, phi)
    cfg.angle_psi = L.oracle_angle_from_degrees(value_type, psi)
    L.oracle_init_scheme(C.byref(cfg))
    cfg.rank, cfg.nranks = 0, 1
    cfg.time_step_start = 0
    return cfg


def sigma_profile(cfg, axis):
    out = np.zeros(cfg.size[axis] + 1, dtype=real_dtype(cfg))
    lib().oracle_init_sigma(C.byref(cfg), axis, out.ctypes.data)
    return out


def eps_sphere(cfg, center=(40.5, 40.5, 40.5), radius=20.0, eps_in=2.0):
    mz = cfg.size[2] + 1 if cfg.dims == 3 else 1
    out = np.zeros((cfg.size[0] + 1) * (cfg.size[1] + 1) * mz, dtype=real_dtype(cfg))
    lib().oracle_init_eps_sphere(C.byref(cfg), out.ctypes.data, *center, radius, eps_in)
    return out


def n_cells(cfg):
    return cfg.size[0] * cfg.size[1] * (cfg.size[2] if cfg.dims == 3 else 1)


def inc_len(cfg):
    return 100 * (cfg.size[0] + cfg.size[1] + (cfg.size[2] if cfg.dims == 3 else 0))


def field_names(cfg):
    names = ["Ex", "Ey", "Ez", "Hx", "Hy", "Hz"] if cfg.dims == 3 else ["Ez", "Hx", "Hy"]
    if cfg.use_pml:
        names += ["Dx", "Dy", "Dz", "Bx", "By", "Bz"] if cfg.dims == 3 else ["Dz", "Bx", "By"]
    return names


def reference_setup(cfg):
    """Inputs the reference's initGrids would produce: eps sphere (3D) / eps=1 (2D), mu=1, sigma profiles."""
    inputs = {}
    mz = cfg.size[2] + 1 if cfg.dims == 3 else 1
    nm = (cfg.size[0] + 1) * (cfg.size[1] + 1) * mz
    inputs["Eps"] = eps_sphere(cfg) if cfg.dims == 3 else np.ones(nm, dtype=real_dtype(cfg))
    inputs["Mu"] = np.ones(nm, dtype=real_dtype(cfg))
    for a in range(cfg.dims):
        inputs["Sigma" + "XYZ"[a]] = sigma_profile(cfg, a) if cfg.use_pml else np.zeros(cfg.size[a] + 1, real_dtype(cfg))
    return inputs


def run(cfg, nsteps, inputs=None, fields=None, nthreads=0, scattered=False):
    """Runs the oracle.  `inputs`: dict with Eps, Mu, SigmaX/Y/Z; `fields`: dict of initial fields
    (missing ones start at zero).  Returns dict of all fields (+EInc/HInc) after nsteps."""
    inputs = inputs or {}
    fields = dict(fields or {})
    arr = capi.HostArrays()
    out = {}
    n = n_cells(cfg)
    for name in field_names(cfg):
        a = np.ascontiguousarray(fields.get(name, np.zeros(n, field_dtype(cfg))), dtype=field_dtype(cfg)).copy()
        out[name] = a
        arr.field[capi.FIELD_IDS[name]] = a.ctypes.data
    keep = []
    for key, attr in (("Eps", "eps"), ("Mu", "mu")):
        if key in inputs:
            a = np.ascontiguousarray(inputs[key], dtype=real_dtype(cfg))
            keep.append(a)
            setattr(arr, attr, a.ctypes.data)
    for ax in range(3):
        key = "Sigma" + "XYZ"[ax]
        if key in inputs:
            a = np.ascontiguousarray(inputs[key], dtype=real_dtype(cfg))
            keep.append(a)
            arr.sigma[ax] = a.ctypes.data
    if cfg.use_tfsf:
        for key, attr in (("EInc", "einc"), ("HInc", "hinc")):
            a = np.ascontiguousarray(fields.get(key, np.zeros(inc_len(cfg), field_dtype(cfg))), dtype=field_dtype(cfg)).copy()
            out[key] = a
            setattr(arr, attr, a.ctypes.data)
    fn = lib().oracle_run_scattered if scattered else lib().oracle_run
    rc = fn(C.byref(cfg), C.byref(arr), nsteps, nthreads)
    if rc != 0:
        raise RuntimeError("oracle_run failed: %d" % rc)
    return out
