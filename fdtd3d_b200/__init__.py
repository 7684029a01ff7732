"""fdtd3d_b200 -- B200-native Yee-grid leapfrog time step of fdtd3d behind a C ABI.

This Python package is a thin ctypes binding over ``libfdtd3d_b200.so`` (CUDA kernels for sm_100a +
the C ABI of ``include/fdtd3d_b200.h``).  It exists for the tests, the benchmark harness and scripting;
the product is the shared library.  There is no CPU fallback: if the library has not been built the
import fails, and on a machine without a CUDA device ``Solver`` raises ``FdtdError(NO_DEVICE)``.
"""
import ctypes as C
import os

import numpy as np

from . import _capi
from ._capi import (ABI_VERSION, Config, HostArrays, FIELD_IDS, FIELD_NAMES, ORDER_NATIVE, ORDER_REFERENCE,
                    OK, ERROR, INVALID_ARGUMENT, NO_DEVICE, COMM_ERROR)

_HERE = os.path.dirname(os.path.abspath(__file__))
# FDTD3D_B200_LIB selects another build of the same library (kernel tuning experiments, tools/)
LIB_PATH = os.environ.get("FDTD3D_B200_LIB") or os.path.join(_HERE, "libfdtd3d_b200.so")

_lib = None


class FdtdError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("fdtd3d_b200 error %d: %s" % (code, message))
        self.code = code


def lib():
    """The loaded C-ABI library; raises if it has not been built (python -m fdtd3d_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: build it with `python -m fdtd3d_b200.build` "
                              "(there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, (restype, argtypes) in _capi.PROTOTYPES.items():
            fn = getattr(L, name)  # AttributeError if the library lacks a declared symbol
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = L
    return _lib


def _check(rc):
    if rc != OK:
        raise FdtdError(rc, lib().fdtd_last_error().decode(errors="replace"))


def device_info(device=0):
    buf = C.create_string_buffer(512)
    _check(lib().fdtd_device_info(device, buf, len(buf)))
    return buf.value.decode()


def real_dtype(cfg):
    return np.float32 if cfg.value_type == 4 else np.float64


def field_dtype(cfg):
    if cfg.complex_values:
        return np.complex64 if cfg.value_type == 4 else np.complex128
    return real_dtype(cfg)


def slab_range(nz, rank, nranks):
    """z-slab owned by `rank`: core = nz // nranks, the last rank takes the remainder
    (reference rule: Source/Layout/ParallelYeeGridLayout.cpp:24-36)."""
    core = nz // nranks
    z0 = rank * core
    return z0, (nz - z0 if rank == nranks - 1 else core)


def make_config(dims=3, size=(100, 100, 100), pml=10, tfsf=20, use_pml=False, use_tfsf=False, value_type=8,
                complex_values=False, dx=0.0005, wavelength=0.02, teta=90.0, phi=0.0, psi=90.0, exact_mode=0,
                device=0, rank=0, nranks=1):
    """fdtd_config from CLI-style parameters (defaults are the reference's, Source/Settings/Settings.inc):
    dt, relPhaseVelocity, frequency, angles and the point-source cell come from fdtd_init_scheme."""
    cfg = Config()
    cfg.abi_version = ABI_VERSION
    cfg.dims, cfg.value_type, cfg.complex_values = dims, value_type, int(complex_values)
    cfg.use_pml, cfg.use_tfsf, cfg.exact_mode, cfg.device = int(use_pml), int(use_tfsf), int(exact_mode), device
    size = tuple(size)[:dims] + (1,) * (3 - dims)
    pml = (pml,) * 3 if np.isscalar(pml) else tuple(pml)
    tfsf = (tfsf,) * 3 if np.isscalar(tfsf) else tuple(tfsf)
    for a in range(3):
        cfg.size[a] = size[a]
        cfg.pml_size[a] = pml[a] if a < dims else 0
        cfg.tfsf_size[a] = tfsf[a] if a < dims else 0
    cfg.dx = dx
    cfg.rank, cfg.nranks = rank, nranks
    rc = lib().fdtd_init_scheme(C.byref(cfg), wavelength, teta, phi, psi)
    if rc != OK:
        raise FdtdError(rc, "fdtd_init_scheme rejected the configuration")
    return cfg


def pml_sigma_profile(cfg, axis):
    out = np.zeros(cfg.size[axis] + 1, dtype=real_dtype(cfg))
    rc = lib().fdtd_pml_sigma_profile(C.byref(cfg), axis, out.ctypes.data, out.size)
    if rc != OK:
        raise FdtdError(rc, "fdtd_pml_sigma_profile rejected the arguments")
    return out


def plan_core_box(cfg, sigmas=None):
    """Box of cells fast mode advances without D/B for `cfg` and the sigma profiles [sx, sy, sz] (None = zero), or None
    (fdtd_plan_core_box; host only)."""
    keep = [None if sg is None else np.ascontiguousarray(sg, dtype=real_dtype(cfg)) for sg in (sigmas or [None] * 3)]
    keep += [None] * (3 - len(keep))
    ptrs = (C.c_void_p * 3)(*[None if a is None else a.ctypes.data for a in keep])
    lo, hi = (C.c_int64 * 3)(), (C.c_int64 * 3)()
    _check(lib().fdtd_plan_core_box(C.byref(cfg), ptrs, lo, hi))
    if any(lo[a] > hi[a] for a in range(3)):
        return None
    return tuple(lo), tuple(hi)


class Solver:
    """One device-resident Scheme (Scheme3D or SchemeTMz) -- wraps an ``fdtd_handle``."""

    def __init__(self, cfg):
        self.cfg = cfg
        self._h = C.c_void_p()
        _check(lib().fdtd_create(C.byref(cfg), C.byref(self._h)))

    def close(self):
        if self._h:
            lib().fdtd_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- geometry helpers
    @property
    def dims(self):
        return self.cfg.dims

    def field_shape(self, field, nz=None):
        c = self.cfg
        fid = FIELD_IDS[field] if isinstance(field, str) else field
        if fid in (_capi.EINC, _capi.HINC):
            return (100 * (c.size[0] + c.size[1] + (c.size[2] if c.dims == 3 else 0)),)
        extra = 1 if fid in (_capi.EPS, _capi.MU) else 0
        if c.dims == 2:
            return (c.size[0] + extra, c.size[1] + extra)
        return (c.size[0] + extra, c.size[1] + extra, (c.size[2] + extra) if nz is None else nz)

    def _dtype(self, fid):
        return real_dtype(self.cfg) if fid in (_capi.EPS, _capi.MU) else field_dtype(self.cfg)

    # -- inputs
    def set_sigma(self, axis, profile):
        a = np.ascontiguousarray(profile, dtype=real_dtype(self.cfg))
        _check(lib().fdtd_set_sigma(self._h, axis, a.ctypes.data, a.size))

    def upload(self, field, array, order=ORDER_REFERENCE, z0=0, nz=None):
        fid = FIELD_IDS[field] if isinstance(field, str) else field
        a = np.ascontiguousarray(array, dtype=self._dtype(fid))
        if nz is None:
            nz = self.field_shape(fid)[2] if self.cfg.dims == 3 and fid < _capi.EINC else 1
        _check(lib().fdtd_upload(self._h, fid, a.ctypes.data, order, z0, nz))

    def download(self, field, order=ORDER_REFERENCE, z0=0, nz=None):
        fid = FIELD_IDS[field] if isinstance(field, str) else field
        if nz is None:
            nz = self.field_shape(fid)[2] if self.cfg.dims == 3 and fid < _capi.EINC else 1
        shape = self.field_shape(fid, nz)
        out = np.zeros(int(np.prod(shape)), dtype=self._dtype(fid))
        _check(lib().fdtd_download(self._h, fid, out.ctypes.data, order, z0, nz))
        return out

    def fill(self, field, value):
        fid = FIELD_IDS[field] if isinstance(field, str) else field
        v = complex(value)
        _check(lib().fdtd_fill(self._h, fid, v.real, v.imag))

    def init_eps_sphere(self, center=(40.5, 40.5, 40.5), radius=20.0, eps_inside=2.0):
        _check(lib().fdtd_init_eps_sphere(self._h, center[0], center[1], center[2], radius, eps_inside))

    # -- time loop
    def run(self, nsteps):
        _check(lib().fdtd_run(self._h, nsteps))

    def subtract_incident(self):
        """Total -> scattered field inside the TF/SF box, in place (the reference's --save-res post-processing)."""
        _check(lib().fdtd_subtract_incident(self._h))

    def run_async(self, nsteps):
        _check(lib().fdtd_run_async(self._h, nsteps))

    def sync(self):
        _check(lib().fdtd_sync(self._h))

    @property
    def busy(self):
        """True while queued steps are still running (fdtd_busy; does not wait)."""
        b = C.c_int()
        _check(lib().fdtd_busy(self._h, C.byref(b)))
        return bool(b.value)

    @property
    def last_run_ms(self):
        ms = C.c_float()
        _check(lib().fdtd_last_run_ms(self._h, C.byref(ms)))
        return ms.value

    def set_profiling(self, enable=True):
        _check(lib().fdtd_set_profiling(self._h, int(enable)))

    def phase_times(self):
        """(E-phase ms, H-phase ms, steps) summed over the last completed run (needs set_profiling)."""
        e, h, n = C.c_float(), C.c_float(), C.c_int64()
        _check(lib().fdtd_phase_times(self._h, C.byref(e), C.byref(h), C.byref(n)))
        return e.value, h.value, n.value

    def step_breakdown(self):
        """(ms[8], steps) of the last profiled multi-rank run (fdtd_step_breakdown)."""
        ms, n = (C.c_float * 8)(), C.c_int64()
        _check(lib().fdtd_step_breakdown(self._h, ms, C.byref(n)))
        return list(ms), n.value

    @property
    def launch_count(self):
        n = C.c_int64()
        _check(lib().fdtd_launch_count(self._h, C.byref(n)))
        return n.value

    @property
    def time_step(self):
        n = C.c_int64()
        _check(lib().fdtd_get_time_step(self._h, C.byref(n)))
        return n.value

    @property
    def core_box(self):
        """((lo_x, lo_y, lo_z), (hi_x, hi_y, hi_z)) of the cells advanced without D/B, or None (fdtd_get_core_box)."""
        lo, hi = (C.c_int64 * 3)(), (C.c_int64 * 3)()
        _check(lib().fdtd_get_core_box(self._h, lo, hi))
        if any(lo[a] > hi[a] for a in range(3)):
            return None
        return tuple(lo), tuple(hi)

    @property
    def fused_box(self):
        """Box advanced with one pass per step (fdtd_get_fused_box), or None."""
        lo, hi = (C.c_int64 * 3)(), (C.c_int64 * 3)()
        _check(lib().fdtd_get_fused_box(self._h, lo, hi))
        if any(lo[a] > hi[a] for a in range(3)):
            return None
        return tuple(lo), tuple(hi)

    @property
    def fused_ms(self):
        ms = C.c_float()
        _check(lib().fdtd_fused_time(self._h, C.byref(ms)))
        return ms.value

    def set_tuning(self, march_planes=0, one_pass_march_planes=0, allow_one_pass=None):
        """allow_one_pass: True / False, or None = the library's default (complex double only)."""
        _check(lib().fdtd_set_tuning(self._h, march_planes, one_pass_march_planes, -1 if allow_one_pass is None else int(allow_one_pass)))

    def sumsq(self, field):
        fid = FIELD_IDS[field] if isinstance(field, str) else field
        out = C.c_double()
        _check(lib().fdtd_field_sumsq(self._h, fid, C.byref(out)))
        return out.value

    # -- multi-GPU
    @staticmethod
    def comm_unique_id():
        buf = (C.c_char * 128)()
        _check(lib().fdtd_comm_unique_id(buf))
        return bytes(buf)

    def comm_init(self, unique_id):
        buf = (C.c_char * 128).from_buffer_copy(unique_id)
        _check(lib().fdtd_comm_init(self._h, buf))


def execute_steps(cfg, arrays, total_steps):
    """One-shot drop-in for cudaExecute3DSteps / cudaExecute2DTMzSteps: `arrays` maps field names
    (Ex.. Bz, Eps, Mu, SigmaX/Y/Z, EInc, HInc) to numpy arrays in reference order; E/H/D/B/EInc/HInc are
    updated in place."""
    arr = HostArrays()
    keep = []
    for name, a in arrays.items():
        if name in ("Eps", "Mu") or name.startswith("Sigma"):
            a = np.ascontiguousarray(a, dtype=real_dtype(cfg))
        else:
            assert a.flags["C_CONTIGUOUS"] and a.dtype == field_dtype(cfg), name
        keep.append(a)
        if name == "Eps":
            arr.eps = a.ctypes.data
        elif name == "Mu":
            arr.mu = a.ctypes.data
        elif name.startswith("Sigma"):
            arr.sigma["XYZ".index(name[-1])] = a.ctypes.data
        elif name == "EInc":
            arr.einc = a.ctypes.data
        elif name == "HInc":
            arr.hinc = a.ctypes.data
        else:
            arr.field[FIELD_IDS[name]] = a.ctypes.data
    status = C.c_int(ERROR)
    rc = lib().fdtd_execute_steps(C.byref(status), C.byref(cfg), C.byref(arr), total_steps)
    _check(rc)
    return status.value
