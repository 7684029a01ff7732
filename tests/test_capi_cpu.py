"""CPU-side checks of the C ABI: the library loads, exports every symbol include/fdtd3d_b200.h declares,
the ctypes struct layouts match the C ones, and -- with no GPU -- it fails loudly instead of falling back."""
import ctypes as C
import os
import re
import subprocess
import tempfile

import pytest

import fdtd3d_b200 as F
from fdtd3d_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "fdtd3d_b200.h")


def _declared_functions():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(fdtd_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    names = _declared_functions()
    assert len(names) >= 24
    L = F.lib()
    for n in names:
        assert hasattr(L, n), n
    assert sorted(_capi.PROTOTYPES) == names


def test_struct_layouts_match_the_header():
    src = r'''
#include <stdio.h>
#include <stddef.h>
#include "fdtd3d_b200.h"
int main (void) {
  printf ("%zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof (fdtd_config), offsetof (fdtd_config, size), offsetof (fdtd_config, dx),
          offsetof (fdtd_config, rank), offsetof (fdtd_config, time_step_start), sizeof (fdtd_host_arrays),
          offsetof (fdtd_host_arrays, eps), offsetof (fdtd_host_arrays, einc));
  return 0; }
'''
    with tempfile.TemporaryDirectory() as tmp:
        c = os.path.join(tmp, "t.c")
        open(c, "w").write(src)
        exe = os.path.join(tmp, "t")
        subprocess.check_call(["/usr/bin/gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        got = [int(x) for x in subprocess.check_output([exe]).split()]
    want = [C.sizeof(_capi.Config), _capi.Config.size.offset, _capi.Config.dx.offset, _capi.Config.rank.offset,
            _capi.Config.time_step_start.offset, C.sizeof(_capi.HostArrays), _capi.HostArrays.eps.offset,
            _capi.HostArrays.einc.offset]
    assert got == want


def test_header_is_plain_c():
    with tempfile.TemporaryDirectory() as tmp:
        c = os.path.join(tmp, "t.c")
        open(c, "w").write('#include "fdtd3d_b200.h"\nint main (void) { return FDTD_FIELD_COUNT == 16 ? 0 : 1; }\n')
        subprocess.check_call(["/usr/bin/gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), c,
                               "-o", os.path.join(tmp, "t")])


def test_argument_validation_needs_no_device():
    import oracle_lib as O
    cfg = O.make_config(size=(16, 16, 16), pml=3, tfsf=4)
    cfg.abi_version = 99
    h = C.c_void_p()
    assert F.lib().fdtd_create(C.byref(cfg), C.byref(h)) == F.INVALID_ARGUMENT
    assert b"abi_version" in F.lib().fdtd_last_error()
    cfg = O.make_config(size=(16, 16, 16), pml=3, tfsf=4)
    cfg.value_type = 16  # long double: no device type
    assert F.lib().fdtd_create(C.byref(cfg), C.byref(h)) == F.INVALID_ARGUMENT


def test_no_cpu_fallback():
    """Without a CUDA device the product refuses to run (the oracle is never used as a fallback)."""
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = os.path.exists("/dev/nvidia0")
    if has_gpu:
        pytest.skip("a GPU is present")
    import oracle_lib as O
    cfg = O.make_config(size=(16, 16, 16), pml=3, tfsf=4)
    with pytest.raises(F.FdtdError) as e:
        F.Solver(cfg)
    assert e.value.code == F.NO_DEVICE
    assert "no CPU fallback" in str(e.value)


def test_product_never_references_the_oracle():
    """Nothing under fdtd3d_b200/ or include/ may include, import or link anything under oracle/."""
    bad = []
    for base in ("fdtd3d_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                    txt = open(os.path.join(dp, f), errors="replace").read()
                    if re.search(r"oracle[/_.]|liboracle|import oracle|from oracle", txt):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_slab_partition_rule():
    """core = Nz / n, the last rank takes the remainder (ParallelYeeGridLayout.cpp:24-36, intended rule)."""
    for nz, n in ((1024, 8), (512, 4), (100, 3), (7, 2)):
        parts = [F.slab_range(nz, r, n) for r in range(n)]
        assert parts[0][0] == 0 and sum(p[1] for p in parts) == nz
        for r in range(1, n):
            assert parts[r][0] == parts[r - 1][0] + parts[r - 1][1]
        assert all(p[1] == nz // n for p in parts[:-1])


def test_host_scheme_init_matches_reference_dumps():
    """fdtd_init_scheme / fdtd_pml_sigma_profile (product host code) reproduce what the unmodified reference
    derived (dt, relPhaseVelocity, frequency, angles, sigma profiles) bit for bit, and agree with the oracle."""
    import numpy as np
    import golden_cases as G
    import oracle_lib as O
    for name in G.case_names():
        meta, data = G.load(name)
        dims, vt, cplx = G.VARIANT[meta["variant"]]
        cfg = F.make_config(dims=dims, size=meta["size"][:dims], pml=meta["pml"], tfsf=meta["tfsf"],
                            use_pml=bool(meta["use_pml"]), use_tfsf=bool(meta["use_tfsf"]), value_type=vt,
                            complex_values=cplx, dx=meta["dx"], wavelength=meta["wavelength"], teta=meta["theta"],
                            phi=meta["phi"], psi=meta["psi"])
        assert cfg.dt == meta["dt"], name
        assert cfg.source_frequency == meta["source_frequency"], name
        assert (cfg.angle_teta, cfg.angle_phi, cfg.angle_psi) == (meta["angle_teta"], meta["angle_phi"], meta["angle_psi"]), name
        if dims == 3:
            assert cfg.rel_phase_velocity == meta["rel_phase_velocity"], name
        ocfg = G.config_from_meta(meta)
        for f in ("dx", "dt", "source_frequency", "rel_phase_velocity", "angle_teta", "angle_phi", "angle_psi"):
            assert getattr(cfg, f) == getattr(ocfg, f), (name, f)
        assert list(cfg.source_pos) == list(ocfg.source_pos), name
        if meta["use_pml"]:
            for a in range(dims):
                assert np.array_equal(F.pml_sigma_profile(cfg, a), data["Sigma" + "XYZ"[a]]), (name, a)


def _cfg(**kw):
    import oracle_lib as O
    return O.make_config(**kw)


def test_core_box_planning_is_host_logic():
    """fdtd_plan_core_box (no GPU): the cells fast mode advances without D/B = per axis the longest zero run of the
    sigma profile (both material samples i and i+1 of a cell inside it), shrunk in x to whole quarter tile rows and in y
    to whole tile rows of the kernels."""
    import numpy as np
    import oracle_lib as O

    def box(cfg, sig="profile"):
        if sig == "profile":
            sig = [O.sigma_profile(cfg, a) for a in range(cfg.dims)]
        return F.plan_core_box(cfg, sig)

    # 512^3 double, PML 10: sigma is zero for material cells 10..502 -> field cells 10..501; x to 16-cell quarters, y to 4 rows
    assert box(_cfg(size=(512, 512, 512), pml=10, exact_mode=0)) == ((16, 12, 10), (495, 499, 501))
    # complex double: one cell per 16 bytes -> 8-cell quarters
    assert box(_cfg(size=(512, 512, 512), pml=10, exact_mode=0, complex_values=True)) == ((16, 12, 10), (495, 499, 501))
    # real float: 32-cell quarters, 16-row tiles
    assert box(_cfg(size=(1024, 1024, 512), pml=10, exact_mode=0, value_type=4)) == ((32, 16, 10), (991, 1007, 501))
    # 2D TMz float: per row, and in x per quarter of a tile row (8 vectors = 32 cells); z is the single plane
    assert box(_cfg(dims=2, size=(4096, 4096), pml=10, exact_mode=0, value_type=4)) == ((32, 10, 0), (4063, 4085, 0))
    # exact mode and the D/B-everywhere mode never skip anything; nor does a run without UPML
    assert box(_cfg(size=(64, 64, 64), pml=6, tfsf=9, exact_mode=1)) is None
    assert box(_cfg(size=(64, 64, 64), pml=6, tfsf=9, exact_mode=2)) is None
    assert box(_cfg(size=(64, 64, 64), pml=6, tfsf=9, exact_mode=0, use_pml=False)) is None
    # too small for one whole quarter row inside the zero run
    assert box(_cfg(size=(24, 64, 64), pml=6, tfsf=9, exact_mode=0)) is None
    # all-zero sigma: the whole grid, tile-aligned
    cfg = _cfg(size=(70, 30, 20), pml=3, tfsf=5, exact_mode=0)
    assert box(cfg, None) == ((0, 0, 0), (63, 27, 19))
    # an absorbing sheet in the middle of x: the longer zero run wins
    sig = [np.zeros(71), np.zeros(31), np.zeros(21)]
    sig[0][20] = 1.0
    assert box(cfg, sig) == ((32, 0, 0), (63, 27, 19))
