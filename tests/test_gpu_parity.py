"""GPU parity tests proper: the CUDA path, called through the C ABI, against

  (1) dumps of the unmodified reference solver (tests/golden/), and
  (2) the pinned Tier-2 oracle on seeded random inputs, ragged sizes and edge cases.

Bar (BASELINE.json north_star): index/halo bookkeeping bit-exact; field values within max relative
error 1e-12 (double) / 1e-5 (float).  exact mode (reference operand order, IEEE division, no FMA
contraction: -fmad=false) is held to BIT-EXACT equality with the reference in double; fast mode
(reciprocal tables + explicit FMA) to the stated tolerances.
"""
import numpy as np
import pytest

import fdtd3d_b200 as F
import golden_cases as G
import oracle_lib as O
from gpu_util import family_rel_errs, field_rel_errs, random_inputs, rel_err, run_gpu, significant

pytestmark = pytest.mark.gpu

TOL = {8: 1e-12, 4: 1e-5}


def _keys(cfg):
    return O.field_names(cfg) + (["EInc", "HInc"] if cfg.use_tfsf else [])


PLAIN_CASES = [n for n in G.case_names() if not n.endswith("_saveres")]


@pytest.mark.parametrize("name", PLAIN_CASES)
def test_exact_mode_matches_reference_dump(name):
    meta, data = G.load(name)
    cfg = G.config_from_meta(meta, exact_mode=1)
    out = run_gpu(cfg, meta["steps"], G.inputs_from_golden(cfg, data))
    assert out["_launches"] > 0
    errs = family_rel_errs(out, data, _keys(cfg))
    for key in _keys(cfg):
        # double, and real float (whose double promotions the kernels reproduce): bit for bit
        assert np.array_equal(out[key], data[key]), "%s not bit-exact: rel err %g" % (key, errs[key])


@pytest.mark.parametrize("name", PLAIN_CASES)
def test_fast_mode_within_tolerance_of_reference_dump(name):
    meta, data = G.load(name)
    cfg = G.config_from_meta(meta, exact_mode=0)
    out = run_gpu(cfg, meta["steps"], G.inputs_from_golden(cfg, data))
    errs = family_rel_errs(out, data, _keys(cfg))
    for key in _keys(cfg):
        assert errs[key] <= TOL[cfg.value_type], (key, errs)
    # per field, against the field's own maximum, for every component that carries signal
    own = field_rel_errs(out, data, _keys(cfg))
    for key in significant(data, _keys(cfg), value_type=cfg.value_type):
        assert own[key] <= TOL[cfg.value_type], (key, own)


RANDOM_CASES = [
    # dims, size, pml, tfsf, use_pml, use_tfsf, value_type, complex, steps, angles
    (3, (37, 21, 19), 4, 6, True, True, 8, False, 12, (90, 0, 90)),
    (3, (18, 35, 23), 4, 6, True, True, 8, False, 12, (50, 70, 20)),
    (3, (33, 17, 40), 4, 6, True, False, 8, False, 10, (90, 0, 90)),
    (3, (20, 20, 20), 3, 5, False, True, 8, False, 10, (90, 90, 0)),
    (3, (16, 19, 17), 3, 5, False, False, 8, True, 10, (90, 0, 90)),
    (3, (17, 18, 19), 3, 5, True, True, 8, True, 10, (30, 60, 45)),
    (3, (35, 22, 18), 4, 6, True, True, 4, False, 12, (90, 0, 90)),
    (3, (21, 22, 23), 4, 6, True, True, 4, True, 10, (90, 0, 90)),
    (3, (5, 4, 3), 1, 1, True, False, 8, False, 6, (90, 0, 90)),
    (2, (75, 40), 5, 8, True, False, 8, False, 20, (90, 0, 90)),
    (2, (131, 77), 6, 9, True, True, 8, False, 20, (90, 0, 90)),
    (2, (90, 33), 5, 8, False, True, 8, True, 20, (90, 0, 90)),
    (2, (200, 45), 5, 8, True, True, 4, False, 20, (90, 0, 90)),
    (2, (100, 50), 5, 8, True, False, 4, True, 20, (90, 0, 90)),
]


@pytest.mark.parametrize("case", RANDOM_CASES, ids=lambda c: "%dd-%s-p%dt%d-vt%d%s" % (c[0], "x".join(map(str, c[1])), c[4], c[5], c[6], "c" if c[7] else ""))
@pytest.mark.parametrize("exact", [1, 0])
def test_random_inputs_against_oracle(case, exact):
    dims, size, pml, tfsf, use_pml, use_tfsf, vt, cplx, steps, ang = case
    src = (size[0] // 2, size[1] // 2, size[2] // 2) if dims == 3 else (min(70, size[0] - 2), size[1] // 2, 0)
    cfg = O.make_config(dims=dims, size=size, pml=pml, tfsf=tfsf, use_pml=use_pml, use_tfsf=use_tfsf, value_type=vt,
                        complex_values=cplx, theta=ang[0], phi=ang[1], psi=ang[2], exact_mode=exact, source_pos=src)
    inputs, fields = random_inputs(cfg, seed=1234 + steps)
    ref = O.run(cfg, steps, inputs, fields)
    out = run_gpu(cfg, steps, inputs, fields)
    errs = family_rel_errs(out, ref, _keys(cfg))
    for key in _keys(cfg):
        if exact and (vt == 8 or not cplx):
            assert np.array_equal(out[key], ref[key]), "%s not bit-exact: rel err %g" % (key, errs[key])
        else:
            assert errs[key] <= TOL[vt], (key, errs)


def test_chunked_runs_equal_one_run():
    """performNSteps (startStep, n): running 7 + 1 + 12 steps is the same as 20 steps (time index carries over)."""
    cfg = O.make_config(size=(24, 20, 22), pml=4, tfsf=6, use_pml=True, use_tfsf=False, exact_mode=1)
    inputs, fields = random_inputs(cfg, seed=7)
    a = run_gpu(cfg, 20, inputs, fields)
    b = run_gpu(cfg, 20, inputs, fields, chunks=[7, 1, 12])
    for key in _keys(cfg):
        assert np.array_equal(a[key], b[key]), key


def test_compute_ranges_leave_border_cells_untouched():
    """Cells outside the compute boxes of YeeGridLayout.h:131-182 keep their initial value bit for bit."""
    cfg = O.make_config(size=(12, 11, 10), pml=2, tfsf=3, use_pml=True, use_tfsf=False, exact_mode=1)
    inputs, fields = random_inputs(cfg, seed=3)
    out = run_gpu(cfg, 3, inputs, fields)
    N = (12, 11, 10)
    lo = {"Ex": (0, 1, 1), "Ey": (1, 0, 1), "Ez": (1, 1, 0), "Hx": (1, 0, 0), "Hy": (0, 1, 0), "Hz": (0, 0, 1)}
    hi = {"Ex": (1, 0, 0), "Ey": (0, 1, 0), "Ez": (0, 0, 1), "Hx": (0, 1, 1), "Hy": (1, 0, 1), "Hz": (1, 1, 0)}
    for name in lo:
        for aux in (name, {"E": "D", "H": "B"}[name[0]] + name[1]):
            a, b = out[aux].reshape(N), fields[aux].reshape(N)
            mask = np.ones(N, bool)
            sl = tuple(slice(lo[name][d], N[d] - hi[name][d]) for d in range(3))
            mask[sl] = False
            if name == "Ez" and aux == "Ez":
                mask[6, 5, 5] = False  # the point source cell
            assert np.array_equal(a[mask], b[mask]), aux
            assert not np.array_equal(a[sl], b[sl]), aux


@pytest.mark.parametrize("vt,cplx", [(8, False), (8, True), (4, False), (4, True)])
def test_upload_download_round_trip(vt, cplx):
    cfg = O.make_config(size=(19, 13, 11), pml=2, tfsf=3, use_pml=True, use_tfsf=True, value_type=vt, complex_values=cplx)
    rng = np.random.default_rng(5)
    with F.Solver(cfg) as s:
        for name in ["Ex", "Hz", "Dy", "Bz", "Eps", "Mu", "EInc"]:
            shape = s.field_shape(name)
            fid = F.FIELD_IDS[name]
            dt = O.real_dtype(cfg) if name in ("Eps", "Mu") else O.field_dtype(cfg)
            a = rng.standard_normal(int(np.prod(shape))).astype(dt)
            s.upload(name, a)
            assert np.array_equal(s.download(name), a), name
            if len(shape) == 3:
                # native order is the transpose of reference order
                nat = s.download(name, order=F.ORDER_NATIVE)
                assert np.array_equal(nat.reshape(shape[::-1]), a.reshape(shape).transpose(2, 1, 0)), name
                # partial plane range, both directions
                part = s.download(name, z0=3, nz=4)
                assert np.array_equal(part.reshape(shape[0], shape[1], 4), a.reshape(shape)[:, :, 3:7]), name
                b = rng.standard_normal(shape[0] * shape[1] * 2).astype(dt)
                s.upload(name, b, z0=5, nz=2)
                full = a.reshape(shape).copy()
                full[:, :, 5:7] = b.reshape(shape[0], shape[1], 2)
                assert np.array_equal(s.download(name).reshape(shape), full), name
                s.upload(name, np.ascontiguousarray(full.transpose(2, 1, 0)), order=F.ORDER_NATIVE)
                assert np.array_equal(s.download(name).reshape(shape), full), name
                sq = float((np.abs(full.astype(np.complex128 if cplx else np.float64)) ** 2).sum())
                assert abs(s.sumsq(name) - sq) <= 1e-12 * sq, name


def test_device_eps_sphere_matches_reference_init():
    """fdtd_init_eps_sphere == Approximation::approximateSphere as run by Scheme3D::initGrids, bit for bit."""
    for vt in (8, 4):
        cfg = O.make_config(size=(47, 52, 45), value_type=vt)
        with F.Solver(cfg) as s:
            s.init_eps_sphere()
            assert np.array_equal(s.download("Eps"), O.eps_sphere(cfg))


def test_one_shot_execute_steps():
    """fdtd_execute_steps, the drop-in for cudaExecute3DSteps / cudaExecute2DTMzSteps."""
    for dims, size in ((3, (22, 20, 18)), (2, (80, 30))):
        cfg = O.make_config(dims=dims, size=size, pml=4, tfsf=6, use_pml=True, use_tfsf=True, exact_mode=1)
        inputs, fields = random_inputs(cfg, seed=11)
        ref = O.run(cfg, 9, inputs, fields)
        arrays = {k: v.copy() for k, v in fields.items()}
        arrays.update(inputs)
        n_inc = O.inc_len(cfg)
        arrays["EInc"] = np.zeros(n_inc, O.field_dtype(cfg))
        arrays["HInc"] = np.zeros(n_inc, O.field_dtype(cfg))
        assert F.execute_steps(cfg, arrays, 9) == F.OK
        for key in _keys(cfg):
            assert np.array_equal(arrays[key], ref[key]), key


def test_errors_are_reported_not_fatal():
    cfg = O.make_config(size=(16, 16, 16), pml=3, tfsf=4)
    with F.Solver(cfg) as s:
        with pytest.raises(F.FdtdError) as e:
            s.set_sigma(0, np.zeros(5))
        assert e.value.code == F.INVALID_ARGUMENT
        with pytest.raises(F.FdtdError):
            s.download("Ex", z0=10, nz=20)
    bad = O.make_config(size=(16, 16, 16), pml=3, tfsf=4)
    bad.tfsf_size[0] = 8
    with pytest.raises(F.FdtdError):
        F.Solver(bad)


def test_full_size_512_cubed_against_oracle():
    """BASELINE.json's headline size (3D 512^3 double, PML + TF/SF), every cell exercised: random fields, random
    materials, 3 steps.  Exact mode: each of the twelve fields is compared with the CPU oracle bit for bit through
    a checksum of per-plane checksums plus full equality of sampled planes, and through |field|^2.  Then the two fast
    modes -- exact_mode 0 is what bench.py times: sigma-free core, 4-row tiles x 3 CTAs per SM, the march length of
    the full-size grid -- on the same inputs: all twelve fields, every cell, within 1e-12 of the same oracle result."""
    n, steps = 512, 3
    cfg = O.make_config(size=(n, n, n), pml=10, tfsf=20, use_pml=True, use_tfsf=True, exact_mode=1)
    rng = np.random.default_rng(2026)
    nm = (n + 1) ** 3
    inputs = {"Eps": rng.uniform(1.0, 4.0, nm), "Mu": rng.uniform(1.0, 2.0, nm)}
    for a in range(3):
        inputs["Sigma" + "XYZ"[a]] = O.sigma_profile(cfg, a)
    scale = {"E": 1.0, "H": 2.6e-3, "D": 8.85e-12, "B": 3.3e-9}
    names = O.field_names(cfg)
    fields = {}
    for name in names:
        fields[name] = rng.standard_normal(n ** 3, dtype=np.float32).astype(np.float64) * scale[name[0]]

    def gpu_run(mode):
        c = O.make_config(size=(n, n, n), pml=10, tfsf=20, use_pml=True, use_tfsf=True, exact_mode=mode)
        s = F.Solver(c)
        for a in range(3):
            s.set_sigma(a, inputs["Sigma" + "XYZ"[a]])
        s.upload("Eps", inputs["Eps"])
        s.upload("Mu", inputs["Mu"])
        for name in names:
            s.upload(name, fields[name])
        s.run(steps)
        return s

    s = gpu_run(1)
    ref = O.run(cfg, steps, inputs, fields)  # ~10 s per step on 16 host cores
    try:
        for name in names + ["EInc", "HInc"]:
            got = s.download(name)
            want = ref[name]
            if name in names:
                # checksum of per-(i)-plane checksums over the raw bits, then sampled planes in full
                gb = got.view(np.uint64).reshape(n, -1)
                wb = want.view(np.uint64).reshape(n, -1)
                gs = np.bitwise_xor.reduce(gb, axis=1)
                ws = np.bitwise_xor.reduce(wb, axis=1)
                assert np.array_equal(gs, ws), "%s: plane checksums differ at i = %s" % (name, np.nonzero(gs != ws)[0][:8])
                for i in (0, 1, 9, 10, 19, 20, 255, 256, 491, 492, 501, 502, 510, 511):
                    assert np.array_equal(gb[i], wb[i]), (name, i)
                sq = float((want ** 2).sum())
                assert abs(s.sumsq(name) - sq) <= 1e-12 * sq, name
            else:
                assert np.array_equal(got, want), name
            del got
    finally:
        s.close()

    fam = {}
    for name in names:
        fam[name[0]] = max(fam.get(name[0], 0.0), float(np.abs(ref[name]).max()))
    for mode in (0, 2):
        s = gpu_run(mode)
        try:
            if mode == 0:
                box = s.core_box
                assert box is not None and (box[1][0] - box[0][0] + 1) * (box[1][1] - box[0][1] + 1) * (box[1][2] - box[0][2] + 1) > 0.8 * n ** 3
            worst = {}
            for name in names:
                got = s.download(name)
                got -= ref[name]
                worst[name] = float(np.abs(got).max()) / fam[name[0]]
                del got
            assert max(worst.values()) <= TOL[8], (mode, worst)
            for name in ("EInc", "HInc"):
                assert rel_err(s.download(name), ref[name]) <= TOL[8], (mode, name)
        finally:
            s.close()


def test_save_res_scattered_field_matches_reference_dump():
    """fdtd_subtract_incident == the block the reference runs before saving with --save-res (Scheme3D.cpp:2767-2921)."""
    meta, data = G.load("3d_d_pml_tfsf_saveres")
    for exact in (1, 0):
        cfg = G.config_from_meta(meta, exact_mode=exact)
        inputs = G.inputs_from_golden(cfg, data)
        with F.Solver(cfg) as s:
            for a in range(3):
                s.set_sigma(a, inputs["Sigma" + "XYZ"[a]])
            s.upload("Eps", inputs["Eps"])
            s.run(meta["steps"])
            s.subtract_incident()
            out = {k: s.download(k) for k in _keys(cfg)}
        errs = family_rel_errs(out, data, _keys(cfg))
        for key in _keys(cfg):
            if exact:
                assert np.array_equal(out[key], data[key]), (key, errs[key])
            else:
                assert errs[key] <= TOL[8], (key, errs)


@pytest.mark.parametrize("vt,cplx", [(8, False), (8, True), (4, False)])
def test_thousand_steps_stay_within_the_north_star_tolerance(vt, cplx):
    """BASELINE.json quotes 1000-step runs.  After 1000 steps:
      - exact mode is still bit-for-bit equal to the reference arithmetic, in double AND in float;
      - fast mode (FMA + reciprocal tables) in double is within 1e-12;
      - fast mode in float: E and H within 1e-5 of the float reference.  The auxiliary D/B fields are running sums
        over time, and there the *reference's own float build* is 3e-5 .. 5e-5 away from exact (double) arithmetic
        after 1000 steps (measured: tools/float_drift.py), so no float implementation with a different rounding
        sequence can stay within 1e-5 of it; the bar for those is: no further from the double result than the float
        reference itself is."""
    n = 56 if cplx else 64
    cfg_args = dict(size=(n, n, n), pml=8, tfsf=14, use_pml=True, use_tfsf=True, complex_values=cplx, theta=65, phi=25, psi=40)
    cfg = O.make_config(exact_mode=0, value_type=vt, **cfg_args)
    inputs = O.reference_setup(cfg)
    ref = O.run(cfg, 1000, inputs)
    assert np.abs(ref["Ez"]).max() > 0.1  # the wave has filled the total-field region
    out = run_gpu(cfg, 1000, inputs)
    errs = family_rel_errs(out, ref, _keys(cfg))
    if vt == 8:
        assert max(errs.values()) <= TOL[8], errs
    else:
        physical = [k for k in _keys(cfg) if k[0] in "EH"]
        assert max(errs[k] for k in physical) <= TOL[4], errs
        cfg_d = O.make_config(exact_mode=0, value_type=8, **cfg_args)
        truth = {k: v.astype(np.float32) for k, v in O.run(cfg_d, 1000, O.reference_setup(cfg_d)).items()}
        ours = family_rel_errs(out, truth, _keys(cfg))
        theirs = family_rel_errs(ref, truth, _keys(cfg))
        for k in _keys(cfg):
            if k[0] in "DB":
                assert ours[k] <= 1.25 * theirs[k] + 1e-6, (k, ours[k], theirs[k])
    cfg_x = O.make_config(exact_mode=1, value_type=vt, **cfg_args)
    out_x = run_gpu(cfg_x, 1000, inputs)
    for key in _keys(cfg):
        assert np.array_equal(out_x[key], ref[key]), key


# ----------------------------------------------------------------------------- sigma-free core (fast mode, 3D, UPML)

CORE_CASES = [
    # size, pml, tfsf, use_tfsf, value_type, complex, steps
    ((40, 24, 20), 3, 5, False, 8, False, 14),   # point source inside the core: its Dz is advanced explicitly
    ((70, 26, 18), 3, 6, True, 8, False, 12),    # two tiles along x, TF/SF faces inside the core
    ((40, 24, 20), 3, 5, True, 8, True, 10),
    ((140, 40, 18), 4, 6, False, 4, False, 12),  # float: 128-cell tiles, 32-cell quarters, 16-row tiles
    ((70, 26, 18), 3, 6, True, 4, True, 10),
    ((90, 40), 5, 8, False, 8, False, 20),       # 2D TMz: per-vector decision, source cell in the core
    ((131, 47), 6, 9, True, 4, False, 20),
    ((64, 50), 5, 8, True, 8, True, 16),
]


@pytest.mark.parametrize("case", CORE_CASES, ids=lambda c: "%s-t%d-vt%d%s" % ("x".join(map(str, c[0])), c[3], c[4], "c" if c[5] else ""))
def test_sigma_free_core_matches_oracle_and_full_formulation(case):
    """exact_mode 0 advances the cells with sigma = 0 without D/B and reconstructs D/B on demand: every field,
    D and B included, is within the north-star tolerance of the oracle, and of exact_mode 2 (same arithmetic, D/B
    carried through every cell), also when D/B are read and E is replaced between two runs."""
    size, pml, tfsf, use_tfsf, vt, cplx, steps = case
    dims = len(size)
    src = (size[0] // 2, size[1] // 2, size[2] // 2 if dims == 3 else 0)
    ang = dict(theta=60, phi=30, psi=20) if dims == 3 else dict(theta=90, phi=0, psi=90)  # TMz allows phi = 0 or 45 only
    kw = dict(dims=dims, size=size, pml=pml, tfsf=tfsf, use_pml=True, use_tfsf=use_tfsf, value_type=vt, complex_values=cplx,
              source_pos=src, **ang)
    cfg = O.make_config(exact_mode=0, **kw)
    inputs, fields = random_inputs(cfg, seed=99)
    keys = _keys(cfg)
    with F.Solver(cfg) as s:
        for a in range(dims):
            s.set_sigma(a, inputs["Sigma" + "XYZ"[a]])
        box = s.core_box
    assert box is not None, "this grid is meant to have a non-empty core"
    lo, hi = box
    assert all(lo[a] >= pml and hi[a] <= size[a] - pml - 1 for a in range(dims)), box
    if not use_tfsf:
        assert all(lo[a] <= src[a] <= hi[a] for a in range(dims)), (box, src)

    ref = O.run(cfg, steps, inputs, fields)
    out = run_gpu(cfg, steps, inputs, fields)
    full = run_gpu(O.make_config(exact_mode=2, **kw), steps, inputs, fields)
    for got in (out, full):
        errs = family_rel_errs(got, ref, keys)
        assert max(errs.values()) <= TOL[vt], errs

    # read D/B (unfold), replace a field (unfold + refold) between two runs
    half = steps // 2
    mid = O.run(cfg, half, inputs, fields)
    probe = ("Dx", "Bz", "Ez") if dims == 3 else ("Dz", "By", "Ez")
    dsum, eup = ("Dy", "Ey") if dims == 3 else ("Bx", "Hx")
    with F.Solver(cfg) as s:
        for a in range(dims):
            s.set_sigma(a, inputs["Sigma" + "XYZ"[a]])
        s.upload("Eps", inputs["Eps"])
        s.upload("Mu", inputs["Mu"])
        for name in O.field_names(cfg):
            s.upload(name, fields[name])
        s.run(half)
        got_mid = {k: s.download(k) for k in probe}
        want = float((np.abs(mid[dsum].astype(np.complex128)) ** 2).sum())
        assert abs(s.sumsq(dsum) - want) <= 1e-4 * want
        s.upload(eup, mid[eup])  # same values the oracle has at this point, through the upload path
        s.run(steps - half)
        again = {k: s.download(k) for k in O.field_names(cfg)}
    errs = family_rel_errs(got_mid, mid, list(probe))
    assert max(errs.values()) <= TOL[vt], errs
    errs = family_rel_errs(again, ref, O.field_names(cfg))
    assert max(errs.values()) <= TOL[vt], errs


def test_sigma_free_core_with_all_zero_sigma():
    """use_pml with sigma = 0 everywhere: the core is the whole (tile-aligned) grid, border cells included."""
    cfg = O.make_config(size=(36, 19, 12), pml=3, tfsf=5, use_pml=True, use_tfsf=True, exact_mode=0)
    inputs, fields = random_inputs(cfg, seed=5)
    for a in range(3):
        inputs["Sigma" + "XYZ"[a]] = np.zeros(cfg.size[a] + 1)
    ref = O.run(cfg, 8, inputs, fields)
    out = run_gpu(cfg, 8, inputs, fields)
    errs = family_rel_errs(out, ref, _keys(cfg))
    assert max(errs.values()) <= TOL[8], errs


# ----------------------------------------------------------------------------- homogeneous mu (fdtd_fill): H kernels without the mu box

@pytest.mark.parametrize("vt,cplx,use_tfsf", [(8, False, True), (8, True, True), (4, False, False), (4, True, True)])
@pytest.mark.parametrize("mode", [1, 0, 2])
def test_homogeneous_mu_set_with_fill(vt, cplx, use_tfsf, mode):
    """A mu array set with fdtd_fill is known to hold one value: the H kernels then skip the material box.  Same
    results as the oracle fed a full array of that value -- bit for bit in exact mode (the 4-point average of equal
    values is exact) -- and the general path returns once mu is uploaded."""
    size = (140, 40, 22) if (vt == 4 and not cplx) else (70, 26, 18)
    cfg = O.make_config(size=size, pml=4, tfsf=6, use_pml=True, use_tfsf=use_tfsf, value_type=vt, complex_values=cplx,
                        theta=60, phi=30, psi=20, exact_mode=mode, source_pos=(size[0] // 2, size[1] // 2, size[2] // 2))
    inputs, fields = random_inputs(cfg, seed=21)
    mu = 1.5
    inputs_ref = dict(inputs, Mu=np.full_like(inputs["Mu"], mu))
    ref = O.run(cfg, 9, inputs_ref, fields)
    out = run_gpu(cfg, 9, {k: v for k, v in inputs.items() if k != "Mu"}, fields, fills={"Mu": mu})
    errs = family_rel_errs(out, ref, _keys(cfg))
    for key in _keys(cfg):
        if mode == 1 and (vt == 8 or not cplx):
            assert np.array_equal(out[key], ref[key]), "%s not bit-exact: rel err %g" % (key, errs[key])
        else:
            assert errs[key] <= TOL[vt], (key, errs)
    # fill, then upload a non-uniform mu: the general kernels must be back
    ref2 = O.run(cfg, 5, inputs, fields)
    out2 = run_gpu(cfg, 5, inputs, fields, fills={"Mu": mu})
    errs2 = family_rel_errs(out2, ref2, _keys(cfg))
    assert max(errs2.values()) <= (0 if mode == 1 and (vt == 8 or not cplx) else TOL[vt]), errs2


@pytest.mark.parametrize("vt,cplx,use_tfsf", [(8, False, True), (4, False, False), (8, True, False)])
@pytest.mark.parametrize("mode", [1, 0, 2])
def test_homogeneous_materials_tmz(vt, cplx, use_tfsf, mode):
    """2D TMz: eps and mu set with fdtd_fill (the reference's own TMz set-up is eps = mu = 1): neither array is loaded."""
    size = (131, 47)
    cfg = O.make_config(dims=2, size=size, pml=6, tfsf=9, use_pml=True, use_tfsf=use_tfsf, value_type=vt, complex_values=cplx,
                        exact_mode=mode, source_pos=(size[0] // 2, size[1] // 2, 0))
    inputs, fields = random_inputs(cfg, seed=22)
    eps, mu = 2.25, 1.5
    inputs_ref = dict(inputs, Eps=np.full_like(inputs["Eps"], eps), Mu=np.full_like(inputs["Mu"], mu))
    ref = O.run(cfg, 15, inputs_ref, fields)
    out = run_gpu(cfg, 15, {k: v for k, v in inputs.items() if k not in ("Eps", "Mu")}, fields, fills={"Eps": eps, "Mu": mu})
    errs = family_rel_errs(out, ref, _keys(cfg))
    for key in _keys(cfg):
        if mode == 1 and (vt == 8 or not cplx):
            assert np.array_equal(out[key], ref[key]), "%s not bit-exact: rel err %g" % (key, errs[key])
        else:
            assert errs[key] <= TOL[vt], (key, errs)


# ----------------------------------------------------------------------------- one pass per step in the core (kernels3d_fused.cu)

FUSED_CASES = [
    # size, pml, tfsf, use_tfsf, value_type, complex, steps
    ((100, 24, 20), 3, 6, True, 8, False, 12),    # two x tiles: the rim lane recomputes E' across the tile boundary
    ((100, 24, 20), 3, 6, False, 8, False, 14),   # point source (its Dz advanced explicitly)
    ((150, 31, 23), 4, 7, True, 8, False, 9),     # three x tiles, ragged y and z
    ((80, 28, 18), 3, 6, True, 8, True, 10),      # complex double: 32-cell tiles
    ((200, 40, 22), 4, 6, False, 4, False, 12),   # float: 128-cell tiles, 8-row fused tiles inside 16-row shell tiles
    ((200, 40, 22), 4, 6, True, 4, False, 12),
    ((100, 36, 20), 3, 6, True, 4, True, 10),     # complex float
    ((48, 36, 28), 5, 9, True, 8, False, 30),     # one x tile (the reference-dump size)
]


@pytest.mark.parametrize("case", FUSED_CASES, ids=lambda c: "%s-t%d-vt%d%s" % ("x".join(map(str, c[0])), c[3], c[4], "c" if c[5] else ""))
def test_one_pass_core_equals_two_pass_and_oracle(case):
    """With fdtd_set_tuning (allow_one_pass) exact_mode 0 with a homogeneous mu advances the core with ONE kernel per step
    (E and H fused, ping-pong buffers).
    Same per-cell arithmetic as the two half-step kernels, so the fields are bit-identical to the two-pass run, and both
    are within the north-star tolerance of the oracle -- also across runs of odd length (buffer parity), D/B reads
    between runs and an upload that switches the one-pass kernel off (non-homogeneous mu)."""
    size, pml, tfsf, use_tfsf, vt, cplx, steps = case
    src = (size[0] // 2, size[1] // 2, size[2] // 2)
    kw = dict(size=size, pml=pml, tfsf=tfsf, use_pml=True, use_tfsf=use_tfsf, value_type=vt, complex_values=cplx, source_pos=src,
              theta=60, phi=30, psi=20)
    cfg = O.make_config(exact_mode=0, **kw)
    inputs, fields = random_inputs(cfg, seed=77)
    mu = 1.25
    inputs_ref = dict(inputs, Mu=np.full_like(inputs["Mu"], mu))
    keys = _keys(cfg)
    ref = O.run(cfg, steps, inputs_ref, fields)

    def run(one_pass, chunks):
        with F.Solver(cfg) as s:
            s.fill("Mu", mu)
            for a in range(3):
                s.set_sigma(a, inputs["Sigma" + "XYZ"[a]])
            s.upload("Eps", inputs["Eps"])
            for name in O.field_names(cfg):
                s.upload(name, fields[name])
            s.set_tuning(allow_one_pass=one_pass)   # off by default
            box = s.fused_box
            mids = []
            for n in chunks:
                s.run(n)
                mids.append(s.download("Ez"))
            out = {k: s.download(k) for k in keys}
        return box, out, mids

    chunks = [1, 2, steps - 3]
    box, out, mids = run(True, chunks)
    assert box is not None, "the one-pass kernel is meant to be in use on this grid"
    assert box[0][0] == 0 and box[1][0] == size[0] - 1 and box[0][1] >= pml and box[1][2] <= size[2] - pml - 1
    box2, two, mids2 = run(False, chunks)
    assert box2 is None
    for k in keys:
        assert np.array_equal(out[k], two[k]), "%s: one-pass and two-pass runs differ, max |d| = %g" % (k, np.abs(out[k] - two[k]).max())
    for a, b in zip(mids, mids2):
        assert np.array_equal(a, b)
    errs = family_rel_errs(out, ref, keys)
    assert max(errs.values()) <= TOL[vt], errs

    # a non-homogeneous mu switches back to the two-pass kernels mid-way (odd parity at that moment), results unchanged
    ref2 = O.run(cfg, steps, inputs, fields)
    with F.Solver(cfg) as s:
        s.fill("Mu", 1.0)
        for a in range(3):
            s.set_sigma(a, inputs["Sigma" + "XYZ"[a]])
        s.upload("Eps", inputs["Eps"])
        for name in O.field_names(cfg):
            s.upload(name, fields[name])
        assert (s.fused_box is not None) == (cplx and vt == 8)   # by default only complex double uses it
        s.set_tuning(allow_one_pass=True)
        assert s.fused_box is not None
        s.upload("Mu", inputs["Mu"])
        assert s.fused_box is None
        s.run(steps)
        got = {k: s.download(k) for k in keys}
    errs = family_rel_errs(got, ref2, keys)
    assert max(errs.values()) <= TOL[vt], errs


FUSED_TMZ_CASES = [
    # size, pml, tfsf, use_tfsf, value_type, complex, steps, uniform eps
    ((600, 60), 5, 8, False, 8, False, 21, True),     # point source, eps and mu homogeneous (the reference's TMz set-up)
    ((600, 60), 5, 8, True, 8, False, 20, False),     # TF/SF, varying eps
    ((1100, 50), 6, 9, False, 4, False, 25, True),    # float: 128-cell tiles
    ((1100, 50), 6, 9, True, 4, False, 20, False),
    ((150, 44), 5, 8, True, 8, True, 16, False),      # complex double: 32-cell tiles
    ((96, 80), 8, 14, False, 8, False, 60, True),     # the reference-dump size
]


@pytest.mark.parametrize("case", FUSED_TMZ_CASES, ids=lambda c: "%s-t%d-vt%d%s%s" % ("x".join(map(str, c[0])), c[3], c[4], "c" if c[5] else "", "-uni" if c[7] else ""))
def test_one_pass_core_tmz_equals_two_pass_and_oracle(case):
    """2D TMz: the one-pass kernel (kernels2d_fused.cu) against the two half-step kernels (bit-identical: same per-cell
    arithmetic) and the oracle, over runs of odd and even length."""
    size, pml, tfsf, use_tfsf, vt, cplx, steps, uni_eps = case
    src = (min(70, size[0] // 2), size[1] // 2, 0)
    kw = dict(dims=2, size=size, pml=pml, tfsf=tfsf, use_pml=True, use_tfsf=use_tfsf, value_type=vt, complex_values=cplx, source_pos=src)
    cfg = O.make_config(exact_mode=0, **kw)
    inputs, fields = random_inputs(cfg, seed=78)
    mu, eps = 1.25, 2.0
    inputs["Mu"] = np.full_like(inputs["Mu"], mu)
    if uni_eps:
        inputs["Eps"] = np.full_like(inputs["Eps"], eps)
    keys = _keys(cfg)
    ref = O.run(cfg, steps, inputs, fields)

    def run(one_pass, chunks):
        with F.Solver(cfg) as s:
            s.fill("Mu", mu)
            for a in range(2):
                s.set_sigma(a, inputs["Sigma" + "XY"[a]])
            if uni_eps:
                s.fill("Eps", eps)
            else:
                s.upload("Eps", inputs["Eps"])
            for name in O.field_names(cfg):
                s.upload(name, fields[name])
            s.set_tuning(allow_one_pass=one_pass)   # off by default
            box = s.fused_box
            mids = []
            for n in chunks:
                s.run(n)
                mids.append(s.download("Hy"))
            out = {k: s.download(k) for k in keys}
        return box, out, mids

    chunks = [1, 2, steps - 3]
    box, out, mids = run(True, chunks)
    assert box is not None, "the one-pass TMz kernel is meant to be in use on this grid"
    assert box[0][0] == 0 and box[1][0] == size[0] - 1 and box[0][1] >= pml and box[1][1] <= size[1] - pml - 1
    box2, two, mids2 = run(False, chunks)
    assert box2 is None
    for k in keys:
        assert np.array_equal(out[k], two[k]), "%s: one-pass and two-pass runs differ, max |d| = %g" % (k, np.abs(out[k] - two[k]).max())
    for a, b in zip(mids, mids2):
        assert np.array_equal(a, b)
    errs = family_rel_errs(out, ref, keys)
    assert max(errs.values()) <= TOL[vt], errs
