"""Pins the Tier-2 oracle (oracle/fdtd_oracle.hpp) against dumps of the unmodified reference solver.

Bit-for-bit: every field, the auxiliary D/B fields and the incident-wave lines, plus the inputs the
reference's initScheme/initGrids derive (dt, relPhaseVelocity, eps sphere, sigma profiles).
"""
import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O


@pytest.mark.parametrize("name", G.case_names())
def test_oracle_matches_reference_dump(name):
    meta, data = G.load(name)
    cfg = G.config_from_meta(meta)

    # initScheme restatement (Scheme3D.cpp:3387-3402, SchemeTMz.cpp:1948-1956)
    assert cfg.dt == meta["dt"]
    assert cfg.source_frequency == meta["source_frequency"]
    assert cfg.angle_teta == meta["angle_teta"] and cfg.angle_phi == meta["angle_phi"] and cfg.angle_psi == meta["angle_psi"]
    if cfg.dims == 3:
        assert cfg.rel_phase_velocity == meta["rel_phase_velocity"]

    # initGrids restatement (Scheme3D.cpp:3405-3818, SchemeTMz.cpp:1959-2332)
    mine = O.reference_setup(cfg)
    assert np.array_equal(mine["Eps"], data["Eps"])
    for a in range(cfg.dims):
        key = "Sigma" + "XYZ"[a]
        if cfg.use_pml:
            assert np.array_equal(mine[key], data[key]), key
    # the reference fills the sigma grids even when PML is off; they are simply unused then

    out = O.run(cfg, meta["steps"], G.inputs_from_golden(cfg, data), scattered=bool(meta.get("save_res")))
    for key in O.field_names(cfg) + (["EInc", "HInc"] if cfg.use_tfsf else []):
        assert out[key].dtype == data[key].dtype, key
        assert np.array_equal(out[key], data[key]), "%s differs: max |d| = %g" % (key, np.abs(out[key] - data[key]).max())
    # the run is not trivially zero
    assert np.abs(out["Ez"]).max() > 0


@pytest.mark.parametrize("name", ["c1_d", "c1_dc"])
def test_oracle_matches_the_reference_on_baseline_config_c1(name):
    """BASELINE.json configs[0] as written (80^3, PML 10, TF/SF 20, 100 steps, real and complex double): the oracle
    against the condensed dump of the unmodified reference (per-plane XOR of the raw bits of all cells, sums, planes)."""
    import sys
    sys.path.insert(0, G.GOLDEN_DIR)
    import make_c1_golden as C1
    meta, gold = G.load_c1(name)
    cfg = O.make_config(size=meta["size"], pml=meta["pml"], tfsf=meta["tfsf"], use_pml=True, use_tfsf=True, value_type=8,
                        complex_values=name.endswith("c"), dx=meta["dx"], wavelength=meta["wavelength"], exact_mode=1)
    assert cfg.dt == meta["dt"] and cfg.rel_phase_velocity == meta["rel_phase_velocity"]
    small = C1.condense(O.run(cfg, meta["steps"], O.reference_setup(cfg)))
    assert sorted(small) == sorted(gold)
    for key, want in gold.items():
        assert np.array_equal(small[key], want), key


def test_dat_files_written_by_the_reference_are_the_raw_cur_layer():
    """DATDumper (Source/File-Management/Dumper/DATDumper.h:84-118) writes sizeof (FieldValue) bytes of the `cur` value per
    cell in linear-index order, named current[<step>]_rank-<p>_<name>.dat: the bytes of two files the reference itself wrote
    (kept in the fixture) are exactly the dumped grids.  The product's --dump-dat is compared with the same bytes on the GPU
    (tests/test_gpu_cli.py)."""
    meta, data = G.load("3d_d_pml_tfsf_dat")
    for comp in ("Ez", "Hy"):
        raw = data["dat_" + comp]
        assert raw.dtype == np.uint8 and raw.size == data[comp].size * 8
        assert np.array_equal(raw.view(np.float64), data[comp]), comp
