"""End to end through the solver binary: the reference's command line in, DAT dumps out, compared with dumps of the
unmodified reference solver (tests/golden/).  The golden runs used the default dx / wavelength, which the CLI keeps
exact (only values *given* on the command line go through std::stof)."""
import os
import subprocess

import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O
from gpu_util import family_rel_errs

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "fdtd3d_b200", "bin", "fdtd3d")


@pytest.mark.parametrize("name", ["3d_d_pml_tfsf", "3d_d_pml_tfsf_angled", "3d_d_vacuum_src", "3d_dc_pml_tfsf", "3d_f_pml_tfsf",
                                  "tmz_d_pml_src", "tmz_d_pml_tfsf", "tmz_f_pml_src"])
def test_cli_run_matches_reference_dump(name, tmp_path):
    meta, data = G.load(name)
    dims, vt, cplx = G.VARIANT[meta["variant"]]
    n = meta["size"]
    argv = ["--3d" if dims == 3 else "--2d", "--sizex", str(n[0]), "--sizey", str(n[1])]
    if dims == 3:
        argv += ["--sizez", str(n[2])]
    argv += ["--pml-sizex", str(meta["pml"]), "--same-size-pml", "--tfsf-sizex", str(meta["tfsf"]), "--same-size-tfsf",
             "--time-steps", str(meta["steps"]), "--angle-teta", str(meta["theta"]), "--angle-phi", str(meta["phi"]),
             "--angle-psi", str(meta["psi"]), "--value-type", "d" if vt == 8 else "f", "--exact", "--dump-dat", str(tmp_path)]
    if meta["use_pml"]:
        argv.append("--use-pml")
    if meta["use_tfsf"]:
        argv.append("--use-tfsf")
    if cplx:
        argv.append("--complex-values")
    r = subprocess.run([EXE] + argv, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout
    assert "Total time =" in r.stdout and "Grid size:" in r.stdout
    cfg = G.config_from_meta(meta)
    keys = O.field_names(cfg)
    out = {}
    for key in keys:
        path = tmp_path / ("current[%d]_rank-0_%s.dat" % (meta["steps"], key))
        out[key] = np.fromfile(str(path), dtype=O.field_dtype(cfg))
    errs = family_rel_errs(out, data, keys)
    for key in keys:
        if vt == 8:
            assert np.array_equal(out[key], data[key]), (key, errs[key])
        else:
            assert errs[key] <= 1e-5, (key, errs)
