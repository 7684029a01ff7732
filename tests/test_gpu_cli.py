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


def _run_cli(name, tmp_path, mode_flags):
    """Runs the binary on a golden case; returns (cfg, dumped fields, reference dump)."""
    meta, data = G.load(name)
    dims, vt, cplx = G.VARIANT[meta["variant"]]
    n = meta["size"]
    argv = ["--3d" if dims == 3 else "--2d", "--sizex", str(n[0]), "--sizey", str(n[1])]
    if dims == 3:
        argv += ["--sizez", str(n[2])]
    argv += ["--pml-sizex", str(meta["pml"]), "--same-size-pml", "--tfsf-sizex", str(meta["tfsf"]), "--same-size-tfsf",
             "--time-steps", str(meta["steps"]), "--angle-teta", str(meta["theta"]), "--angle-phi", str(meta["phi"]),
             "--angle-psi", str(meta["psi"]), "--value-type", "d" if vt == 8 else "f", "--dump-dat", str(tmp_path)] + mode_flags
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
    out = {}
    for key in O.field_names(cfg):
        path = tmp_path / ("current[%d]_rank-0_%s.dat" % (meta["steps"], key))
        out[key] = np.fromfile(str(path), dtype=O.field_dtype(cfg))
    return cfg, out, data


@pytest.mark.parametrize("name", ["3d_d_pml_tfsf", "3d_d_pml_tfsf_angled", "3d_d_vacuum_src", "3d_dc_pml_tfsf", "3d_f_pml_tfsf",
                                  "tmz_d_pml_src", "tmz_d_pml_tfsf", "tmz_f_pml_src"])
def test_cli_run_matches_reference_dump(name, tmp_path):
    cfg, out, data = _run_cli(name, tmp_path, ["--exact"])
    keys = O.field_names(cfg)
    errs = family_rel_errs(out, data, keys)
    for key in keys:
        assert np.array_equal(out[key], data[key]), (key, errs[key])  # exact mode: double and real float alike


@pytest.mark.parametrize("name", ["3d_dc_pml_tfsf", "tmz_d_pml_tfsf", "tmz_f_pml_src"])
@pytest.mark.parametrize("flags", [["--fast"], ["--fast", "--keep-aux"], []])
def test_cli_fast_modes_within_tolerance_of_reference_dump(name, flags, tmp_path):
    """The default (fast) arithmetic through the binary: D/B are not carried where sigma = 0 (these grids have such a
    core) unless --keep-aux is given; every dumped field, D and B included, is within tolerance of the reference's dump."""
    cfg, out, data = _run_cli(name, tmp_path, flags)
    keys = O.field_names(cfg)
    errs = family_rel_errs(out, data, keys)
    assert max(errs.values()) <= (1e-12 if cfg.value_type == 8 else 1e-5), errs


def test_cli_save_res_writes_the_scattered_field(tmp_path):
    """--save-res: like Scheme3D::performNSteps, subtract the incident field inside the TF/SF box, then save the grids
    under the reference's names (3D-in-time-Ex ...), as DAT instead of BMP."""
    meta, data = G.load("3d_d_pml_tfsf_saveres")
    n = meta["size"]
    argv = ["--3d", "--sizex", str(n[0]), "--sizey", str(n[1]), "--sizez", str(n[2]), "--pml-sizex", str(meta["pml"]), "--same-size-pml",
            "--tfsf-sizex", str(meta["tfsf"]), "--same-size-tfsf", "--time-steps", str(meta["steps"]), "--angle-teta", str(meta["theta"]),
            "--angle-phi", str(meta["phi"]), "--angle-psi", str(meta["psi"]), "--use-pml", "--use-tfsf", "--exact", "--save-res",
            "--dump-dat", str(tmp_path)]
    r = subprocess.run([EXE] + argv, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout
    for key in ("Ex", "Ey", "Ez", "Hx", "Hy", "Hz"):
        got = np.fromfile(str(tmp_path / ("current[%d]_rank-0_3D-in-time-%s.dat" % (meta["steps"], key))), dtype=np.float64)
        assert np.array_equal(got, data[key]), key


@pytest.mark.parametrize("dims", [3, 2])
def test_cli_checkpoint_and_resume(dims, tmp_path):
    """--dump-dat after 14 steps, --load-dat/--start-step for 9 more == 23 steps in one go, bit for bit (fields,
    auxiliary fields and the incident lines are all part of the checkpoint)."""
    base = (["--3d", "--sizex", "26", "--sizey", "22", "--sizez", "24"] if dims == 3 else ["--2d", "--sizex", "90", "--sizey", "40"]) + \
        ["--pml-sizex", "4", "--same-size-pml", "--tfsf-sizex", "6", "--same-size-tfsf", "--use-pml", "--use-tfsf", "--exact"]
    a, b, c = tmp_path / "a", tmp_path / "b", tmp_path / "c"
    for d in (a, b, c):
        d.mkdir()
    for argv in (["--time-steps", "23", "--dump-dat", str(a)],
                 ["--time-steps", "14", "--dump-dat", str(b)],
                 ["--time-steps", "9", "--load-dat", str(b), "--start-step", "14", "--dump-dat", str(c)]):
        r = subprocess.run([EXE] + base + argv, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
        assert r.returncode == 0, r.stdout
    names = sorted(f for f in os.listdir(a))
    assert len(names) == (14 if dims == 3 else 8) and all(f.startswith("current[23]_rank-0_") for f in names)
    for f in names:
        x, y = np.fromfile(str(a / f), dtype=np.uint8), np.fromfile(str(c / f), dtype=np.uint8)
        assert x.size > 0 and np.array_equal(x, y), f


@pytest.mark.parametrize("ngpus", [2, 4])
@pytest.mark.parametrize("flags", [["--exact"], []])
def test_cli_num_cuda_gpus_is_byte_identical_to_one_gpu(ngpus, flags, tmp_path):
    """--num-cuda-gpus N (Source/Settings/Settings.inc:81, Source/main.cpp:88-96): N z-slabs in one process, NCCL halo
    planes, slabs gathered into the reference's whole-grid DAT files -- byte for byte the single-GPU run, in exact mode
    and in the default fast mode."""
    import torch
    if torch.cuda.device_count() < ngpus:
        pytest.skip("needs %d GPUs" % ngpus)
    base = ["--3d", "--sizex", "100", "--sizey", "28", "--sizez", "46", "--pml-sizex", "4", "--same-size-pml", "--tfsf-sizex", "8",
            "--same-size-tfsf", "--time-steps", "21", "--angle-teta", "60", "--angle-phi", "30", "--angle-psi", "20",
            "--use-pml", "--use-tfsf"] + flags
    dumps = {}
    for n in (1, ngpus):
        d = tmp_path / ("g%d" % n)
        d.mkdir()
        r = subprocess.run([EXE] + base + ["--num-cuda-gpus", str(n), "--dump-dat", str(d)], stdout=subprocess.PIPE,
                           stderr=subprocess.STDOUT, text=True, timeout=300)
        assert r.returncode == 0, r.stdout
        assert "GPUs: %d" % n in r.stdout
        dumps[n] = {f: open(str(d / f), "rb").read() for f in sorted(os.listdir(str(d)))}
    assert sorted(dumps[1]) == sorted(dumps[ngpus]) and len(dumps[1]) == 14
    for f in dumps[1]:
        assert dumps[1][f] == dumps[ngpus][f], f
    ez = np.frombuffer(dumps[1]["current[21]_rank-0_Ez.dat"], dtype=np.float64)
    assert np.abs(ez).max() > 1e-3


def test_cli_dat_files_equal_files_written_by_the_reference(tmp_path):
    """--dump-dat against DAT files the reference's own DATDumper wrote for the same run (bytes kept in the fixture)."""
    cfg, out, data = _run_cli("3d_d_pml_tfsf_dat", tmp_path, ["--exact"])
    meta, _ = G.load("3d_d_pml_tfsf_dat")
    for comp in ("Ez", "Hy"):
        mine = open(str(tmp_path / ("current[%d]_rank-0_%s.dat" % (meta["steps"], comp))), "rb").read()
        assert mine == data["dat_" + comp].tobytes(), comp
