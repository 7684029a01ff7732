"""Semantics of the C ABI beyond the arithmetic: asynchronous runs, the time index, error reporting, edge sizes."""
import ctypes as C

import numpy as np
import pytest

import fdtd3d_b200 as F
import oracle_lib as O
from gpu_util import random_inputs, run_gpu

pytestmark = pytest.mark.gpu


def _setup(s, cfg, inputs, fields):
    for a in range(cfg.dims):
        s.set_sigma(a, inputs["Sigma" + "XYZ"[a]])
    s.upload("Eps", inputs["Eps"])
    s.upload("Mu", inputs["Mu"])
    for name in O.field_names(cfg):
        s.upload(name, fields[name])


def test_async_run_then_implicit_sync_on_download():
    cfg = O.make_config(size=(28, 24, 26), pml=4, tfsf=6, use_pml=True, use_tfsf=True, exact_mode=1)
    inputs, fields = random_inputs(cfg, seed=21)
    ref = O.run(cfg, 11, inputs, fields)
    with F.Solver(cfg) as s:
        _setup(s, cfg, inputs, fields)
        s.run_async(4)
        s.run_async(7)            # queues behind the first run
        assert s.time_step == 11  # the time index is advanced when the steps are queued
        got = s.download("Ez")    # every transfer waits for queued steps first
        assert np.array_equal(got, ref["Ez"])
        s.sync()
        assert s.last_run_ms > 0
        s.run(0)
        assert s.last_run_ms == 0 and s.time_step == 11


def test_run_async_queues_runs_without_waiting():
    """Three fdtd_run_async calls in a row: the host is back while the first is still running (fdtd_busy), each has its
    own event pair, and the result is the one run of the same number of steps."""
    import time
    kw = dict(size=(256, 256, 192), pml=10, tfsf=20, use_pml=True, use_tfsf=True, exact_mode=0)
    cfg = O.make_config(**kw)
    inputs = O.reference_setup(cfg)
    chunks = (60, 60, 60)
    with F.Solver(cfg) as s:
        for a in range(3):
            s.set_sigma(a, inputs["Sigma" + "XYZ"[a]])
        s.upload("Eps", inputs["Eps"])
        s.run(5)                      # warm-up: module load, first-launch set-up
        t0 = time.perf_counter()
        for n in chunks:
            s.run_async(n)
        host_s = time.perf_counter() - t0
        busy_after_queueing = s.busy
        s.sync()
        total_s = time.perf_counter() - t0
        last_ms = s.last_run_ms
        assert not s.busy
        a = {k: s.download(k) for k in ("Ez", "Hy", "Dz")}
    assert busy_after_queueing, "the queued runs had already finished when fdtd_run_async returned"
    assert host_s < 0.5 * total_s, (host_s, total_s)
    assert 0 < last_ms * 1e-3 < 0.6 * total_s      # the time of the run queued last, not of all three
    with F.Solver(cfg) as s:
        for a_ in range(3):
            s.set_sigma(a_, inputs["Sigma" + "XYZ"[a_]])
        s.upload("Eps", inputs["Eps"])
        s.run(5)
        s.run(sum(chunks))
        for k in a:
            assert np.array_equal(s.download(k), a[k]), k


def test_time_step_start_is_performNSteps_start_step():
    """cfg.time_step_start = the reference's startStep: the source waveform continues from t, not from 0."""
    cfg = O.make_config(size=(20, 18, 16), pml=3, tfsf=5, use_pml=True, use_tfsf=False, exact_mode=1)
    inputs, fields = random_inputs(cfg, seed=5)
    cfg.time_step_start = 37
    ref = O.run(cfg, 6, inputs, fields)
    out = run_gpu(cfg, 6, inputs, fields)
    for key in O.field_names(cfg):
        assert np.array_equal(out[key], ref[key]), key
    cfg0 = O.make_config(size=(20, 18, 16), pml=3, tfsf=5, use_pml=True, use_tfsf=False, exact_mode=1)
    assert not np.array_equal(O.run(cfg0, 6, inputs, fields)["Ez"], ref["Ez"])


@pytest.mark.parametrize("size", [(2, 2, 2), (3, 70, 2), (65, 2, 3), (129, 9, 5)])
def test_tiny_and_lopsided_grids(size):
    """Grids smaller than one CTA tile in some direction, and just past a tile edge in x."""
    cfg = O.make_config(size=size, pml=1, tfsf=1, use_pml=True, use_tfsf=False, exact_mode=1,
                        source_pos=(size[0] // 2, size[1] // 2, size[2] // 2))
    inputs, fields = random_inputs(cfg, seed=size[0])
    ref = O.run(cfg, 5, inputs, fields)
    out = run_gpu(cfg, 5, inputs, fields)
    for key in O.field_names(cfg):
        assert np.array_equal(out[key], ref[key]), key


def test_multi_rank_config_without_communicator_is_an_error_not_a_hang():
    cfg = O.make_config(size=(16, 16, 16), pml=3, tfsf=4)
    cfg.rank, cfg.nranks = 0, 2
    with F.Solver(cfg) as s:
        with pytest.raises(F.FdtdError) as e:
            s.run(1)
        assert e.value.code == F.COMM_ERROR


def test_invalid_requests_are_rejected_with_a_message():
    cfg = O.make_config(dims=2, size=(80, 40), pml=4, tfsf=6, use_pml=False, use_tfsf=False)
    with F.Solver(cfg) as s:
        for field in ("Ex", "Dz", "EInc"):   # not part of a TMz scheme without PML / TF-SF
            with pytest.raises(F.FdtdError) as e:
                s.download(field)
            assert e.value.code == F.INVALID_ARGUMENT
        with pytest.raises(F.FdtdError):
            s.init_eps_sphere()
        with pytest.raises(F.FdtdError):
            s.subtract_incident()
        assert "3D" in F.lib().fdtd_last_error().decode()
        h = C.c_void_p()
        assert F.lib().fdtd_run(h, 1) == F.INVALID_ARGUMENT


def test_profiling_reports_both_half_steps():
    cfg = O.make_config(size=(64, 64, 64), pml=6, tfsf=10)
    with F.Solver(cfg) as s:
        s.set_profiling(True)
        n0 = s.launch_count
        s.run(5)
        e_ms, h_ms, steps = s.phase_times()
        assert steps == 5 and e_ms > 0 and h_ms > 0
        # per step: incident E, E face tables, E phase, incident H, H face tables, H phase
        assert s.launch_count - n0 == 5 * 6
        assert e_ms + h_ms <= s.last_run_ms * 1.05
        # the eight parts of a MULTI-rank step (fdtd_step_breakdown): nothing to report for a single rank
        parts, bsteps = s.step_breakdown()
        assert bsteps == 0 and len(parts) == 8 and not any(parts)
