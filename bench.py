#!/usr/bin/env python
"""Benchmark of the fdtd3d_b200 hot path: Mcell-updates/s of the Yee leapfrog step.

Workload (BASELINE.json configs[2], the roofline headline): 3D 512x512x512, double, real, UPML + TF/SF
plane wave, eps = the reference's dielectric sphere.  With N > 1 GPUs the grid is 512 x 512 x (512*N),
split into N z-slabs of 512 planes (weak scaling), halo planes exchanged with NCCL send/recv.
One bench "step" = one leapfrog time step (E half step + H half step) of the whole grid.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line (see the task contract): value = whole-job Mcell-updates/s from device time (CUDA
events on the solver's stream, max over ranks), e2e = the same metric through the C ABI with host
buffers (upload + K steps + download inside the timed region), roofline for the dominant kernel,
cpu_baseline = the CPU oracle timed on this box's host cores (a reported baseline, not the target).

--impl reference times the reference's own CPU implementation of the path: the unmodified reference
solver built into oracle/_ref (single-threaded: the reference has no OpenMP) on a bounded sub-grid of
the same workload; if that binary is absent, the oracle port.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mcell-updates/s"
NX = NY = 512
NZ_PER_GPU = 512
PML, TFSF = 10, 20
DX, WAVELENGTH = 0.0005, 0.02

# BASELINE.json configs; "c3" is the configuration the metric is quoted on (roofline headline) and the default.
# nz: planes per GPU (weak scaling) unless "nz_total" is given (the grid is then split over the ranks).
WORKLOADS = {
    "c3": dict(dims=3, nx=512, ny=512, nz=512, vt=8, cplx=False, pml=True, tfsf=True, sphere=True,
               label="3D 512x512x%(nz)d double real, UPML(10)+TF/SF(20) plane wave, eps sphere; BASELINE.json configs[2]"),
    "c1": dict(dims=3, nx=80, ny=80, nz_total=80, vt=8, cplx=False, pml=True, tfsf=True, sphere=True,
               label="3D 80x80x%(nz)d double real, UPML(10)+TF/SF(20), eps sphere; BASELINE.json configs[0] (README launch example)"),
    "c2": dict(dims=2, nx=4096, ny=4096, nz_total=1, vt=4, cplx=False, pml=True, tfsf=False, sphere=False,
               label="2D TMz 4096x4096 float real, UPML(10), point source at Ez(70, 2048); BASELINE.json configs[1]"),
    "c4": dict(dims=3, nx=1024, ny=1024, nz_total=1024, vt=8, cplx=True, pml=True, tfsf=True, sphere=True,
               label="3D 1024x1024x%(nz)d complex double, UPML(10)+TF/SF(20), z-slab sharded; BASELINE.json configs[3]"),
    "c4w": dict(dims=3, nx=1024, ny=1024, nz=128, vt=8, cplx=True, pml=True, tfsf=True, sphere=True,
                label="3D 1024x1024x%(nz)d complex double, UPML(10)+TF/SF(20); 128 planes per GPU = configs[3]'s slab at 8 GPUs (tuning aid)"),
    "c4s": dict(dims=3, nx=512, ny=512, nz=256, vt=8, cplx=True, pml=True, tfsf=True, sphere=True,
                label="3D 512x512x%(nz)d complex double, UPML(10)+TF/SF(20); one-GPU sized replica of configs[3] (tuning aid)"),
    "c5": dict(dims=3, nx=1024, ny=1024, nz=512, vt=4, cplx=False, pml=True, tfsf=False, sphere=True,
               label="3D 1024x1024x%(nz)d float real, UPML(10), point source; BASELINE.json configs[4] (weak scaling)"),
}


# ----------------------------------------------------------------------------- clocks

class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.proc, self.rows = device, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = sorted(int(float(r[1])) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit())
        mx = [int(float(r[2])) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 8 for i in range(4) if r[4 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU legs (oracle = checker / baseline only)

def cpu_oracle_baseline(workload="c3", n=None, steps=None, budget_s=15.0):
    """The Tier-2 oracle (OpenMP port of the reference step) on a bounded sub-grid sample of the workload."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    w = WORKLOADS[workload]
    if n is None:
        n = min(256, w["nx"]) if w["dims"] == 3 else min(2048, w["nx"])
    size = (n, n, n)[:w["dims"]]
    cfg = O.make_config(dims=w["dims"], size=size, pml=PML, tfsf=TFSF, use_pml=w["pml"], use_tfsf=w["tfsf"], value_type=w["vt"],
                        complex_values=w["cplx"], dx=DX, wavelength=WAVELENGTH)
    inputs = O.reference_setup(cfg)
    cores = O.lib().oracle_max_threads()
    t0 = time.perf_counter()
    O.run(cfg, 1, inputs)
    one = time.perf_counter() - t0
    if steps is None:
        steps = max(2, min(50, int(budget_s / max(one, 1e-3))))
    t0 = time.perf_counter()
    O.run(cfg, steps, inputs)
    dt = time.perf_counter() - t0
    return {"value": n ** w["dims"] * steps / dt / 1e6, "unit": METRIC, "cores": cores, "kind": "port",
            "sample": "%s sub-grid of the workload (same flags), %d steps, OpenMP oracle, alloc+init included" % ("x".join([str(n)] * w["dims"]), steps)}


def run_reference_arm(args):
    """--impl reference: the unmodified reference CPU solver (oracle/_ref) on a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    steps, warmup = args.steps, args.warmup
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_3d_d")
    n = 64
    if os.path.exists(exe):
        # the reference has no warm-up notion and runs all steps in one call: W + K steps, whole time counted
        total = steps + warmup
        # keep the run within a few minutes: ~0.5 s per step at 64^3 on one core
        total = min(total, 120)
        with tempfile.TemporaryDirectory() as tmp:
            out = subprocess.run([exe, str(n), str(n), str(n), str(PML), str(TFSF), str(total), "1", "1", repr(DX), repr(WAVELENGTH),
                                  "90", "0", "90", os.path.join(tmp, "dump.bin")], check=True, stdout=subprocess.PIPE, text=True).stdout
        secs = float(out.split("Total time =")[1].split()[0])
        value = n ** 3 * total / secs / 1e6
        cb = {"value": value, "unit": METRIC, "cores": 1, "kind": "reference",
              "sample": "%d^3 sub-grid of the workload (same flags: double, PML, TF/SF, sphere), %d steps, unmodified reference "
                        "solver from oracle/_ref, single thread (the reference has no OpenMP; MPI absent)" % (n, total)}
        ms = secs / total * 1e3
        timed = total
    else:
        n = 128
        cb = cpu_oracle_baseline("c3", n=n, steps=max(2, min(steps, 20)))
        value = cb["value"]
        ms = n ** 3 / (value * 1e6) * 1e3
        timed = max(2, min(steps, 20))
    # The line says what ran: a sub-grid sample of the headline workload.  The reference cannot reach 512^3 (about
    # 250 GB and days), so the per-cell rate of the sample is the only comparison there is.
    cfg = workload_config(args.gpus)
    cfg["workload"] = "%d^3 sample of: %s" % (n, cfg["workload"])
    cfg["full_size"] = cfg["size"]
    cfg["size"] = [n, n, n]
    cfg["parallelism"] = "1 host thread"
    cfg["l2"] = "CPU run"
    cfg["arithmetic"] = "the reference's own (x86-64 -O2, no FMA contraction)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": args.gpus, "steps": steps,
            "warmup": warmup, "timed_steps": timed, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": cfg,
            "note": "the reference runs warm-up and timed steps in one performSteps call: its own 'Total time' over all "
                    "timed_steps is what value and ms_per_step are computed from",
            "cpu_baseline": cb, "e2e": {"value": value, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def workload_nz(w, n_gpus):
    return w["nz_total"] if "nz_total" in w else w["nz"] * n_gpus


def workload_config(n_gpus, name="c3", exact=0):
    w = WORKLOADS[name]
    nz = workload_nz(w, n_gpus)
    fv = w["vt"] * (2 if w["cplx"] else 1)
    nfields = (12 if w["pml"] else 6) if w["dims"] == 3 else (6 if w["pml"] else 3)
    gb = nfields * w["nx"] * w["ny"] * nz * fv / n_gpus / 1e9
    return {"workload": w["label"] % {"nz": nz} + ("" if n_gpus == 1 else "; %d z-slabs" % n_gpus),
            "name": name, "size": [w["nx"], w["ny"], nz][:w["dims"]], "pml": PML if w["pml"] else 0, "tfsf": TFSF if w["tfsf"] else 0,
            "dx": DX, "wavelength": WAVELENGTH,
            "parallelism": "1 GPU" if n_gpus == 1 else "z-slab x%d, NCCL send/recv halo planes" % n_gpus,
            "l2": "inputs larger than L2 (%.2f GB of fields per GPU vs 126 MB L2); no flush" % gb if gb > 0.25 else
                  "fields (%.3f GB) fit in the 126 MB L2: numbers are L2-resident, not an HBM roofline" % gb,
            "arithmetic": {1: "exact mode (reference operand order, IEEE division, no FMA contraction): bit-for-bit with the reference",
                           2: "fast mode (1-D reciprocal tables + explicit FMA; -fmad=false elsewhere) with D/B carried through every cell as the "
                              "reference does, within 1e-12 (double) / 1e-5 (float) of the reference",
                           0: "fast mode (1-D reciprocal tables + explicit FMA; -fmad=false elsewhere), within 1e-12 (double) / 1e-5 (float) of the "
                              "reference" + ("; UPML: cells with sigma = 0 are advanced without their D/B values (reconstructed when read)"
                                             if w["pml"] else "")}[exact]}


# ----------------------------------------------------------------------------- measuring one configuration

def _peak():
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(peaks["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    except Exception:
        return 6650.0, "fallback 6650 GB/s (of fallback)"


def byte_model(w, nzl, box, z0, mu_uniform=True, eps_uniform=False, fbox=None):
    """SURVEY.md §8d bytes of one rank's slab per half step: algorithmic (D/B through every cell, a material value per
    cell, as the reference formulates the step) and needed (what the default mode has to move: no D/B in the sigma-free
    core, no material array that was set with fdtd_fill).  With the one-pass kernel (fbox = its box) the cells inside
    the box cost 12 field values + eps per whole step (+ 12 for D/B in the x absorber lanes); `need_f` is that, and
    need_e / need_h are then the two half steps of the remaining shell."""
    dims, nx, ny = w["dims"], w["nx"], w["ny"]
    fp = w["vt"]
    fv = fp * (2 if w["cplx"] else 1)
    ncell = nx * ny * nzl
    if dims == 3:
        per_e = per_h = (15 if w["pml"] else 9) * fv + fp
        aux_e = aux_h = 3
    else:
        per_e, per_h = (6 if w["pml"] else 4) * fv + fp, (9 if w["pml"] else 5) * fv + fp
        aux_e, aux_h = 1, 2
    alg_e, alg_h = per_e * ncell, per_h * ncell

    def inter(lo, hi, lo2, hi2):
        return max(0, min(hi, hi2) - max(lo, lo2) + 1)

    core_cells = 0
    if box:
        core_cells = (box[1][0] - box[0][0] + 1) * (box[1][1] - box[0][1] + 1) * (inter(box[0][2], box[1][2], z0, z0 + nzl - 1) if dims == 3 else 1)
    mat_e = 0 if eps_uniform else fp
    mat_h = 0 if mu_uniform else fp
    out = {"alg_e": alg_e, "alg_h": alg_h, "core_cells": core_cells, "cells": ncell, "fused_cells": 0, "alg_f": 0, "need_f": 0}
    if not fbox:
        out["need_e"] = alg_e - 2 * aux_e * fv * core_cells - (fp - mat_e) * ncell
        out["need_h"] = alg_h - 2 * aux_h * fv * core_cells - (fp - mat_h) * ncell
        return out
    # one-pass box: whole x range, rows and planes of fbox (global indices)
    rows = fbox[1][1] - fbox[0][1] + 1
    planes = inter(fbox[0][2], fbox[1][2], z0, z0 + nzl - 1) if dims == 3 else 1
    fcells = nx * rows * planes
    fcore = (box[1][0] - box[0][0] + 1) * rows * planes       # sigma-free cells inside the box
    out["fused_cells"] = fcells
    out["alg_f"] = (per_e + per_h) * fcells
    nf = 12 if dims == 3 else 6    # E and H components read once + written once; as many again for D / B in the x absorber
    out["need_f"] = (nf * fv + mat_e + mat_h) * fcells + nf * fv * (fcells - fcore)
    shell = ncell - fcells
    shell_core = core_cells - fcore                               # sigma-free cells the half-step kernels still advance
    out["alg_e"], out["alg_h"] = per_e * shell, per_h * shell
    out["need_e"] = per_e * shell - 2 * aux_e * fv * shell_core - (fp - mat_e) * shell
    out["need_h"] = per_h * shell - 2 * aux_h * fv * shell_core - (fp - mat_h) * shell
    return out


def make_solver(F, name, exact, rank, world, local, dist):
    w = WORKLOADS[name]
    dims = w["dims"]
    nz_total = workload_nz(w, world)
    cfg = F.make_config(dims=dims, size=(w["nx"], w["ny"], nz_total)[:dims], pml=PML, tfsf=TFSF, use_pml=w["pml"], use_tfsf=w["tfsf"],
                        value_type=w["vt"], complex_values=w["cplx"], dx=DX, wavelength=WAVELENGTH, exact_mode=exact,
                        device=local, rank=rank, nranks=world)
    s = F.Solver(cfg)
    if world > 1:
        # the 128-byte NCCL id travels over torch.distributed; the exchange itself is inside the library
        from fdtd3d_b200 import multi
        s.comm_init(multi.share_unique_id(dist, rank, "cuda"))
    if w["pml"]:
        for a in range(dims):
            s.set_sigma(a, F.pml_sigma_profile(cfg, a))
    if w["sphere"]:
        s.init_eps_sphere()  # Scheme3D.cpp:3437: centre (40.5,40.5,40.5), r = 20, eps = 2
    return s, cfg, nz_total


TUNING = None  # (march, one_pass_march, allow_one_pass) from the command line, applied to every solver bench.py creates


def measure(F, env, name, exact, steps, warmup, sample_clocks=False, keep=False, tuning=None):
    """Device-timed throughput of one configuration on the ranks of this job: W untimed steps, then K steps between
    barriers, CUDA events on the solver's stream, max over ranks.  Returns the numbers of the bench line for it."""
    rank, world, local, dist, barrier, max_over_ranks = env
    w = WORKLOADS[name]
    s, cfg, nz_total = make_solver(F, name, exact, rank, world, local, dist)
    if tuning or TUNING:
        s.set_tuning(*(tuning or TUNING))
    z0, nzl = F.slab_range(nz_total, rank, world)
    cells = w["nx"] * w["ny"] * nz_total
    s.run(warmup)
    barrier()
    launches0 = s.launch_count
    # Large 3D grids: the per-kernel CUDA events are recorded inside the timed region (4 events per 3.6 ms step).  2D TMz
    # and small 3D grids are launch-bound (a step is a few launches of 5-50 microseconds, replayed as a CUDA graph): there
    # 4 events per step would be measured themselves and would switch the graph off, so the timed region runs without
    # them and the same K steps are repeated right afterwards with events for the per-kernel times.
    prof_inline = world == 1 and w["dims"] == 3 and w["nx"] * w["ny"] * (nz_total // world) > (4 << 20)  # small 3D grids: as TMz
    if prof_inline:
        s.set_profiling(True)
    sampler = ClockSampler(local) if (sample_clocks and rank == 0) else None
    if sampler:
        sampler.start()
    barrier()
    t_wall = time.perf_counter()
    s.run(steps)                       # the timed region
    barrier()
    t_wall = time.perf_counter() - t_wall
    clocks = sampler.stop() if sampler else None
    ms = max_over_ranks(s.last_run_ms)
    launches = s.launch_count - launches0
    if not prof_inline:
        s.set_profiling(True)
        s.run(steps if world == 1 else min(steps, 50))
    e_ms, h_ms, psteps = s.phase_times()
    breakdown = None
    if world > 1:
        # N > 1: the same K steps (50 at most) once more with nine events per step on the compute stream (fdtd_step_breakdown): where
        # a rank's step goes -- the two waits for a neighbour's halo plane against the kernels.  Outside the timed region.
        import torch
        part_ms, bsteps = s.step_breakdown()
        t = torch.tensor(part_ms if bsteps else [0.0] * 8, dtype=torch.float64, device="cuda:%d" % local) / max(bsteps, 1)
        allp = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allp, t)
        if bsteps:
            breakdown = {"parts": ["incident_e", "wait_h_halo", "e_boundary_plane", "e_interior", "incident_h", "wait_e_halo",
                                   "h_boundary_plane", "h_interior"],
                         "ms_per_step_by_rank": [[round(float(v), 5) for v in a.tolist()] for a in allp],
                         "note": "CUDA events on each rank's compute stream, a separate run of the same K steps right after "
                                 "the timed one; a wait_* part is the time the stream sat behind the neighbour's plane"}
    fused_ms_total = s.fused_ms if psteps else 0.0
    s.set_profiling(False)
    peak, peak_source = _peak()
    box = s.core_box
    fbox = s.fused_box
    bm = byte_model(w, nzl, box, z0, mu_uniform=True, eps_uniform=(w["dims"] == 2 and not w["sphere"]), fbox=fbox)
    step_s = ms / steps * 1e-3
    alg_step = bm["alg_e"] + bm["alg_h"] + bm["alg_f"]
    need_step = bm["need_e"] + bm["need_h"] + bm["need_f"]
    res = {"name": name, "exact_mode": exact, "n_gpus": world, "steps": steps, "warmup": warmup,
           "value": cells * steps / (ms * 1e-3) / 1e6, "ms_per_step": ms / steps, "gpu_launches": launches,
           "whole_step_frac": alg_step / step_s / 1e9 / peak,
           "whole_step_frac_of_needed_bytes": need_step / step_s / 1e9 / peak,
           "core_cell_fraction": bm["core_cells"] / float(bm["cells"]), "one_pass_cell_fraction": bm["fused_cells"] / float(bm["cells"]),
           "host_wall_s": t_wall, "clocks": clocks, "peak": peak, "peak_source": peak_source, "bytes": bm}
    if psteps:
        kn = ("e_phase_3d", "h_phase_3d") if w["dims"] == 3 else ("ez_phase_tmz", "h_phase_tmz")

        def frac(nbytes, t_ms):
            return nbytes / (t_ms * 1e-3) / 1e9 / peak if t_ms > 0 and nbytes > 0 else None

        res.update({"e_phase_ms": e_ms / psteps, "h_phase_ms": h_ms / psteps, "kernels": kn,
                    "e_frac": frac(bm["alg_e"], e_ms / psteps), "h_frac": frac(bm["alg_h"], h_ms / psteps),
                    "e_frac_of_needed_bytes": frac(bm["need_e"], e_ms / psteps),
                    "h_frac_of_needed_bytes": frac(bm["need_h"], h_ms / psteps)})
        if fbox:
            f_ms = fused_ms_total / psteps
            res.update({"one_pass_ms": f_ms, "one_pass_frac": frac(bm["alg_f"], f_ms),
                        "one_pass_frac_of_needed_bytes": frac(bm["need_f"], f_ms)})
    if breakdown:
        res["step_breakdown"] = breakdown
    if keep:
        return res, s, cfg
    s.close()
    return res


def brief(res):
    """The part of a measure() result that goes into roofline.modes / other_configs."""
    keys = ("name", "exact_mode", "n_gpus", "steps", "warmup", "value", "ms_per_step", "e_phase_ms", "h_phase_ms", "e_frac", "h_frac",
            "e_frac_of_needed_bytes", "h_frac_of_needed_bytes", "one_pass_ms", "one_pass_frac", "one_pass_frac_of_needed_bytes",
            "whole_step_frac", "whole_step_frac_of_needed_bytes", "core_cell_fraction", "one_pass_cell_fraction", "gpu_launches")
    out = {k: res[k] for k in keys if k in res}
    out["unit"] = METRIC
    out["workload"] = workload_config(res["n_gpus"], res["name"], res["exact_mode"])["workload"]
    return out


def parity_check(F, env):
    """Before anything is timed: one small case, sharded over the ranks of this job exactly as the timed runs are
    (z-slabs, NCCL halo planes), against the CPU oracle (checker only) -- exact mode bit for bit, the default fast mode
    (sigma-free core) within 1e-12.  Every rank checks its own slab; the verdict is reduced over the ranks."""
    import numpy as np
    rank, world, local, dist, barrier, max_over_ranks = env
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    from fdtd3d_b200 import multi
    size, steps = (72, 40, 12 * world + 5), 10
    out = {"case": "3D %dx%dx%d double, UPML(4)+TF/SF(7), theta 60 phi 30 psi 20, eps sphere, %d steps, %d z-slab(s), vs the CPU oracle"
                   % (size + (steps, world)), "tol": 1e-12, "ranks": world}
    worst = 0.0
    exact_equal = True
    for mode in (1, 0):
        cfg1 = O.make_config(size=size, pml=4, tfsf=7, use_pml=True, use_tfsf=True, theta=60, phi=30, psi=20, exact_mode=mode)
        inputs = O.reference_setup(cfg1)
        ref = O.run(cfg1, steps, inputs)
        cfg = O.make_config(size=size, pml=4, tfsf=7, use_pml=True, use_tfsf=True, theta=60, phi=30, psi=20, exact_mode=mode)
        cfg.rank, cfg.nranks, cfg.device = rank, world, local
        z0, nzl = F.slab_range(size[2], rank, world)
        with F.Solver(cfg) as s:
            if world > 1:
                s.comm_init(multi.share_unique_id(dist, rank, "cuda"))
            for a in range(3):
                s.set_sigma(a, inputs["Sigma" + "XYZ"[a]])
            s.upload("Eps", inputs["Eps"])
            s.run(steps - 3)
            s.run(3)
            if mode == 0:
                out["core_box"] = s.core_box
            fam = {}
            for nme in O.field_names(cfg):
                fam[nme[0]] = max(fam.get(nme[0], 0.0), float(np.abs(ref[nme]).max()))
            for nme in O.field_names(cfg):
                got = s.download(nme, z0=z0, nz=nzl)
                want = multi.slab_of_array(ref[nme], size, z0, nzl)
                if mode == 1:
                    exact_equal = exact_equal and bool(np.array_equal(got, want))
                else:
                    worst = max(worst, float(np.abs(got - want).max()) / fam[nme[0]])
    worst = max_over_ranks(worst)
    bad = max_over_ranks(0.0 if exact_equal else 1.0)
    out.update({"exact_mode_bit_equal": bad == 0.0, "fast_mode_max_rel_err": worst, "ok": bad == 0.0 and worst <= 1e-12})
    return out


# ----------------------------------------------------------------------------- our arm

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--exact", type=int, default=0, choices=[0, 1, 2],
                    help="fdtd_config.exact_mode: 0 fast (default), 1 bit-exact kernels, 2 fast with D/B carried through every cell")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the small sharded parity run against the CPU oracle")
    ap.add_argument("--quick", action="store_true", help="headline line only: no roofline.modes, other_configs, sustained")
    ap.add_argument("--march", type=int, default=0, help="tuning: planes (2D: rows) marched per CTA by the half-step kernels (0 = library rule)")
    ap.add_argument("--one-pass-march", type=int, default=0, help="tuning: planes (2D: tiles) marched per CTA by the one-pass kernel")
    ap.add_argument("--one-pass", action="store_true", help="experimental: one kernel per step in the core (fdtd_set_tuning; slower today, DESIGN 3.1c)")
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS), help="BASELINE.json config (default c3, the headline)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import numpy as np
    import torch
    import fdtd3d_b200 as F

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("WORLD_SIZE %d != --gpus %d" % (world, args.gpus))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: fdtd3d_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    w = WORKLOADS[args.workload]
    dims, nx, ny = w["dims"], w["nx"], w["ny"]
    nz_total = workload_nz(w, world)
    if dims == 2 and world > 1:
        raise SystemExit("the 2D TMz workload runs on one GPU")
    env = (rank, world, local, dist, barrier, max_over_ranks)
    global TUNING
    if args.march or args.one_pass_march or args.one_pass:
        TUNING = (args.march, args.one_pass_march, True if args.one_pass else None)
    headline = args.workload == "c3" and args.exact == 0 and not args.quick and not (args.march or args.one_pass_march or args.one_pass)

    # ---- parity of the sharded path against the CPU oracle, before anything is timed
    parity = None
    if not args.no_parity:
        try:
            parity = parity_check(F, env)
        except Exception as ex:
            parity = {"ok": False, "error": repr(ex)}

    # ---- device-resident throughput of the configuration the metric is quoted on
    res, s, cfg = measure(F, env, args.workload, args.exact, args.steps, args.warmup, sample_clocks=True, keep=True)
    value, ms, launches, clocks, t_wall = res["value"], res["ms_per_step"] * args.steps, res["gpu_launches"], res["clocks"], res["host_wall_s"]
    cells = nx * ny * nz_total
    z0, nzl = F.slab_range(nz_total, rank, world)
    fp = w["vt"]
    fv = fp * (2 if w["cplx"] else 1)
    rdt = torch.float64 if fp == 8 else torch.float32
    fdt = {(8, False): torch.float64, (4, False): torch.float32, (8, True): torch.complex128, (4, True): torch.complex64}[(fp, w["cplx"])]
    peak, peak_source = res["peak"], res["peak_source"]
    bm = res["bytes"]
    box = s.core_box

    # ---- roofline of the dominant kernel (single-rank run: per-kernel CUDA events inside the timed region)
    if "e_phase_ms" in res:
        e_dom = res["e_phase_ms"] >= res["h_phase_ms"]
        name = res["kernels"][0 if e_dom else 1]
        k_ms = res["e_phase_ms"] if e_dom else res["h_phase_ms"]
        alg, need = (bm["alg_e"], bm["need_e"]) if e_dom else (bm["alg_h"], bm["need_h"])
        if res.get("one_pass_ms"):
            # the one-pass core kernel dominates the step: its algorithmic bytes are both half steps of its cells
            name = "step_core_3d_fused" if dims == 3 else "step_core_tmz_fused"
            k_ms, alg, need = res["one_pass_ms"], bm["alg_f"], bm["need_f"]
        achieved = alg / (k_ms * 1e-3) / 1e9
        traffic = None
        try:
            prof = json.load(open(os.path.join(ROOT, "profiles", "ncu_summary.json")))
            key = name if name.startswith("step_core") else name + ("_core" if box else "")
            # ncu captures exist for c3 (default mode and exact_mode 2) and for c2's default mode
            have = (args.workload == "c3" and args.exact != 1) or (args.workload == "c2" and args.exact == 0)
            traffic = prof.get(key, {}).get("dram_bytes_per_launch") if have else None
            traffic_tag = prof.get(key, {}).get("tag") if have else None
        except Exception:
            traffic_tag = None
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                    "traffic_source": None if traffic is None else
                    "constant read from profiles/ncu_summary.json (dram__bytes_read.sum + dram__bytes_write.sum of one "
                    "`ncu --set full` capture of this kernel, tag %s), NOT measured by this run" % traffic_tag,
                    "kernel": name, "kernel_ms": k_ms, "e_phase_ms": res["e_phase_ms"], "h_phase_ms": res["h_phase_ms"],
                    "algorithmic_bytes_per_launch": alg, "peak_source": peak_source, "whole_step_frac": res["whole_step_frac"],
                    "whole_step_frac_of_needed_bytes": res["whole_step_frac_of_needed_bytes"]}
        if res.get("one_pass_ms"):
            roofline.update({"one_pass_cell_fraction": res["one_pass_cell_fraction"], "one_pass_ms": res["one_pass_ms"],
                             "shell_e_phase_ms": res["e_phase_ms"], "shell_h_phase_ms": res["h_phase_ms"],
                             "one_pass_note": "inside its box the step is ONE kernel that reads E, H, eps and writes E, H once: 12 (TMz: 6) "
                                              "field values + eps per cell-update instead of the 30 + 2 (15 + 2) of the reference's formulation; "
                                              "e_phase_ms / h_phase_ms are the half-step kernels on the remaining shell"})
        if need != alg:
            roofline.update({"core_cell_fraction": res["core_cell_fraction"], "needed_bytes_per_launch": need,
                             "frac_of_needed_bytes": need / (k_ms * 1e-3) / 1e9 / peak,
                             "note": "achieved/frac use SURVEY 8(d)'s algorithmic bytes (D/B through every cell and a material value "
                                     "per cell, as the reference formulates the step); the sigma-free core skips D/B and a material "
                                     "set with fdtd_fill is not loaded, so frac > 1 is bytes not moved, not bandwidth above the peak "
                                     "-- frac_of_needed_bytes is the kernel against its own traffic"})
    else:
        whole = (bm["alg_e"] + bm["alg_h"] + bm["alg_f"]) / (res["ms_per_step"] * 1e-3) / 1e9
        roofline = {"bound": "hbm", "achieved": whole, "peak": peak, "unit": "GB/s", "frac": whole / peak, "traffic": None,
                    "peak_source": peak_source, "whole_step_frac_of_needed_bytes": res["whole_step_frac_of_needed_bytes"],
                    "kernel": "whole step per GPU (multi-rank run: boundary + interior launches overlap with the exchange)"}
        if "step_breakdown" in res:
            roofline["step_breakdown"] = res["step_breakdown"]

    # ---- end to end through the C ABI with host buffers (upload + K steps + download in the timed region)
    e2e = None
    if not args.no_e2e:
        names = ["Ex", "Ey", "Ez", "Hx", "Hy", "Hz"] if dims == 3 else ["Ez", "Hx", "Hy"]
        if w["pml"]:
            names += ["Dx", "Dy", "Dz", "Bx", "By", "Bz"] if dims == 3 else ["Dz", "Bx", "By"]
        n_slab = nx * ny * nzl
        nzm = nzl + 1 if dims == 3 else 1
        host = torch.zeros(n_slab, dtype=fdt).pin_memory().numpy()  # all fields start at zero (no RNG)
        eps_host = s.download("Eps", z0=z0, nz=nzm)
        eps_pin = torch.from_numpy(eps_host).pin_memory().numpy()
        mu_pin = torch.ones(eps_host.size, dtype=rdt).pin_memory().numpy()
        h2d = len(names) * n_slab * fv + 2 * eps_host.size * fp
        d2h = len(names) * n_slab * fv
        outs = host  # results land in the same pinned buffer, field after field
        barrier()
        t0 = time.perf_counter()
        s.upload("Eps", eps_pin, z0=z0, nz=nzm)
        s.upload("Mu", mu_pin, z0=z0, nz=nzm)
        for nme in names:
            s.upload(nme, host, z0=z0, nz=nzl)
        s.run(args.steps)
        for nme in names:
            F._check(F.lib().fdtd_download(s._h, F.FIELD_IDS[nme], outs.ctypes.data, F.ORDER_REFERENCE, z0, nzl))
        barrier()
        t_e2e = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": cells * args.steps / t_e2e / 1e6, "unit": METRIC, "h2d_bytes_per_step": h2d * world / args.steps,
               "d2h_bytes_per_step": d2h * world / args.steps, "seconds": t_e2e,
               "what": "per rank: upload eps, mu and %d fields from pinned host memory in reference (z-fastest) order, %d steps, "
                       "download %d fields; host wall clock, max over ranks" % (len(names), args.steps, len(names))}
    s.close()

    # ---- everything else the round claims, measured by this same run (headline invocation only)
    modes = other = sustained = one_pass = None
    if headline:
        short = max(5, min(args.steps, 12))
        # the like-for-like figure (D/B through every cell) and the bit-exact one, same workload, same process
        modes = {}
        for mode, label in ((2, "fast_db_everywhere"), (1, "exact_bit_for_bit")):
            try:
                modes[label] = brief(measure(F, env, "c3", mode, short, 3))
            except Exception as ex:
                modes[label] = {"error": repr(ex)}
        # the other BASELINE.json configurations this job's GPUs can hold
        other = {}
        todo = [("c5", 0)] + ([("c2", 0), ("c2", 2), ("c1", 0)] if world == 1 else [("c4", 0)])
        for nme, mode in todo:
            key = nme if mode == 0 else "%s_exact%d" % (nme, mode)
            try:
                other[key] = brief(measure(F, env, nme, mode, 100 if nme in ("c2", "c1") else short, 5 if nme in ("c2", "c1") else 3))
            except Exception as ex:
                other[key] = {"error": repr(ex)}
        # the experimental one-pass core kernels (off by default), measured beside the shipped two-pass path
        one_pass = {}
        for nme in ["c3", "c5"] + (["c2"] if world == 1 else []):
            try:
                one_pass[nme] = brief(measure(F, env, nme, 0, 100 if nme == "c2" else short, 3, tuning=(0, 0, True)))
            except Exception as ex:
                one_pass[nme] = {"error": repr(ex)}
        # BASELINE.json quotes 1000-step runs: the sustained rate of the headline configuration, with its own clock sample
        try:
            r1000 = measure(F, env, "c3", 0, 1000, 3, sample_clocks=True)
            sustained = {"steps": 1000, "value": r1000["value"], "unit": METRIC, "ms_per_step": r1000["ms_per_step"],
                         "whole_step_frac": r1000["whole_step_frac"], "clocks": r1000["clocks"],
                         "note": "no per-kernel events in this run's timed region would change it: profiling events are recorded "
                                 "between launches on the same stream"}
        except Exception as ex:
            sustained = {"error": repr(ex)}

    # ---- CPU baseline beside it (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu = cpu_oracle_baseline(args.workload)
        except Exception as ex:  # the baseline must not take the GPU numbers down with it
            cpu = {"value": None, "unit": METRIC, "cores": None, "kind": "port", "sample": "failed: %r" % (ex,)}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": {8: "f64", 4: "f32"}[fp] + ("c" if w["cplx"] else ""), "data": "synthetic",
                "config": workload_config(world, args.workload, args.exact), "roofline": roofline,
                "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
                "per_gpu_value": value / world, "host_wall_s": t_wall, "exact_mode": args.exact, "parity": parity}
        if modes is not None:
            line["roofline"]["modes"] = modes
        if other is not None:
            line["other_configs"] = other
        if sustained is not None:
            line["sustained"] = sustained
        if one_pass is not None:
            line["one_pass_experiment"] = {"note": "fdtd_set_tuning (allow_one_pass = 1): E and H of the core in ONE kernel per step (12 instead of 18 "
                                                   "field values per cell, bit-identical results); latency-bound today, hence not the default "
                                                   "(DESIGN.md 3.1c, profiles/r2_one_pass.md)", "configs": one_pass}
        print(json.dumps(line))
    if dist:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
