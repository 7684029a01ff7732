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
        # keep the run within a few minutes: ~1.2 s per step at 64^3 on one core
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
        steps_run = total
    else:
        cb = cpu_oracle_baseline("c3", n=128, steps=max(2, min(steps, 20)))
        value = cb["value"]
        ms = 128 ** 3 / (value * 1e6) * 1e3
        steps_run = steps
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": args.gpus, "steps": steps_run,
            "warmup": 0, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args.gpus),
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
    cfg = F.make_config(dims=dims, size=(nx, ny, nz_total)[:dims], pml=PML, tfsf=TFSF, use_pml=w["pml"], use_tfsf=w["tfsf"],
                        value_type=w["vt"], complex_values=w["cplx"], dx=DX, wavelength=WAVELENGTH, exact_mode=args.exact,
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
    cells = nx * ny * nz_total
    z0, nzl = F.slab_range(nz_total, rank, world)
    fp = w["vt"]
    fv = fp * (2 if w["cplx"] else 1)
    rdt = torch.float64 if fp == 8 else torch.float32
    fdt = {(8, False): torch.float64, (4, False): torch.float32, (8, True): torch.complex128, (4, True): torch.complex64}[(fp, w["cplx"])]

    # ---- device-resident throughput
    s.run(args.warmup)
    barrier()
    launches0 = s.launch_count
    s.set_profiling(True)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    t_wall = time.perf_counter()
    s.run(args.steps)
    barrier()
    t_wall = time.perf_counter() - t_wall
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(s.last_run_ms)
    e_ms, h_ms, psteps = s.phase_times()
    launches = s.launch_count - launches0
    s.set_profiling(False)
    value = cells * args.steps / (ms * 1e-3) / 1e6

    # ---- roofline of the dominant kernel (single-rank run: per-kernel CUDA events inside the timed region)
    # SURVEY.md §8d, per cell per half step: 3D: (9 or 15) FieldValue + 1 FPValue; TMz E: (4 or 6) FV + FP, H: (5 or 9) FV + FP
    if dims == 3:
        bytes_e = bytes_h = (15 if w["pml"] else 9) * fv + fp
        kn = ("e_phase_3d", "h_phase_3d")
    else:
        bytes_e, bytes_h = (6 if w["pml"] else 4) * fv + fp, (9 if w["pml"] else 5) * fv + fp
        kn = ("ez_phase_tmz", "h_phase_tmz")
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # cells of this rank that the time loop advances without D/B (exact_mode 0, 3D UPML): their D/B values are neither
    # read nor written, so the bytes actually needed are below the algorithmic figure and frac can exceed 1
    box = s.core_box
    core_cells = 0
    if box:
        zlo, zhi = max(box[0][2], z0), min(box[1][2], z0 + nzl - 1)
        core_cells = (box[1][0] - box[0][0] + 1) * (box[1][1] - box[0][1] + 1) * (max(0, zhi - zlo + 1) if dims == 3 else 1)
    roofline = None
    if psteps:
        name, k_ms, bytes_k = (kn[1], h_ms / psteps, bytes_h) if h_ms >= e_ms else (kn[0], e_ms / psteps, bytes_e)
        achieved = bytes_k * (nx * ny * nzl) / (k_ms * 1e-3) / 1e9
        traffic = None
        try:
            prof = json.load(open(os.path.join(ROOT, "profiles", "ncu_summary.json")))
            key = name + ("_core" if box else "")
            # ncu captures exist for c3 (default mode and exact_mode 2) and for c2's default mode
            have = (args.workload == "c3" and args.exact != 1) or (args.workload == "c2" and args.exact == 0)
            traffic = prof.get(key, {}).get("dram_bytes_per_launch") if have else None
        except Exception:
            pass
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                    "kernel": name, "kernel_ms": k_ms, "e_phase_ms": e_ms / psteps, "h_phase_ms": h_ms / psteps,
                    "algorithmic_bytes_per_launch": bytes_k * nx * ny * nzl,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                    "whole_step_frac": ((bytes_e + bytes_h) * nx * ny * nzl / (ms / args.steps * 1e-3) / 1e9) / peak}
        # materials set with fdtd_fill are known to be homogeneous and are not loaded: mu always here (never uploaded
        # before the timed run), eps unless the sphere was initialised; the 3D kernels use this for mu only
        skip_mat = fp * nx * ny * nzl if ((name == kn[1]) or (dims == 2 and not w["sphere"])) else 0
        if box or skip_mat:
            # skipped in the core: D (or B) of each component, read + written (3D: 3 components; TMz: Dz / Bx, By)
            n_aux = 3 if dims == 3 else (2 if name == kn[1] else 1)
            needed = bytes_k * nx * ny * nzl - 2 * n_aux * fv * core_cells - skip_mat
            roofline.update({"core_cell_fraction": core_cells / float(nx * ny * nzl), "needed_bytes_per_launch": needed,
                             "frac_of_needed_bytes": needed / (k_ms * 1e-3) / 1e9 / peak,
                             "note": "achieved/frac use SURVEY 8(d)'s algorithmic bytes (D/B through every cell and a material value "
                                     "per cell, as the reference formulates the step); the sigma-free core skips D/B and a material "
                                     "set with fdtd_fill is not loaded, so frac > 1 is bytes not moved, not bandwidth above the peak "
                                     "-- frac_of_needed_bytes is the kernel against its own traffic"})
    else:
        whole = (bytes_e + bytes_h) * nx * ny * nzl / (ms / args.steps * 1e-3) / 1e9
        roofline = {"bound": "hbm", "achieved": whole, "peak": peak, "unit": "GB/s", "frac": whole / peak, "traffic": None,
                    "kernel": "whole step per GPU (multi-rank run: boundary + interior launches overlap with the exchange)"}

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
        launches_e2e = s.launch_count
    s.close()

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
                "per_gpu_value": value / world, "host_wall_s": t_wall, "exact_mode": args.exact}
        print(json.dumps(line))
    if dist:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
