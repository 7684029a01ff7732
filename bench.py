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

def cpu_oracle_baseline(n=256, steps=None, budget_s=15.0):
    """The Tier-2 oracle (OpenMP port of the reference step) on a bounded n^3 sample of the workload."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    cfg = O.make_config(size=(n, n, n), pml=PML, tfsf=TFSF, use_pml=True, use_tfsf=True, dx=DX, wavelength=WAVELENGTH)
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
    return {"value": n ** 3 * steps / dt / 1e6, "unit": METRIC, "cores": cores, "kind": "port",
            "sample": "%d^3 sub-grid of the workload (same flags), %d steps, OpenMP oracle, alloc+init included" % (n, steps)}


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
        cb = cpu_oracle_baseline(n=128, steps=max(2, min(steps, 20)))
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


def workload_config(n_gpus):
    return {"workload": "3D %dx%dx%d double real, UPML(10)+TF/SF(20) plane wave, eps sphere; BASELINE.json configs[2]%s"
                        % (NX, NY, NZ_PER_GPU * n_gpus, "" if n_gpus == 1 else " extended to %d z-slabs of 512 planes" % n_gpus),
            "size": [NX, NY, NZ_PER_GPU * n_gpus], "pml": PML, "tfsf": TFSF, "dx": DX, "wavelength": WAVELENGTH,
            "parallelism": "1 GPU" if n_gpus == 1 else "z-slab x%d, NCCL send/recv halo planes" % n_gpus,
            "l2": "inputs larger than L2 (>= 15 GB of fields per GPU vs 126 MB L2); no flush needed",
            "arithmetic": "fast mode (1-D reciprocal tables + explicit FMA; -fmad=false elsewhere), within 1e-12 of the reference"}


# ----------------------------------------------------------------------------- our arm

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--exact", type=int, default=0, help="1: bench the exact-mode kernels instead")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--nz", type=int, default=NZ_PER_GPU, help=argparse.SUPPRESS)
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

    nz_total = args.nz * world
    cfg = F.make_config(dims=3, size=(NX, NY, nz_total), pml=PML, tfsf=TFSF, use_pml=True, use_tfsf=True, value_type=8,
                        complex_values=False, dx=DX, wavelength=WAVELENGTH, exact_mode=args.exact, device=local, rank=rank,
                        nranks=world)
    s = F.Solver(cfg)
    if world > 1:
        # the 128-byte NCCL id travels over torch.distributed; the exchange itself is inside the library
        from fdtd3d_b200 import multi
        s.comm_init(multi.share_unique_id(dist, rank, "cuda"))
    for a in range(3):
        s.set_sigma(a, F.pml_sigma_profile(cfg, a))
    s.init_eps_sphere()  # Scheme3D.cpp:3437: centre (40.5,40.5,40.5), r = 20, eps = 2
    cells = NX * NY * nz_total
    z0, nzl = F.slab_range(nz_total, rank, world)

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
    fv = 8
    bytes_per_cell_half_step = 15 * fv + 1 * 8  # 15 FieldValue + 1 FPValue per cell per half step
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    roofline = None
    if psteps:
        name, k_ms = ("h_phase_3d", h_ms / psteps) if h_ms >= e_ms else ("e_phase_3d", e_ms / psteps)
        achieved = bytes_per_cell_half_step * (NX * NY * nzl) / (k_ms * 1e-3) / 1e9
        traffic = None
        try:
            prof = json.load(open(os.path.join(ROOT, "profiles", "ncu_summary.json")))
            traffic = prof.get(name, {}).get("dram_bytes_per_launch")
        except Exception:
            pass
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                    "kernel": name, "kernel_ms": k_ms, "e_phase_ms": e_ms / psteps, "h_phase_ms": h_ms / psteps,
                    "algorithmic_bytes_per_launch": bytes_per_cell_half_step * NX * NY * nzl,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                    "whole_step_frac": (2 * bytes_per_cell_half_step * NX * NY * nzl / (ms / args.steps * 1e-3) / 1e9) / peak}
    else:
        whole = 2 * bytes_per_cell_half_step * NX * NY * nzl / (ms / args.steps * 1e-3) / 1e9
        roofline = {"bound": "hbm", "achieved": whole, "peak": peak, "unit": "GB/s", "frac": whole / peak, "traffic": None,
                    "kernel": "whole step per GPU (multi-rank run: boundary + interior launches overlap with the exchange)"}

    # ---- end to end through the C ABI with host buffers (upload + K steps + download in the timed region)
    e2e = None
    if not args.no_e2e:
        names = ["Ex", "Ey", "Ez", "Hx", "Hy", "Hz", "Dx", "Dy", "Dz", "Bx", "By", "Bz"]
        n_slab = NX * NY * nzl
        host = torch.zeros(n_slab, dtype=torch.float64).pin_memory().numpy()  # all fields start at zero (no RNG)
        eps_host = s.download("Eps", z0=z0, nz=nzl + 1)
        eps_pin = torch.from_numpy(eps_host).pin_memory().numpy()
        mu_pin = torch.ones(eps_host.size, dtype=torch.float64).pin_memory().numpy()
        h2d = (len(names) * n_slab + 2 * eps_host.size) * 8
        d2h = len(names) * n_slab * 8
        outs = host  # results land in the same pinned buffer, field after field
        barrier()
        t0 = time.perf_counter()
        s.upload("Eps", eps_pin, z0=z0, nz=nzl + 1)
        s.upload("Mu", mu_pin, z0=z0, nz=nzl + 1)
        for nme in names:
            s.upload(nme, host, z0=z0, nz=nzl)
        s.run(args.steps)
        for nme in names:
            F._check(F.lib().fdtd_download(s._h, F.FIELD_IDS[nme], outs.ctypes.data, F.ORDER_REFERENCE, z0, nzl))
        barrier()
        t_e2e = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": cells * args.steps / t_e2e / 1e6, "unit": METRIC, "h2d_bytes_per_step": h2d * world / args.steps,
               "d2h_bytes_per_step": d2h * world / args.steps, "seconds": t_e2e,
               "what": "per rank: upload eps, mu and 12 fields from pinned host memory in reference (z-fastest) order, %d steps, "
                       "download 12 fields; host wall clock, max over ranks" % args.steps}
        launches_e2e = s.launch_count
    s.close()

    # ---- CPU baseline beside it (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu = cpu_oracle_baseline()
        except Exception as ex:  # the baseline must not take the GPU numbers down with it
            cpu = {"value": None, "unit": METRIC, "cores": None, "kind": "port", "sample": "failed: %r" % (ex,)}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": workload_config(world), "roofline": roofline,
                "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
                "per_gpu_value": value / world, "host_wall_s": t_wall, "exact_mode": args.exact}
        print(json.dumps(line))
    if dist:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
