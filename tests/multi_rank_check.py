"""Launched by torchrun (one rank per GPU) from tests/test_gpu_multi.py: z-slab sharded runs, every rank's slab compared
with the oracle's whole-grid result -- bit for bit in exact mode, within the north-star tolerance in fast mode (where
the sigma-free core, whose box spans the slabs, skips D/B).  Exit code 0 = parity."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import fdtd3d_b200 as F  # noqa: E402
from fdtd3d_b200 import multi  # noqa: E402
import oracle_lib as O  # noqa: E402
from gpu_util import random_inputs  # noqa: E402


def main():
    rank, world, local = multi.env_rank()
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    failures = 0
    cases = [
        # size (nz not divisible by world on purpose), pml, tfsf, use_pml, use_tfsf, value_type, complex, steps, angles
        ((24, 20, 27), 4, 6, True, True, 8, False, 9, (90, 0, 90)),
        ((21, 18, 26), 3, 5, True, True, 8, True, 7, (40, 50, 30)),
        ((20, 22, 24), 4, 6, True, False, 8, False, 8, (90, 0, 90)),
        ((18, 16, 22), 3, 5, False, True, 4, False, 8, (90, 0, 90)),
        # large enough in x and y for a non-empty core box in fast mode
        ((40, 24, 27), 3, 5, True, True, 8, False, 9, (60, 30, 20)),
        ((40, 24, 26), 3, 5, True, False, 8, False, 9, (90, 0, 90)),
        ((140, 40, 22), 4, 6, True, True, 4, False, 8, (90, 0, 90)),
    ]
    tol = {8: 1e-12, 4: 1e-5}
    # (case, homogeneous mu set with fdtd_fill): with it, fast mode advances the core of every slab with the one-pass
    # kernel (ping-pong buffers; the halo planes then travel between the `nxt` buffers of neighbouring ranks)
    runs = [(c, False) for c in cases] + [(c, True) for c in cases[4:]] + [(((100, 24, 31), 3, 6, True, True, 8, False, 9, (60, 30, 20)), True)]
    for exact_mode in (1, 0):
      for (size, pml, tfsf, use_pml, use_tfsf, vt, cplx, steps, ang), fill_mu in runs:
        kw = dict(size=size, pml=pml, tfsf=tfsf, use_pml=use_pml, use_tfsf=use_tfsf, value_type=vt, complex_values=cplx,
                  theta=ang[0], phi=ang[1], psi=ang[2], exact_mode=exact_mode,
                  source_pos=(size[0] // 2, size[1] // 2, size[2] // 2))
        cfg1 = O.make_config(**kw)
        inputs, fields = random_inputs(cfg1, seed=99)
        if fill_mu:
            inputs["Mu"] = np.full_like(inputs["Mu"], 1.5)
        ref = O.run(cfg1, steps, inputs, fields)  # whole grid, on every rank's CPU
        cfg = O.make_config(**kw)
        cfg.rank, cfg.nranks, cfg.device = rank, world, local
        z0, nzl = F.slab_range(size[2], rank, world)
        with F.Solver(cfg) as s:
            s.comm_init(multi.share_unique_id(dist, rank, "cuda"))
            for a in range(3):
                s.set_sigma(a, inputs["Sigma" + "XYZ"[a]])
            s.upload("Eps", inputs["Eps"])   # whole-grid arrays: planes a rank does not own are ignored
            if fill_mu:
                s.fill("Mu", 1.5)
            else:
                s.upload("Mu", inputs["Mu"])
            for name in O.field_names(cfg):
                s.upload(name, fields[name])
            if fill_mu:
                s.set_tuning(allow_one_pass=True)
            x_ok = size[0] in (40, 100)   # grids whose x tile boundaries lie inside the x core (one tile, or 64 of 100)
            cb = s.core_box
            # planes of the z core on this rank, without the plane next to another rank (it travels mid-step)
            z_ok = cb is not None and (min(cb[1][2], z0 + nzl - 1 - (1 if rank + 1 < world else 0))
                                       >= max(cb[0][2], z0 + (1 if rank > 0 else 0)))
            if fill_mu and x_ok and z_ok and exact_mode == 0 and s.fused_box is None:
                failures += 1
                print("rank %d case %s: the one-pass kernel was expected to be in use" % (rank, size), flush=True)
            s.run(steps - 3)
            s.set_profiling(True)
            s.run(3)                          # a second run exercises the halo refresh at run start
            parts, bsteps = s.step_breakdown()   # the eight parts of a rank's step (two half-step kernels only)
            s.set_profiling(False)
            if not (bsteps in (0, 3) and (bsteps == 3 or s.fused_box is not None) and all(v >= 0 for v in parts)
                    and (bsteps == 0 or 0 < sum(parts) <= 1.5 * s.last_run_ms)):
                failures += 1
                print("rank %d case %s: step breakdown %s over %d steps (run %.3f ms)" % (rank, size, parts, bsteps, s.last_run_ms),
                      flush=True)
            scale = {}
            for name in O.field_names(cfg):
                scale[name[0]] = max(scale.get(name[0], 0.0), float(np.abs(ref[name]).max()))
            for name in O.field_names(cfg):
                got = s.download(name, z0=z0, nz=nzl)
                want = multi.slab_of_array(ref[name], size, z0, nzl)
                if exact_mode == 1 and (vt == 8 or not cplx):
                    ok = np.array_equal(got, want)
                else:
                    ok = float(np.abs(got - want).max()) <= tol[vt] * scale[name[0]]
                if not ok:
                    failures += 1
                    print("rank %d mode %d case %s field %s: max |d| = %g (family scale %g)"
                          % (rank, exact_mode, size, name, np.abs(got - want).max(), scale[name[0]]), flush=True)
            # a whole-grid host buffer that the ranks fill side by side: a rank's download of [0, nz) must write its own
            # planes only (the others keep the sentinel), and an upload from such a buffer must not touch the halos
            for name in ("Ez", "Eps"):
                shape = s.field_shape(name)
                dt = O.real_dtype(cfg) if name == "Eps" else O.field_dtype(cfg)
                whole = np.full(int(np.prod(shape)), -7.25, dtype=dt)
                F._check(F.lib().fdtd_download(s._h, F.FIELD_IDS[name], whole.ctypes.data, F.ORDER_REFERENCE, 0, shape[2]))
                w3 = whole.reshape(shape)
                own = nzl + (1 if name == "Eps" else 0)   # material slabs carry one more plane
                mine = np.ascontiguousarray(w3[:, :, z0:z0 + own]).reshape(-1)
                if not (np.array_equal(mine, s.download(name, z0=z0, nz=own))
                        and (w3[:, :, :z0] == dt(-7.25)).all() and (w3[:, :, z0 + own:] == dt(-7.25)).all()):
                    failures += 1
                    print("rank %d mode %d case %s: whole-grid download of %s wrote outside the slab or differs"
                          % (rank, exact_mode, size, name), flush=True)
            if use_tfsf and exact_mode == 1:
                for name in ("EInc", "HInc"):
                    if vt == 8 and not np.array_equal(s.download(name), ref[name]):
                        failures += 1
                        print("rank %d: %s differs" % (rank, name), flush=True)
    t = torch.tensor([failures], device="cuda")
    dist.all_reduce(t)
    if rank == 0:
        print("multi_rank_check: world %d, total failures %d" % (world, int(t.item())), flush=True)
    dist.destroy_process_group()
    return 1 if int(t.item()) else 0


if __name__ == "__main__":
    sys.exit(main())
