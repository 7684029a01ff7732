"""fdtd_sumsq_noalloc against fdtd_field_sumsq.

fdtd_sumsq_noalloc is the reduction for callers that may not allocate on the GPU (a hook between queued
chunks of steps, an application whose memory pool owns the device, a process that has run out of HBM) or
that have no solver handle.  Its scratch is a fixed array inside the library, shared by all calls, so the
interesting failure is a second call being polluted by what the first one left there: the test makes two
calls in a row, the second on a much shorter array, and holds both to the existing entry point's answer.

Collected last on purpose (file name), so a run that stops at the first failure has finished everything else.
"""
import ctypes as C

import numpy as np
import pytest

import fdtd3d_b200 as F
import oracle_lib as O

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12  # the tolerance test_upload_download_round_trip uses for fdtd_field_sumsq


def _noalloc_sumsq(t_in, t_out, value_type):
    import torch
    err = C.create_string_buffer(512)
    rc = F.lib().fdtd_sumsq_noalloc(t_in.data_ptr(), value_type, t_in.numel(), t_out.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream, err, len(err))
    assert rc == F.OK, err.value.decode()
    torch.cuda.synchronize()
    return float(t_out.item())


@pytest.mark.parametrize("value_type", [8, 4])
def test_two_calls_in_a_row_match_the_handle_based_reduction(value_type):
    import torch
    torch.cuda.set_device(0)
    rng = np.random.default_rng(17)
    dt = np.float64 if value_type == 8 else np.float32
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    # first a large field (many partial sums), then a small one (few): stale partials must not leak
    for size in ((96, 80, 72), (9, 7, 5)):
        cfg = O.make_config(size=size, pml=2, tfsf=3, use_pml=False, use_tfsf=False, value_type=value_type)
        field = rng.standard_normal(size[0] * size[1] * size[2]).astype(dt)
        with F.Solver(cfg) as s:
            s.upload("Ez", field)
            want = s.sumsq("Ez")  # the existing entry point, scratch from the handle's cudaMalloc
        got = _noalloc_sumsq(torch.from_numpy(field).cuda(), out, value_type)
        assert abs(got - want) <= REL_TOL * want, (size, got, want)
        # and against the sum computed on the host, so the check does not lean on the other entry point alone
        host = float((field.astype(np.float64) ** 2).sum())
        assert abs(got - host) <= REL_TOL * host, (size, got, host)


def test_sizes_beyond_the_fixed_scratch_are_refused():
    import torch
    torch.cuda.set_device(0)
    x = torch.ones(16, dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    err = C.create_string_buffer(512)
    limit = 268435456
    rc = F.lib().fdtd_sumsq_noalloc(x.data_ptr(), 8, limit + 1, out.data_ptr(), None, err, len(err))
    assert rc == F.INVALID_ARGUMENT
    assert str(limit).encode() in err.value and b"split" in err.value
    # nothing was launched: the output is untouched, and a valid call still works afterwards
    torch.cuda.synchronize()
    assert float(out.item()) == 0.0
    assert _noalloc_sumsq(x, out, 8) == 16.0
    assert F.lib().fdtd_sumsq_noalloc(x.data_ptr(), 2, 16, out.data_ptr(), None, err, len(err)) == F.INVALID_ARGUMENT
