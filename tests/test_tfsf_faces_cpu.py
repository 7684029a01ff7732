"""Index bookkeeping of the TF/SF correction, bit-exact: the integer face boxes the kernels evaluate
(kernels3d_common.cuh, tfsf_e_faces / tfsf_h_faces, run on the host through fdtd_tfsf_face_masks) against the
reference's floating-point predicates YeeGridLayout::doNeedTFSFUpdate{Ex..Hz}Border (YeeGridLayout.cpp:327-809),
evaluated by the reference's own code for every cell (tests/golden/make_faces_golden.py -> tfsf_faces.faces.npz).
SURVEY.md §7 step 2 / §8c: N = 32, L = 8 gives 24 full boxes of 272 cells."""
import os

import numpy as np
import pytest

import fdtd3d_b200 as F
import oracle_lib as O

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tfsf_faces.faces.npz"))
NAMES = ["Ex", "Ey", "Ez", "Hx", "Hy", "Hz"]


@pytest.mark.parametrize("case", ["n32_l8", "ragged", "thin"])
def test_face_boxes_equal_the_reference_predicates(case):
    want = GOLD[case]
    n, l = tuple(int(v) for v in GOLD[case + "_cfg"][:3]), tuple(int(v) for v in GOLD[case + "_cfg"][3:])
    cfg = O.make_config(size=n, pml=1, tfsf=1, use_pml=False, use_tfsf=True)
    for a in range(3):
        cfg.tfsf_size[a] = l[a]
    for c, name in enumerate(NAMES):
        got = np.zeros(n, dtype=np.uint8)
        F._check(F.lib().fdtd_tfsf_face_masks(cfg, F.FIELD_IDS[name], got.ctypes.data, got.size))
        assert np.array_equal(got, want[c]), (case, name, np.argwhere(got != want[c])[:5])
    if case == "n32_l8":
        # the known-answer boxes quoted in SURVEY.md §8c
        assert all(int((want[c] & (1 << d) != 0).sum()) in (0, 272) for c in range(6) for d in range(6))
        ex_down = np.argwhere(want[0] & 4)
        assert ex_down.min(0).tolist() == [7, 7, 8] and ex_down.max(0).tolist() == [23, 7, 23]
        hy_right = np.argwhere(want[4] & 2)
        assert (hy_right[:, 0] == 23).all()
