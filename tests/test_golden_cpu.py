"""Both CPU oracles against the committed fixtures (tests/golden/).  The plain-C oracle shares no
code with the torch oracle that generated the .npz files, so its agreement is an independent check;
the torch oracle's agreement is a regression pin."""
import numpy as np
import pytest
import torch

import golden_io as G
import parity as P
from oracle import c_ref
from oracle import torch_ref as T


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64).reshape(-1), np.asarray(b, dtype=np.float64).reshape(-1)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("name", G.LOSS_CASES)
def test_c_oracle_reproduces_golden(name):
    inp, gold, simple = G.load_loss_case(name)
    got = c_ref.train_loss(inp.x, inp.disparities, inp.rvecs, inp.tvecs, inp.auto_loss, inp.K, inp.invK,
                           target_id=inp.target_id, source_ids=inp.source_ids, scales=inp.scales,
                           automasking=inp.automasking, normalize=not simple, smooth_mode=1 if simple else 0)
    assert abs(got["loss"] - gold["loss"]) < 1e-11 * abs(gold["loss"])
    for a, b in zip(got["d_disp"] + got["d_rvec"] + got["d_tvec"], gold["d_disp"] + gold["d_rvec"] + gold["d_tvec"]):
        assert _rel(a, b) < 1e-9


@pytest.mark.parametrize("name", G.LOSS_CASES)
def test_torch_oracle_reproduces_golden(name):
    inp, gold, simple = G.load_loss_case(name)
    ref = P.oracle_train_loss(inp, normalize=not simple, smooth_mode=1 if simple else 0)
    assert abs(ref["loss"] - gold["loss"]) < 1e-12 * abs(gold["loss"])
    for a, b in zip(ref["d_disp"], gold["d_disp"]):
        assert _rel(a.numpy(), b) < 1e-10


def test_torch_oracle_reproduces_render_golden():
    z = G.load_render_case()
    f = lambda a: torch.tensor(a, dtype=torch.float64)
    ro, do, mo = T.render_tgt_rgb_depth(f(z["rgb"]), f(z["sigma"]), f(z["disparity_src"]), f(z["xyz_tgt"]),
                                        T.Pose(f(z["rvec"]), f(z["tvec"])), f(z["K_inv"]), f(z["K"]))
    assert _rel(ro.numpy(), z["rgb_out"]) < 1e-12 and _rel(do.numpy(), z["depth_out"]) < 1e-12
    assert np.array_equal(mo.numpy(), z["mask_out"])


def test_known_answers_file():
    ka = G.known_answers()
    assert ka["smooth_loss"]["value"] == 0.2542            # test/runtests.jl:82
    pd = ka["pose_derivative"]                              # test/runtests.jl:124-142 scenario, oracle values
    r = np.array([pd["r"]], dtype=np.float64)
    R = c_ref.so3_exp_map(r)[0].T
    l = np.linalg.norm(R @ np.array(pd["x"], dtype=np.float64) - np.array(pd["target"], dtype=np.float64))
    assert abs(l - pd["loss"]) < 1e-12
    # d l / d t = unit residual; d l / d r through the C pullback of so3_exp_map
    res = (R @ np.array(pd["x"], dtype=np.float64) - np.array(pd["target"], dtype=np.float64)) / l
    assert np.allclose(res, pd["d_t"], atol=1e-12)
    dR = np.outer(res, np.array(pd["x"], dtype=np.float64))          # math-natural dR[i][j]
    gr, _ = c_ref.compose_T_bwd(r, np.zeros((1, 3)), False, dR.T[None], np.zeros((1, 3)))
    assert np.allclose(gr[0], pd["d_r"], atol=1e-12)
