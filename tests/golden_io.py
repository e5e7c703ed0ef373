"""Loads the fixtures under tests/golden/ (see tests/golden/make_golden.py for their provenance)."""
import json
import os

import numpy as np

from mdb200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LOSS_CASES = ["loss_48x20_b2_automask", "loss_40x24_b1", "simple_depth_32x16"]


def known_answers():
    with open(os.path.join(GOLDEN, "runtests_known_answers.json")) as f:
        return json.load(f)


def load_loss_case(name):
    """-> (synth.LossInputs, dict of golden outputs, simple: bool)"""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    x = z["x"]
    N, _, C, H, W = x.shape
    Ls = len([k for k in z.files if k.startswith("disp")])
    inp = synth.LossInputs(W=W, H=H, C=C, N=N, x=x, disparities=[z[f"disp{k}"] for k in range(Ls)],
                           scales=[float(s) for s in z["scales"]], rvecs=[z["rvec0"], z["rvec1"]],
                           tvecs=[z["tvec0"], z["tvec1"]], K=z["K"], invK=z["invK"], target_id=int(z["target_id"]),
                           source_ids=[int(s) for s in z["source_ids"]], automasking=bool(int(z["automasking"])),
                           auto_loss=z["auto_loss"] if "auto_loss" in z.files else None)
    gold = dict(loss=float(z["loss"]), d_disp=[z[f"d_disp{k}"] for k in range(Ls)],
                d_rvec=[z["d_rvec0"], z["d_rvec1"]], d_tvec=[z["d_tvec0"], z["d_tvec1"]])
    return inp, gold, bool(int(z["simple"]))


def load_render_case(name="render_24x16_p6_b2"):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz")))
