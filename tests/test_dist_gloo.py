"""World-size-2 test of the multi-GPU contract on CPU (gloo): contiguous batch shards evaluated
with n_global = N, one all-reduce of [loss, pose grads] -- checked with the float64 oracle playing
the role of each rank's kernel (the CUDA path's n_global handling has its own GPU test)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _shard_loss(inp, lo, hi, n_global):
    """Oracle value/gradients of items [lo,hi) normalised by the GLOBAL batch size."""
    import copy
    import parity as P
    sh = copy.copy(inp)
    sh.N = hi - lo
    sh.x = inp.x[lo:hi]
    sh.disparities = [d[lo:hi] for d in inp.disparities]
    sh.rvecs = [r[lo:hi] for r in inp.rvecs]
    sh.tvecs = [t[lo:hi] for t in inp.tvecs]
    r = P.oracle_train_loss(sh)
    f = (hi - lo) / n_global            # every mean in train_loss is over the batch axis too
    return r["loss"] * f, [g * f for g in r["d_rvec"]], [g * f for g in r["d_tvec"]], [g * f for g in r["d_disp"]]


def _worker(rank, world, port, q):
    for p in (ROOT, os.path.join(ROOT, "monodepth2.jl_b200"), os.path.join(ROOT, "tests")):
        sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from mdb200 import dist as mdist, synth
    N, S = 3, 2                                     # ragged on purpose: shards of 2 and 1 items
    inp = synth.make_loss_inputs(40, 24, N, seed=4, scales=(0.5, 1.0))
    lo, hi = mdist.shard_range(N, world, rank)
    loss, drv, dtv, _ = _shard_loss(inp, lo, hi, N)
    packed = mdist.pack(torch.tensor(loss, dtype=torch.float32), [g.float() for g in drv], [g.float() for g in dtv], N, lo)
    mdist.allreduce_loss_posegrads(packed)
    gl, grv, gtv = mdist.unpack(packed, N, S)
    if rank == 0:
        q.put((float(gl), [g.numpy() for g in grv], [g.numpy() for g in gtv]))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_ranges():
    sys.path.insert(0, os.path.join(ROOT, "monodepth2.jl_b200"))
    from mdb200 import dist as mdist
    for N in (1, 3, 8, 64):
        for G in (1, 2, 4, 8):
            r = [mdist.shard_range(N, G, k) for k in range(G)]
            assert r[0][0] == 0 and r[-1][1] == N and all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


def test_two_rank_allreduce_matches_single_process():
    import parity as P
    sys.path.insert(0, os.path.join(ROOT, "monodepth2.jl_b200"))
    from mdb200 import synth
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    gl, grv, gtv = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    inp = synth.make_loss_inputs(40, 24, 3, seed=4, scales=(0.5, 1.0))
    ref = P.oracle_train_loss(inp)
    assert abs(gl - ref["loss"]) <= 1e-6 * abs(ref["loss"])
    for s in range(2):
        assert np.allclose(grv[s], ref["d_rvec"][s].numpy(), rtol=1e-5, atol=1e-8)
        assert np.allclose(gtv[s], ref["d_tvec"][s].numpy(), rtol=1e-5, atol=1e-8)


def test_bind_host_to_device_is_advisory():
    """No NVML / no GPU here: the call reports why it did nothing and leaves the affinity alone."""
    import os
    from mdb200 import dist as mdist
    before = os.sched_getaffinity(0)
    info = mdist.bind_host_to_device("cuda:0")
    assert isinstance(info, dict) and "bound" in info
    if not info["bound"]:
        assert os.sched_getaffinity(0) == before
    os.sched_setaffinity(0, before)
