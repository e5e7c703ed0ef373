"""Batch sharding of the loss path across the GPUs of one box.

Batch items are independent, N is the slowest axis, so rank r owns the contiguous item range
[r N/G, (r+1) N/G) of every array -- no repacking.  Each rank evaluates its shard with
`n_global = N` (the kernels divide by the GLOBAL batch), so shard losses add.  The only exchange
is ONE all-reduce (sum) of a packed float32 buffer

    [ loss | d_rvec[s][n] (3 floats) for s, n | d_tvec[s][n] (3 floats) for s, n ]

of 1 + 6 S N floats in which a rank fills only its own items' slots (others zero); disparity
gradients stay sharded.  `torch.distributed` is the plumbing (NCCL on GPUs, gloo in CPU tests).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(N: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous, balanced: the first N % world ranks get one extra item."""
    base, rem = divmod(N, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def packed_size(N: int, S: int) -> int:
    return 1 + 6 * S * N


def pack(loss: torch.Tensor, d_rvec: Sequence[torch.Tensor], d_tvec: Sequence[torch.Tensor], N: int, lo: int,
         out: torch.Tensor = None) -> torch.Tensor:
    """Fill this rank's slots of the packed buffer.  d_rvec[s] (n_local,3), d_tvec[s] (n_local,1,3)."""
    S = len(d_rvec)
    n_local = d_rvec[0].shape[0]
    if out is None:
        out = torch.zeros(packed_size(N, S), dtype=torch.float32, device=loss.device)
    else:
        out.zero_()
    out[0] = loss.reshape(())
    rv = out[1:1 + 3 * S * N].view(S, N, 3)
    tv = out[1 + 3 * S * N:].view(S, N, 3)
    for s in range(S):
        rv[s, lo:lo + n_local] = d_rvec[s].reshape(n_local, 3)
        tv[s, lo:lo + n_local] = d_tvec[s].reshape(n_local, 3)
    return out


def unpack(packed: torch.Tensor, N: int, S: int):
    """-> loss (0-dim), d_rvec [S x (N,3)], d_tvec [S x (N,1,3)] for the GLOBAL batch."""
    rv = packed[1:1 + 3 * S * N].view(S, N, 3)
    tv = packed[1 + 3 * S * N:].view(S, N, 3)
    return packed[0], [rv[s] for s in range(S)], [tv[s].reshape(N, 1, 3) for s in range(S)]


def allreduce_loss_posegrads(packed: torch.Tensor, group=None) -> torch.Tensor:
    """The path's single collective (sum).  Enqueued on the current stream for NCCL."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group)
    return packed


def bind_host_to_device(device) -> dict:
    """Pin this rank's host threads to the CPUs NVML reports as local to `device` (its NUMA node), BEFORE the
    pinned staging buffers are allocated: first-touch then places them on the node the GPU's PCIe root hangs
    off, and the H2D / D2H copies of the ranks stop crossing the socket interconnect.  One process per GPU
    makes this a per-process setting.  Returns what was done; never raises (a cpuset that excludes the ideal
    CPUs, or no NVML, leaves the affinity alone)."""
    import os
    info = {"bound": False}
    try:
        import pynvml as nv
        nv.nvmlInit()
        dev = torch.device(device)
        uuid = str(torch.cuda.get_device_properties(dev).uuid)
        h = nv.nvmlDeviceGetHandleByUUID(uuid if uuid.startswith("GPU-") else "GPU-" + uuid)
        ncpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        ideal = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        cpus = ideal & allowed
        info.update(ideal=len(ideal), allowed=len(allowed))
        if cpus and cpus != allowed:
            os.sched_setaffinity(0, cpus)
            info.update(bound=True, cpus=len(cpus))
    except Exception as e:   # noqa: BLE001 -- advisory only
        info["error"] = f"{type(e).__name__}: {e}"
    return info
