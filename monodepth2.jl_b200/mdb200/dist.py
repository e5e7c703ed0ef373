"""Batch sharding of the loss path across the GPUs of one box.

Batch items are independent, N is the slowest axis, so rank r owns the contiguous item range
[r N/G, (r+1) N/G) of every array -- no repacking.  Each rank evaluates its shard with
`n_global = N` (the kernels divide by the GLOBAL batch), so shard losses add.  The only exchange
is ONE all-reduce (sum) of a packed float32 buffer

    [ loss | d_rvec[s][n] (3 floats) for s, n | d_tvec[s][n] (3 floats) for s, n ]

of 1 + 6 S N floats in which a rank fills only its own items' slots (others zero); disparity
gradients stay sharded.

On the GPU the packed buffer is not assembled by copies: `PackedGrads.out()` hands the fused loss
views INTO the rank's slots of a zeroed send buffer (`mdb_packed_bind` in the C ABI), so the
finalize kernel writes the pose gradients and the loss where the collective reads them, and the
step is the five loss kernels + one out-of-place all-reduce.  The collective itself is
`mdb_allreduce_loss_posegrad` -- NCCL resolved with dlopen, or the library's own one-shot
peer-write kernel over NVLink (`init_comm(transport=...)`); `torch.distributed` only carries the
bootstrap (unique id / IPC handles) and is the transport of the CPU (gloo) tests.
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


class PackedGrads:
    """The rank's send buffer (zero outside its own slots, for good: the all-reduce is out of place) and the
    receive buffer every rank ends up with.  `out()` is the `out=` argument of `mdb200.FusedLoss`."""

    def __init__(self, n_global: int, S: int, lo: int, n_local: int, device):
        self.N, self.S, self.lo, self.n = int(n_global), int(S), int(lo), int(n_local)
        size = packed_size(self.N, self.S)
        self.send = torch.zeros(size, dtype=torch.float32, device=device)
        self.recv = torch.zeros(size, dtype=torch.float32, device=device)

    def out(self) -> dict:
        N, S, lo, n = self.N, self.S, self.lo, self.n
        rv = self.send[1:1 + 3 * S * N].view(S, N, 3)
        tv = self.send[1 + 3 * S * N:].view(S, N, 3)
        return dict(loss=self.send[0:1], d_rvec=[rv[s, lo:lo + n] for s in range(S)],
                    d_tvec=[tv[s, lo:lo + n].view(n, 1, 3) for s in range(S)])

    def result(self):
        """-> loss (0-dim), d_rvec [S x (N,3)], d_tvec [S x (N,1,3)] of the GLOBAL batch (after allreduce())."""
        return unpack(self.recv, self.N, self.S)


_comm = {"transport": None}


def init_comm(device, max_count: int, transport: str = "peer") -> str:
    """Set up the C-ABI communicator of this rank's handle; `torch.distributed` (already initialised) carries
    the bootstrap.  transport "peer": the one-shot peer-write reduce (falls back to "nccl" when the peers'
    buffers cannot be mapped through CUDA IPC); "nccl": ncclAllReduce through the library (falls back to "torch" when
    the communicator cannot be created); "torch": torch.distributed.all_reduce.  Returns the transport in use."""
    import ctypes as ct
    from . import api, _lib
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = torch.device(device)
    h = api.handle(dev.index)
    if world == 1 or transport == "torch":
        _comm["transport"] = "torch"
        return "torch"
    if transport == "peer":
        mine = (ct.c_char * 64)()
        rc = _lib.lib.mdb_comm_peers_export(h, world, int(max_count), mine, 64)
        ok = torch.tensor([1 if rc == 0 else 0], device=dev)
        handles = [None] * world
        dist.all_gather_object(handles, bytes(mine))
        if rc == 0:
            blob = b"".join(handles)
            rc = _lib.lib.mdb_comm_peers_attach(h, ct.c_char_p(blob), rank)
            ok[0] = 1 if rc == 0 else 0
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)   # also the barrier the transport needs before its first use
        if int(ok) == 1:
            _comm["transport"] = "peer"
            return "peer"
        _lib.lib.mdb_comm_destroy(h)
        transport = "nccl"
    uid = (ct.c_char * 128)()
    if rank == 0:
        _lib.check(h, _lib.lib.mdb_comm_get_unique_id(uid, 128))
    box = [bytes(uid)]
    dist.broadcast_object_list(box, src=0)
    rc = _lib.lib.mdb_comm_init(h, ct.c_char_p(box[0]), rank, world)
    ok = torch.tensor([1 if rc == 0 else 0], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)          # every rank must take the same transport
    if int(ok) != 1:                                    # e.g. no libnccl to dlopen: torch.distributed carries the reduction
        _lib.lib.mdb_comm_destroy(h)
        _comm["transport"] = "torch"
        return "torch"
    _comm["transport"] = "nccl"
    return "nccl"


def allreduce(pg: PackedGrads, group=None) -> torch.Tensor:
    """recv = sum over ranks of send, enqueued on the current stream."""
    import ctypes as ct
    if _comm["transport"] in ("peer", "nccl") and pg.send.is_cuda:
        from . import api, _lib
        h = api.handle(pg.send.device.index)
        _lib.check(h, _lib.lib.mdb_allreduce_loss_posegrad(h, ct.c_void_p(pg.send.data_ptr()), ct.c_void_p(pg.recv.data_ptr()),
                                                         pg.send.numel(), api._stream()))
        return pg.recv
    pg.recv.copy_(pg.send)
    return allreduce_loss_posegrads(pg.recv, group)


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
