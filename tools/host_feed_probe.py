"""is the 8-GPU end-to-end step bound by the host side of the VM?  Every rank copies its shard of config 5's
inputs (866 MB / world) host->device at the same time, four ways: torch-pinned vs cudaHostRegister-ed memory that was
allocated AFTER the rank bound itself to the GPU's NUMA node, each on one and on two copy streams.  Prints the aggregate
GB/s over all ranks.  torchrun --nproc-per-node N tools/host_feed_probe.py"""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "monodepth2.jl_b200"))
import numpy as np, torch, torch.distributed as dist
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
dev = torch.device("cuda", torch.cuda.current_device())
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
from mdb200 import dist as mdist
info = mdist.bind_host_to_device(dev)
nbytes = 866388992 // world // 4096 * 4096
d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
h_pin = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
h_pin.fill_(1)
raw = np.empty(nbytes + 4096, dtype=np.uint8)          # allocated and first-touched after the affinity call
off = (-raw.ctypes.data) % 4096
reg = raw[off:off + nbytes]
reg[:] = 1
cudart = ctypes.CDLL("libcudart.so.12")
rc = cudart.cudaHostRegister(ctypes.c_void_p(reg.ctypes.data), ctypes.c_size_t(nbytes), 0)
h_reg = torch.from_numpy(reg) if rc == 0 else None
streams = [torch.cuda.Stream(), torch.cuda.Stream()]

def run(h, nstreams, reps=5):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        step = nbytes // nstreams
        for k in range(nstreams):
            with torch.cuda.stream(streams[k]):
                d[k * step:(k + 1) * step].copy_(h[k * step:(k + 1) * step], non_blocking=True)
        torch.cuda.synchronize()
        t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t))
    return best

rows = []
for name, h in (("torch pin_memory", h_pin), ("cudaHostRegister after sched_setaffinity", h_reg)):
    if h is None:
        rows.append((name, None, None)); continue
    for ns in (1, 2):
        t = run(h, ns)
        rows.append((name, ns, nbytes * world / t / 1e9))
if rank == 0:
    print(f"world {world}, {nbytes >> 20} MiB per rank, affinity: {info}")
    for name, ns, gbs in rows:
        print(f"  {name:42s} {ns} copy stream(s): {gbs:7.1f} GB/s aggregate ({gbs / world:5.1f} per GPU)" if gbs else f"  {name}: cudaHostRegister failed rc={rc}")
if rc == 0:
    cudart.cudaHostUnregister(ctypes.c_void_p(reg.ctypes.data))
if world > 1:
    dist.destroy_process_group()
