"""CUDA-event timing of the HBM-bound side kernels against their algorithmic bytes (L2 flushed between calls)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monodepth2.jl_b200")):
    sys.path.insert(0, p)
import torch, numpy as np
import mdb200
dev = torch.device("cuda:0")
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6550.0) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6550.0
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, reps=10):
    ts = []
    for _ in range(reps + 3):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts[3:]))
W, H, N = 1024, 320, 64
strips = torch.randint(0, 256, (N, H, 3 * W, 3), dtype=torch.uint8, device=dev)
out = torch.empty((N, 3, 3, H, W), device=dev)
for gray in (False, True):
    o = out if not gray else torch.empty((N, 3, 1, H, W), device=dev)
    ms = timeit(lambda: mdb200.ingest_triplet_u8(strips, W, None, grayscale=gray, out=o))
    b = strips.numel() + o.numel() * 4
    print(f"ingest_triplet_u8 gray={gray}: {ms*1e3:.1f} us, {b/1e6:.0f} MB, {b/ms/1e6:.0f} GB/s = {b/ms/1e6/peak*100:.1f} % of {peak:.0f}")
x = out
auto = None
ms = timeit(lambda: mdb200.automasking_loss(mdb200.SSIM(), x, x[:, 1], source_ids=[1, 3]))
b = x.numel() * 4 + N * H * W * 4
print(f"automasking_loss: {ms*1e3:.1f} us, {b/1e6:.0f} MB, {b/ms/1e6:.0f} GB/s = {b/ms/1e6/peak*100:.1f} %")
m = torch.rand((N, 1, H, W), device=dev)
ms = timeit(lambda: mdb200.item_mean(m))
print(f"item_mean: {ms*1e3:.1f} us, {m.numel()*4/ms/1e6:.0f} GB/s = {m.numel()*4/ms/1e6/peak*100:.1f} %")
d = torch.rand((N, H, W), device=dev); img = x[:, 1].contiguous()
ms = timeit(lambda: mdb200.smooth_loss(d, img, with_grad=True))
b = (d.numel() * 2 + img.numel()) * 4
print(f"smooth_loss+grad: {ms*1e3:.1f} us, {b/1e6:.0f} MB, {b/ms/1e6:.0f} GB/s = {b/ms/1e6/peak*100:.1f} %")
