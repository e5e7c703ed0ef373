#!/usr/bin/env python
"""bench.py -- warped px/s of the view-synthesis photometric loss, forward + backward.

Metric (BASELINE.json): warped px/s (loss fwd+bwd).  One *warped pixel* = one (item, scale,
source, full-res target pixel); one step = one `train_loss` value+gradient evaluation
(src/training.jl:21-78) over one batch: N * Ls * S * W * H warped pixels.

Workload at every GPU count: BASELINE config 5 as stated -- 1024x320, GLOBAL batch 64, 3-frame
stacks, 2 sources, 4 scales -- sharded over the ranks ("scaling": "strong": rank r owns the
contiguous items shard_range(64, world, r) of ONE 64-item input, 32 / 16 / 8 items per GPU at
2 / 4 / 8 GPUs), one all-reduce of [loss, pose gradients] per step.  Before timing, rank 0
evaluates the whole batch on its own GPU and the all-reduced result is checked against it
("shard_parity").  `weak_aux` repeats the run with 64 items PER GPU (round 1's line).
Synthetic images / disparities / poses from mdb200.synth (no dataset or checkpoint is
reachable).  At one GPU the step's inputs are larger than L2 (866 MB vs 126 MB); from 8 items
per GPU down they are not, and a buffer larger than L2 is written between the timed steps
OUTSIDE the event pairs (per-step events, `config.l2`).

  value     device-resident inputs, CUDA-event timed on the launching stream, max over ranks
  e2e       the same call made through mdb200.train_loss_host with PINNED HOST buffers: every
            step copies the inputs H2D and the loss + all gradients D2H inside the timed region
  roofline  the dominant kernel (photo_strip_kernel) timed live with CUDA events recorded
            inside the C-ABI call (mdb_profile_*), against the algorithmic bytes of SURVEY 8d
  cpu_baseline / --impl reference
            the reference's algorithm on the host cores: the plain-C float32 port
            (oracle/c_port/mdb_oracle.c, hand-written pullbacks), one item per thread on all
            cores, on a bounded sample (one item per core).
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "monodepth2.jl_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "warped px/s (loss fwd+bwd)"
UNIT = "warped_px/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="config5", choices=["config1", "config2", "config3", "config5", "config4"])
    ap.add_argument("--render-batch", type=int, default=2, help="config4: render batch B (reference scripts use 2 and 4)")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch override (default: the config's)")
    ap.add_argument("--cpu-sample-items", type=int, default=0, help="items in the CPU sample (0 = one per host core)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--disp-noise", type=float, default=None,
                    help="diagnostic: amplitude of the per-pixel disparity noise of the synthetic inputs (default 0.05 "
                         "as SURVEY 8d specifies; 0 = smooth disparities, what a network predicts)")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong (default): the config's global batch sharded over the ranks; weak: the config's batch per GPU")
    ap.add_argument("--transport", default="peer", choices=["peer", "nccl", "torch"],
                    help="the all-reduce of [loss, pose grads]: the library's one-shot peer-write kernel, ncclAllReduce "
                         "through the C ABI, or torch.distributed")
    ap.add_argument("--no-weak-aux", action="store_true")
    ap.add_argument("--no-aux-configs", action="store_true", help="skip the compact records of configs 1-4 (N = 1 only)")
    ap.add_argument("--e2e-chunk", type=int, default=0, help="items per pipeline chunk of the host-buffer call (0 = HostLoss.default_chunk: N/8, at least ~16 MB)")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def algorithmic_bytes(inp, N):
    """every distinct input byte once + every output byte once, fp32:
    4 N [ W H C (1+S) + 2 sum_k w_k h_k + (automask ? W H : 0) ]."""
    S = len(inp.source_ids)
    per_item = inp.W * inp.H * inp.C * (1 + S) + 2 * sum(d.shape[-1] * d.shape[-2] for d in inp.disparities)
    if inp.automasking:
        per_item += inp.W * inp.H
    return 4 * N * per_item


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md's clocks line).  NVML -- what
    nvidia-smi itself reads -- polled from a thread of this process: the first sample lands milliseconds after
    start() (an `nvidia-smi -lms` child needs over a second to come up on an 8-GPU box, longer than a short
    timed region), and sample_now() takes one more from the main thread while the timed steps are queued."""
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, device):
        self.h, self.nv, self.samples, self.thread, self.run = None, None, [], None, False
        try:
            import torch
            import pynvml as nv
            nv.nvmlInit()
            self.nv = nv
            try:
                uuid = str(torch.cuda.get_device_properties(device).uuid)
                self.h = nv.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                idx = device.index or 0
                self.h = nv.nvmlDeviceGetHandleByIndex(int(vis.split(",")[idx]) if vis else idx)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
            self.bits = [nv.nvmlClocksEventReasonHwSlowdown, nv.nvmlClocksEventReasonHwThermalSlowdown,
                         nv.nvmlClocksEventReasonSwThermalSlowdown, nv.nvmlClocksEventReasonSwPowerCap]
        except Exception as e:   # no NVML: say so in the line rather than fail the bench
            self.h, self.err = None, f"{type(e).__name__}: {e}"

    def sample_now(self):
        if self.h is None:
            return
        nv = self.nv
        try:
            mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
            self.samples.append((mhz, mask))
        except Exception:
            pass

    def _loop(self):
        while self.run:
            self.sample_now()
            time.sleep(0.005)

    def start(self):
        if self.h is None:
            return
        import threading
        self.samples, self.run = [], True
        self.thread = threading.Thread(target=self._loop, daemon=True)
        self.thread.start()

    def stop(self):
        if self.h is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [f"NVML unavailable ({self.err})"]}
        self.run = False
        self.thread.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["no samples"]}
        sm = sorted(m for m, _ in self.samples)
        reasons = sorted({nm for _, mask in self.samples for nm, bit in zip(self.NAMES, self.bits) if mask & bit})
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(sm),
                "source": "NVML, polled during the timed region"}


def load_synth():
    """mdb200/synth.py (numpy only) loaded by path: the reference arm must not import the mdb200 package, whose
    __init__ dlopens the product library."""
    import importlib.util
    if "mdb200_synth" in sys.modules:
        return sys.modules["mdb200_synth"]
    spec = importlib.util.spec_from_file_location("mdb200_synth", os.path.join(ROOT, "monodepth2.jl_b200", "mdb200", "synth.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["mdb200_synth"] = mod
    spec.loader.exec_module(mod)
    return mod


def make_inputs(args):
    synth = load_synth()
    kw = dict(synth.CONFIGS[args.config])
    if args.batch:
        kw["N"] = args.batch
    if getattr(args, "disp_noise", None) is not None:   # diagnostic only: the bench line's workload is SURVEY 8d's (noise 0.05)
        kw["disp_noise"] = args.disp_noise
    return synth.make_loss_inputs(**kw), kw


def cpu_reference_rate(inp, items, steps, warmup):
    """The reference algorithm on the host cores: the plain-C float32 port (oracle/c_port/mdb_oracle.c,
    unfused, op by op in the reference's order, hand-written pullbacks), one item per thread on all
    cores.  `items` <= 0 means one item per core.  Returns (warped px/s, threads, seconds per step, items)."""
    import numpy as np
    from oracle import c_ref
    cores = os.cpu_count() or 1
    items = min(items if items > 0 else cores, inp.N)
    threads = min(cores, items)
    a = (inp.x[:items], [d[:items] for d in inp.disparities], [r[:items] for r in inp.rvecs],
         [t[:items] for t in inp.tvecs], None, inp.K, inp.invK)
    kw = dict(target_id=inp.target_id, source_ids=inp.source_ids, scales=inp.scales, automasking=False,
              dtype=np.float32, threads=threads)
    units = items * len(inp.disparities) * len(inp.source_ids) * inp.W * inp.H
    for _ in range(warmup):
        c_ref.train_loss(*a, **kw)
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        c_ref.train_loss(*a, **kw)
        ts.append(time.perf_counter() - t0)
    t = sum(ts) / len(ts)
    # one item on one thread (the closest stand-in for the reference's single-threaded CPU broadcasts)
    a1 = (inp.x[:1], [d[:1] for d in inp.disparities], [r[:1] for r in inp.rvecs], [t_[:1] for t_ in inp.tvecs], None, inp.K, inp.invK)
    kw1 = dict(kw); kw1["threads"] = 1
    t0 = time.perf_counter()
    c_ref.train_loss(*a1, **kw1)
    cpu_reference_rate.one_thread = (units / items) / (time.perf_counter() - t0)
    return units / t, threads, t, items


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    ncpu = os.cpu_count() or 1
    want = args.cpu_sample_items if args.cpu_sample_items > 0 else ncpu
    inp, _ = make_inputs(argparse.Namespace(config=args.config, batch=min(args.batch or want, want),
                                            disp_noise=getattr(args, "disp_noise", None)))
    steps, warm = max(1, min(args.steps, 20)), max(1, min(args.warmup, 5))
    rate, cores, t, items = cpu_reference_rate(inp, args.cpu_sample_items, steps, warm)
    sample = (f"{items} of the per-GPU batch's items at {inp.W}x{inp.H}, {len(inp.disparities)} scales, "
              f"{steps} steps after {warm} warm-up (plain-C float32 port of the Julia path, oracle/c_port, one item per thread, "
              f"{cores} threads)")
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.config}: {inp.W}x{inp.H}, CPU sample of {items} items", "timing": "host wall clock"},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "one_thread_value": cpu_reference_rate.one_thread},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import mdb200
    from mdb200 import dist as mdist

    rank, local_rank, world = dist_env()
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback for the product path)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    # host threads (and, by first touch, the pinned staging buffers below) onto the GPU's own NUMA node
    numa = mdist.bind_host_to_device(dev)

    inp, kw = make_inputs(args)                       # the config's WHOLE batch, identical on every rank (seeded)
    NB, W, H, C = inp.N, inp.W, inp.H, inp.C
    S, Ls = len(inp.source_ids), len(inp.disparities)
    strong = args.scaling == "strong"
    Ng = NB if strong else NB * world                 # batch the means divide by
    lo, hi = mdist.shard_range(NB, world, rank) if strong else (0, NB)
    first = lo if strong else rank * NB               # this rank's first slot in the packed vector
    N = hi - lo
    units_total = Ng * Ls * S * W * H
    sizes = [(d.shape[-1], d.shape[-2]) for d in inp.disparities]

    # pinned host copies of the rank's shard (e2e) and device-resident copies (value)
    hx = torch.from_numpy(inp.x[lo:hi]).pin_memory()
    hd = [torch.from_numpy(d[lo:hi]).pin_memory() for d in inp.disparities]
    hr = [torch.from_numpy(r[lo:hi]).pin_memory() for r in inp.rvecs]
    ht = [torch.from_numpy(t[lo:hi]).pin_memory() for t in inp.tvecs]
    x = hx.to(dev); disps = [d.to(dev) for d in hd]
    poses = [mdb200.Pose(r.to(dev), t.to(dev)) for r, t in zip(hr, ht)]
    auto = mdb200.automasking_loss(mdb200.SSIM(), x, x[:, inp.target_id - 1], source_ids=inp.source_ids) if inp.automasking else None
    hauto = auto.cpu().pin_memory() if auto is not None else None

    common = dict(W=W, H=H, C=C, L=3, disp_sizes=sizes, scales=inp.scales, target_id=inp.target_id,
                  source_ids=inp.source_ids, K=inp.K, invK=inp.invK, automasking=inp.automasking)
    # [loss, d_rvec (S,Ng,3), d_tvec (S,Ng,3)]: the finalize kernel writes this rank's slots of the send buffer in
    # place (mdb_packed_bind); one out-of-place all-reduce per step
    pg = mdist.PackedGrads(Ng, S, first, N, dev)
    plan = mdb200.FusedLoss(N=N, n_global=Ng, device=dev, out=pg.out(), **common)
    # the peer buffers are sized for the larger of the two packed vectors this run reduces (weak_aux: 64 items per rank)
    transport = mdist.init_comm(dev, mdist.packed_size(NB * world, S), args.transport) if world > 1 else "none"

    mid_events = []     # per step: recorded between the loss kernels and the all-reduce (what part of a step is exchange + waiting for the slowest rank)

    def step():
        res = plan(x, disps, poses, auto)
        if world > 1:
            if mid_events is not None and len(mid_events) < 4096:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                mid_events.append(ev)
            mdist.allreduce(pg)
        return res

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- shard parity: the all-reduced [loss, pose grads] of the sharded batch == the whole batch on one GPU ----
    shard_parity = None
    if world > 1 and strong:
        step()
        torch.cuda.synchronize()
        got = pg.recv.clone()
        worst = torch.zeros(3, device=dev, dtype=torch.float64)
        fx = torch.from_numpy(inp.x).to(dev)
        fd = [torch.from_numpy(d).to(dev) for d in inp.disparities]
        fposes = [mdb200.Pose(torch.from_numpy(r).to(dev), torch.from_numpy(t).to(dev)) for r, t in zip(inp.rvecs, inp.tvecs)]
        fauto = mdb200.automasking_loss(mdb200.SSIM(), fx, fx[:, inp.target_id - 1], source_ids=inp.source_ids) if inp.automasking else None
        fpg = mdist.PackedGrads(Ng, S, 0, NB, dev)
        fplan = mdb200.FusedLoss(N=NB, n_global=Ng, device=dev, out=fpg.out(), **common)
        fres = fplan(fx, fd, fposes, fauto)
        torch.cuda.synchronize()
        ref = fpg.send.double()
        g = got.double()
        nr = 1 + 3 * S * Ng
        worst[0] = (g[0] - ref[0]).abs() / ref[0].abs()
        worst[1] = (g[1:nr] - ref[1:nr]).abs().max() / ref[1:nr].abs().max()
        worst[2] = (g[nr:] - ref[nr:]).abs().max() / ref[nr:].abs().max()
        # the shard's disparity gradients (which stay sharded) against the same rows of the whole-batch run
        dd = max(float((a.double() - b[lo:hi].double()).abs().max() / b.double().abs().max())
                 for a, b in zip(plan.d_disp, fres.d_disparities))
        worst = torch.cat([worst, torch.tensor([dd], device=dev, dtype=torch.float64)])
        dist.all_reduce(worst, op=dist.ReduceOp.MAX)
        shard_parity = {"ok": bool(float(worst.max()) <= 1e-6), "tol": 1e-6,
                        "loss_rel": float(worst[0]), "d_rvec_rel_maxabs": float(worst[1]), "d_tvec_rel_maxabs": float(worst[2]),
                        "d_disparity_rel_maxabs": float(worst[3]),
                        "how": "every rank: all-reduced [loss, pose grads] and its own shard's disparity gradients vs the whole "
                               f"{NB}-item batch evaluated on that rank's GPU alone; max over ranks"}
        assert shard_parity["ok"], f"sharded result differs from the single-GPU result: {shard_parity}"
        del fx, fd, fposes, fauto, fplan, fres, fpg
        torch.cuda.empty_cache()

    # ---- value: device-resident inputs, one CUDA-event pair per step on the launching stream --------------------
    in_bytes = x.numel() * 4 + sum(d.numel() for d in disps) * 4
    l2_bytes = torch.cuda.get_device_properties(dev).L2_cache_size
    flush = torch.empty(2 * l2_bytes, dtype=torch.uint8, device=dev) if in_bytes < 2 * l2_bytes else None

    def timed(step_fn, steps, warmup, profile=False):
        for _ in range(warmup):
            step_fn()
        sync_all()
        sampler = ClockSampler(dev)
        if profile:
            mdb200.profile_begin(steps)
        l0 = mdb200.launch_count()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        sampler.start()
        sync_all()
        del mid_events[:]
        t0 = time.perf_counter()
        for e0, e1 in evs:
            if flush is not None:
                flush.zero_()          # outside the event pair
            e0.record()
            r = step_fn()
            e1.record()
        sampler.sample_now()   # the timed steps are queued or running: at least one sample under load
        sync_all()
        wall = time.perf_counter() - t0
        clocks = sampler.stop()
        launches = mdb200.launch_count() - l0
        ms = torch.tensor([sum(e0.elapsed_time(e1) for e0, e1 in evs)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        seg = None
        if profile:
            seg = np.array([mdb200.profile_read(i) for i in range(steps)])  # [steps][4] ms
            mdb200.profile_begin(0)
        timed.split = None
        if world > 1 and len(mid_events) == steps:
            comp = sum(e0.elapsed_time(m) for (e0, _), m in zip(evs, mid_events)) / steps
            exch = sum(m.elapsed_time(e1) for (_, e1), m in zip(evs, mid_events)) / steps
            both = torch.tensor([comp, exch], device=dev, dtype=torch.float64)
            allr = [torch.zeros_like(both) for _ in range(world)]
            dist.all_gather(allr, both)
            timed.split = {"loss_kernels_ms_per_rank": [float(t[0]) for t in allr],
                           "allreduce_incl_wait_for_slowest_rank_ms_per_rank": [float(t[1]) for t in allr]}
        return float(ms), clocks, launches, seg, r, wall

    warm = max(args.warmup, 3)
    ms_total, clocks, launches, seg, res, wall = timed(step, args.steps, warm, profile=True)
    step_split = timed.split
    loss_val = float(pg.recv[0]) if world > 1 else float(res.loss)

    # ---- weak_aux: 64 items per GPU (round 1's line), same kernels, global batch 64 x world -------------------
    weak_aux = None
    if world > 1 and strong and not args.no_weak_aux:
        wx = torch.from_numpy(inp.x).to(dev)
        wd = [torch.from_numpy(d).to(dev) for d in inp.disparities]
        wposes = [mdb200.Pose(torch.from_numpy(r).to(dev), torch.from_numpy(t).to(dev)) for r, t in zip(inp.rvecs, inp.tvecs)]
        wauto = mdb200.automasking_loss(mdb200.SSIM(), wx, wx[:, inp.target_id - 1], source_ids=inp.source_ids) if inp.automasking else None
        wpg = mdist.PackedGrads(NB * world, S, rank * NB, NB, dev)
        wplan = mdb200.FusedLoss(N=NB, n_global=NB * world, device=dev, out=wpg.out(), **common)
        keep_flush, flush = flush, None     # 866 MB of inputs per rank: larger than L2

        def wstep():
            r = wplan(wx, wd, wposes, wauto)
            mdist.allreduce(wpg)
            return r
        wms, _, _, _, _, _ = timed(wstep, args.steps, warm)
        flush = keep_flush
        weak_aux = {"scaling": "weak", "batch_per_gpu": NB, "global_batch": NB * world, "ms_per_step": wms / args.steps,
                    "value": NB * world * Ls * S * W * H * args.steps / (wms * 1e-3), "unit": UNIT}
        del wx, wd, wposes, wauto, wplan, wpg
        torch.cuda.empty_cache()

    # ---- e2e: host buffers, copies inside the timed region ---------------------------------
    e2e = None
    if not args.no_e2e:
        e_chunk = args.e2e_chunk if args.e2e_chunk > 0 else None   # None: HostLoss.default_chunk
        hplan = mdb200.HostLoss(N=N, n_global=Ng, device=dev, chunk=e_chunk, packed=pg if world > 1 else None, **common)
        hposes = [mdb200.Pose(r, t) for r, t in zip(hr, ht)]
        e_steps = max(3, min(args.steps, 10))
        for _ in range(3):
            mdb200.train_loss_host(hplan, hx, hd, hposes, hauto)
        sync_all()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            hres = mdb200.train_loss_host(hplan, hx, hd, hposes, hauto)   # returns after the D2H of the (all-reduced) result
        sync_all()
        dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        h2d = torch.tensor([hplan.h2d_bytes, hplan.d2h_bytes], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(h2d)
        e2e = {"value": units_total * e_steps / float(dt), "unit": UNIT,
               "h2d_bytes_per_step": int(h2d[0]), "d2h_bytes_per_step": int(h2d[1]),
               "ms_per_step": float(dt) / e_steps * 1e3, "steps": e_steps, "host_numa_binding": numa,
               "loss_matches_device_path": bool(abs(hres.loss - loss_val) <= 1e-5 * abs(loss_val))}
        # the same call fed the dataset's own format (src/dtk.jl:29-46): uint8 triplet strips, converted on the
        # device (mdb_ingest_triplet_u8) -- a quarter of the image bytes over PCIe.  Auxiliary: the headline
        # `e2e` stays the Float32 call above, which is what the reference's `transfer(x)` ships.
        if not inp.automasking and inp.C == 3:
            del hplan
            torch.cuda.empty_cache()
            synth = load_synth()
            hu8 = torch.tensor(synth.to_u8_strips(inp.x[lo:hi])).pin_memory()
            uplan = mdb200.HostLoss(N=N, n_global=Ng, device=dev, chunk=e_chunk, x_format="u8_strip",
                                    packed=pg if world > 1 else None, **common)
            for _ in range(3):
                mdb200.train_loss_host(uplan, hu8, hd, hposes, None)
            sync_all()
            t0 = time.perf_counter()
            for _ in range(e_steps):
                mdb200.train_loss_host(uplan, hu8, hd, hposes, None)
            sync_all()
            dtu = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(dtu, op=dist.ReduceOp.MAX)
            hu = torch.tensor([uplan.h2d_bytes, uplan.d2h_bytes], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(hu)
            e2e["u8_strip_input"] = {"value": units_total * e_steps / float(dtu), "unit": UNIT,
                                     "h2d_bytes_per_step": int(hu[0]), "d2h_bytes_per_step": int(hu[1]),
                                     "ms_per_step": float(dtu) / e_steps * 1e3,
                                     "note": "images as uint8 triplet strips (8-bit quantised copies of the same synthetic frames), "
                                             "disparities / poses / gradients Float32 as above; on the device the strips are "
                                             "transposed to a planar uint8 stack that the fused kernel reads directly "
                                             "(mdb_loss_io.x_format = 1: no Float32 stack is ever written)"}
            # the same kernel with the uint8 stack resident on the device: what reading bytes costs / saves in the kernel itself
            if world == 1:
                del uplan
                torch.cuda.empty_cache()
                xu8 = mdb200.ingest_triplet_u8_planar(hu8.to(dev), W)
                for _ in range(3):
                    plan(xu8, disps, poses, None)
                mdb200.profile_begin(e_steps)
                evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(e_steps)]
                for a_, b_ in evs:
                    a_.record(); plan(xu8, disps, poses, None); b_.record()
                torch.cuda.synchronize()
                fused = sum(mdb200.profile_read(i)[1] for i in range(e_steps)) / e_steps
                mdb200.profile_begin(0)
                e2e["u8_strip_input"]["device_resident"] = {
                    "ms_per_step": sum(a_.elapsed_time(b_) for a_, b_ in evs) / e_steps, "fused_kernel_ms": fused,
                    "image_bytes": int(xu8.numel()), "image_bytes_float32": int(x.numel() * 4)}
                del xu8

    if rank == 0:
        line = {
            "metric": METRIC, "value": units_total * args.steps / (ms_total * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.config}: {W}x{H} global batch {Ng} ({N} items on this rank), 3-frame stacks, {S} sources, "
                                   f"{Ls} scales, SSIM+L1 min-reprojection{' + automask' if inp.automasking else ''} "
                                   f"+ smoothness, fwd+bwd", "global_batch": Ng, "batch_per_gpu": N,
                       "l2": "inputs_larger_than_l2" if flush is None else
                             f"flushed: {flush.numel() >> 20} MiB written between the timed steps, outside the per-step event pairs",
                       "timing": "sum of one CUDA-event pair per step on the launching stream, max over ranks",
                       "parallelism": f"batch-sharded x{world}, 1 all-reduce of [loss, pose grads] per step "
                                      f"(transport: {transport})",
                       "units_per_step": units_total, "loss": loss_val, "wall_s_timed_region": wall,
                       **({"disp_noise": args.disp_noise, "diagnostic": "not SURVEY 8d's inputs"}
                          if args.disp_noise is not None else {})},
            "clocks": clocks,
            "e2e": e2e,
            "gpu_launches": int(launches),
            "roofline": roofline_block(args, inp, N, seg, plan),
        }
        if step_split is not None:
            line["step_split"] = step_split
        if shard_parity is not None:
            line["shard_parity"] = shard_parity["ok"]
            line["shard_parity_detail"] = shard_parity
        if weak_aux is not None:
            line["weak_aux"] = weak_aux
        if not args.no_cpu_baseline and world == 1:
            rate, cores, t, items = cpu_reference_rate(inp, args.cpu_sample_items, 10, 2)
            line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                                    "one_thread_value": cpu_reference_rate.one_thread,
                                    "sample": f"{items} of the batch's {NB} items, 10 steps after 2 warm-up, plain-C float32 port "
                                              f"(oracle/c_port), one item per thread, {cores} threads ({t:.2f} s/step)"}
        if world == 1 and not args.no_aux_configs and args.config == "config5" and not args.batch:
            del plan, x, disps
            torch.cuda.empty_cache()
            line["aux_configs"] = aux_configs(args)
        print(json.dumps(line), flush=True)
    if world > 1:
        sync_all()
        mdb200._lib.lib.mdb_comm_destroy(mdb200.handle())
        dist.destroy_process_group()


def _blob_hash(rel):
    import hashlib
    d = open(os.path.join(ROOT, rel), "rb").read()
    return hashlib.sha1(b"blob %d\0" % len(d) + d).hexdigest()


def _profile_record(name):
    """A record under profiles/ (SASS histogram, ncu DRAM traffic) counts only for the kernel source it was captured
    on: every record carries the git blob hash of the files it depends on; a stale record reads as None."""
    path = os.path.join(ROOT, "profiles", name)
    try:
        rec = json.load(open(path))
        for rel, want in rec["source_hash"].items():
            if _blob_hash(rel) != want:
                return None
        return rec
    except Exception:
        return None


def roofline_block(args, inp, N, seg, plan):
    """The dominant kernel (photo_strip_kernel) of THIS rank's launch: timed live with CUDA events recorded inside the
    C-ABI call (mdb_profile_*) on the launching stream.  The fused loss is bound by FP32 issue, not by HBM (SURVEY
    F10, 8d "Honest roofline"): T_roof = max(B_alg / BW_hbm, F_alg / P_fp32) with P_fp32 MEASURED in this run
    (mdb_measure_fp32_peak) and F_alg counted from the committed SASS histogram of the row loop."""
    import mdb200
    peak, peak_src = load_peaks()
    S, Ls = len(inp.source_ids), len(inp.disparities)
    units = N * Ls * S * inp.W * inp.H
    b_alg = algorithmic_bytes(inp, N)
    k_ms = float(seg[:, 1].mean())
    hbm_gbs = b_alg / (k_ms * 1e-3) / 1e9
    p_fp32, p_ms = mdb200.measure_fp32_peak(5)
    geo = plan.strip_geometry()
    sass = _profile_record("r2_strip_sass.json")
    ncu = _profile_record("r2_strip_ncu.json")
    traffic = None
    if ncu is not None and ncu.get("config") == args.config and ncu.get("items") == N:
        traffic = ncu["dram_bytes_read"] + ncu["dram_bytes_write"]
    blk = {"kernel": "photo_strip_kernel<3,2>", "kernel_ms": k_ms, "traffic": traffic,
           "segments_ms": {"prep_stats": float(seg[:, 0].mean()), "fused": k_ms,
                           "finalize": float(seg[:, 2].mean()), "upsample_bwd": float(seg[:, 3].mean())},
           "strip_geometry": geo,
           "hbm": {"achieved": hbm_gbs, "peak": peak, "unit": "GB/s", "frac": hbm_gbs / peak, "peak_source": peak_src,
                   "algorithmic_bytes_per_launch": b_alg, "T_mem_ms": b_alg / (peak * 1e9) * 1e3},
           "P_fp32_tflops": p_fp32, "P_fp32_source": f"measured in this run: independent FFMA chains on every SM, best of 5 ({p_ms:.3f} ms)"}
    t_mem = b_alg / (peak * 1e9) * 1e3
    if sass is not None:
        # executed fp32 flops per warped pixel: per lane and row from the SASS histogram, 32 lanes per 28-pixel strip
        # (ragged last strip included), rows + 4 halo rows per segment, two sources per lane
        lanes_per_px = 32.0 * geo["strips"] / inp.W
        rows_f = (inp.H + 4.0 * geo["segments"]) / inp.H      # 4 halo rows per row segment (segments may differ in length)
        f_wpx = sass["fp32_flops_per_lane_row"] * lanes_per_px * rows_f / S
        t_fp32 = units * f_wpx / (p_fp32 * 1e12) * 1e3
        t_roof = max(t_mem, t_fp32)
        blk.update({"bound": "fp32" if t_fp32 >= t_mem else "hbm", "F_alg_flop_per_warped_px": f_wpx,
                    "F_alg_source": "profiles/r2_strip_sass.json (static SASS histogram of the row loop, source hash checked): "
                                    f"{sass['fp32_flops_per_lane_row']} flop per lane and row x {lanes_per_px:.3f} lanes per pixel "
                                    f"x {rows_f:.3f} (halo rows) / {S} sources",
                    "T_fp32_ms": t_fp32, "T_mem_ms": t_mem,
                    "achieved": units * f_wpx / (k_ms * 1e-3) / 1e12, "peak": p_fp32, "unit": "TFLOP/s", "frac": t_roof / k_ms})
    else:   # the kernel source changed since the SASS record was made: no flop count to stand on
        blk.update({"bound": "hbm", "achieved": hbm_gbs, "peak": peak, "unit": "GB/s", "frac": hbm_gbs / peak,
                    "F_alg_flop_per_warped_px": None, "note": "profiles/r2_strip_sass.json is stale for this kernel source"})
    return blk


def run_render(args, emit=True):
    """BASELINE config 4: render_tgt_rgb_depth forward + backward (src/render.jl:92-114), 32 planes at
    416x128.  Not the headline line (bench.py's default stays config 5); prints its own JSON line with
    the HBM roofline of the two render kernels.  The working set (110 MB at B=2) fits L2, so a buffer
    larger than L2 is written between the timed calls."""
    import ctypes as ct
    import torch
    import mdb200
    from mdb200 import synth, api
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    ri = synth.make_render_inputs(B=args.render_batch)
    W, H, Np, B = ri.W, ri.H, ri.Np, ri.B
    t = lambda a: torch.tensor(a, device=dev)
    rgb, sigma, disp, xyz, rvec, tvec = t(ri.rgb), t(ri.sigma), t(ri.disparity_src), t(ri.xyz_tgt), t(ri.rvec), t(ri.tvec)
    g_rgb = torch.rand((B, 3, H, W), device=dev)
    g_dep = torch.rand((B, Np, 1, H, W), device=dev) * 0.1
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    pose = mdb200.Pose(rvec, tvec)

    def time_plan(plan, xyz_in, g_z=None):
        f_ms, b_ms, l0 = [], [], None
        for it in range(args.warmup + args.steps):
            if it == args.warmup:
                l0 = mdb200.launch_count()
            flush.zero_()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            e[0].record()
            plan.forward(rgb, sigma, disp, xyz_in, pose)
            e[1].record()
            flush.zero_()
            e[2].record()
            plan.backward(g_rgb, g_dep, g_z)
            e[3].record()
            torch.cuda.synchronize()
            if it >= args.warmup:
                f_ms.append(e[0].elapsed_time(e[1])); b_ms.append(e[2].elapsed_time(e[3]))
        return sum(f_ms) / len(f_ms), sum(b_ms) / len(b_ms), mdb200.launch_count() - l0

    fm, bm, launches = time_plan(mdb200.RenderPlan(W, H, Np, B, ri.K_inv, ri.K, device=dev), xyz)
    # SURVEY 8(f) rank 2 variants, same inputs: (a) as written but xyz_tgt never materialised; (b) the sampler MINE
    # intends (samples land inside the image instead of on the border), its mask, the expected depth, no xyz_tgt
    g_z = torch.rand((B, 1, H, W), device=dev) * 0.01
    fa, ba, _ = time_plan(mdb200.RenderPlan(W, H, Np, B, ri.K_inv, ri.K, device=dev, fused_xyz=True), None)
    fb, bb, _ = time_plan(mdb200.RenderPlan(W, H, Np, B, ri.K_inv, ri.K, device=dev, fused_xyz=True, mine_grid=True,
                                            mine_mask=True, expected_depth=True), None, g_z)
    fc, bc, _ = time_plan(mdb200.RenderPlan(W, H, Np, B, ri.K_inv, ri.K, device=dev, mine_grid=True, mine_mask=True), xyz)
    units = W * H * Np * B
    inp = 7 * units * 4
    outb = (3 * W * H * B + units) * 4
    peak, peak_src = load_peaks()
    fwd_b, bwd_b = inp + outb, 2 * inp + outb
    variants = {
        "fused_xyz (as written, xyz_tgt not an input: 4/7 of the input bytes)": {"fwd_ms": fa, "bwd_ms": ba},
        "mine_grid + mine_mask (xyz_tgt given)": {"fwd_ms": fc, "bwd_ms": bc},
        "mine_grid + mine_mask + expected_depth + fused_xyz": {"fwd_ms": fb, "bwd_ms": bb}}
    line = {"metric": "plane-px/s (render_tgt_rgb_depth fwd+bwd)", "value": units / ((fm + bm) * 1e-3), "unit": "plane_px/s",
            "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": fm + bm, "higher_is_better": True,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"config4: {W}x{H}, {Np} planes, B={B}, render_tgt_rgb_depth fwd+bwd", "l2": "flushed between calls",
                       "timing": "CUDA events around the pre-planned C-ABI calls (mdb200.RenderPlan: buffers allocated once)"},
            "gpu_launches": int(launches), "variants": variants,
            "roofline": {"bound": "hbm", "kernel": "render_fwd_kernel + render_bwd_kernel", "unit": "GB/s", "peak": peak, "peak_source": peak_src,
                         "note": "ncu (profiles/r2_render_ncu.json): the forward issues 259 warp-instructions per warp-sample and reads 25 MB "
                                 "of DRAM at B=2 (the reference's grid normalisation clamps 3/4 of the samples onto the border): "
                                 "issue- and latency-bound; the pullback (767 per sample, one thread per pixel, 32 planes in series) "
                                 "is bound by that serial chain at 33 % occupancy",
                         "fwd": {"ms": fm, "algorithmic_bytes": fwd_b, "achieved": fwd_b / (fm * 1e-3) / 1e9, "frac": fwd_b / (fm * 1e-3) / 1e9 / peak},
                         "bwd": {"ms": bm, "algorithmic_bytes": bwd_b, "achieved": bwd_b / (bm * 1e-3) / 1e9, "frac": bwd_b / (bm * 1e-3) / 1e9 / peak},
                         "achieved": (fwd_b + bwd_b) / ((fm + bm) * 1e-3) / 1e9, "frac": (fwd_b + bwd_b) / ((fm + bm) * 1e-3) / 1e9 / peak}}
    if not args.no_cpu_baseline:
        from oracle import c_ref
        import numpy as np
        cores = os.cpu_count() or 1
        gr, gd = g_rgb.cpu().numpy(), g_dep.cpu().numpy()
        call = lambda: c_ref.render_tgt_rgb_depth(ri.rgb, ri.sigma, ri.disparity_src, ri.xyz_tgt, ri.rvec, ri.tvec, ri.K_inv, ri.K,
                                                  g_rgb=gr, g_depth=gd, dtype=np.float32, threads=cores)
        call()
        t0 = time.perf_counter()
        for _ in range(3):
            call()
        t_cpu = (time.perf_counter() - t0) / 3
        line["cpu_baseline"] = {"value": units / t_cpu, "unit": "plane_px/s", "cores": cores, "kind": "port",
                                "sample": f"the whole config (B={B}), 3 calls after 1 warm-up of the plain-C float32 port "
                                          f"(oracle/c_port: forward + hand-written pullback, rows / planes over {cores} threads, "
                                          f"{t_cpu * 1e3:.0f} ms per call)"}
    if emit:
        print(json.dumps(line), flush=True)
    return line


def run_slow_depth(args, emit=True):
    """BASELINE config 1: the `slow_depth` loop (src/simple_depth.jl:1-62) at 416x128, one triplet:
    500 iterations of {loss fwd+bwd, ADAM(3e-4)} replayed from one CUDA graph.  Prints its own JSON line
    (not the headline; the default stays config 5).  Compared with (a) the same kernels launched one by
    one from the host mirror (what a Julia caller without the graph would do) and (b) the plain-C CPU
    port doing the same iteration on one host core."""
    import numpy as np
    import torch
    import mdb200
    from mdb200 import synth
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    inp = synth.make_config("config1")
    x = torch.tensor(inp.x, device=dev)
    iters = 500
    units = inp.W * inp.H * 2                    # warped px per iteration (1 item, 1 scale, 2 sources)
    ms = []
    l0 = None
    for it in range(args.warmup + args.steps):
        sd = mdb200.SlowDepth(x, inp.K, inp.invK, max_iters=iters)
        if it == args.warmup:
            l0 = mdb200.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sd.run(iters)
        e1.record()
        torch.cuda.synchronize()
        if it >= args.warmup:
            ms.append(e0.elapsed_time(e1))
        hist = sd.loss_history.cpu().numpy()
        sd.close()
    launches = mdb200.launch_count() - l0
    t_graph = sum(ms) / len(ms)
    # (a) per-kernel launches from the host mirror, no graph
    plan = mdb200.simple_depth_plan(x, inp.K, inp.invK)
    disp = torch.full((1, 1, inp.H, inp.W), 0.5, device=dev)
    poses = [mdb200.Pose(torch.tensor([[0.0, 0.0, 0.01]], device=dev), torch.zeros((1, 1, 3), device=dev)) for _ in range(2)]
    opt = mdb200.ADAM(3e-4)
    params = [disp] + [t for pz in poses for t in (pz.rvec, pz.tvec)]
    def eager(n):
        for _ in range(n):
            r = plan(x, [disp], poses, None)
            opt.update(params, [r.d_disparities[0]] + [t for pz in r.d_poses for t in (pz.rvec, pz.tvec)])
    eager(20)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); eager(iters); torch.cuda.synchronize(); t_eager = (time.perf_counter() - t0) * 1e3
    line = {"metric": "slow_depth iterations/s (loss fwd+bwd + ADAM)", "value": iters / (t_graph * 1e-3), "unit": "iter/s",
            "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_graph, "higher_is_better": True,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"config1: slow_depth, {inp.W}x{inp.H}, 1 triplet, {iters} iterations per step",
                       "timing": "CUDA events around mdb_slow_depth_run (500 graph replays)", "l2": "working set 2 MB: L2-resident by design"},
            "us_per_iteration": t_graph * 1e3 / iters, "warped_px_per_s": units * iters / (t_graph * 1e-3),
            "eager_ms_per_500": t_eager, "eager_us_per_iteration": t_eager * 1e3 / iters,
            "loss_first_last": [float(hist[0]), float(hist[-1])], "gpu_launches": int(launches)}
    if not args.no_cpu_baseline:
        from oracle import c_ref
        d = np.full((1, 1, inp.H, inp.W), 0.5, dtype=np.float32)
        rv = [np.array([[0, 0, 0.01]], dtype=np.float32) for _ in range(2)]
        tv = [np.zeros((1, 1, 3), dtype=np.float32) for _ in range(2)]
        n_cpu = 10
        t0 = time.perf_counter()
        for _ in range(n_cpu):          # the loss + gradient dominates; the ADAM arithmetic on 53 K numbers is noise
            c_ref.train_loss(inp.x, [d], rv, tv, None, inp.K, inp.invK, scales=(1.0,), automasking=False, normalize=False,
                             smooth_mode=1, dtype=np.float32, threads=1)
        t_cpu = (time.perf_counter() - t0) / n_cpu
        line["cpu_baseline"] = {"value": 1.0 / t_cpu, "unit": "iter/s", "cores": 1, "kind": "port",
                                "sample": f"{n_cpu} loss+gradient evaluations of the plain-C float32 port on one core "
                                          f"({t_cpu * 1e3:.1f} ms each; one item cannot be spread over threads)"}
    if emit:
        print(json.dumps(line), flush=True)
    return line


def run_loss_aux(name, steps=20, warmup=5, cpu=True):
    """One of the other loss configurations of BASELINE.json (config 2: 416x128 batch 12; config 3: 640x192 batch 32
    with automasking) on one GPU, device-resident, L2 flushed between the timed steps: a compact record for the
    `aux_configs` key of the headline line."""
    import numpy as np
    import torch
    import mdb200
    synth = load_synth()
    dev = torch.device("cuda", torch.cuda.current_device())
    inp = synth.make_config(name)
    N, W, H = inp.N, inp.W, inp.H
    S, Ls = len(inp.source_ids), len(inp.disparities)
    x = torch.from_numpy(inp.x).to(dev)
    disps = [torch.from_numpy(d).to(dev) for d in inp.disparities]
    poses = [mdb200.Pose(torch.from_numpy(r).to(dev), torch.from_numpy(t).to(dev)) for r, t in zip(inp.rvecs, inp.tvecs)]
    auto = mdb200.automasking_loss(mdb200.SSIM(), x, x[:, inp.target_id - 1], source_ids=inp.source_ids) if inp.automasking else None
    plan = mdb200.FusedLoss(W=W, H=H, C=inp.C, L=3, N=N, disp_sizes=[(d.shape[-1], d.shape[-2]) for d in disps], scales=inp.scales,
                            target_id=inp.target_id, source_ids=inp.source_ids, K=inp.K, invK=inp.invK, automasking=inp.automasking,
                            device=dev)
    flush = torch.empty(2 * torch.cuda.get_device_properties(dev).L2_cache_size, dtype=torch.uint8, device=dev)
    for _ in range(warmup):
        plan(x, disps, poses, auto)
    mdb200.profile_begin(steps)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize()
    for e0, e1 in evs:
        flush.zero_()
        e0.record()
        res = plan(x, disps, poses, auto)
        e1.record()
    torch.cuda.synchronize()
    ms = sum(e0.elapsed_time(e1) for e0, e1 in evs) / steps
    k_ms = float(np.mean([mdb200.profile_read(i)[1] for i in range(steps)]))
    mdb200.profile_begin(0)
    units = N * Ls * S * W * H
    peak, _ = load_peaks()
    b_alg = algorithmic_bytes(inp, N)
    rec = {"workload": f"{name}: {W}x{H} batch {N}, {Ls} scales, {S} sources{', automask' if inp.automasking else ''}, fwd+bwd",
           "ms_per_step": ms, "value": units / (ms * 1e-3), "unit": UNIT, "fused_kernel_ms": k_ms, "loss": float(res.loss),
           "hbm_frac_of_step": b_alg / (ms * 1e-3) / 1e9 / peak, "algorithmic_bytes": b_alg, "l2": "flushed between steps"}
    if cpu:
        auto_np = auto.cpu().numpy() if auto is not None else None
        from oracle import c_ref
        cores = os.cpu_count() or 1
        a = (inp.x, inp.disparities, inp.rvecs, inp.tvecs, auto_np, inp.K, inp.invK)
        kw = dict(target_id=inp.target_id, source_ids=inp.source_ids, scales=inp.scales, automasking=inp.automasking,
                  dtype=np.float32, threads=min(cores, N))
        c_ref.train_loss(*a, **kw)
        t0 = time.perf_counter()
        for _ in range(3):
            c_ref.train_loss(*a, **kw)
        t = (time.perf_counter() - t0) / 3
        rec["cpu_port"] = {"value": units / t, "unit": UNIT, "cores": min(cores, N), "ms_per_step": t * 1e3,
                           "sample": "the whole batch, 3 steps after 1 warm-up, plain-C float32 port (oracle/c_port)"}
    return rec


def aux_configs(args):
    """Configs 1-4 of BASELINE.json beside the headline (config 5), so that the driver's record carries them."""
    import argparse as ap
    a2 = ap.Namespace(**vars(args))
    a2.steps, a2.warmup = min(args.steps, 10), 3
    out = {}
    c1 = run_slow_depth(a2, emit=False)
    out["config1"] = {k: c1[k] for k in ("metric", "value", "unit", "ms_per_step", "us_per_iteration", "warped_px_per_s",
                                         "eager_us_per_iteration", "loss_first_last", "cpu_baseline") if k in c1}
    out["config1"]["workload"] = c1["config"]["workload"]
    out["config2"] = run_loss_aux("config2", cpu=not args.no_cpu_baseline)
    out["config3"] = run_loss_aux("config3", cpu=not args.no_cpu_baseline)
    c4 = run_render(a2, emit=False)
    out["config4"] = {"workload": c4["config"]["workload"], "metric": c4["metric"], "value": c4["value"], "unit": c4["unit"],
                      "ms_per_step": c4["ms_per_step"], "fwd": c4["roofline"]["fwd"], "bwd": c4["roofline"]["bwd"],
                      "variants": c4["variants"], "cpu_baseline": c4.get("cpu_baseline")}
    return out


if __name__ == "__main__":
    a = parse()
    if a.config == "config1" and a.impl != "reference":
        run_slow_depth(a)
    elif a.config == "config4" and a.impl != "reference":
        run_render(a)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
