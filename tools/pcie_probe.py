"""PCIe ceiling for the host-buffer step: one large pinned H2D copy alone, and with a concurrent D2H."""
import time, torch
dev = torch.device("cuda:0")
n_in, n_out = 866388992, 111414304
h_in = torch.empty(n_in, dtype=torch.uint8).pin_memory(); d_in = torch.empty(n_in, dtype=torch.uint8, device=dev)
h_out = torch.empty(n_out, dtype=torch.uint8).pin_memory(); d_out = torch.empty(n_out, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(both, chunks=1, reps=5):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        with torch.cuda.stream(s1):
            step = n_in // chunks
            for c in range(chunks):
                d_in[c * step:(c + 1) * step].copy_(h_in[c * step:(c + 1) * step], non_blocking=True)
        if both:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    return best
for both in (False, True):
    for chunks in (1, 8, 72):
        t = run(both, chunks)
        print(f"H2D 866 MB in {chunks} copies{' + concurrent D2H 111 MB' if both else ''}: {t*1e3:.2f} ms = {n_in/t/1e9:.1f} GB/s")
