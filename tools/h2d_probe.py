"""GPU box: pinned host -> device copy rate for the e2e input size (context for bench.py's e2e line)."""
import time
import torch
n = 279_714_356
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda:0")
for chunks in (1, 2, 4):
    best = 1e9
    for _ in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        step = (n + chunks - 1) // chunks
        for c in range(chunks):
            d[c * step:(c + 1) * step].copy_(h[c * step:(c + 1) * step], non_blocking=True)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    print("chunks %d: %.3f ms  %.1f GB/s" % (chunks, best * 1e3, n / best / 1e9))
