"""GPU box: does a small device -> host readback on one stream wait for a large device -> host copy
queued on another stream?  Measures the latency of `.item()` on the current stream while a 238 MB copy
runs on a side stream, for several chunk sizes of the large copy.
usage: python tools/d2h_interleave.py"""
import json
import time

import torch

dev = torch.device("cuda", 0)
n = 59_556_379
src = torch.zeros(n, dtype=torch.int32, device=dev)
dst = torch.empty(n, dtype=torch.int32, pin_memory=True)
small = torch.ones(1, dtype=torch.int64, device=dev)
side = torch.cuda.Stream(device=dev)
for chunk_mb in (0, 64, 16, 4, 1):
    lat, tot = [], []
    for rep in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        with torch.cuda.stream(side):
            if chunk_mb == 0:
                dst.copy_(src, non_blocking=True)
            else:
                step = chunk_mb * (1 << 20) // 4
                for s in range(0, n, step):
                    dst[s:s + step].copy_(src[s:s + step], non_blocking=True)
        t1 = time.perf_counter()
        small.item()                       # cudaMemcpyAsync + synchronize on the current stream
        t2 = time.perf_counter()
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        lat.append((t2 - t1) * 1e3)
        tot.append((t3 - t0) * 1e3)
    print(json.dumps({"chunk_mb": chunk_mb, "queue_ms": round((t1 - t0) * 1e3, 3), "item_latency_ms_min": round(min(lat), 3),
                      "item_latency_ms_max": round(max(lat), 3), "whole_copy_ms_min": round(min(tot), 3)}))
