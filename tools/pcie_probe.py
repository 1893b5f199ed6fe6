#!/usr/bin/env python
"""Host <-> device bandwidth table for the end-to-end leg of bench.py: raw pinned-memory copies, k = 1..N devices at once, from one
process (what the library's host driver does) -- names the link that caps `e2e` when several GPUs are driven from one host.
    python tools/pcie_probe.py [--gpus N] [--mb 1024]      -> one JSON line per k"""
import argparse, json, time
import torch

ap = argparse.ArgumentParser(); ap.add_argument("--gpus", type=int, default=0); ap.add_argument("--mb", type=int, default=1024)
a = ap.parse_args()
N = a.gpus or torch.cuda.device_count()
nbytes = a.mb << 20
host_in = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(N)]
host_out = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(N)]
dev_a = [torch.empty(nbytes, dtype=torch.uint8, device="cuda:%d" % d) for d in range(N)]
dev_b = [torch.empty(nbytes, dtype=torch.uint8, device="cuda:%d" % d) for d in range(N)]
s_in = [torch.cuda.Stream(device=d) for d in range(N)]
s_out = [torch.cuda.Stream(device=d) for d in range(N)]


def run(k, h2d, d2h, reps=4):
    def once():
        for d in range(k):
            if h2d:
                with torch.cuda.stream(s_in[d]):
                    dev_a[d].copy_(host_in[d], non_blocking=True)
            if d2h:
                with torch.cuda.stream(s_out[d]):
                    host_out[d].copy_(dev_b[d], non_blocking=True)
        for d in range(k):
            s_in[d].synchronize(); s_out[d].synchronize()
    once()
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    dt = (time.perf_counter() - t0) / reps
    return k * nbytes * (int(h2d) + int(d2h)) / dt / 1e9


for k in [n for n in (1, 2, 4, 8) if n <= N]:
    print(json.dumps({"devices": k, "h2d_GBs": round(run(k, True, False), 1), "d2h_GBs": round(run(k, False, True), 1),
                      "bidirectional_GBs": round(run(k, True, True), 1), "buffer_MB_per_device": a.mb}), flush=True)
