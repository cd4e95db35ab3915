"""Streaming write / copy bandwidth of the box's GPU with torch ops (CUDA events), for the store floor of the corpus:
    python tools/bw_probe.py"""
import json

import torch

dev = torch.device("cuda:0")
out = {}
for gb in (2, 16):
    a = torch.empty(gb << 30, dtype=torch.uint8, device=dev)
    b = torch.empty(gb << 30, dtype=torch.uint8, device=dev)
    for name, fn, nbytes in (("fill", lambda: a.fill_(3), gb << 30), ("zero", lambda: a.zero_(), gb << 30), ("copy", lambda: b.copy_(a), 2 * (gb << 30))):
        best = 1e9
        for _ in range(6):
            s, t = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); t.record(); torch.cuda.synchronize()
            best = min(best, s.elapsed_time(t))
        out[f"{name}_{gb}GiB_GBs"] = round(nbytes / best / 1e6, 1)
    del a, b
print(json.dumps(out))
