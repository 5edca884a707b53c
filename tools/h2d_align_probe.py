#!/usr/bin/env python
"""Does the alignment / size of the host source matter for cudaMemcpyAsync H2D?  (diagnosis of the packed56 path)"""
import json, time, torch
dev = torch.device("cuda", 0)
host = torch.empty(4 << 30, dtype=torch.uint8, pin_memory=True)
host.fill_(3)
dst = torch.empty(160 << 20, dtype=torch.uint8, device=dev)
s = torch.cuda.Stream()
def run(chunk, stride, start, n=28):
    with torch.cuda.stream(s):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(2):
            dst[:chunk].copy_(host[start:start + chunk], non_blocking=True)
        e0.record(s)
        for i in range(n):
            o = start + i * stride
            dst[:chunk].copy_(host[o:o + chunk], non_blocking=True)
        e1.record(s)
        s.synchronize()
    ms = e0.elapsed_time(e1)
    return chunk * n / (ms * 1e-3) / 1e9
out = {}
out["u64 slices 134000000 B, 128-B aligned"] = run(134000000, 134000000, 0)
out["packed slices 117250000 B, 16-B aligned starts"] = run(117250000, 117250000, 0)
out["packed slices 117250000 B, starts +1 B"] = run(117250000, 117250000, 1)
out["117250048 B (64-B multiple), 64-B aligned starts"] = run(117250048, 117250048, 0)
out["117440512 B (2^20 multiple), 4096-aligned"] = run(117440512, 117440512, 0)
print(json.dumps(out, indent=1))
