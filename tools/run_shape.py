#!/usr/bin/env python
"""Device-resident projection of one uniform synthetic shape, a few launches -- the target command for ncu captures of
kernels other than the bench default (e.g. project_kernel<2> at d = 256 of BASELINE configs[2]).

    ncu --set full --clock-control none --import-source on -k regex:project_kernel -s 3 -c 1 -o gpurun_out/x \\
        python tools/run_shape.py --samples 1024 --per 100000 --d 256 --steps 5
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import metagenome_vector_sketches_b200 as mvs                      # noqa: E402
from metagenome_vector_sketches_b200 import synthetic as syn       # noqa: E402


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=1024)
    ap.add_argument("--per", type=int, default=100000)
    ap.add_argument("--d", type=int, default=256)
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    sk = mvs.Sketcher(0)
    dh, off = syn.uniform_batch_torch(a.samples, a.per, dev, seed=0)
    dout = torch.empty((a.samples, a.d), dtype=torch.int32, device=dev)
    ts = torch.cuda.Stream(device=dev)           # stream 0 would make every call synchronous (the context's own stream)
    torch.cuda.set_stream(ts)
    stream = ts.cuda_stream
    for _ in range(3):
        sk.project_device(dh.data_ptr(), off, a.d, dout.data_ptr(), stream=stream)
    torch.cuda.synchronize()
    sk.timing()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        sk.project_device(dh.data_ptr(), off, a.d, dout.data_ptr(), stream=stream)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    t = sk.timing()
    hps = a.samples * a.per / (ms * 1e-3)
    print(json.dumps({"samples": a.samples, "per": a.per, "d": a.d, "ms_per_step": ms, "kernel_ms": t["kernel_ms"],
                      "reduce_ms": t["reduce_ms"], "warp_tasks": int(t["warp_tasks"]), "partial_rows": int(t["partial_rows"]),
                      "hashes_per_s": hps, "entries_per_s": hps * a.d}))


if __name__ == "__main__":
    main()
