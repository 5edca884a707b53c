#!/usr/bin/env python
"""Throughput of the projection kernel on BASELINE configs 3 and 4 (GPU box only).

  config 3  dimension sweep d in {256, 512, 1024, 2048, 4096} on 1024 samples x 100 000 hashes
  config 4  heavy-tailed batch: log-normal sample sizes clip(exp(N(ln 2e4, 1.5^2)), 1e3, 1e7),
            d = 2048 (--c4-samples, default 100 000 = the full config; ~6e9 hashes, ~48 GB in HBM)

Device-resident timing (CUDA events around mvs_project_device, as bench.py's `value`), one JSON
object per line.  These are parity-test configurations, not bench lines; the numbers go to
profiles/ and DESIGN.md to show that small d / ragged sizes have no performance cliff.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import metagenome_vector_sketches_b200 as mvs                      # noqa: E402
from metagenome_vector_sketches_b200 import synthetic as syn       # noqa: E402


def work_per_hash(d):
    return d + 20 * ((d + 63) // 64)


def time_device(sk, torch, dh, off, d, steps=5, warmup=2):
    n = len(off) - 1
    dout = torch.empty((n, d), dtype=torch.int32, device=dh.device)
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(warmup):
        sk.project_device(dh.data_ptr(), off, d, dout.data_ptr(), stream=stream)
    torch.cuda.synchronize()
    sk.timing()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        sk.project_device(dh.data_ptr(), off, d, dout.data_ptr(), stream=stream)
    e1.record()
    torch.cuda.synchronize()
    t = sk.timing()
    return e0.elapsed_time(e1) / steps, t, dout


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--c3-samples", type=int, default=1024)
    ap.add_argument("--c4-samples", type=int, default=100000)
    ap.add_argument("--skip-c4", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    sk = mvs.Sketcher(0)
    p_int = sk.int_issue_peak(2)

    # ---- config 3 ------------------------------------------------------------------------
    per = 100000
    dh, off = syn.uniform_batch_torch(args.c3_samples, per, dev, seed=0)
    total = args.c3_samples * per
    for d in (256, 512, 1024, 2048, 4096):
        ms, t, _ = time_device(sk, torch, dh, off, d)
        hps = total / (ms * 1e-3)
        print(json.dumps({"config": 3, "d": d, "samples": args.c3_samples, "hashes_per_sample": per,
                          "ms_per_step": ms, "kernel_ms": t["kernel_ms"], "hashes_per_s": hps,
                          "entries_per_s": hps * d, "warp_tasks": t["warp_tasks"],
                          "roofline_frac": hps * work_per_hash(d) / p_int}), flush=True)
    del dh
    torch.cuda.empty_cache()

    # ---- config 4 ------------------------------------------------------------------------
    if not args.skip_c4:
        sizes = syn.lognormal_sizes(args.c4_samples, seed=0)
        dh, off = syn.ragged_batch_torch(sizes, dev, seed=0)
        total = int(off[-1])
        d = 2048
        ms, t, dout = time_device(sk, torch, dh, off, d, steps=3, warmup=1)
        hps = total / (ms * 1e-3)
        # size-independent property at full size: column 0..d of every row has the parity of |H_s|
        par = (dout[:, :8] & 1).to(torch.int64).cpu().numpy()
        ok = bool((par == (sizes[:, None] & 1)).all())
        print(json.dumps({"config": 4, "d": d, "samples": args.c4_samples, "total_hashes": total,
                          "min_size": int(sizes.min()), "max_size": int(sizes.max()),
                          "median_size": float(np.median(sizes)), "ms_per_step": ms,
                          "kernel_ms": t["kernel_ms"], "reduce_ms": t["reduce_ms"], "hashes_per_s": hps,
                          "samples_per_s": args.c4_samples / (ms * 1e-3), "warp_tasks": t["warp_tasks"],
                          "partial_rows": t["partial_rows"], "parity_property_ok": ok,
                          "roofline_frac": hps * work_per_hash(d) / p_int}), flush=True)


if __name__ == "__main__":
    main()
