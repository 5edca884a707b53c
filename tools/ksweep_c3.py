#!/usr/bin/env python
"""Kernel time of BASELINE configs[2] (1024 x 100 000 hashes) at every d versus the number of chunks k a sample is split
into (experiment build, option task_cap), next to the planner's own choice (cap 0).  How launches of about one wave of
CTAs should be cut: DESIGN section 4.   python tools/ksweep_c3.py > gpurun_out/ksweep_c3.jsonl"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import metagenome_vector_sketches_b200 as mvs
from metagenome_vector_sketches_b200 import synthetic as syn
from sweep import time_device

dev = torch.device("cuda", 0)
sk = mvs.Sketcher(0, experiments=True)
n, per = 1024, 100000
dh, off = syn.uniform_batch_torch(n, per, dev, seed=0)
for d in (256, 512, 1024, 2048, 4096):
    nb = (d + 63) // 64
    S = max(1, 32 // nb) if nb < 32 else 1
    while S & (S - 1):
        S -= 1                                    # lanes per hash is a power of two
    rows = []
    for k in [0] + list(range(1, 17)):
        cap = 0 if k == 0 else ((per + k * S - 1) // (k * S) + 63) // 64 * 64
        sk.set_option("task_cap", cap)
        ms, t, _ = time_device(sk, torch, dh, off, d, steps=6, warmup=2)
        rows.append({"d": d, "k": k, "cap_rounds": cap, "tasks": int(t["warp_tasks"]), "kernel_ms": t["kernel_ms"],
                     "reduce_ms": t["reduce_ms"], "ms": ms})
        print(json.dumps(rows[-1]), flush=True)
sk.set_option("task_cap", 0)
