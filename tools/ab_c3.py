#!/usr/bin/env python
"""A/B of projection-kernel builds on BASELINE configs[2] with the planner's own task caps (MVS_EXP_LIB selects the
experiment library): kernel time at d = 64..1024.  One JSON line per d."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import metagenome_vector_sketches_b200 as mvs
from metagenome_vector_sketches_b200 import synthetic as syn
from sweep import time_device

dev = torch.device("cuda", 0)
sk = mvs.Sketcher(0, experiments=True)
tag = os.path.basename(os.environ.get("MVS_EXP_LIB", "default"))
dh, off = syn.uniform_batch_torch(1024, 100000, dev, seed=0)
for d in (64, 128, 256, 512, 1024):
    best = None
    for rep in range(3):
        ms, t, _ = time_device(sk, torch, dh, off, d, steps=6, warmup=2)
        best = t["kernel_ms"] if best is None else min(best, t["kernel_ms"])
    print(json.dumps({"lib": tag, "d": d, "kernel_ms": round(best, 4), "tasks": t["warp_tasks"]}), flush=True)
