#!/usr/bin/env python
"""Kernel time of config 2 / config 3 (d=256) / a slice of config 4 versus the per-task round cap (experiment)."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import metagenome_vector_sketches_b200 as mvs
from metagenome_vector_sketches_b200 import synthetic as syn
from sweep import time_device

dev = torch.device("cuda", 0)
sk = mvs.Sketcher(0, experiments=True)       # knobs of the experiment build: make -C metagenome_vector_sketches_b200/csrc experiments
cases = []
dh, off = syn.uniform_batch_torch(10000, 50000, dev, seed=0)
cases.append(("c2", dh, off, 2048))
dh3, off3 = syn.uniform_batch_torch(1024, 100000, dev, seed=0)
cases.append(("c3d256", dh3, off3, 256))
cases.append(("c3d2048", dh3, off3, 2048))
sizes = syn.lognormal_sizes(20000, seed=0)
dh4, off4 = syn.ragged_batch_torch(sizes, dev, seed=0)
cases.append(("c4_20k", dh4, off4, 2048))
for name, h, o, d in cases:
    for cap in (0, 1024, 2048, 3072, 4096, 6144, 8192, 12288, 16368):
        sk.set_option("task_cap", cap)
        ms, t, _ = time_device(sk, torch, h, o, d, steps=4, warmup=1)
        print(json.dumps({"case": name, "cap": cap, "ms": ms, "kernel_ms": t["kernel_ms"], "reduce_ms": t["reduce_ms"],
                          "tasks": t["warp_tasks"]}), flush=True)
