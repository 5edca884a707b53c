#!/usr/bin/env python
"""A/B of projection-kernel builds at FIXED task caps (experiment builds; MVS_EXP_LIB selects the library):
kernel time of config 2 and of the config-3 shape at d = 2048 / 256.  One JSON line per (case, cap)."""
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
dh, off = syn.uniform_batch_torch(10000, 50000, dev, seed=0)
dh3, off3 = syn.uniform_batch_torch(1024, 100000, dev, seed=0)
for name, h, o, d, caps in (("c2", dh, off, 2048, (4096, 8192, 16320)), ("c3d2048", dh3, off3, 2048, (4096, 8192)),
                            ("c3d256", dh3, off3, 256, (2048, 4096, 8128))):
    for cap in caps:
        sk.set_option("task_cap", cap)
        ms, t, _ = time_device(sk, torch, h, o, d, steps=4, warmup=1)
        print(json.dumps({"lib": tag, "case": name, "cap": cap, "ms": round(ms, 4), "kernel_ms": round(t["kernel_ms"], 4),
                          "reduce_ms": round(t["reduce_ms"], 4), "tasks": t["warp_tasks"]}), flush=True)
