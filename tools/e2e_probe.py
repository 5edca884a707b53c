#!/usr/bin/env python
"""End-to-end (pinned host buffers) time of config 2 through the u64 and the packed56 entry points, for a few slice
sizes: where does the host link's time go?  One JSON line per case."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import metagenome_vector_sketches_b200 as mvs
from metagenome_vector_sketches_b200 import synthetic as syn

dev = torch.device("cuda", 0)
n, per, d = 10000, 50000, 2048
sk = mvs.Sketcher(0)
dh, off = syn.uniform_batch_torch(n, per, dev, seed=0)
hh = torch.empty(n * per, dtype=torch.int64, pin_memory=True)
hh.copy_(dh)
del dh
h_np = hh.numpy().view(np.uint64)
hout = torch.empty((n, d), dtype=torch.int32, pin_memory=True)
out_np = hout.numpy()
pk = torch.empty(7 * n * per, dtype=torch.uint8, pin_memory=True)
pk_np = pk.numpy()
mvs.pack56(h_np, out=pk_np)
for slice_hashes in (16 << 20, 8 << 20, 32 << 20, 64 << 20):
    sk.set_option("slice_hashes", slice_hashes)
    sk.set_option("slice_rows_bytes", 1 << 30)
    for name, fn in (("u64", lambda: sk.project_csr(h_np, off, d, out=out_np)),
                     ("packed56", lambda: sk.project_csr_packed56(pk_np, off, d, out=out_np))):
        fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3 / 3
        t = sk.timing()
        print(json.dumps({"path": name, "slice_hashes": slice_hashes, "ms_per_step": round(ms, 2),
                          "h2d_GBps": round(t["h2d_bytes"] / ms / 1e6, 2), "checksum": int(out_np[:64].sum())}), flush=True)
