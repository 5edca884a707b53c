#!/usr/bin/env python
"""BASELINE config 4 on 1/2/4/8 GPUs: the heavy-tailed batch (100 000 samples, log-normal sizes 1e3..1e7, d = 2048)
sharded over the ranks by the library's static size-balanced partition (mvs_partition_samples) -- STRONG
scaling, no data-path collective.  Each rank generates its own contiguous shard on its device, projects it
device-resident (CUDA events around mvs_project_device, as bench.py's `value`) and the job's step time is the slowest
rank's.  One JSON line from rank 0.  A parity-test configuration, not a bench line.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \\
        tools/sweep_c4_ranks.py [--samples 100000]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import metagenome_vector_sketches_b200 as mvs                      # noqa: E402
from metagenome_vector_sketches_b200 import sharding, synthetic as syn   # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=100000)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    d = 2048
    sizes = syn.lognormal_sizes(args.samples, seed=0)               # every rank computes the same sizes
    offsets = np.zeros(args.samples + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum(sizes).astype(np.uint64)
    bounds = sharding.partition_bounds(offsets, world)
    a, e = int(bounds[rank]), int(bounds[rank + 1])
    dh, off = syn.ragged_batch_torch(sizes[a:e], dev, seed=0, first_sample=a)
    mine = int(off[-1])
    sk = mvs.Sketcher(local)
    dout = torch.empty((e - a, d), dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    sk.project_device(dh.data_ptr(), off, d, dout.data_ptr(), stream=stream)
    torch.cuda.synchronize()
    sk.timing()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        sk.project_device(dh.data_ptr(), off, d, dout.data_ptr(), stream=stream)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    # size-independent property on every row of the shard: each column has the parity of |H_s|
    ok = bool(((dout[:, :8] & 1).to(torch.int64).cpu().numpy() == (sizes[a:e, None] & 1)).all())
    t = torch.tensor([ms, float(mine), float(e - a), 1.0 if ok else 0.0], dtype=torch.float64, device=dev)
    allt = [t.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(allt, t)
    if rank == 0:
        per = [[float(x[0]), int(x[1]), int(x[2])] for x in allt]
        step = max(p[0] for p in per)
        total = int(offsets[-1])
        print(json.dumps({"config": 4, "d": d, "n_gpus": world, "scaling": "strong", "samples": args.samples,
                          "total_hashes": total, "ms_per_step": step, "hashes_per_s": total / (step * 1e-3),
                          "samples_per_s": args.samples / (step * 1e-3),
                          "per_rank_ms_hashes_samples": per,
                          "hash_imbalance_max_over_mean": max(p[1] for p in per) / (total / world),
                          "parity_property_ok": all(bool(x[3]) for x in allt)}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
