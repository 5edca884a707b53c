#!/usr/bin/env python
"""Host-to-device bandwidth of every GPU alone, of pairs, of each half of the box and of all GPUs at once, with the
pinned source buffers first-touched from different vCPUs -- the evidence behind DESIGN.md section 6 (where the 8-GPU
end-to-end number is lost: the host memory path, not the devices).  One process drives all GPUs, one stream per GPU,
CUDA events per GPU, wall clock for the aggregate.  Plain cudaMemcpyAsync (torch copy_), no batch-copy API.

    python tools/h2d_matrix.py [--mb 1024] [--reps 6] > profiles/r02_h2d_matrix_8gpu.json
"""
import argparse
import json
import os
import threading
import time

import torch


def first_touch(t, cpu):
    """Write the pinned buffer from a thread bound to `cpu` (first touch decides the page's NUMA node, if any)."""
    def work():
        try:
            os.sched_setaffinity(0, {cpu})
        except OSError:
            pass
        t.fill_(1)
    th = threading.Thread(target=work)
    th.start()
    th.join()


def measure(gpus, host, dev, streams, reps):
    for g in gpus:                                                      # warm-up
        with torch.cuda.stream(streams[g]):
            dev[g].copy_(host[g], non_blocking=True)
    for g in gpus:
        streams[g].synchronize()
    ev = {g: (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for g in gpus}
    t0 = time.perf_counter()
    for g in gpus:
        with torch.cuda.stream(streams[g]):
            ev[g][0].record(streams[g])
            for _ in range(reps):
                dev[g].copy_(host[g], non_blocking=True)
            ev[g][1].record(streams[g])
    for g in gpus:
        streams[g].synchronize()
    wall = time.perf_counter() - t0
    nbytes = host[gpus[0]].numel() * host[gpus[0]].element_size() * reps
    per = {str(g): nbytes / (ev[g][0].elapsed_time(ev[g][1]) * 1e-3) / 1e9 for g in gpus}
    return {"gpus": gpus, "per_gpu_GBps": per, "aggregate_GBps_wall": nbytes * len(gpus) / wall / 1e9}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=6)
    a = ap.parse_args()
    n = torch.cuda.device_count()
    cpus = sorted(os.sched_getaffinity(0))
    host, dev, streams = {}, {}, {}
    for g in range(n):
        torch.cuda.set_device(g)
        host[g] = torch.empty(a.mb << 20, dtype=torch.uint8, pin_memory=True)
        first_touch(host[g], cpus[(g * len(cpus)) // n])               # spread the touching threads over the vCPUs
        dev[g] = torch.empty(a.mb << 20, dtype=torch.uint8, device="cuda:%d" % g)
        streams[g] = torch.cuda.Stream(device=g)
    sets = [[g] for g in range(n)]
    if n >= 2:
        sets += [[0, 1], [0, n - 1]]
    if n >= 4:
        sets += [list(range(n // 2)), list(range(n // 2, n)), [0, 1, n // 2, n // 2 + 1], list(range(n))]
    out = {"n_gpus": n, "mb_per_copy": a.mb, "reps": a.reps, "vcpus": len(cpus),
           "numa_nodes": sorted(x for x in os.listdir("/sys/devices/system/node") if x.startswith("node")) if os.path.isdir("/sys/devices/system/node") else [],
           "rows": [measure(s, host, dev, streams, a.reps) for s in sets]}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
