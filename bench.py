#!/usr/bin/env python
"""bench.py -- hashes projected / s at d=2048 (BASELINE.json metric) on N B200s.

A "step" is one pass of the hot path over one CSR batch: BASELINE config 2, 10 000 samples
x 50 000 hashes, d = 2048, seed 0, synthetic (`mvs-synth-v1`, generated on the device).  With
--gpus N every rank owns its own batch of that shape (samples are independent: weak scaling,
no data-path collective).

  value  device-resident throughput: K calls of mvs_project_device, CUDA events on the
         launching stream, barrier + synchronize on both sides, max over ranks.
  e2e    the same batch through mvs_project_batch with PINNED HOST buffers: H2D of the
         hashes, kernels and D2H of the vectors all inside the timed region.
  roofline     issue-slot roofline of the projection kernel: warp instructions the
         kernel issues per launch (from the committed ncu capture of the SAME cubin, profiles/
         project_kernel_profile.json, keyed by tools/cubin_key.py) / live kernel time, against the SM sub-partitions'
         issue slots (1 warp instruction per clock each) -- a fraction that cannot exceed 1.  The algorithmic view
         (W(d) = d + 20*ceil(d/64) int32 ops per hash against the live LOP3+IMAD issue peak) is `frac_algorithmic`,
         the HBM view (8 B/hash in, 4d B/sample out) is reported beside it.
  config3      BASELINE configs[2]: dimension sweep d in {256..4096} on 1024 x 100 000 hashes (N = 1 only).
  config4      BASELINE configs[3]: heavy-tailed batch (100 000 samples, log-normal 1e3..1e7, d = 2048) STRONG-scaled
         over the N ranks with the library's own mvs_partition_samples; per-rank ms, imbalance, reduce time, a parity
         property on every row and an in-process cross-check through mvs_sketch_batch_multi.
  cpu_baseline the reference's own transform_set_into_vector (oracle/_ref, compiled from
         /root/reference) under OpenMP on the host cores, on a bounded prefix of the batch.

`--impl reference` times only that CPU arm (rank 0; other ranks exit).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "hashes_projected_per_sec_d2048"
UNIT = "hashes/s"
WORKLOAD = "configs[1]: synthetic 10k samples x 50k hashes, d=2048, seed 0"


PROFILE_JSON = os.path.join(ROOT, "profiles", "project_kernel_profile.json")


def load_kernel_profile(n, per, d):
    """The committed ncu capture of the dominant kernel (tools/ncu_summary.py --json) if -- and only if -- it was
    taken on the cubin that is loaded now and on this workload.  Returns (profile or None, why)."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    try:
        from cubin_key import projection_cubin_key
        prof = json.load(open(PROFILE_JSON))
    except Exception as e:
        return None, "no profile (%s)" % repr(e)[:80]
    key = projection_cubin_key()
    if key is None:
        return None, "cubin key unavailable (cuobjdump missing)"
    if key != prof.get("kernel_sass_sha256"):
        return None, "stale: profile is of cubin %s..., library is %s..." % (str(prof.get("kernel_sass_sha256"))[:12], key[:12])
    if prof.get("workload") != "%dx%dx%d" % (n, per, d):
        return None, "profile is of workload %s" % prof.get("workload")
    return prof, "cubin %s..., %s" % (key[:12], prof.get("source"))


def balanced_relay_share(p0, g0, p1, g1):
    """Share (percent) of a relaying rank's bytes that should take the relay, from two measured steps: g = (slowest
    relaying rank's time - slowest direct rank's time) at shares p0 and p1.  The gap falls linearly with the share
    (more relayed = relaying ranks faster, direct ranks slower); its root is where both groups finish together.  A gap
    that does not move keeps p0; the result is clamped to [0, 100]."""
    if abs(g1 - g0) <= 1e-3 or p1 == p0:
        return int(p0)
    return int(round(min(100.0, max(0.0, p0 - g0 * (p1 - p0) / (g1 - g0)))))


def work_per_hash(d):
    """algorithmic int32 ops per hash, d sign-accumulates + 20 per splitmix64."""
    return d + 20 * ((d + 63) // 64)


# ------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.proc = None
        self.th = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.th = threading.Thread(target=self._read, daemon=True)
        self.th.start()

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 6:
                self.samples.append(parts)

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for p in self.samples:
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's own kernel (oracle/_ref) or, if it cannot run here, the oracle port
# ------------------------------------------------------------------------------------------
def _load_cpu_arm():
    """Returns (kind, fn(hashes, offsets, d) -> seconds_of_the_sample_loop)."""
    import ctypes as C
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_util
    lib = oracle_util.ref_lib()
    if lib is not None:
        def run_ref(h, off, d):
            out = np.empty((len(off) - 1, d), dtype=np.int32)
            sec = C.c_double()
            lib.ref_project_csr(h.ctypes.data, off.ctypes.data, len(off) - 1, d, out.ctypes.data, C.byref(sec))
            return sec.value
        return "reference", run_ref
    orc = oracle_util.oracle()

    def run_port(h, off, d):
        out = np.empty((len(off) - 1, d), dtype=np.int32)
        t0 = time.perf_counter()
        orc.oracle_project_csr(h.ctypes.data, off.ctypes.data, len(off) - 1, d, C.c_uint64(0), out.ctypes.data)
        return time.perf_counter() - t0
    return "port", run_port


def cpu_arm(per_sample, d, steps, warmup, target_s):
    """Time the CPU arm on a bounded prefix of the workload.  Returns dict(value, cores, kind, sample, ms_per_step)."""
    import metagenome_vector_sketches_b200.synthetic as syn
    try:
        cores = len(os.sched_getaffinity(0)) or 1
    except AttributeError:  # pragma: no cover
        cores = os.cpu_count() or 1
    # The CPU arm uses every host thread it may run on.  torchrun exports OMP_NUM_THREADS=1 to its workers, so the
    # variable is overwritten (before libgomp is loaded) and the thread count is also set through libgomp itself.
    os.environ["OMP_NUM_THREADS"] = str(cores)
    kind, fn = _load_cpu_arm()
    try:
        import ctypes as C
        C.CDLL("libgomp.so.1").omp_set_num_threads(int(cores))
    except OSError:  # pragma: no cover
        pass
    # calibrate on one sample per core, then size a step to ~target_s seconds
    h, off = syn.uniform_batch(cores, per_sample, seed=0)
    t_cal = max(fn(h, off, d), 1e-4)
    mult = int(max(1, min(16, round(target_s / t_cal))))
    n = cores * mult
    h, off = syn.uniform_batch(n, per_sample, seed=0)
    for _ in range(warmup):
        fn(h, off, d)
    times = [fn(h, off, d) for _ in range(max(1, steps))]
    sec = sum(times) / len(times)
    return {"value": n * per_sample / sec, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "first %d of 10000 samples x %d hashes (d=%d), OpenMP schedule(dynamic) sample loop, "
                      "%d threads, %d timed passes" % (n, per_sample, d, cores, len(times)),
            "ms_per_step": sec * 1e3}


def run_reference_impl(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    r = cpu_arm(args.hashes_per_sample, args.d, args.steps, min(args.warmup, 1), target_s=3.0)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "d": args.d, "hashes_per_sample": args.hashes_per_sample},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                             "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    _emit(line)
    return 0


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    import torch.distributed as dist
    import metagenome_vector_sketches_b200 as mvs
    from metagenome_vector_sketches_b200 import synthetic as syn

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: no CUDA device and there is no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # host locality: this rank's pinned buffers should live on the NUMA node its GPU hangs off (matters at N > 1,
    # where the ranks' uploads otherwise meet on one socket's memory controllers / the socket interconnect)
    binding = {"numa_node": None, "applied": 0}
    if not args.no_bind:
        node, applied = mvs.bind_host_to_device(local)
        binding = {"numa_node": node, "applied": applied}
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n, per, d = args.samples, args.hashes_per_sample, args.d
    sk = mvs.Sketcher(local)
    # every rank owns its own batch (weak scaling): sample ids are offset by rank
    dh, off = syn.uniform_batch_torch(n, per, dev, seed=0, first_sample=rank * n)
    dout = torch.zeros((n, d), dtype=torch.int32, device=dev)
    total = n * per
    # a real stream: stream 0 (torch's default) means "the context's own stream, synchronous" to mvs_project_device
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream

    def step_device():
        sk.project_device(dh.data_ptr(), off, d, dout.data_ptr(), seed=0, stream=stream)

    # ---- value: device-resident ---------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    torch.cuda.synchronize()
    sk.timing()                                   # drop warm-up launches from the kernel average
    launches0 = sk.timing()["launches"]
    clocks = ClockSampler(local)
    clocks.start()
    time.sleep(0.3)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    tm = sk.timing()
    kernel_ms = max_over_ranks(tm["kernel_ms"])
    launches = tm["launches"] - launches0
    ms_step = ms_total / args.steps
    value = world * total / (ms_step * 1e-3)
    checksum_dev = int(dout.sum(dtype=torch.int64).item())

    # ---- e2e: pinned host buffers through mvs_project_batch -------------------------------
    if args.wc_input:                             # experiment: write-combined input buffer (no CPU cache snoops)
        from metagenome_vector_sketches_b200.projection import pinned_empty
        wc = pinned_empty(total, np.int64, write_combined=True)
        hh = torch.from_numpy(wc.a)
    else:
        hh = torch.empty(total, dtype=torch.int64, pin_memory=True)
    hh.copy_(dh)
    # ---- host-link probe and upload relay (N > 1): every rank copies 256 MB x 4 to its GPU at the same time; if the
    # ranks' rates differ by more than 25 % (unequal PCIe roots, profiles/r02l_h2d_matrix_8gpu.json), the slow half
    # routes its hash slices through a partner of the fast half (option relay_device: host -> partner GPU -> NVLink)
    relay = {"mode": args.relay, "probe_GBps": None, "via": None}
    if world > 1 and args.relay != "off":
        probe_dst = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        probe_src = hh[: (256 << 20) // 8].view(torch.uint8)
        probe_dst.copy_(probe_src, non_blocking=True)
        barrier()
        pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        pe0.record()
        for _ in range(4):
            probe_dst.copy_(probe_src, non_blocking=True)
        pe1.record()
        torch.cuda.synchronize()
        rate = 4 * (256 << 20) / (pe0.elapsed_time(pe1) * 1e-3) / 1e9
        rt = torch.tensor([rate], dtype=torch.float64, device=dev)
        allr = [torch.zeros_like(rt) for _ in range(world)]
        dist.all_gather(allr, rt)
        rates = [float(x.item()) for x in allr]
        relay["probe_GBps"] = rates
        order = sorted(range(world), key=lambda r: rates[r])             # slowest first
        via = None
        if args.relay.startswith("force:"):
            via = int(args.relay.split(":")[1])
        elif min(rates) < 0.75 * max(rates) and world % 2 == 0:
            pos = order.index(rank)
            if pos < world // 2:
                via = order[world - 1 - pos]                              # slowest pairs with fastest
        if via is not None and via != local:
            sk.set_option("relay_device", via)
            relay["via"] = via
            if args.relay_percent >= 0:
                pct = args.relay_percent
            else:
                # balanced split for rates that stay as probed: a slow rank (rate s_) keeps x = 2 s_/(f_ + s_) of its
                # bytes on its own link and relays 1 - x, so that its link and its partner's (own 4 GB + relayed) finish together
                s_, f_ = rates[rank], rates[via] if via < world else max(rates)
                pct = int(round(100.0 * max(0.0, 1.0 - 2.0 * s_ / (f_ + s_))))
            sk.set_option("relay_percent", pct)
            relay["percent"] = pct
        viat = torch.tensor([-1.0 if relay["via"] is None else float(relay["via"])], dtype=torch.float64, device=dev)
        allv = [torch.zeros_like(viat) for _ in range(world)]
        dist.all_gather(allv, viat)
        relay["via_per_rank"] = [int(x.item()) for x in allv]
        del probe_dst
    hout = torch.empty((n, d), dtype=torch.int32, pin_memory=True)
    h_np = hh.numpy().view(np.uint64)
    out_np = hout.numpy()
    for _ in range(min(args.warmup, 3)):
        sk.project_csr(h_np, off, d, out=out_np)
    # ---- relay share: two more untimed steps balance it.  The probe's rates do not stay as probed once the relay runs
    # (the host side of both PCIe roots is shared), so the split is corrected from what the steps themselves show: one
    # step at the probe's share p0, one at p0 + 15, both all-gathered as (relaying ranks' time - direct ranks' time);
    # that difference is linear in the share (profiles/r02s_*: 166.8 | 127.6 ms at 10 %, 155.6 | 141.7 at 21 %,
    # 143.0 | 152.1 at 35 %, 114.3 | 152.1 at 50 %), its root is the share at which both groups finish together.
    if world > 1 and args.relay_percent < 0 and any(v >= 0 for v in relay.get("via_per_rank", [])):
        def one_step_gap(pct):
            if relay["via"] is not None:
                sk.set_option("relay_percent", pct)
            barrier()
            t0 = time.perf_counter()
            sk.project_csr(h_np, off, d, out=out_np)
            ms = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=dev)
            allm = [torch.zeros_like(ms) for _ in range(world)]
            dist.all_gather(allm, ms)
            allm = [float(x.item()) for x in allm]
            slow = max(m for m, v in zip(allm, relay["via_per_rank"]) if v >= 0)
            fast = max(m for m, v in zip(allm, relay["via_per_rank"]) if v < 0)
            return slow - fast
        pcts = torch.tensor([float(relay.get("percent", -1))], dtype=torch.float64, device=dev)
        allp = [torch.zeros_like(pcts) for _ in range(world)]
        dist.all_gather(allp, pcts)
        p0 = int(round(max(float(x.item()) for x in allp)))              # the same for every relaying rank up to probe noise
        p1 = min(100, p0 + 15) if p0 <= 85 else p0 - 15
        g0, g1 = one_step_gap(p0), one_step_gap(p1)
        tuned = balanced_relay_share(p0, g0, p1, g1)
        if relay["via"] is not None:
            sk.set_option("relay_percent", tuned)
        relay["percent_probe"] = p0
        relay["percent"] = tuned
        relay["tuning_gap_ms"] = [[p0, g0], [p1, g1]]
        sk.project_csr(h_np, off, d, out=out_np)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sk.project_csr(h_np, off, d, out=out_np)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    barrier()
    clk = clocks.stop()
    my_e2e_ms = (t1 - t0) * 1e3 / args.steps
    e2e_ms = max_over_ranks((t1 - t0) * 1e3) / args.steps
    per_rank = [[my_e2e_ms, binding["numa_node"], binding["applied"]]]
    if world > 1:
        t = torch.tensor([my_e2e_ms, -1.0 if binding["numa_node"] is None else float(binding["numa_node"]),
                          float(binding["applied"])], dtype=torch.float64, device=dev)
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        per_rank = [[float(x[0]), int(x[1]), int(x[2])] for x in allt]
    tb = sk.timing()
    e2e_value = world * total / (e2e_ms * 1e-3)
    checksum_host = int(out_np.sum(dtype=np.int64))
    if checksum_host != checksum_dev:
        raise SystemExit("bench: device-resident and host paths disagree (%d vs %d)" % (checksum_dev, checksum_host))

    # ---- the same batch through mvs_sketch_batch_packed56: 7 bytes per hash on the host link (all synthetic hashes
    # are < 2^54, like FracMinHash hashes at scaled = 1000).  Packing is the producer's job (a parser would emit it
    # directly); its cost here is reported, not timed into the step.  N = 1 only (another 3.5 GB of pinned memory).
    packed_ms = pack_s = None
    if world == 1 and not args.no_packed:
        pk = torch.empty(7 * total, dtype=torch.uint8, pin_memory=True)
        pk_np = pk.numpy()
        t0 = time.perf_counter()
        mvs.pack56(h_np, out=pk_np)
        pack_s = time.perf_counter() - t0
        for _ in range(2):
            sk.project_csr_packed56(pk_np, off, d, out=out_np)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            sk.project_csr_packed56(pk_np, off, d, out=out_np)
        torch.cuda.synchronize()
        packed_ms = (time.perf_counter() - t0) * 1e3 / args.steps
        tp = sk.timing()
        packed_h2d = int(tp["h2d_bytes"])
        if int(out_np.sum(dtype=np.int64)) != checksum_dev:
            raise SystemExit("bench: packed56 host path disagrees with the device-resident path")
        del pk, pk_np

    # ---- the same call with PAGEABLE host buffers (what a caller with a std::vector has); extra information ----
    pageable_ms = None                           # single-GPU runs only: N ranks would hold N more 4 GB host copies
    if world == 1:
        pg_in = dh.cpu().numpy().view(np.uint64) if args.wc_input else np.array(h_np, copy=True)
        pg_out = np.empty((n, d), dtype=np.int32)
        sk.project_csr(pg_in, off, d, out=pg_out)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(2):
            sk.project_csr(pg_in, off, d, out=pg_out)
        torch.cuda.synchronize()
        pageable_ms = (time.perf_counter() - t0) * 1e3 / 2
        if int(pg_out.sum(dtype=np.int64)) != checksum_dev:
            raise SystemExit("bench: pageable host path disagrees with the device-resident path")
        del pg_in, pg_out

    # ---- BASELINE configs[2]: dimension sweep (N = 1 only; device-resident, same timing rules as `value`) ----
    config3 = None
    if world == 1 and not args.no_config3:
        config3 = run_config3(sk, torch, syn, dev)
    # ---- BASELINE configs[3]: heavy-tailed batch, strong-scaled over the ranks ----
    config4 = None
    if args.c4_samples > 0:
        del hh, hout, h_np, out_np
        config4 = run_config4(sk, mvs, torch, dist if world > 1 else None, syn, dev, rank, world, local, args.c4_samples,
                              barrier)
    return locals()


def run_config3(sk, torch, syn, dev, ns=1024, per=100000):
    dh, off = syn.uniform_batch_torch(ns, per, dev, seed=0)
    stream = torch.cuda.current_stream().cuda_stream
    rows = []
    for d in (256, 512, 1024, 2048, 4096):
        dout = torch.empty((ns, d), dtype=torch.int32, device=dev)
        for _ in range(3):
            sk.project_device(dh.data_ptr(), off, d, dout.data_ptr(), stream=stream)
        torch.cuda.synchronize()
        sk.timing()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            sk.project_device(dh.data_ptr(), off, d, dout.data_ptr(), stream=stream)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        t = sk.timing()
        hps = ns * per / (ms * 1e-3)
        # size-independent property: every column of a row has the parity of |H_s|
        ok = bool(((dout & 1) == (per & 1)).all().item())
        rows.append({"d": d, "ms_per_step": ms, "kernel_ms": t["kernel_ms"], "hashes_per_s": hps, "entries_per_s": hps * d,
                     "warp_tasks": int(t["warp_tasks"]), "parity_property_ok": ok})
        del dout
    del dh
    torch.cuda.empty_cache()
    return {"workload": "configs[2]: d in {256,512,1024,2048,4096}, %d samples x %d hashes, 1 B200, device-resident "
                        "(0.82 GB of hashes per step > L2)" % (ns, per), "steps": 5, "warmup": 3, "rows": rows}


def run_config4(sk, mvs, torch, dist, syn, dev, rank, world, local, samples, barrier, d=2048, steps=3):
    from metagenome_vector_sketches_b200 import sharding
    sizes = syn.lognormal_sizes(samples, seed=0)                    # every rank computes the same sizes
    offsets = np.zeros(samples + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum(sizes).astype(np.uint64)
    bounds = sharding.partition_bounds(offsets, world)              # the library's own mvs_partition_samples
    a, e = int(bounds[rank]), int(bounds[rank + 1])
    dh, off = syn.ragged_batch_torch(sizes[a:e], dev, seed=0, first_sample=a)
    mine = int(off[-1])
    dout = torch.empty((e - a, d), dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(3):                            # one warm-up per plan slot of the context (buffers are allocated on first use)
        sk.project_device(dh.data_ptr(), off, d, dout.data_ptr(), stream=stream)
    torch.cuda.synchronize()
    sk.timing()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        sk.project_device(dh.data_ptr(), off, d, dout.data_ptr(), stream=stream)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / steps
    tm = sk.timing()
    # size-independent property on every row of the shard: each column has the parity of |H_s|
    ok = bool(((dout[:, :8] & 1).to(torch.int64).cpu().numpy() == (sizes[a:e, None] & 1)).all())
    # in-process cross-check of the torchrun sharding against the library's one-call multi-context path: a prefix of
    # this rank's shard (<= 2e7 hashes) through mvs_sketch_batch_multi on two contexts of this rank's device
    k = int(np.searchsorted(off, np.uint64(20_000_000), side="right")) - 1
    xok = None
    if k >= 1:
        h_host = dh[: int(off[k])].cpu().numpy().view(np.uint64)
        sk2 = mvs.Sketcher(local)
        try:
            got = mvs.sketch_csr_multi([sk, sk2], h_host, off[: k + 1], d)
        finally:
            sk2.close()
        xok = bool(np.array_equal(got["vec"], dout[:k].cpu().numpy()))
    t = torch.tensor([ms, float(mine), float(e - a), 1.0 if ok else 0.0, -1.0 if xok is None else float(xok),
                      float(tm["reduce_ms"]), float(tm["kernel_ms"]), float(tm["partial_rows"])], dtype=torch.float64, device=dev)
    allt = [t.clone() for _ in range(world)]
    if dist is not None:
        dist.all_gather(allt, t)
    per = [[float(x[0]), int(x[1]), int(x[2])] for x in allt]
    step = max(p[0] for p in per)
    total = int(offsets[-1])
    del dh, dout
    torch.cuda.empty_cache()
    return {"workload": "configs[3]: heavy-tailed synthetic, %d samples, log-normal sizes 1e3..1e7 hashes, d=%d, seed 0, "
                        "device-resident, sharded by mvs_partition_samples" % (samples, d),
            "scaling": "strong", "n_gpus": world, "steps": steps, "warmup": 3, "total_hashes": total,
            "ms_per_step": step, "hashes_per_s": total / (step * 1e-3), "samples_per_s": samples / (step * 1e-3),
            "per_rank_ms_hashes_samples": per,
            "hash_imbalance_max_over_mean": max(p[1] for p in per) / (total / world),
            "reduce_ms_max": max(float(x[5]) for x in allt), "kernel_ms_max": max(float(x[6]) for x in allt),
            "partial_rows": int(sum(float(x[7]) for x in allt)),
            "parity_property_ok": all(bool(x[3]) for x in allt),
            "multi_context_crosscheck_ok": (None if any(float(x[4]) < 0 for x in allt) else all(bool(x[4]) for x in allt)),
            "collective": "none on the data path (NCCL only for the timing barrier / gathering these numbers)"}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        j = json.load(open(p))
        return float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


_REAL_STDOUT = None


def _quiet_stdout():
    """Route fd 1 to stderr while the job runs (NCCL / the CUDA runtime print banners on stdout) and keep
    the real stdout for the ONE JSON line."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def _emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    _quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--samples", type=int, default=10000)
    ap.add_argument("--hashes-per-sample", type=int, default=50000)
    ap.add_argument("--d", type=int, default=2048)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--wc-input", action="store_true", help="experiment: write-combined pinned input buffer")
    ap.add_argument("--no-bind", action="store_true", help="skip mvs_bind_host_to_device (A/B of the host locality)")
    ap.add_argument("--relay", default="auto",
                    help="N > 1 end-to-end leg: auto = probe the ranks' concurrent H2D rates and route the slow half through "
                         "partners of the fast half (relay_device); off; force:<device> = every rank relays through that device")
    ap.add_argument("--relay-percent", type=int, default=-1,
                    help="share of a relaying rank's slice bytes that takes the relay (default -1: 1 - 2s/(f+s) from the probe, then "
                         "balanced over two untimed steps)")
    ap.add_argument("--no-packed", action="store_true", help="skip the packed56 end-to-end leg (N = 1)")
    ap.add_argument("--no-config3", action="store_true", help="skip the BASELINE configs[2] dimension sweep (N = 1 leg)")
    ap.add_argument("--c4-samples", type=int, default=100000,
                    help="samples of the BASELINE configs[3] heavy-tailed leg (0 = skip; 100000 = the full config)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference_impl(args)

    g = run_gpu(args)
    rank, world, sk = g["rank"], g["world"], g["sk"]
    d, n, per, total = g["d"], g["n"], g["per"], g["total"]

    # ---- roofline of the projection kernel (rank 0's device; kernel time = max over ranks) ----
    peaks = {name: sk.int_issue_peak(which) for which, name in
             ((0, "lop3"), (1, "imad"), (2, "lop3_imad_mix"), (3, "iadd"), (4, "shf"))}
    p_int = peaks["lop3_imad_mix"]               # both integer pipes busy: the most any int32 stream can issue
    w = work_per_hash(d)
    k_s = g["kernel_ms"] * 1e-3
    achieved_ops = total * w / k_s
    hbm_peak, hbm_how = measured_peaks()
    alg_bytes = total * 8 + n * d * 4
    # Issue-slot view, the one that cannot exceed 1.  A B200 SM has 4 sub-partitions that each issue at most one warp
    # instruction per clock: peak = 4 x SMs x clock.  The warp instructions of one launch come from the committed ncu
    # capture of THIS cubin on THIS workload (smsp__inst_executed.sum, a property of code + input, not of timing);
    # a capture of another cubin is refused (frac = null) rather than silently reused.
    prof, prof_why = load_kernel_profile(n, per, d)
    clk_mhz = g["clk"].get("sm_mhz") or 1965.0
    sm_count = g["torch"].cuda.get_device_properties(g["local"]).multi_processor_count
    slots_per_s = 4.0 * sm_count * clk_mhz * 1e6
    if prof is not None:
        inst_per_s = prof["inst_executed"] / k_s
        scale = (prof["duration_ms_under_ncu"] * prof["sm_clock_ghz"] * 1e3) / (g["kernel_ms"] * clk_mhz)   # cycles under ncu / cycles now
        issue = {"achieved": inst_per_s / 1e9, "frac": inst_per_s / slots_per_s,
                 "warp_instructions_per_launch": prof["inst_executed"],
                 "warp_instructions_per_hash": prof["inst_executed"] / total,
                 "pipes_pct_live": {"alu": prof["pipe_alu_pct"] * scale, "fmaheavy": prof["pipe_fmaheavy_pct"] * scale,
                                    "xu": prof["pipe_xu_pct"] * scale, "lsu": prof["pipe_lsu_pct"] * scale,
                                    "how": "ncu pipe utilisation of the capture x (cycles under ncu / cycles of the live launch)"},
                 "ncu": {"issue_active_pct": prof["issue_active_pct"], "pipe_alu_pct": prof["pipe_alu_pct"],
                         "pipe_fmaheavy_pct": prof["pipe_fmaheavy_pct"], "pipe_xu_pct": prof["pipe_xu_pct"],
                         "kernel_ms_under_ncu": prof["duration_ms_under_ncu"],
                         "shared_bank_conflicts": prof["shared_bank_conflicts"], "registers": prof["registers_per_thread"]},
                 "traffic": prof["dram_bytes_read"] + prof["dram_bytes_write"]}
    else:
        issue = {"achieved": None, "frac": None, "traffic": None}
    roofline = {
        "bound": "issue slots of the SM sub-partitions (integer ALU + FMA-heavy pipes); not hbm, not tensor -- DESIGN.md section 4",
        "achieved": issue["achieved"], "peak": slots_per_s / 1e9, "unit": "Gwarp-instr/s", "frac": issue["frac"],
        "traffic": issue["traffic"],
        "how": "warp instructions of one launch (smsp__inst_executed.sum of the committed ncu capture, " + prof_why + ") / live "
               "kernel time, over 4 x %d SMs x %.0f MHz (median SM clock sampled during the timed region)" % (sm_count, clk_mhz),
        "profile": {k: issue[k] for k in issue if k not in ("achieved", "frac", "traffic")},
        "kernel_ms": g["kernel_ms"], "kernel_hashes_per_s": total / k_s,
        "frac_algorithmic": achieved_ops / p_int,
        "algorithmic": {"achieved": achieved_ops / 1e12, "peak": p_int / 1e12, "unit": "Tint32op/s", "work_per_hash_ops": w,
                        "peak_how": "live microbenchmark mvs_int_issue_peak(2): LOP3 (ALU pipe) and IMAD (FMA pipe) interleaved 1:1",
                        "issue_peaks_Top_s": {k: v / 1e12 for k, v in peaks.items()},
                        "note": "W(d)=d+20*ceil(d/64) int32 ops per hash (SURVEY 8d) charges one op per sign-accumulate; the "
                                "bit-sliced kernel does 32 columns per LOP3, so this ratio exceeds 1 by construction and is "
                                "not a roofline fraction"},
        "hbm": {"bound": "hbm", "achieved": alg_bytes / k_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": alg_bytes / k_s / 1e9 / hbm_peak, "peak_how": hbm_how,
                "algorithmic_bytes_per_launch": alg_bytes},
    }
    # Mix limit: the loop's own opcode mix without data dependencies (mvs_loop_mix_peak, live).  One warp round = one
    # hash x 32 lanes = 2048 columns, so a launch of this workload is `total * ceil(d/2048)` rounds.
    mix_rounds_per_s = sk.loop_mix_peak()
    rounds_per_launch = total * ((d + 2047) // 2048) if d >= 2048 else None
    if rounds_per_launch:
        roofline["mix_limit"] = {
            "achieved": rounds_per_launch / k_s / 1e9, "peak": mix_rounds_per_s / 1e9, "unit": "Gwarp-round/s",
            "frac": rounds_per_launch / k_s / mix_rounds_per_s,
            "clk_per_round": {"kernel": 4.0 * sm_count * clk_mhz * 1e6 * k_s / rounds_per_launch,
                              "mix_without_dependencies": 4.0 * sm_count * clk_mhz * 1e6 / mix_rounds_per_s},
            "how": "live microbenchmark mvs_loop_mix_peak: the projection loop's opcodes in the same proportions (8.5 LOP3, "
                   "3 SHF, 4 IMAD, 3 IMAD add, 2 IMAD.WIDE, 1 POPC, 1 LDS.128 per round) on independent registers at the "
                   "same residency -- what the issue ports and integer pipes sustain for this mix; the kernel sits on it, so "
                   "the idle issue slots of `frac` are the machine's for this mix, not stalls of the kernel (DESIGN.md 4)"}
    if g["config3"] is not None:
        for row in g["config3"]["rows"]:
            row["frac_algorithmic"] = row["hashes_per_s"] * work_per_hash(row["d"]) / p_int

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            c = cpu_arm(per, d, steps=2, warmup=0, target_s=6.0)
            cpu = {"value": c["value"], "unit": UNIT, "cores": c["cores"], "kind": c["kind"], "sample": c["sample"]}
        except Exception as e:  # the checker being absent must not void the GPU numbers
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": repr(e)[:200]}

    if rank == 0:
        tb = g["tb"]
        line = {
            "metric": METRIC, "value": g["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": g["ms_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "samples_per_gpu": n, "hashes_per_sample": per, "d": d, "seed": 0,
                       "generator": "mvs-synth-v1 (54-bit bijection, unique hashes, on-device)",
                       "l2": "inputs (%.1f GB per step) larger than the 126 MB L2" % (total * 8 / 1e9),
                       "parallelism": "samples sharded by rank, no collective"},
            "samples_per_s": g["value"] / per,
            "e2e": {"value": g["e2e_value"], "unit": UNIT, "ms_per_step": g["e2e_ms"],
                    "h2d_bytes_per_step": int(tb["h2d_bytes"]), "d2h_bytes_per_step": int(tb["d2h_bytes"]),
                    "api": "mvs_project_batch (%s host CSR in, pinned host int32 rows out)"
                           % ("write-combined pinned" if args.wc_input else "pinned"),
                    "relay": g["relay"],
                    "host_binding": {"how": "mvs_bind_host_to_device per rank before the pinned buffers are allocated "
                                            "(applied: 1 = CPU affinity, 2 = preferred memory node)",
                                     "per_rank_ms_node_applied": g["per_rank"]},
                    "packed56_host_buffers": None if g["packed_ms"] is None else {
                        "value": g["total"] / (g["packed_ms"] * 1e-3), "unit": UNIT, "ms_per_step": g["packed_ms"],
                        "h2d_bytes_per_step": g["packed_h2d"], "pack_seconds_once": g["pack_s"],
                        "how": "mvs_sketch_batch_packed56: the same batch as 7 little-endian bytes per hash in pinned memory "
                               "(12.5 % fewer bytes on the PCIe link that bounds e2e), widened on the device; packing "
                               "(mvs_pack56, host threads) happens once outside the timed steps"},
                    "pageable_host_buffers": None if g["pageable_ms"] is None else {
                        "value": g["total"] / (g["pageable_ms"] * 1e-3), "unit": UNIT, "ms_per_step": g["pageable_ms"],
                        "how": "same call, numpy (pageable) arrays in and out: the library stages them through a pinned "
                               "ring with 8 copy threads using streaming stores (measured at N = 1 only)"}},
            "gpu_launches": int(g["launches"]),
            "clocks": g["clk"],
            "roofline": roofline,
            "cpu_baseline": cpu,
            "config3": g["config3"],
            "config4": g["config4"],
            "checksum": g["checksum_dev"],
        }
        _emit(line)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
