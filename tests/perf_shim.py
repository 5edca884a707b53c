"""perf_shim.py -- throughput of the reference's in-process interface transform_set_into_vector
(include/random_projection.h:11) called from the sample loop of sketch() (an OpenMP `parallel for
schedule(dynamic)` over in-memory unordered_sets, src/project_everything.cpp:287-298): the reference's own
function on the host cores next to our link-time replacement (csrc/shim/random_projection_b200.cpp) on the GPU.
Run on the GPU box:

    python tests/perf_shim.py [--samples 1024] [--ref-samples 128] [--hashes 50000] [--d 2048] > gpurun_out/shim_bench.json

Both binaries are tests/a4_loop_harness.cpp, linked by oracle/Makefile (they include the reference's header, so
they are built where the reference tree is and travel as oracle/_ref/v3/a4_loop_*).  The first --ref-samples
sets are the same in both runs; a third run of the shim binary on exactly that prefix must give the reference's
checksum.  Lives under tests/ because it executes oracle/_ref; not collected by pytest."""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFV3 = os.path.join(ROOT, "oracle", "_ref", "v3")


def run(exe, n, h, d, passes, threads, extra_env=None):
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    env.update(extra_env or {})
    r = subprocess.run([os.path.join(REFV3, exe), str(n), str(h), str(d), str(passes)], capture_output=True, text=True, env=env)
    if r.returncode != 0:
        sys.exit("%s failed: %s" % (exe, r.stderr[-2000:]))
    return json.loads(r.stdout.strip().splitlines()[-1])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=1024)
    ap.add_argument("--ref-samples", type=int, default=128)
    ap.add_argument("--hashes", type=int, default=50000)
    ap.add_argument("--d", type=int, default=2048)
    ap.add_argument("--threads", default="1,4,16")
    ap.add_argument("--gpus", type=int, default=1)
    a = ap.parse_args()
    cores = os.cpu_count() or 1
    ref = run("a4_loop_ref", a.ref_samples, a.hashes, a.d, 1, cores)
    same = run("a4_loop_b200shim", a.ref_samples, a.hashes, a.d, 1, cores, {"MVS_SHIM_DEVICES": str(a.gpus)})
    out = {"what": "transform_set_into_vector in the a4 sample loop over in-memory unordered_sets",
           "hashes_per_sample": a.hashes, "d": a.d, "host_cores": cores, "gpus": a.gpus,
           "reference_cpu": ref, "checksum_equal_on_reference_prefix": ref["checksum"] == same["checksum"],
           "b200_shim": []}
    for t in [int(x) for x in a.threads.split(",")]:
        out["b200_shim"].append(run("a4_loop_b200shim", a.samples, a.hashes, a.d, 3, t, {"MVS_SHIM_DEVICES": str(a.gpus)}))
    out["b200_shim_one_lock_around_the_device_call"] = run("a4_loop_b200shim", a.samples, a.hashes, a.d, 3, cores,
                                                           {"MVS_SHIM_DEVICES": str(a.gpus), "MVS_SHIM_SERIAL": "1"})
    best = max(out["b200_shim"], key=lambda r: r["hashes_per_s"])
    out["speedup_best_vs_reference_all_cores"] = best["hashes_per_s"] / ref["hashes_per_s"]
    print(json.dumps(out))
    if not out["checksum_equal_on_reference_prefix"]:
        sys.exit("checksum mismatch")


if __name__ == "__main__":
    main()
