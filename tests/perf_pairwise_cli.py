"""perf_pairwise_cli.py -- wall-clock of the all-pairs comparison CLI on a synthetic int32 db: ours
(bin/pairwise_comp_b200, GPU) next to the reference binary (oracle/_ref/v3/pairwise_comp_optimized, the unmodified
reference sources on the host cores) with the same arguments, and a byte comparison of the three shard files.
Run on the GPU box:

    python tests/perf_pairwise_cli.py [--n 10000] [--ref-n 2000] [--d 2048] > gpurun_out/pairwise_cli_bench.json

The reference run is bounded: it compares the first --ref-n vectors (its cost grows with n^2); ours runs both sizes, so
the files are compared on the same db.  Lives under tests/ because it executes oracle/_ref; not collected by pytest."""
import argparse
import filecmp
import json
import os
import re
import shutil
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OURS = os.path.join(ROOT, "metagenome_vector_sketches_b200", "bin", "pairwise_comp_b200")
REF = os.path.join(ROOT, "oracle", "_ref", "v3", "pairwise_comp_optimized")
FILES = ("matrix.bin", "row_index.bin", "neighbor_start.bin")


def make_db(path, n, d, seed=7):
    """Families of 20 related vectors (entries ~ N(0, 220)) like tests/perf_pairwise.py; norms printed with %g as
    sketch() does."""
    os.makedirs(path, exist_ok=True)
    rng = np.random.default_rng(seed)
    fam = np.rint(rng.normal(0.0, 190.0, ((n + 19) // 20, d)))
    v = (fam[np.arange(n) // 20] + np.rint(rng.normal(0.0, 110.0, (n, d)))).astype(np.int32)
    v.tofile(os.path.join(path, "vectors.bin"))
    nrm = np.sqrt((v.astype(np.float64) ** 2).sum(axis=1) / d)
    with open(os.path.join(path, "vector_norms.txt"), "w") as f:
        for i, x in enumerate(nrm):
            f.write("s%06d %g\n" % (i, x))
    open(os.path.join(path, "dimension.txt"), "w").write("%d\n" % d)
    open(os.path.join(path, "dtype.txt"), "w").write("int32\n")
    return path + "/"


def run(exe, db, out, env=None):
    shutil.rmtree(out, ignore_errors=True)
    os.makedirs(out)
    cmd = [exe, "--db", db, "--max_memory_gb", "12", "--num_threads", str(os.cpu_count()), "--output_folder", out + "/",
           "--num_shards", "1", "--shard_idx", "0"]
    t = time.perf_counter()
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    return time.perf_counter() - t, r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=10000)
    ap.add_argument("--ref-n", type=int, default=2000)
    ap.add_argument("--d", type=int, default=2048)
    ap.add_argument("--work", default="/tmp/mvs_pairwise_cli_bench")
    a = ap.parse_args()
    shutil.rmtree(a.work, ignore_errors=True)
    full = make_db(os.path.join(a.work, "db_full"), a.n, a.d)
    small = make_db(os.path.join(a.work, "db_small"), a.ref_n, a.d)
    env = dict(os.environ, MVS_TIMING="1")
    out = {"metric": "pairwise_cli_wall_seconds", "n_vectors": a.n, "d": a.d, "ref_n_vectors": a.ref_n, "host_cores": os.cpu_count()}
    run(OURS, small, os.path.join(a.work, "warm"), env)               # page cache, driver
    walls = []
    for _ in range(3):
        w, r = run(OURS, full, os.path.join(a.work, "ours_full"), env)
        if r.returncode != 0:
            print(json.dumps({"error": "pairwise_comp_b200 failed", "stderr": r.stderr[-400:]}))
            return 1
        stages = {k: float(v) for k, v in re.findall(r"(\w+)=([0-9.eE+-]+)", r.stderr)}
        walls.append((w, stages))
    w, stages = min(walls, key=lambda x: x[0])
    out["ours"] = {"wall_s": w, "pairs_per_s": float(a.n) ** 2 / w, "stages": stages}
    w_small, r = run(OURS, small, os.path.join(a.work, "ours_small"), env)
    out["ours_on_ref_db"] = {"wall_s": w_small}
    if os.path.exists(REF):
        w_ref, r = run(REF, small, os.path.join(a.work, "ref_small"))
        same = r.returncode == 0 and all(
            filecmp.cmp(os.path.join(a.work, "ours_small", "shard_0", f), os.path.join(a.work, "ref_small", "shard_0", f), shallow=False)
            for f in FILES)
        out["reference"] = {"wall_s": w_ref, "pairs_per_s": float(a.ref_n) ** 2 / w_ref, "n_vectors": a.ref_n,
                            "binary": "oracle/_ref/v3/pairwise_comp_optimized (unmodified reference sources)",
                            "extrapolated_wall_s_at_n": w_ref * (float(a.n) / a.ref_n) ** 2}
        out["shard_files_byte_identical_on_ref_db"] = bool(same)
    else:
        out["reference"] = None
    print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
