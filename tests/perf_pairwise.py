"""perf_pairwise.py -- throughput of the all-pairs comparison on a synthetic db, next to the
reference function (oracle/_ref/*/libref_pairwise.so = the unmodified reference source) on the host cores.

    python tests/perf_pairwise.py [--n 10000] [--d 2048] [--ref-n 768] > gpurun_out/pairwise_bench.json

db: n vectors in families of 20 that share a base vector (so that neighbours exist), entries ~ N(0, 220) like the
projection of ~5e4-hash samples (two base-256 limbs).  value = pairs compared per second (kernel time, db resident in HBM);
e2e = the same through mvs_pairwise_shard with host buffers (upload + result download inside the timed region).
Roofline: one IMAD per (pair, dimension) against the live IMAD issue peak (mvs_int_issue_peak(1)).
Lives under tests/ because it executes oracle/_ref (the compiled reference) as checker and CPU baseline; not collected by pytest."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))   # oracle_util: the checker / CPU baseline

import metagenome_vector_sketches_b200 as mvs  # noqa: E402
from metagenome_vector_sketches_b200 import pairwise as pw  # noqa: E402


def make_db(n, d, seed=0):
    """Entries like the projection of ~5e4-hash samples (sigma ~ 220, |v| up to ~1200: two base-256 limbs)."""
    rng = np.random.default_rng(seed)
    fam = np.rint(rng.normal(0.0, 190.0, ((n + 19) // 20, d)))
    v = (fam[np.arange(n) // 20] + np.rint(rng.normal(0.0, 110.0, (n, d)))).astype(np.int32)
    nrm = np.array([float("%g" % x) ** 2 for x in np.sqrt((v.astype(np.float64) ** 2).sum(axis=1) / d)])
    return v, nrm


def roofline(a, pairs, gemm_ms, imad_peak, max_abs):
    """Dominant kernel = the limb GEMM.  tcgen05 / mma.sync paths: tensor bound, int8 MACs = pairs x d x limb products
    (4 for two limbs per side, 10 for four), against the measured dense peak of this pool's B200 (MEASURED_PEAKS.json
    holds cuBLAS bf16; kind::i8 has twice the bf16 rate: 8192 vs 4096 MAC/clk/SM).  IMAD path: FMA-heavy pipe issue."""
    if a.path == 1:
        return {"bound": "int32 IMAD issue (FMA-heavy pipe)", "unit": "TMAC/s", "achieved": pairs * a.d / (gemm_ms * 1e-3) / 1e12,
                "peak": imad_peak / 1e12, "frac": pairs * a.d / (gemm_ms * 1e-3) / imad_peak,
                "peak_how": "live microbenchmark mvs_int_issue_peak(1) (IMAD only)", "traffic": None}
    limbs = max(a.limbs, 1 if max_abs <= 127 else 2 if max_abs <= 32639 else 3 if max_abs <= 8355711 else 4)
    products = sum(1 for p in range(limbs) for q in range(limbs) if p + q <= 3)
    ops = 2.0 * pairs * a.d * products
    peak_tf, how = 2250.0, "nominal dense bf16 (no MEASURED_PEAKS.json)"
    try:
        mp = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")))
        peak_tf, how = float(mp["bf16_tflops"]), "MEASURED_PEAKS.json bf16_tflops (burst, cuBLAS)"
    except (OSError, KeyError, ValueError):
        pass
    return {"bound": "tensor", "unit": "TOP/s (int8)", "achieved": ops / (gemm_ms * 1e-3) / 1e12, "peak": 2.0 * peak_tf,
            "frac": ops / (gemm_ms * 1e-3) / 1e12 / (2.0 * peak_tf), "peak_how": "2 x " + how + ": kind::i8 runs at twice the bf16 rate",
            "limb_products": products, "int32_mac_equivalent_TMAC_s": pairs * a.d / (gemm_ms * 1e-3) / 1e12,
            "imad_issue_peak_TMAC_s": imad_peak / 1e12,
            "algorithmic": "2 x pairs x d x limb products int8 operations; bytes: operands re-read from L2, C stored only for "
                           "segments that may keep a pair (fused pre-filter)", "traffic": None}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=10000)
    ap.add_argument("--d", type=int, default=2048)
    ap.add_argument("--ref-n", type=int, default=768)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--path", type=int, default=0, help="0 = tcgen05 limb GEMM, 1 = IMAD GEMM, 2 = mma.sync limb GEMM")
    ap.add_argument("--cluster", type=int, default=0, help="tcgen05 path, CTAs per cluster: 0/1 = 1x1 (default), 2 = 1x2, 4 = 2x2 with TMA multicast")
    ap.add_argument("--pair", type=int, default=0, help="tcgen05 path: 1 = CTA pairs (cta_group::2), 2 = single CTAs, 0 = library default")
    ap.add_argument("--fuse", type=int, default=0, help="tcgen05 path: 0 = keep-test in the GEMM epilogue (default), 2 = separate count pass")
    ap.add_argument("--stream", type=int, default=0, help="1 = tcgen05 path streams limb pairs instead of fused stages")
    ap.add_argument("--limbs", type=int, default=0, help="force at least this many base-256 limbs (0 = from the data)")
    a = ap.parse_args()
    v, nrm = make_db(a.n, a.d)
    ap2 = None
    knobs = (("pairwise_path", a.path), ("pairwise_limbs", a.limbs), ("pairwise_stream", a.stream),
             ("pairwise_cluster", a.cluster), ("pairwise_pair", a.pair), ("pairwise_fuse", a.fuse))
    exp = any(v for _, v in knobs)                # non-default variants live in the experiment build only
    sk = mvs.Sketcher(0, experiments=exp)
    if exp:
        for k, val in knobs:
            sk.set_option(k, val)
    pw.pairwise_shard(sk, v[:256], nrm[:256])                         # warm-up (module load, clocks)
    runs = []
    for _ in range(a.steps):
        t = time.perf_counter()
        r = pw.pairwise_shard(sk, v, nrm)
        runs.append((time.perf_counter() - t, r.gemm_ms, r.filter_ms, r.wall_ms, len(r.rows), r.h2d_bytes, r.d2h_bytes, r.prep_ms))
    best = min(runs, key=lambda x: x[1] + x[2])
    pairs = float(a.n) * a.n
    imad_peak = sk.int_issue_peak(1)
    k_s = (best[1] + best[2] + (best[7] if a.path != 1 else 0.0)) * 1e-3
    out = {
        "metric": "pairs_compared_per_sec_d%d" % a.d, "unit": "pairs/s", "n_vectors": a.n, "d": a.d,
        "kept_pairs": best[4], "value": pairs / k_s, "kernel_ms": k_s * 1e3, "gemm_ms": best[1],
        "filter_ms": best[2], "prep_ms": best[7], "path": {0: "tcgen05-limbs", 1: "imad", 2: "mma.sync-limbs"}[a.path], "forced_limbs": a.limbs, "stream": a.stream, "cluster": a.cluster, "pair": a.pair, "fuse": a.fuse,
        "max_abs_entry": int(np.abs(v).max()),
        "e2e": {"value": pairs / (min(x[3] for x in runs) * 1e-3), "unit": "pairs/s", "wall_ms": min(x[3] for x in runs),
                "h2d_bytes": best[5], "d2h_bytes": best[6], "api": "mvs_pairwise_shard (host vectors in, host triples out)"},
        "roofline": roofline(a, pairs, best[1], imad_peak, int(np.abs(v).max())),
    }
    # parity spot check + CPU baseline on a bounded prefix
    import oracle_util as ou
    m = min(a.ref_n, a.n)
    got = pw.compute_sparse_dot_products(sk, v[:m], v[:m], nrm[:m], nrm[:m])
    if ou.ref_pairwise_lib() is not None:
        rr, rc, rv, sec = ou.ref_pairwise(v[:m], v[:m], a.d, nrm[:m], nrm[:m])
        sec = min(sec, ou.ref_pairwise(v[:m], v[:m], a.d, nrm[:m], nrm[:m])[3])
        kind = "reference"
    else:
        t = time.perf_counter()
        rr, rc, rv = ou.oracle_pairwise(v[:m], v[:m], a.d, nrm[:m], nrm[:m])
        sec = time.perf_counter() - t
        kind = "port"
    out["parity_on_prefix"] = bool(np.array_equal(got.rows, rr) and np.array_equal(got.cols, rc) and
                                   np.array_equal(got.values, rv))
    out["cpu_baseline"] = {"value": float(m) * m / sec, "unit": "pairs/s", "cores": os.cpu_count(), "kind": kind,
                           "sample": "%d x %d block of the same db through compute_sparse_dot_products_optimized "
                                     "(Eigen int32 product, OMP_NUM_THREADS unset)" % (m, m), "seconds": sec}
    print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
