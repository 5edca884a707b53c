"""Generates tests/golden/pairwise_ref.npz with the UNMODIFIED reference function
compute_sparse_dot_products_optimized (src/pairwise_comp_optimized.cpp:57-160) compiled by oracle/Makefile
into oracle/_ref/*/libref_pairwise.so.  Run in the build container (needs /root/reference for `make ref`):

    make -C oracle ref && python tests/golden/make_golden_pairwise.py

Cases: the toy db of the reference (61 vectors, d = 2048, norms as printed in its vector_norms.txt) and three
seeded random block pairs, one of them with values large enough for the int32 product to wrap."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_util as ou  # noqa: E402


def random_case(seed, mi, mj, d, scale, zero_norms):
    rng = np.random.default_rng(seed)
    a = rng.integers(-scale, scale + 1, (mi, d), dtype=np.int64).astype(np.int32)
    b = rng.integers(-scale, scale + 1, (mj, d), dtype=np.int64).astype(np.int32)
    b[: min(mi, mj) // 2] = a[: min(mi, mj) // 2]
    ni = np.zeros(mi) if zero_norms else ou.norms_sq_like_reference(a, d)
    nj = np.zeros(mj) if zero_norms else ou.norms_sq_like_reference(b, d)
    return a, b, ni, nj


def main():
    assert ou.ref_pairwise_lib() is not None, "build oracle/_ref first (make -C oracle ref)"
    out = {}
    ref = ou.load_toy_ref("d2048")
    v = ref["vectors"].reshape(-1, 2048)
    nrm = np.array([float(l.rsplit(" ", 1)[1]) ** 2 for l in ref["norms_txt"].strip().split("\n")])
    r, c, val, _ = ou.ref_pairwise(v, v, 2048, nrm, nrm)
    out["toy_rows"], out["toy_cols"], out["toy_values"] = r, c, val
    specs = [(101, 37, 53, 300, 60, False), (102, 64, 40, 96, 2 ** 20, True), (103, 9, 130, 17, 5, False)]
    out["specs"] = np.array(specs, dtype=np.int64)
    for k, (seed, mi, mj, d, scale, zn) in enumerate(specs):
        a, b, ni, nj = random_case(seed, mi, mj, d, scale, bool(zn))
        r, c, val, _ = ou.ref_pairwise(a, b, d, ni, nj)
        out["r%d_rows" % k], out["r%d_cols" % k], out["r%d_values" % k] = r, c, val
    np.savez_compressed(os.path.join(HERE, "pairwise_ref.npz"), **out)
    print({k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items()})


if __name__ == "__main__":
    main()
