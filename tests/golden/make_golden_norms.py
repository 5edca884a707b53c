#!/usr/bin/env python
"""Golden vectors for the norm column of vector_norms.txt, produced by the REFERENCE ITSELF.

Run in the build container only (needs `make -C oracle ref`):   python tests/golden/make_golden_norms.py

norms_ref.npz holds, for a list of dimensions d (1 ... 4100, multiples of the SSE packet and not),
  rows_<d>   random int32 rows at several magnitudes,
  norms_<d>  the float32 norm of every row computed by the two statements of
             /root/reference/src/project_everything.cpp:328-329 compiled with the reference's Eigen and flags
             (oracle/ref_shim.cpp ref_norms, oracle/_ref/v3/libref_projection.so),
  inv_<d>    the bit pattern of the 1/sqrt(float(d)) factor of the generating host (rsqrtss + Newton-Raphson;
             rsqrtss is CPU-specific, see csrc/mvs_norms.cpp) so that the test can tell whether it runs on a
             host with the same approximation table,
and `family_*`: the head of a live run of the reference BINARY (project_everything sketch) on a 300-sample
family hash file at d = 2048 -- vectors.bin and vector_norms.txt as written by sketch() itself.
"""
import ctypes as C
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_util import oracle, write_family_hash_file          # noqa: E402

REFDIR = os.path.join(ROOT, "oracle", "_ref", "v3")
DIMS = (1, 2, 3, 4, 5, 7, 8, 9, 11, 12, 13, 16, 63, 64, 65, 100, 1000, 2048, 4100)


def main():
    ref = C.CDLL(os.path.join(REFDIR, "libref_projection.so"))
    ref.ref_norms.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
    orc = oracle()
    rng = np.random.default_rng(20261105)
    out = {}
    for d in DIMS:
        rows = np.concatenate([rng.integers(-s, s + 1, size=(12, d)) for s in (1, 40, 3000, 2 ** 21, 2 ** 30)]).astype(np.int32)
        rows[0] = 0
        nr = np.empty(len(rows), np.float64)
        ref.ref_norms(rows.ctypes.data, len(rows), d, nr.ctypes.data)
        assert np.array_equal(nr.astype(np.float32).astype(np.float64), nr)
        out["rows_%d" % d] = rows
        out["norms_%d" % d] = nr.astype(np.float32)
        out["inv_%d" % d] = np.array([orc.oracle_inv_sqrt_d(d)], np.float32).view(np.uint32)
    with tempfile.TemporaryDirectory() as td:
        hf = os.path.join(td, "fam.txt")
        write_family_hash_file(hf, 300, seed=9, mean_size=150)
        db = os.path.join(td, "db")
        subprocess.run([os.path.join(REFDIR, "project_everything"), "sketch", hf, db, "-d", "2048"], check=True,
                       stdout=subprocess.DEVNULL)
        out["family_vectors"] = np.fromfile(os.path.join(db, "vectors.bin"), dtype=np.int32).reshape(300, 2048).astype(np.int16)
        assert np.array_equal(out["family_vectors"].astype(np.int32).ravel(), np.fromfile(os.path.join(db, "vectors.bin"), dtype=np.int32))
        out["family_norms_txt"] = np.array(open(os.path.join(db, "vector_norms.txt")).read())
    np.savez_compressed(os.path.join(HERE, "norms_ref.npz"), **out)
    print("wrote norms_ref.npz:", {k: getattr(v, "shape", None) for k, v in out.items() if k.startswith(("family", "inv_2048"))})


if __name__ == "__main__":
    main()
