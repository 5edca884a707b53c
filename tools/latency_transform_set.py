"""Per-call latency of the one-sample entry point mvs_transform_set (what the a1-signature shim calls once per sample):
run on the GPU box.  Prints microseconds per call for several set sizes."""
import ctypes as C
import sys
import time

import numpy as np

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import metagenome_vector_sketches_b200 as mvs
from metagenome_vector_sketches_b200.projection import _ptr, check

sk = mvs.Sketcher(0)
rng = np.random.default_rng(3)
d = 2048
out = np.empty(d, dtype=np.int32)
for n in (100, 5000, 50000, 500000, 5000000):
    h = np.unique(rng.integers(0, 2 ** 54, n, dtype=np.uint64))
    call = lambda: check(sk._lib.mvs_transform_set(sk._h, _ptr(h), len(h), d, C.c_uint64(0), _ptr(out)), "mvs_transform_set", sk._h)
    for _ in range(5):
        call()
    reps = 200 if n <= 50000 else 20
    t = time.perf_counter()
    for _ in range(reps):
        call()
    us = (time.perf_counter() - t) / reps * 1e6
    print("%8d hashes: %9.1f us per call = %8.1f M hashes/s" % (len(h), us, len(h) / us))
