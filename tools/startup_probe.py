"""startup_probe.py -- where the cold-start second of the CLI goes: cuInit, context + mvs_create,
first call (module load, allocations), second call.  Fresh process per measurement.  python tools/startup_probe.py"""
import ctypes as C
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import ctypes as C, json, sys, time
sys.path.insert(0, %r)
t0 = time.perf_counter()
cuda = C.CDLL("libcuda.so.1")
t1 = time.perf_counter()
rc = cuda.cuInit(0)
t2 = time.perf_counter()
from metagenome_vector_sketches_b200 import _lib
lib = _lib.load_library()
t3 = time.perf_counter()
h = C.c_void_p()
assert lib.mvs_create(C.byref(h), 0) == 0
t4 = time.perf_counter()
import numpy as np
hs = np.arange(1, 1001, dtype=np.uint64)
out = np.empty(2048, dtype=np.int32)
def call():
    t = time.perf_counter()
    assert lib.mvs_transform_set(h, hs.ctypes.data_as(C.c_void_p), 1000, 2048, C.c_uint64(0), out.ctypes.data_as(C.c_void_p)) == 0
    return time.perf_counter() - t
c1 = call(); c2 = call()
p = lib.mvs_host_alloc(1 << 30)
t5 = time.perf_counter()
p2 = lib.mvs_host_alloc(1 << 30)
t6 = time.perf_counter()
print(json.dumps({"dlopen_libcuda_s": t1 - t0, "cuInit_s": t2 - t1, "dlopen_libmvs_s": t3 - t2, "mvs_create_s": t4 - t3,
                  "first_call_s": c1, "second_call_s": c2, "pinned_alloc_1GiB_s": t6 - t5}))
''' % ROOT

if __name__ == "__main__":
    runs = []
    for _ in range(3):
        r = subprocess.run([sys.executable, "-c", CHILD], capture_output=True, text=True)
        if r.returncode != 0:
            print(json.dumps({"error": r.stderr[-400:]}))
            sys.exit(1)
        runs.append(json.loads(r.stdout.strip().split("\n")[-1]))
    print(json.dumps({"runs": runs}))
