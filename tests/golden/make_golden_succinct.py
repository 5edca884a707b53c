"""Golden digests of the reference's `bits` containers for the seeded inputs of tests/succinct_cases.py.

Run in the build container (needs oracle/_ref/v3/libref_bits.so = the reference's own headers behind a C wrapper,
`make -C oracle ref`):   python tests/golden/make_golden_succinct.py
Writes tests/golden/succinct_ref.json: case -> [serialised length, num_bytes(), sha256 of the serialised bytes].
"""
import ctypes as C
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from succinct_cases import compact_cases, rice_cases      # noqa: E402


def encode(fn, v):
    v = np.ascontiguousarray(v, dtype=np.uint64)
    nb = C.c_uint64(0)
    need = fn(v.ctypes.data_as(C.c_void_p), len(v), None, 0, C.byref(nb))
    buf = np.empty(need, dtype=np.uint8)
    got = fn(v.ctypes.data_as(C.c_void_p), len(v), buf.ctypes.data_as(C.c_void_p), need, C.byref(nb))
    assert got == need
    return buf.tobytes(), int(nb.value)


def bind(lib, name):
    fn = getattr(lib, name)
    fn.restype = C.c_uint64
    fn.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
    return fn


if __name__ == "__main__":
    lib = C.CDLL(os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle", "_ref", "v3", "libref_bits.so"))
    out = {"compact_vector": {}, "rice_sequence": {}}
    for kind, fn, cases in (("compact_vector", bind(lib, "ref_compact_vector"), compact_cases()),
                            ("rice_sequence", bind(lib, "ref_rice_sequence"), rice_cases())):
        for name, v in cases.items():
            data, nb = encode(fn, v)
            out[kind][name] = [len(data), nb, hashlib.sha256(data).hexdigest()]
    with open(os.path.join(HERE, "succinct_ref.json"), "w") as f:
        json.dump(out, f, indent=0, sort_keys=True)
    print("wrote %d + %d cases" % (len(out["compact_vector"]), len(out["rice_sequence"])))
