"""The in-tree writers of the all-pairs shard containers (csrc/mvs_succinct.h) against the reference's vendored `bits`
library: byte for byte against golden digests made from the reference's headers (tests/golden/succinct_ref.json, see
make_golden_succinct.py) and, where oracle/_ref/v3/libref_bits.so exists, against those headers live on fresh random
inputs.  Host code only; the harness is compiled here with the host compiler."""
import ctypes as C
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

from succinct_cases import compact_cases, rice_cases

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REFLIB = os.path.join(ROOT, "oracle", "_ref", "v3", "libref_bits.so")
SIG = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]


def _bind(lib, name):
    fn = getattr(lib, name)
    fn.restype = C.c_uint64
    fn.argtypes = SIG
    return fn


@pytest.fixture(scope="module")
def ours(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("succinct") / "libsuccinct_harness.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-Wall", "-Werror", "-fPIC", "-shared",
                    os.path.join(HERE, "succinct_harness.cpp"), "-o", so], check=True)
    lib = C.CDLL(so)
    return {k: _bind(lib, "mvs_" + k) for k in ("compact_vector", "compact_vector_u16", "rice_sequence")}


def _encode(fn, v, dtype=np.uint64):
    v = np.ascontiguousarray(v, dtype=dtype)
    nb = C.c_uint64(0)
    need = fn(v.ctypes.data_as(C.c_void_p), len(v), None, 0, C.byref(nb))
    buf = np.empty(need, dtype=np.uint8)
    assert fn(v.ctypes.data_as(C.c_void_p), len(v), buf.ctypes.data_as(C.c_void_p), need, C.byref(nb)) == need
    return buf.tobytes(), int(nb.value)


def test_containers_match_the_golden_digests_of_the_reference_headers(ours):
    with open(os.path.join(HERE, "golden", "succinct_ref.json")) as f:
        gold = json.load(f)
    for kind, cases in (("compact_vector", compact_cases()), ("rice_sequence", rice_cases())):
        assert sorted(cases) == sorted(gold[kind])
        for name, v in cases.items():
            data, nb = _encode(ours[kind], v)
            assert [len(data), nb, hashlib.sha256(data).hexdigest()] == gold[kind][name], (kind, name)


def test_compact_vector_layout_restated(ours):
    """The format itself, independent of any library: 3 header words, the word count, value i at bit i * width."""
    v = np.array([5, 0, 7, 1, 6, 2, 3, 4] * 9, dtype=np.uint64)                    # 72 values of 3 bits = 216 bits
    data, nb = _encode(ours["compact_vector"], v)
    w = np.frombuffer(data, dtype=np.uint64)
    assert list(w[:4]) == [72, 3, 7, 5] and len(w) == 4 + 5 and nb == len(data)    # ceil(216 / 64) + 1 spare word
    bits = int.from_bytes(w[4:].tobytes(), "little")
    assert [(bits >> (3 * i)) & 7 for i in range(72)] == list(v)
    data16, _ = _encode(ours["compact_vector_u16"], v, np.uint16)                  # the u16 instantiation (Jaccard codes)
    assert data16 == data


def test_rice_sequence_decodes_back(ours):
    """Round trip through a decoder written from the format description: unary high parts in the bit vector, low parts
    of `width` bits in the compact vector, select index consistent with the bit vector."""
    rng = np.random.default_rng(5)
    for v in (rng.geometric(0.01, 3000).astype(np.uint64), np.zeros(10, dtype=np.uint64),
              rice_cases()["sparse_tail_block"]):
        data, nb = _encode(ours["rice_sequence"], v)
        assert nb == len(data)
        w = np.frombuffer(data[:len(data) // 8 * 8], dtype=np.uint64)
        num_bits, nwords = int(w[0]), int(w[1])
        high = int.from_bytes(w[2:2 + nwords].tobytes(), "little")
        at = 8 * (2 + nwords)
        ones = int.from_bytes(data[at:at + 8], "little")
        assert ones == len(v) and bin(high).count("1") == ones and high >> num_bits == 0
        nblk = int.from_bytes(data[at + 8:at + 16], "little")
        assert nblk == (ones + 1023) // 1024
        blocks = np.frombuffer(data[at + 16:at + 16 + 8 * nblk], dtype=np.int64)
        at += 16 + 8 * nblk
        nsub = int.from_bytes(data[at:at + 8], "little")
        assert nsub == sum((min(1024, ones - 1024 * b) + 31) // 32 for b in range(nblk))
        at += 8 + 2 * nsub
        nover = int.from_bytes(data[at:at + 8], "little")
        assert nover == sum(min(1024, ones - 1024 * b) for b in range(nblk) if blocks[b] < 0)
        at += 8 + 8 * nover
        size, width, mask, lw = (int.from_bytes(data[at + 8 * k:at + 8 * k + 8], "little") for k in range(4))
        assert size == len(v) and mask == (1 << width) - 1 and lw == (size * width + 63) // 64 + 1
        low = int.from_bytes(data[at + 32:at + 32 + 8 * lw], "little")
        got, pos = [], 0
        for i in range(len(v)):
            run = 0
            while not (high >> pos) & 1:
                run += 1
                pos += 1
            pos += 1
            got.append((run << width) | ((low >> (i * width)) & mask))
        assert got == [int(x) for x in v]


@pytest.mark.skipif(not os.path.exists(REFLIB), reason="oracle/_ref/v3/libref_bits.so not built (make -C oracle ref)")
def test_containers_match_the_reference_headers_live(ours):
    ref = C.CDLL(REFLIB)
    refs = {"compact_vector": _bind(ref, "ref_compact_vector"), "rice_sequence": _bind(ref, "ref_rice_sequence")}
    rng = np.random.default_rng()                                                  # fresh inputs every run
    for _ in range(300):
        n = int(rng.integers(1, 4000))
        w = int(rng.integers(1, 50))
        v = rng.integers(0, 1 << w, n, dtype=np.uint64)
        assert _encode(ours["compact_vector"], v) == _encode(refs["compact_vector"], v), (n, w)
        if rng.random() < 0.5:
            v = rng.geometric(1.0 / (1.0 + 2.0 ** rng.uniform(-3, 30)), n).astype(np.uint64) - np.uint64(1)
        if rng.random() < 0.2:
            v[rng.integers(0, n, max(1, n // 50))] = np.uint64(rng.integers(0, 1 << 20))   # heavy tail: sparse blocks
        assert _encode(ours["rice_sequence"], v) == _encode(refs["rice_sequence"], v), (n, w)
