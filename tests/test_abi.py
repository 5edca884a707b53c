"""CPU tests of the drop-in boundary: the library loads without a GPU, exports every
symbol include/mvs_b200.h declares, and fails loudly (no CPU fallback) when asked to compute."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import metagenome_vector_sketches_b200 as mvs
from metagenome_vector_sketches_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "mvs_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mvs_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported_and_bound():
    lib = mvs.load_library()
    names = _declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), "libmvs_b200.so does not export %s" % n
        assert n in _lib.SIGNATURES, "ctypes binding misses %s" % n
    assert sorted(_lib.SIGNATURES) == names
    assert lib.mvs_abi_version() >= 1


def test_every_entry_point_is_documented_for_integrators():
    """INTEGRATION.md names the reference interface each entry point replaces (or says that it has none)."""
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    text = text.replace("`mvs_finalize(_device)`", "`mvs_finalize` `mvs_finalize_device`")
    missing = [n for n in _declared_symbols() if n not in text]
    assert not missing, missing
    assert lib_status_texts_are_distinct()


def lib_status_texts_are_distinct():
    lib = mvs.load_library()
    texts = [lib.mvs_strerror(c) for c in range(0, -8, -1)]
    return len(set(texts)) == 8 and b"unknown" not in b" ".join(texts)


def test_strerror_and_argument_checks_need_no_gpu():
    lib = mvs.load_library()
    assert lib.mvs_strerror(0) == b"ok"
    assert b"no CPU fallback" in lib.mvs_strerror(-2)
    assert lib.mvs_create(None, 0) == -1
    need = C.c_uint64(7)
    assert lib.mvs_project_device_static(None, None, None, 0, 2048, C.c_uint64(0), None, None, C.byref(need)) == -1
    assert need.value == 0                                   # cleared even when the call is refused
    assert b"MVS_STATIC_SCRATCH_BYTES" in lib.mvs_strerror(-7)
    assert lib.mvs_write_hash_batch(None, b"/nonexistent/x") == -1
    a = np.array([5, 3, 5, 1, 3], dtype=np.uint64)
    n = lib.mvs_sort_unique(a.ctypes.data, len(a))
    assert n == 3 and a[:3].tolist() == [1, 3, 5]


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(mvs.MvsError) as e:
        mvs.Sketcher(0)
    assert e.value.status == -2          # MVS_ERR_NO_DEVICE
    with pytest.raises(mvs.MvsError):
        mvs.transform_set_into_vector([1, 2, 3], 64)


def test_product_does_not_reference_oracle():
    pkg = os.path.join(ROOT, "metagenome_vector_sketches_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle/" not in txt and "liboracle" not in txt and "oracle_util" not in txt, f


def test_synthetic_generator_is_injective_and_in_range():
    h, off = mvs.synthetic.uniform_batch(64, 1000, seed=0)
    assert len(np.unique(h)) == len(h) and int(h.max()) < 2 ** 54
    h2, _ = mvs.synthetic.uniform_batch(64, 1000, seed=1)
    assert not np.array_equal(h, h2)
    sizes = mvs.synthetic.lognormal_sizes(2000, seed=0)
    assert sizes.min() >= 1000 and sizes.max() <= 10 ** 7 and 1e4 < np.median(sizes) < 4e4
    hr, offr = mvs.synthetic.ragged_batch([3, 0, 5], seed=0, first_sample=10)
    assert offr.tolist() == [0, 3, 3, 8] and len(np.unique(hr)) == 8
    assert np.array_equal(hr[:3], mvs.synthetic.synth_hashes([10, 10, 10], [0, 1, 2], 0))


def test_streaming_copy_handles_every_alignment_and_tail():
    """The copy that fills the pinned staging ring (streaming stores after an aligned head) is a memcpy."""
    lib = mvs.load_library()
    fn = getattr(lib, "_Z27mvs_internal_copy_streamingPvPKvm")          # internal C++ symbol, not part of the C ABI
    fn.restype, fn.argtypes = None, [C.c_void_p, C.c_void_p, C.c_size_t]
    rng = np.random.default_rng(3)
    src = rng.integers(0, 256, 4096 + 700, dtype=np.uint8)
    for doff in (0, 1, 31, 32, 33, 63):
        for soff in (0, 5, 64):
            for n in (0, 1, 31, 127, 128, 129, 1000, 4096):
                dst = np.full(4096 + 800, 0xAB, dtype=np.uint8)
                fn(dst.ctypes.data + doff, src.ctypes.data + soff, n)
                assert np.array_equal(dst[doff:doff + n], src[soff:soff + n])
                assert (dst[:doff] == 0xAB).all() and (dst[doff + n:] == 0xAB).all()
