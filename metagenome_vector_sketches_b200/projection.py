"""Host-side mirror of the reference's projection interface over the C ABI.

reference                                              here
---------------------------------------------------   ---------------------------------------
transform_set_into_vector(set, d)                      transform_set_into_vector(hashes, d)
  include/random_projection.h:11                          (dedupes like the unordered_set)
sample loop of sketch()                                Sketcher.project_csr / project_sets
  src/project_everything.cpp:289-298
norm + int16 epilogue of sketch()                      Sketcher.finalize
  src/project_everything.cpp:327-347

All arithmetic happens in libmvs_b200.so on the GPU; numpy is only the container.
"""
import ctypes as C
import math

import numpy as np

from . import _lib
from ._lib import Timing, check, load_library, load_experiments_library


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Sketcher:
    """One context = one GPU (mvs_create).  Not thread-safe; one per process/thread."""

    def __init__(self, device=0, experiments=False):
        """experiments=True binds the context to the experiment build (libmvs_b200_exp.so) whose set_option knows the
        round-1 A/B knobs; the default is the shipped library."""
        self._lib = load_experiments_library() if experiments else load_library()
        h = C.c_void_p()
        check(self._lib.mvs_create(C.byref(h), int(device)), "mvs_create")
        self._h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.mvs_destroy(self._h)
            self._h = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- hot path -------------------------------------------------------------------
    def project_csr(self, hashes, offsets, d, seed=0, out=None):
        """CSR batch (host numpy arrays) -> (n_samples, d) int32.  Hashes must be unique per sample."""
        hashes = np.ascontiguousarray(hashes, dtype=np.uint64)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = len(offsets) - 1
        if n < 0:
            raise ValueError("offsets needs n_samples+1 entries")
        if n and int(offsets[-1]) > len(hashes):
            raise ValueError("offsets index past the end of hashes")
        if out is None:
            out = np.empty((n, d), dtype=np.int32)
        assert out.dtype == np.int32 and out.flags.c_contiguous and out.size == n * d
        check(self._lib.mvs_project_batch(self._h, _ptr(hashes), _ptr(offsets), n, int(d),
                                          C.c_uint64(seed), _ptr(out)), "mvs_project_batch", self._h)
        return out

    def project_csr_packed56(self, packed, offsets, d, seed=0, out=None):
        """mvs_sketch_batch_packed56: the CSR payload as 7 little-endian bytes per hash (see pack56)."""
        packed = np.ascontiguousarray(packed, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = len(offsets) - 1
        if n and 7 * int(offsets[-1]) > len(packed):
            raise ValueError("offsets index past the end of the packed hashes")
        if out is None:
            out = np.empty((n, d), dtype=np.int32)
        assert out.dtype == np.int32 and out.flags.c_contiguous and out.size == n * d
        check(self._lib.mvs_sketch_batch_packed56(self._h, _ptr(packed), _ptr(offsets), n, int(d), C.c_uint64(seed),
                                                  _ptr(out), None, None, None), "mvs_sketch_batch_packed56", self._h)
        return out

    def project_sets(self, sets, d, seed=0):
        """List of iterables of hashes -> (len(sets), d) int32.  Dedupes each set (reference: unordered_set)."""
        uniq = [np.unique(np.asarray(list(s) if not isinstance(s, np.ndarray) else s, dtype=np.uint64))
                for s in sets]
        offsets = np.zeros(len(uniq) + 1, dtype=np.uint64)
        if uniq:
            offsets[1:] = np.cumsum([len(u) for u in uniq])
        hashes = np.concatenate(uniq) if uniq else np.zeros(0, dtype=np.uint64)
        return self.project_csr(hashes, offsets, d, seed)

    def transform_set(self, hashes, d, seed=0):
        h = np.unique(np.asarray(hashes, dtype=np.uint64))
        out = np.empty(d, dtype=np.int32)
        check(self._lib.mvs_transform_set(self._h, _ptr(h), len(h), int(d), C.c_uint64(seed), _ptr(out)),
              "mvs_transform_set", self._h)
        return out

    def project_device(self, d_hashes_ptr, h_offsets, d, d_out_ptr, seed=0, stream=None):
        """Device-resident batch.  d_hashes_ptr/d_out_ptr are raw device addresses (e.g. tensor.data_ptr())."""
        h_offsets = np.ascontiguousarray(h_offsets, dtype=np.uint64)
        n = len(h_offsets) - 1
        check(self._lib.mvs_project_device(self._h, C.c_void_p(d_hashes_ptr), _ptr(h_offsets), n, int(d),
                                           C.c_uint64(seed), C.c_void_p(d_out_ptr),
                                           C.c_void_p(stream) if stream else None),
              "mvs_project_device", self._h)

    def project_device_static(self, d_hashes_ptr, h_offsets, d, d_out_ptr, seed=0, stream=None):
        """mvs_project_device_static: project_device with the scratch in the library's static device array -- no
        allocation in the call.  Returns the scratch bytes the call needed; MvsError (status -7) beyond the limit."""
        h_offsets = np.ascontiguousarray(h_offsets, dtype=np.uint64)
        n = len(h_offsets) - 1
        needed = C.c_uint64(0)
        check(self._lib.mvs_project_device_static(self._h, C.c_void_p(d_hashes_ptr), _ptr(h_offsets), n, int(d),
                                                  C.c_uint64(seed), C.c_void_p(d_out_ptr),
                                                  C.c_void_p(stream) if stream else None, C.byref(needed)),
              "mvs_project_device_static", self._h)
        return int(needed.value)

    def finalize_device(self, d_vec_ptr, n, d, d_sumsq_ptr, d_out16_ptr=None, stream=None, d_ref_norms_ptr=None):
        check(self._lib.mvs_finalize_device(self._h, C.c_void_p(d_vec_ptr), int(n), int(d),
                                            C.c_void_p(d_sumsq_ptr) if d_sumsq_ptr else None,
                                            C.c_void_p(d_out16_ptr) if d_out16_ptr else None,
                                            C.c_void_p(d_ref_norms_ptr) if d_ref_norms_ptr else None,
                                            C.c_void_p(stream) if stream else None),
              "mvs_finalize_device", self._h)

    def finalize(self, vec, want_int16=False, want_ref_norms=False):
        """(n, d) int32 host array -> (sumsq uint64[n], int16 copy or None[, reference float32 norms])."""
        vec = np.ascontiguousarray(vec, dtype=np.int32)
        n, d = vec.shape
        sumsq = np.empty(n, dtype=np.uint64)
        out16 = np.empty((n, d), dtype=np.int16) if want_int16 else None
        ref = np.empty(n, dtype=np.float32) if want_ref_norms else None
        check(self._lib.mvs_finalize(self._h, _ptr(vec), n, d, _ptr(sumsq), _ptr(out16), _ptr(ref)),
              "mvs_finalize", self._h)
        return (sumsq, out16, ref) if want_ref_norms else (sumsq, out16)

    def sketch_csr(self, hashes, offsets, d, seed=0, int16=False):
        """mvs_sketch_batch: projection + epilogue.  Returns dict(vec | vec16, sumsq, ref_norms)."""
        return sketch_csr_multi([self], hashes, offsets, d, seed, int16)

    # ---- measurement ----------------------------------------------------------------
    def timing(self):
        t = Timing()
        check(self._lib.mvs_get_timing(self._h, C.byref(t)), "mvs_get_timing", self._h)
        return {k: getattr(t, k) for k, _ in Timing._fields_}

    def int_issue_peak(self, which=0):
        v = C.c_double()
        check(self._lib.mvs_int_issue_peak(self._h, int(which), C.byref(v)), "mvs_int_issue_peak", self._h)
        return v.value

    def loop_mix_peak(self):
        v = C.c_double()
        check(self._lib.mvs_loop_mix_peak(self._h, C.byref(v)), "mvs_loop_mix_peak", self._h)
        return v.value

    def set_option(self, key, value):
        check(self._lib.mvs_set_option(self._h, key.encode(), int(value)), "mvs_set_option", self._h)


def sketch_csr_multi(sketchers, hashes, offsets, d, seed=0, int16=False):
    """mvs_sketch_batch_multi: ONE call over several contexts (one per GPU; they may also share a device).
    Returns dict(vec | vec16, sumsq, ref_norms); rows are in input order."""
    lib = sketchers[0]._lib
    hashes = np.ascontiguousarray(hashes, dtype=np.uint64)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    n = len(offsets) - 1
    vec = None if int16 else np.empty((n, d), dtype=np.int32)
    vec16 = np.empty((n, d), dtype=np.int16) if int16 else None
    sumsq = np.empty(n, dtype=np.uint64)
    ref = np.empty(n, dtype=np.float32)
    handles = (C.c_void_p * len(sketchers))(*[s._h for s in sketchers])
    check(lib.mvs_sketch_batch_multi(handles, len(sketchers), _ptr(hashes), _ptr(offsets), n, int(d), C.c_uint64(seed),
                                     _ptr(vec), _ptr(vec16), _ptr(sumsq), _ptr(ref)), "mvs_sketch_batch_multi",
          sketchers[0]._h)
    out = {"sumsq": sumsq, "ref_norms": ref}
    out["vec16" if int16 else "vec"] = vec16 if int16 else vec
    return out


def pack56(hashes, out=None):
    """mvs_pack56: uint64 hashes (all < 2^56) -> 7 little-endian bytes each."""
    h = np.ascontiguousarray(hashes, dtype=np.uint64)
    if out is None:
        out = np.empty(7 * len(h), dtype=np.uint8)
    check(load_library().mvs_pack56(_ptr(h), len(h), _ptr(out)), "mvs_pack56")
    return out


def reference_norms(vec):
    """mvs_reference_norms: the reference's float32 norm of (n, d) int32 rows on the HOST (no GPU needed)."""
    vec = np.ascontiguousarray(vec, dtype=np.int32)
    n, d = vec.shape
    out = np.empty(n, dtype=np.float32)
    check(load_library().mvs_reference_norms(_ptr(vec), n, d, _ptr(out)), "mvs_reference_norms")
    return out


_default = {}


def transform_set_into_vector(hashes, d, seed=0, device=0):
    """Drop-in for the reference function (include/random_projection.h:11): set -> int32[d]."""
    sk = _default.get(device)
    if sk is None:
        sk = _default[device] = Sketcher(device)
    return sk.transform_set(hashes, d, seed)


def bind_host_to_device(device):
    """mvs_bind_host_to_device: pin the calling host thread (CPU affinity + preferred memory node) to the NUMA
    node of `device`; call it once per rank before allocating pinned buffers.  Returns (numa_node, applied)."""
    node, applied = C.c_int(-1), C.c_int(0)
    check(load_library().mvs_bind_host_to_device(int(device), C.byref(node), C.byref(applied)), "mvs_bind_host_to_device")
    return node.value, applied.value


def norm_from_sumsq(sumsq, d):
    """norm written to vector_norms.txt: ||v / sqrt(d)|| (reference project_everything.cpp:327-329)."""
    return math.sqrt(float(int(sumsq)) / float(d))


def format_norm(x):
    """Default ostream formatting of a double: 6 significant digits, %g style (project_everything.cpp:330)."""
    return "%g" % x


def pinned_empty(n, dtype, write_combined=False):
    """numpy array over pinned host memory from mvs_host_alloc (kept alive by the array's base);
    write_combined: mvs_host_alloc_write_combined (input buffers the host only writes)."""
    lib = load_library()
    dt = np.dtype(dtype)
    nbytes = max(1, int(n) * dt.itemsize)
    p = (lib.mvs_host_alloc_write_combined if write_combined else lib.mvs_host_alloc)(nbytes)
    if not p:
        raise MemoryError("mvs_host_alloc(%d)" % nbytes)
    buf = (C.c_char * nbytes).from_address(p)

    class _Owner:
        def __init__(self, ptr):
            self.ptr = ptr

        def __del__(self):
            try:
                lib.mvs_host_free(self.ptr)
            except Exception:  # pragma: no cover
                pass

    arr = np.frombuffer(buf, dtype=dt, count=int(n))
    holder = _PinnedArray(arr, _Owner(p))
    return holder


class _PinnedArray:
    """Tiny holder: `.a` is the numpy view, the pinned allocation lives as long as this object."""

    def __init__(self, a, owner):
        self.a = a
        self._owner = owner
