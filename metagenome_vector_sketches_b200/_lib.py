"""ctypes binding of include/mvs_b200.h (the C-ABI drop-in boundary).

Nothing here computes: every function forwards to libmvs_b200.so, which is built
in-tree by `__graft_entry__.build()` / `make -C metagenome_vector_sketches_b200/csrc`.
If the library is missing the import of any compute entry point fails loudly --
there is no Python or CPU fallback.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBNAME = "libmvs_b200.so"
_EXPLIBNAME = "libmvs_b200_exp.so"      # -DMVS_EXPERIMENTS build (round-1 A/B kernels and knobs); tools/ and one test module only
_lib = None
_explib = None


class MvsError(RuntimeError):
    def __init__(self, status, what, cuda_error=0):
        self.status = status
        self.cuda_error = cuda_error
        super().__init__("%s failed: mvs status %d (%s)%s" % (
            what, status, _strerror(status),
            ", cudaError %d" % cuda_error if cuda_error else ""))


class Timing(C.Structure):
    _fields_ = [("kernel_ms", C.c_float), ("reduce_ms", C.c_float),
                ("calls", C.c_uint32), ("reserved", C.c_uint32),
                ("launches", C.c_uint64), ("warp_tasks", C.c_uint64),
                ("partial_rows", C.c_uint64), ("h2d_bytes", C.c_uint64),
                ("d2h_bytes", C.c_uint64)]


class HashBatch(C.Structure):
    _fields_ = [("n_samples", C.c_int64), ("n_hashes", C.c_uint64), ("hashes", C.POINTER(C.c_uint64)),
                ("offsets", C.POINTER(C.c_uint64)), ("names", C.POINTER(C.c_char_p)), ("hashes_pinned", C.c_int)]


# name -> (restype, argtypes); must list every symbol of include/mvs_b200.h
SIGNATURES = {
    "mvs_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "mvs_destroy": (None, [C.c_void_p]),
    "mvs_strerror": (C.c_char_p, [C.c_int]),
    "mvs_last_cuda_error": (C.c_int, [C.c_void_p]),
    "mvs_abi_version": (C.c_int, []),
    "mvs_transform_set": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_uint64, C.c_void_p]),
    "mvs_project_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_uint64, C.c_void_p]),
    "mvs_sketch_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_uint64, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p]),
    "mvs_sketch_batch_packed56": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_uint64, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_void_p]),
    "mvs_pack56": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p]),
    "mvs_sketch_batch_multi": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int,
                                         C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "mvs_project_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_uint64,
                                     C.c_void_p, C.c_void_p]),
    "mvs_project_device_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_uint64,
                                            C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]),
    "mvs_finalize_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p]),
    "mvs_finalize": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "mvs_reference_norms": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_void_p]),
    "mvs_reference_inv_sqrt_d": (C.c_float, [C.c_int]),
    "mvs_set_norms_mode": (C.c_int, [C.c_int]),
    "mvs_get_norms_mode": (C.c_int, []),
    "mvs_parse_hash_file": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(C.POINTER(HashBatch))]),
    "mvs_free_hash_batch": (None, [C.POINTER(HashBatch)]),
    "mvs_write_hash_batch": (C.c_int, [C.POINTER(HashBatch), C.c_char_p]),
    "mvs_write_db": (C.c_int, [C.c_char_p, C.c_int64, C.c_int, C.POINTER(C.c_char_p), C.c_void_p, C.c_void_p,
                               C.c_void_p, C.c_void_p]),
    "mvs_partition_samples": (None, [C.c_void_p, C.c_int64, C.c_int, C.c_void_p]),
    "mvs_sketch_file": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_char_p, C.c_char_p, C.c_int, C.c_int,
                                  C.c_uint64, C.c_int]),
    "mvs_sketch_file_on_devices": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.c_char_p, C.c_char_p, C.c_int, C.c_int,
                                             C.c_uint64, C.c_int]),
    "mvs_pairwise_block": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_void_p,
                                     C.c_void_p, C.c_void_p]),
    "mvs_pairwise_shard": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64, C.c_int64,
                                     C.c_void_p]),
    "mvs_free_sparse_result": (None, [C.c_void_p]),
    "mvs_standalone_file": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_uint64, C.c_char_p]),
    "mvs_convert_signatures": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int, C.c_int]),
    "mvs_sort_unique": (C.c_int64, [C.c_void_p, C.c_int64]),
    "mvs_host_alloc": (C.c_void_p, [C.c_size_t]),
    "mvs_host_alloc_write_combined": (C.c_void_p, [C.c_size_t]),
    "mvs_host_free": (None, [C.c_void_p]),
    "mvs_host_register": (C.c_int, [C.c_void_p, C.c_size_t]),
    "mvs_host_unregister": (C.c_int, [C.c_void_p]),
    "mvs_bind_host_to_device": (C.c_int, [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "mvs_get_timing": (C.c_int, [C.c_void_p, C.POINTER(Timing)]),
    "mvs_int_issue_peak": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_double)]),
    "mvs_plan_tasks": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                 C.POINTER(C.c_uint64), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]),
    "mvs_loop_mix_peak": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "mvs_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "mvs_has_experiments": (C.c_int, []),
}


def library_path():
    return os.path.join(_HERE, _LIBNAME)


def _bind(path):
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    return lib


def load_library():
    """Load the in-tree native library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise ImportError(
            "%s is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C metagenome_vector_sketches_b200/csrc`).  There is no CPU fallback." % path)
    _lib = _bind(path)
    return _lib


def load_experiments_library():
    """The experiment build (make -C csrc experiments), for tools/ and tests/test_gpu_experiments.py only."""
    global _explib
    if _explib is None:
        path = os.environ.get("MVS_EXP_LIB") or os.path.join(_HERE, _EXPLIBNAME)
        if not os.path.exists(path):
            raise ImportError("%s is not built: make -C metagenome_vector_sketches_b200/csrc experiments" % path)
        _explib = _bind(path)
        assert _explib.mvs_has_experiments() == 1
    return _explib


def _strerror(status):
    try:
        return load_library().mvs_strerror(int(status)).decode()
    except Exception:  # pragma: no cover
        return "?"


def check(status, what, ctx=None):
    if status != 0:
        cuda = load_library().mvs_last_cuda_error(ctx) if ctx else 0
        raise MvsError(status, what, cuda)
