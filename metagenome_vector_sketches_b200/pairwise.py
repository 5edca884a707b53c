"""Host-side mirror of the reference's all-pairs comparison interface over the C ABI.

reference                                                   here
--------------------------------------------------------   -----------------------------------------------
compute_sparse_dot_products_optimized(block_i, block_j,     compute_sparse_dot_products(sk, block_i, block_j,
    norms_i, norms_j, dimension) -> SparseResult                norms_i, norms_j) -> SparseResult
  src/pairwise_comp_optimized.cpp:57-160
block loops of main over one shard (:936-983)               pairwise_shard(sk, vectors, norms_sq, begin, end)
all_norms: stod(printed norm)^2 (:889-901)                  norms_squared_from_db(norm_lines)

Blocks are numpy int32 [m, d] (row-major = the memory image of the reference's column-major d x m MatrixXi).
All arithmetic happens in libmvs_b200.so on the GPU; numpy is only the container."""
import ctypes as C
from collections import namedtuple

import numpy as np

from ._lib import check

SparseResult = namedtuple("SparseResult", "rows cols values gemm_ms filter_ms prep_ms wall_ms h2d_bytes d2h_bytes")


class _CSparse(C.Structure):
    _fields_ = [("n", C.c_int64), ("rows", C.POINTER(C.c_int32)), ("cols", C.POINTER(C.c_int32)),
                ("values", C.POINTER(C.c_int64)), ("gemm_ms", C.c_double), ("filter_ms", C.c_double),
                ("prep_ms", C.c_double), ("wall_ms", C.c_double), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64)]


def _take(lib, ptr):
    r = C.cast(ptr, C.POINTER(_CSparse)).contents
    n = int(r.n)
    if n:
        rows = np.ctypeslib.as_array(r.rows, shape=(n,)).copy()
        cols = np.ctypeslib.as_array(r.cols, shape=(n,)).copy()
        vals = np.ctypeslib.as_array(r.values, shape=(n,)).copy()
    else:
        rows, cols, vals = np.empty(0, np.int32), np.empty(0, np.int32), np.empty(0, np.int64)
    out = SparseResult(rows, cols, vals, r.gemm_ms, r.filter_ms, r.prep_ms, r.wall_ms, int(r.h2d_bytes), int(r.d2h_bytes))
    lib.mvs_free_sparse_result(ptr)
    return out


def compute_sparse_dot_products(sk, block_i, block_j, norms_i, norms_j):
    """One block pair on the GPU of Sketcher `sk`; pairs in (row, col) order like the reference's loop."""
    a = np.ascontiguousarray(block_i, dtype=np.int32)
    b = np.ascontiguousarray(block_j, dtype=np.int32)
    if a.ndim != 2 or b.ndim != 2 or a.shape[1] != b.shape[1]:
        raise ValueError("blocks must be [m, d] with the same d")
    ni = np.ascontiguousarray(norms_i, dtype=np.float64)
    nj = np.ascontiguousarray(norms_j, dtype=np.float64)
    if len(ni) != a.shape[0] or len(nj) != b.shape[0]:
        raise ValueError("one squared norm per vector")
    out = C.c_void_p()
    check(sk._lib.mvs_pairwise_block(sk._h, a.ctypes.data, a.shape[0], b.ctypes.data, b.shape[0], a.shape[1],
                                     ni.ctypes.data, nj.ctypes.data, C.byref(out)), "mvs_pairwise_block")
    return _take(sk._lib, out)


def pairwise_shard(sk, vectors, norms_sq, begin_row=0, end_row=None):
    """Rows [begin_row, end_row) of the db against all vectors; global indices, (row, col) order."""
    v = np.ascontiguousarray(vectors, dtype=np.int32)
    if v.ndim != 2:
        raise ValueError("vectors must be [n, d]")
    nrm = np.ascontiguousarray(norms_sq, dtype=np.float64)
    if len(nrm) != v.shape[0]:
        raise ValueError("one squared norm per vector")
    end_row = v.shape[0] if end_row is None else int(end_row)
    out = C.c_void_p()
    check(sk._lib.mvs_pairwise_shard(sk._h, v.ctypes.data, v.shape[0], v.shape[1], nrm.ctypes.data, int(begin_row),
                                     end_row, C.byref(out)), "mvs_pairwise_shard")
    return _take(sk._lib, out)


def norms_squared_from_db(norm_lines):
    """all_norms of pairwise_comp_optimized.cpp:889-901 from the lines of vector_norms.txt."""
    out = []
    for line in norm_lines:
        pos = line.find(" ")
        if pos < 0:
            continue
        x = float(line[pos + 1:])
        out.append(x * x)
    return np.array(out, dtype=np.float64)
