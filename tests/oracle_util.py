"""Loader for the CPU oracle (oracle/liboracle.so) and, when prebuilt, the compiled
reference (oracle/_ref/<variant>/libref_projection.so).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
GOLDEN = os.path.join(ROOT, "tests", "golden")

_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        so = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(so):
            subprocess.run(["make", "-C", ORACLE_DIR, "oracle"], check=True, stdout=subprocess.DEVNULL)
        lib = C.CDLL(so)
        lib.oracle_transform_set.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_uint64, C.c_void_p]
        lib.oracle_project_csr.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_uint64, C.c_void_p]
        lib.oracle_saturate_int16.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
        lib.oracle_norm_exact.argtypes = [C.c_void_p, C.c_int]
        lib.oracle_norm_exact.restype = C.c_double
        lib.oracle_sumsq.argtypes = [C.c_void_p, C.c_int]
        lib.oracle_sumsq.restype = C.c_uint64
        lib.oracle_splitmix64.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64]
        lib.oracle_splitmix64.restype = C.c_uint64
        lib.oracle_norm_reference.argtypes = [C.c_void_p, C.c_int, C.c_float]
        lib.oracle_norm_reference.restype = C.c_float
        lib.oracle_inv_sqrt_d.argtypes = [C.c_int]
        lib.oracle_inv_sqrt_d.restype = C.c_float
        _oracle = lib
    return _oracle


def ref_dir():
    """Directory of a reference build that runs on this host (native first, then v3), or None."""
    for variant in ("native", "v3"):
        d = os.path.join(ORACLE_DIR, "_ref", variant)
        exe = os.path.join(d, "standalone_projection")
        if not os.path.exists(exe):
            continue
        try:
            r = subprocess.run([exe], capture_output=True, timeout=20)
            if r.returncode == 1:          # usage + exit 1 (standalone_projection.cpp:12-15)
                return d
        except Exception:
            pass
    return None


def ref_lib():
    global _ref
    if _ref is None:
        d = ref_dir()
        if d is None:
            return None
        lib = C.CDLL(os.path.join(d, "libref_projection.so"))
        lib.ref_transform_set.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
        lib.ref_project_csr.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p,
                                        C.POINTER(C.c_double)]
        lib.ref_norms.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
        _ref = lib
    return _ref


def oracle_set(hashes, d, seed=0):
    h = np.ascontiguousarray(hashes, dtype=np.uint64)
    out = np.empty(d, dtype=np.int32)
    oracle().oracle_transform_set(h.ctypes.data, len(h), d, C.c_uint64(seed), out.ctypes.data)
    return out


def oracle_csr(hashes, offsets, d, seed=0):
    h = np.ascontiguousarray(hashes, dtype=np.uint64)
    o = np.ascontiguousarray(offsets, dtype=np.uint64)
    n = len(o) - 1
    out = np.empty((n, d), dtype=np.int32)
    oracle().oracle_project_csr(h.ctypes.data, o.ctypes.data, n, d, C.c_uint64(seed), out.ctypes.data)
    return out


def ref_set(hashes, d):
    h = np.ascontiguousarray(hashes, dtype=np.uint64)
    out = np.empty(d, dtype=np.int32)
    ref_lib().ref_transform_set(h.ctypes.data, len(h), d, out.ctypes.data)
    return out


def load_toy():
    z = np.load(os.path.join(GOLDEN, "toy_hashes.npz"))
    return [str(n) for n in z["names"]], z["offsets"].astype(np.uint64), z["hashes"].astype(np.uint64)


def load_toy_ref(tag):
    z = np.load(os.path.join(GOLDEN, "toy_ref_%s.npz" % tag))
    return {k: (z[k] if z[k].shape else z[k].item()) for k in z.files}


def load_kat():
    return json.load(open(os.path.join(GOLDEN, "kat.json")))


def parse_line_like_reference(text):
    """`while (iss >> hash)` for unsigned long (project_everything.cpp:275-279,
    standalone_projection.cpp:30-34), character by character: skip whitespace, optional sign, at
    least one digit (else the loop ends), magnitude > 2^64-1 ends the loop without inserting,
    '-N' wraps modulo 2^64; the next extraction starts right after the last digit, so "5x" gives
    [5] and "5-3" gives [5, 2^64-3].  Small-case oracle for the parser tests, pinned by kat.json."""
    out = []
    i, n = 0, len(text)
    ws = " \t\n\v\f\r"
    while True:
        while i < n and text[i] in ws:
            i += 1
        if i >= n:
            break
        neg = False
        if text[i] in "+-":
            neg = text[i] == "-"
            i += 1
        j = i
        while j < n and text[j] in "0123456789":
            j += 1
        if j == i:
            break
        v = int(text[i:j])
        i = j
        if v > 0xFFFFFFFFFFFFFFFF:
            break
        out.append((-v) & 0xFFFFFFFFFFFFFFFF if neg else v)
    return out


# ---- pairwise comparison (reference src/pairwise_comp_optimized.cpp:57-160) ----
_ref_pw = None


def _pairwise_call(fn, a, b, d, ni, nj, with_seconds):
    a = np.ascontiguousarray(a, dtype=np.int32)
    b = np.ascontiguousarray(b, dtype=np.int32)
    ni = np.ascontiguousarray(ni, dtype=np.float64)
    nj = np.ascontiguousarray(nj, dtype=np.float64)
    mi, mj = a.shape[0], b.shape[0]
    cap = mi * mj
    rows = np.empty(cap, dtype=np.int32)
    cols = np.empty(cap, dtype=np.int32)
    vals = np.empty(cap, dtype=np.int64)
    args = [a.ctypes.data, mi, b.ctypes.data, mj, d, ni.ctypes.data, nj.ctypes.data, rows.ctypes.data,
            cols.ctypes.data, vals.ctypes.data, cap]
    sec = C.c_double(0)
    if with_seconds:
        args.append(C.byref(sec))
    n = fn(*args)
    return rows[:n].copy(), cols[:n].copy(), vals[:n].copy(), sec.value


def oracle_pairwise(a, b, d, ni, nj):
    lib = oracle()
    lib.oracle_pairwise_block.restype = C.c_int64
    lib.oracle_pairwise_block.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
    return _pairwise_call(lib.oracle_pairwise_block, a, b, d, ni, nj, False)[:3]


def oracle_quantize_jaccard(value, d, norm_row, norm_col):
    lib = oracle()
    lib.oracle_quantize_jaccard.restype = C.c_uint16
    lib.oracle_quantize_jaccard.argtypes = [C.c_int64, C.c_int, C.c_double, C.c_double]
    return int(lib.oracle_quantize_jaccard(int(value), int(d), float(norm_row), float(norm_col)))


def ref_pairwise_lib():
    """The unmodified reference function behind oracle/ref_pairwise_shim.cpp, or None if not built."""
    global _ref_pw
    if _ref_pw is None:
        d = ref_dir()
        if d is None or not os.path.exists(os.path.join(d, "libref_pairwise.so")):
            return None
        lib = C.CDLL(os.path.join(d, "libref_pairwise.so"))
        lib.ref_pairwise_block.restype = C.c_int64
        lib.ref_pairwise_block.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                           C.POINTER(C.c_double)]
        _ref_pw = lib
    return _ref_pw


def ref_pairwise(a, b, d, ni, nj):
    return _pairwise_call(ref_pairwise_lib().ref_pairwise_block, a, b, d, ni, nj, True)


def norms_sq_like_reference(vectors, d):
    """all_norms of pairwise_comp_optimized.cpp:889-901: the norm PRINTED by sketch() with 6 significant
    digits (project_everything.cpp:327-330), parsed back with stod and squared."""
    v = np.asarray(vectors, dtype=np.float64)
    norm = np.sqrt((v * v).sum(axis=1) / float(d))
    return np.array([float("%g" % x) ** 2 for x in norm], dtype=np.float64)


# ---- the norm column of vector_norms.txt (reference src/project_everything.cpp:328-330) ----
def oracle_reference_norms(vectors, inv=None):
    """float32 norms of (n, d) int32 rows by oracle_norm_reference; inv defaults to this host's rsqrtss value."""
    v = np.ascontiguousarray(vectors, dtype=np.int32)
    n, d = v.shape
    lib = oracle()
    if inv is None:
        inv = lib.oracle_inv_sqrt_d(d)
    return np.array([lib.oracle_norm_reference(v[i].ctypes.data, d, C.c_float(inv)) for i in range(n)], dtype=np.float32)


def ref_norms_live(vectors):
    """The reference's two statements compiled with its own Eigen and flags (oracle/ref_shim.cpp), float32 widened."""
    v = np.ascontiguousarray(vectors, dtype=np.int32)
    n, d = v.shape
    out = np.empty(n, dtype=np.float64)
    ref_lib().ref_norms(v.ctypes.data, n, d, out.ctypes.data)
    return out


def load_norms_golden():
    z = np.load(os.path.join(GOLDEN, "norms_ref.npz"))
    return {k: z[k] for k in z.files}


def golden_host_has_same_rsqrt(d, golden=None):
    """rsqrtss is an approximation whose table is the CPU's: the committed norm TEXT fixtures were printed on the
    build container's CPU and only bind hosts whose 1/sqrt(d) factor has the same bits."""
    g = golden or load_norms_golden()
    key = "inv_%d" % d
    if key not in g:
        return False
    mine = np.array([oracle().oracle_inv_sqrt_d(d)], np.float32).view(np.uint32)[0]
    return int(mine) == int(g[key][0])


def norms_text(names, norms_f32):
    """vector_norms.txt as sketch() prints it: default ostream formatting of the float norm widened to double."""
    return "".join("%s %s\n" % (nm, "%g" % float(np.float64(x))) for nm, x in zip(names, norms_f32))


def write_family_hash_file(path, n, seed=1, mean_size=200, family=10):
    """A hash file whose samples come in families sharing most of their hashes (so that the all-pairs
    comparison keeps several neighbours per row), sizes ragged, one empty and one tiny sample included.
    Returns (names, offsets, hashes) with hashes sorted + unique per sample."""
    rng = np.random.default_rng(seed)
    names, sets = [], []
    base = None
    for i in range(n):
        if i % family == 0:
            size = max(8, int(rng.lognormal(np.log(mean_size), 0.6)))
            base = rng.integers(0, 2 ** 54, size=size, dtype=np.uint64)
        keep = base[rng.random(len(base)) < rng.uniform(0.55, 0.95)]
        extra = rng.integers(0, 2 ** 54, size=int(len(base) * rng.uniform(0.02, 0.3)), dtype=np.uint64)
        s = np.unique(np.concatenate([keep, extra]))
        if i == 7:
            s = s[:0]
        if i == 8:
            s = s[:3]
        names.append("fam%05d" % i)
        sets.append(s)
    with open(path, "w") as f:
        for nm, s in zip(names, sets):
            f.write(nm + ":" + "".join(" %d" % int(h) for h in s) + "\n")
    offsets = np.zeros(n + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum([len(s) for s in sets])
    return names, offsets, (np.concatenate(sets) if sets else np.zeros(0, np.uint64))
