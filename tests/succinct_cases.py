"""Seeded inputs for the shard-file container tests (tests/test_succinct.py, tests/golden/make_golden_succinct.py)."""
import numpy as np


def compact_cases():
    """name -> uint64 values for compact_vector::build(begin, n) (width from the maximum)."""
    rng = np.random.default_rng(20240607)
    cases = {}
    for w in list(range(1, 65)):
        hi = (1 << w) - 1
        for n in (1, 2, 63, 64, 65, 1000):
            v = rng.integers(0, hi, n, dtype=np.uint64, endpoint=True)
            if w <= 53:                                   # widths follow ceil(log2(max + 1)) in double precision:
                v[rng.integers(0, n)] = hi                # force the full width where the double is exact
            cases["w%d_n%d" % (w, n)] = v
    for k in (1, 2, 7, 8, 15, 16, 31, 32, 33, 52):
        cases["pow2_%d" % k] = np.array([3, 1 << k, 0], dtype=np.uint64)           # max = 2^k needs k + 1 bits
        cases["pow2m1_%d" % k] = np.array([(1 << k) - 1, 0, 1], dtype=np.uint64)
    cases["zeros"] = np.zeros(77, dtype=np.uint64)
    cases["one_zero"] = np.zeros(1, dtype=np.uint64)
    cases["jaccard_codes"] = rng.integers(0, 256, 5000, dtype=np.uint64)            # what matrix.bin holds per row
    cases["row_ids"] = np.sort(rng.choice(1 << 20, 30000, replace=False)).astype(np.uint64)
    return cases


def rice_cases():
    """name -> uint64 values for rice_sequence<>::encode(begin, n) (Rice parameter by Kiely's rule)."""
    rng = np.random.default_rng(977)
    cases = {}
    for mean in (0.2, 1, 3, 10, 100, 1e3, 1e4, 1e6, 1e9):
        for n in (1, 2, 31, 32, 33, 1023, 1024, 1025, 5000):
            cases["geom_%g_n%d" % (mean, n)] = rng.geometric(1.0 / (1.0 + mean), n).astype(np.uint64) - np.uint64(1)
    cases["zeros"] = np.zeros(3000, dtype=np.uint64)                               # p = 1: parameter 0, one bit each
    cases["one_zero"] = np.zeros(1, dtype=np.uint64)
    cases["ones"] = np.ones(2049, dtype=np.uint64)
    cases["gaps_of_columns"] = np.diff(np.sort(rng.choice(1 << 22, 40000, replace=False))).astype(np.uint64)
    cases["first_neighbours"] = rng.integers(0, 1 << 20, 20000, dtype=np.uint64)
    # sparse super-blocks of the select index: 1024 ones spread over >= 2^16 bits while the global parameter stays 0
    v = np.zeros(100000, dtype=np.uint64)
    v[-1024:] = 100
    cases["sparse_tail_block"] = v
    v = np.zeros(120000, dtype=np.uint64)
    v[3000:4024] = 90
    v[70000:70500] = 200
    cases["sparse_middle_blocks"] = v
    cases["single_large"] = np.array([123456789012], dtype=np.uint64)
    cases["uniform_u32"] = rng.integers(0, 1 << 32, 7000, dtype=np.uint64)
    return cases
