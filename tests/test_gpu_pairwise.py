"""GPU parity of the all-pairs comparison through the C ABI, against the oracle
(oracle/mvs_oracle.c oracle_pairwise_block, pinned to the reference function by tests/test_oracle.py)
and, where the compiled reference travels with the repo, against the reference function itself.
Bit-exact: dot products are int32 arithmetic modulo 2^32, the keep-test is integer division + one
double comparison, the order is the reference's (row, col) loop order."""
import numpy as np
import pytest

import metagenome_vector_sketches_b200 as mvs
from metagenome_vector_sketches_b200 import pairwise as pw
from oracle_util import (load_toy_ref, norms_sq_like_reference, oracle_pairwise, ref_pairwise, ref_pairwise_lib)

pytestmark = pytest.mark.gpu


# The shipped library has ONE path per shape class: tcgen05 CTA pairs (cta_group::2, 256 x 128 tiles) when a block has
# two tile rows, a single CTA otherwise, keep-test fused into the GEMM epilogue, limb count from the data (the cases
# below reach 2, 3 and 4 limbs and both shape classes through their inputs).  The round-1 A/B variants (forced limbs,
# clusters with TMA multicast, separate count pass, mma.sync and IMAD GEMMs) live in the experiment build and are
# covered by tests/test_gpu_experiments.py, which re-runs every case of this module on them.
@pytest.fixture(scope="module")
def sk():
    s = mvs.Sketcher(0)
    yield s
    s.close()


def same(got, want):
    return all(np.array_equal(a, b) for a, b in zip(got[:3], want[:3]))


def toy():
    ref = load_toy_ref("d2048")
    v = ref["vectors"].reshape(-1, 2048)
    return v, pw.norms_squared_from_db(ref["norms_txt"].strip().split("\n"))


def test_toy_db_pairs_match_oracle_and_reference(sk):
    v, nrm = toy()
    got = pw.compute_sparse_dot_products(sk, v, v, nrm, nrm)
    want = oracle_pairwise(v, v, 2048, nrm, nrm)
    assert len(want[0]) == 1291 and same(got, want)
    if ref_pairwise_lib() is not None:
        assert same(got, ref_pairwise(v, v, 2048, nrm, nrm))
    sh = pw.pairwise_shard(sk, v, nrm)
    assert same(sh, want)
    part = pw.pairwise_shard(sk, v, nrm, 20, 45)                       # a shard of rows: global indices
    keep = (want[0] >= 20) & (want[0] < 45)
    assert same(part, [w[keep] for w in want])


@pytest.mark.parametrize("mi,mj,d", [(1, 1, 1), (3, 5, 7), (37, 53, 300), (128, 128, 64), (129, 127, 2048),
                                     (200, 333, 100), (257, 1, 4096), (64, 300, 18)])
def test_ragged_blocks_all_dimensions(sk, mi, mj, d):
    rng = np.random.default_rng(mi * 1000 + mj + d)
    a = rng.integers(-60, 61, (mi, d), dtype=np.int64).astype(np.int32)
    b = rng.integers(-60, 61, (mj, d), dtype=np.int64).astype(np.int32)
    k = min(mi, mj) // 2
    b[:k] = a[:k]                                                      # related vectors pass the threshold
    if k > 2:
        b[1] = -a[1]
        b[2] = a[2] // 2
    ni, nj = norms_sq_like_reference(a, d), norms_sq_like_reference(b, d)
    got = pw.compute_sparse_dot_products(sk, a, b, ni, nj)
    assert same(got, oracle_pairwise(a, b, d, ni, nj))


def test_int32_wraparound_is_the_references(sk):
    """Eigen's MatrixXi product wraps modulo 2^32; with tiny norms every positive wrapped value is kept,
    so the triples expose the wrapped dot products themselves."""
    rng = np.random.default_rng(5)
    for scale in (40000, 2 ** 20, 2 ** 31 - 1):
        a = rng.integers(-scale, scale + 1, (70, 96), dtype=np.int64).astype(np.int32)
        b = rng.integers(-scale, scale + 1, (45, 96), dtype=np.int64).astype(np.int32)
        a[0] = np.int32(-2 ** 31)
        ni, nj = np.zeros(70), np.zeros(45)
        got = pw.compute_sparse_dot_products(sk, a, b, ni, nj)
        want = oracle_pairwise(a, b, 96, ni, nj)
        assert len(want[0]) > 500 and same(got, want)
        if ref_pairwise_lib() is not None:
            assert same(got, ref_pairwise(a, b, 96, ni, nj))


def test_threshold_edges_and_negative_quotients(sk):
    # dot / d truncates toward zero: -1 / 8 == 0 > threshold only if threshold < 0 (negative "norms")
    a = np.array([[1, 0, 0, 0, 0, 0, 0, 0], [8, 0, 0, 0, 0, 0, 0, 0], [-1, 0, 0, 0, 0, 0, 0, 0]], dtype=np.int32)
    b = np.array([[1, 0, 0, 0, 0, 0, 0, 0], [7, 0, 0, 0, 0, 0, 0, 0], [9, 0, 0, 0, 0, 0, 0, 0]], dtype=np.int32)
    for ni, nj in ((np.zeros(3), np.zeros(3)), (np.full(3, -1.0), np.zeros(3)), (np.full(3, 10.0), np.full(3, 10.0)),
                   (np.array([0.0, 20.0, 0.0]), np.array([0.0, 0.0, 100.0])),
                   # thresholds the division-free test does not cover: NaN, +-inf, >= 2^31, -0.0
                   (np.array([np.nan, np.inf, -np.inf]), np.zeros(3)), (np.array([1e12, 4.2e10, -0.0]), np.array([-0.0, 4.3e10, -0.0])),
                   (np.array([-0.0, 4.2949e10, 4.29496e10]), np.array([-0.0, 0.0, 1.0]))):
        got = pw.compute_sparse_dot_products(sk, a, b, ni, nj)
        assert same(got, oracle_pairwise(a, b, 8, ni, nj))


def test_empty_blocks_and_errors(sk):
    z = np.zeros((0, 16), dtype=np.int32)
    v = np.ones((4, 16), dtype=np.int32)
    assert len(pw.compute_sparse_dot_products(sk, z, v, np.zeros(0), np.zeros(4)).rows) == 0
    assert len(pw.compute_sparse_dot_products(sk, v, z, np.zeros(4), np.zeros(0)).rows) == 0
    assert len(pw.pairwise_shard(sk, v, np.zeros(4), 2, 2).rows) == 0
    with pytest.raises(mvs.MvsError):
        pw.pairwise_shard(sk, v, np.zeros(4), 3, 9)
    with pytest.raises(ValueError):
        pw.compute_sparse_dot_products(sk, v, np.ones((4, 8), dtype=np.int32), np.zeros(4), np.zeros(4))


def test_shard_of_projected_synthetic_db_matches_blockwise_oracle(sk):
    """End to end on our own projection output: 600 samples in 30 families sharing most hashes (so that
    neighbours exist), d = 2048; the shard call (several GEMM row blocks) equals the oracle."""
    rng = np.random.default_rng(11)
    fams = [rng.integers(0, 2 ** 54, 3000, dtype=np.uint64) for _ in range(30)]
    sets = []
    for s in range(600):
        base = fams[s % 30]
        keepn = int(rng.integers(1500, 3000))
        extra = rng.integers(0, 2 ** 54, int(rng.integers(10, 800)), dtype=np.uint64)
        sets.append(np.unique(np.concatenate([base[:keepn], extra])))
    vec = sk.project_sets(sets, 2048)
    nrm = norms_sq_like_reference(vec, 2048)
    got = pw.pairwise_shard(sk, vec, nrm)
    want = oracle_pairwise(vec, vec, 2048, nrm, nrm)
    assert len(want[0]) > 600 * 5 and same(got, want)
    assert got.gemm_ms > 0 and got.h2d_bytes == vec.nbytes + 8 * 600


def test_more_triples_than_the_initial_output_buffers(sk):
    """2200 identical vectors: every pair is kept (4.84 M triples > the 4 Mi the device buffers start with), so the
    emit kernel first declines, the host grows the buffers and launches it again."""
    n, d = 2200, 8
    v = np.tile(np.arange(1, d + 1, dtype=np.int32) * 3, (n, 1))
    nrm = norms_sq_like_reference(v, d)
    got = pw.compute_sparse_dot_products(sk, v, v, nrm, nrm)
    want = oracle_pairwise(v, v, d, nrm, nrm)
    assert len(want[0]) == n * n and same(got, want)


def test_workspace_can_be_released_between_calls(sk):
    """The context keeps its device buffers between calls; option release_workspace frees them and the next call
    simply allocates again."""
    v, nrm = toy()
    want = oracle_pairwise(v, v, 2048, nrm, nrm)
    assert same(pw.pairwise_shard(sk, v, nrm), want)
    sk.set_option("release_workspace", 1)
    assert same(pw.pairwise_shard(sk, v, nrm), want)
    assert same(pw.compute_sparse_dot_products(sk, v[:40], v, nrm[:40], nrm), [w[want[0] < 40] for w in want])


def test_full_size_shard_properties():
    """10 000 vectors x d = 2048 (two GEMM row blocks, CTA pairs, fused pre-filter) is too big for the oracle; check what
    does not depend on size: the kept set is symmetric with equal values, the diagonal is kept with value = sum v^2,
    rows and columns come out in (row, col) order, every kept value equals an int64 numpy dot product on a sample, the
    keep-test holds for all kept pairs and fails for a sample of dropped ones, and a 300-row shard equals the oracle."""
    n, d = 10000, 2048
    rng = np.random.default_rng(5)
    fam = np.rint(rng.normal(0.0, 190.0, ((n + 19) // 20, d)))
    v = (fam[np.arange(n) // 20] + np.rint(rng.normal(0.0, 110.0, (n, d)))).astype(np.int32)
    nrm = norms_sq_like_reference(v, d)
    with mvs.Sketcher(0) as s:
        r = pw.pairwise_shard(s, v, nrm)
        part = pw.pairwise_shard(s, v, nrm, 4990, 5290)
    rows, cols, vals = r.rows.astype(np.int64), r.cols.astype(np.int64), r.values
    key = rows * n + cols
    assert np.all(np.diff(key) > 0)                                                  # (row, col) order, no duplicates
    tkey = cols * n + rows
    order = np.argsort(tkey)
    assert np.array_equal(tkey[order], key) and np.array_equal(vals[order], vals)    # symmetric
    diag = rows == cols
    assert diag.sum() == n and np.array_equal(vals[diag], (v.astype(np.int64) ** 2).sum(axis=1))
    pick = rng.choice(len(rows), 2000, replace=False)
    exact = np.einsum("ij,ij->i", v[rows[pick]].astype(np.int64), v[cols[pick]].astype(np.int64))
    assert np.array_equal(vals[pick], exact)                                         # |dot| < 2^31 here: no wrap
    q = np.trunc(vals / d)                                                           # the reference's keep-test on every kept pair
    assert np.all(q > 0.05 * (nrm[rows] + nrm[cols]))
    ri, ci = rng.integers(0, n, 20000), rng.integers(0, n, 20000)
    dots = np.einsum("ij,ij->i", v[ri].astype(np.int64), v[ci].astype(np.int64))
    keep = np.trunc(dots / d) > 0.05 * (nrm[ri] + nrm[ci])
    assert np.array_equal(keep, np.isin(ri * n + ci, key))                           # dropped pairs really fail it
    want = oracle_pairwise(v[4990:5290], v, d, nrm[4990:5290], nrm)
    assert same(part, (want[0] + 4990, want[1], want[2]))


def test_seeded_fuzz_of_shards(sk):
    """25 random dbs: any n, d, value scale (1 .. 4 limbs, wrapped products), shard range and a few impossible norms
    (zero, negative, NaN), through mvs_pairwise_shard against the oracle on the same rows."""
    rng = np.random.default_rng(20261104)
    for case in range(25):
        n, d = int(rng.integers(1, 700)), int(rng.integers(1, 600))
        scale = float(rng.choice([3.0, 60.0, 900.0, 2.0e4, 3.0e6, 7.0e8]))
        v = np.clip(np.rint(rng.normal(0.0, scale, (n, d))), -2 ** 31, 2 ** 31 - 1).astype(np.int32)
        if n > 4:
            v[rng.integers(0, n, n // 3)] = v[int(rng.integers(0, n))]                 # duplicates: pairs that are kept
        nrm = norms_sq_like_reference(v, d) * float(rng.choice([1.0, 1.0, 0.3, 3.0]))
        for k in rng.integers(0, n, 3):
            nrm[k] = float(rng.choice([0.0, -1.0, -1e9, np.nan, 1e300]))
        b = int(rng.integers(0, n))
        e = int(rng.integers(b, n + 1))
        got = pw.pairwise_shard(sk, v, nrm, b, e)
        want = oracle_pairwise(v[b:e], v, d, nrm[b:e], nrm)
        assert same(got, (want[0] + b, want[1], want[2])), (case, n, d, scale, b, e)
