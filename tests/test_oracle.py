"""CPU tests: pin the oracle (oracle/mvs_oracle.c) against the reference's own outputs.

Golden fixtures in tests/golden were produced by the unmodified reference binaries
(tests/golden/make_golden.py); when oracle/_ref is prebuilt it is also called live."""
import hashlib

import numpy as np
import pytest

from oracle_util import (oracle, oracle_set, oracle_csr, ref_lib, ref_set, load_toy, load_toy_ref, load_kat,
                         parse_line_like_reference, oracle_reference_norms, ref_norms_live, load_norms_golden,
                         golden_host_has_same_rsqrt, norms_text)


def test_splitmix_kat1():
    # splitmix64(0) = 0xE220A8397B1DCDAF
    assert oracle().oracle_splitmix64(0, 0, 0) == 0xE220A8397B1DCDAF


def test_standalone_kats():
    kat = load_kat()
    for case in kat["standalone"]:
        d = case["d"]
        got_lines = case["stdout"].split("\n")[:-1]
        assert len(got_lines) == len(case["lines"])
        for line, ref_out in zip(case["lines"], got_lines):
            hs = sorted(set(parse_line_like_reference(line)))
            vec = oracle_set(np.array(hs, dtype=np.uint64), d)
            # standalone_projection prints static_cast<float>(v) with default ostream formatting
            assert " ".join("%g" % float(v) for v in vec) == ref_out


def test_sketch_parser_kat():
    kat = load_kat()["sketch_parser"]
    d = kat["d"]
    rows = []
    for line in kat["text"].split("\n"):
        if ":" not in line:
            continue
        rest = line[line.index(":") + 1:]
        hs = sorted(set(parse_line_like_reference(rest)))
        rows.append(oracle_set(np.array(hs, dtype=np.uint64), d))
    assert np.array_equal(np.concatenate(rows), np.array(kat["vectors"], dtype=np.int32))


@pytest.mark.parametrize("tag,d", [("d2048", 2048), ("d100", 100)])
def test_toy_golden(tag, d):
    names, offsets, hashes = load_toy()
    ref = load_toy_ref(tag)
    got = oracle_csr(hashes, offsets, d)
    assert np.array_equal(got, ref["vectors"].astype(np.int32))
    assert hashlib.sha256(got.tobytes()).hexdigest() == ref["vectors_sha256"]
    if d == 2048:
        assert ref["vectors_sha256"] == "93fb4358e2774ad31cd50fe8f00a1eb6a97aa0ad029f166654bbf3385df74cb8"
        assert got[0, :8].tolist() == [88, -118, 22, 30, -10, -22, 56, -72]


def test_toy_norms_within_print_precision():
    names, offsets, hashes = load_toy()
    ref = load_toy_ref("d2048")
    got = oracle_csr(hashes, offsets, 2048)
    lines = ref["norms_txt"].split("\n")[:-1]
    assert [l.rsplit(" ", 1)[0] for l in lines] == names
    for row, line in zip(got, lines):
        exact = oracle().oracle_norm_exact(row.ctypes.data, 2048)
        printed = float(line.rsplit(" ", 1)[1])
        # reference prints a float32 Eigen norm with 6 significant digits
        # half a unit of the 6th significant digit + 2e-6 relative for the float32 evaluation
        ulp6 = 10.0 ** (np.floor(np.log10(exact)) - 5) if exact > 0 else 0.0
        assert abs(printed - exact) <= 0.5 * ulp6 + 2e-6 * exact


def test_reference_norm_restatement_matches_golden_rows():
    """oracle_norm_reference against norms the reference's own expression produced (tests/golden/
    make_golden_norms.py), bit for bit, with the generating host's 1/sqrt(d) factor: every packet-count
    case of the Eigen reduction (d < 4, one packet, odd / even packet counts, scalar tails) and magnitudes
    up to 2^30."""
    g = load_norms_golden()
    dims = sorted(int(k[5:]) for k in g if k.startswith("rows_"))
    assert dims[0] == 1 and 2048 in dims and len(dims) >= 15
    for d in dims:
        inv = g["inv_%d" % d].view(np.float32)[0]
        got = oracle_reference_norms(g["rows_%d" % d], inv)
        assert np.array_equal(got.view(np.uint32), g["norms_%d" % d].view(np.uint32)), d


def test_reference_norm_restatement_prints_the_reference_binary_file():
    """vectors.bin + vector_norms.txt written by the reference BINARY (sketch(), 300-sample family file):
    the oracle's norm, printed like an ostream prints a double, reproduces the file byte for byte."""
    g = load_norms_golden()
    vec = g["family_vectors"].astype(np.int32)
    inv = g["inv_2048"].view(np.float32)[0]
    names = ["fam%05d" % i for i in range(len(vec))]
    assert norms_text(names, oracle_reference_norms(vec, inv)) == str(g["family_norms_txt"])
    # and the toy db of SURVEY KAT-3 (61/61 lines), the file the round-1 verdict found one digit off
    names, _, _ = load_toy()
    for tag, d in (("d2048", 2048), ("d100", 100)):
        ref = load_toy_ref(tag)
        inv = g["inv_%d" % d].view(np.float32)[0]
        assert norms_text(names, oracle_reference_norms(ref["vectors"].astype(np.int32), inv)) == ref["norms_txt"]


def test_reference_norm_restatement_matches_live_reference():
    if ref_lib() is None:
        pytest.skip("oracle/_ref not built on this host")
    rng = np.random.default_rng(77)
    for d in (1, 3, 4, 6, 8, 10, 12, 15, 16, 17, 100, 257, 2048, 2051):
        for scale in (1, 50, 5000, 2 ** 24):
            v = rng.integers(-scale, scale + 1, size=(25, d)).astype(np.int32)
            live = ref_norms_live(v)
            assert np.array_equal(oracle_reference_norms(v).astype(np.float64), live), (d, scale)


def test_int16_golden_matches_saturation():
    ref32 = load_toy_ref("d2048")["vectors"].astype(np.int32)
    ref16 = load_toy_ref("d2048_i16")["vectors"]
    out = np.empty(ref32.size, dtype=np.int16)
    oracle().oracle_saturate_int16(ref32.ctypes.data, ref32.size, out.ctypes.data)
    assert np.array_equal(out.reshape(ref32.shape), ref16)
    big = np.array([40000, -40000, 32767, -32768, 32768, -32769, 0], dtype=np.int32)
    o = np.empty(big.size, dtype=np.int16)
    oracle().oracle_saturate_int16(big.ctypes.data, big.size, o.ctypes.data)
    assert o.tolist() == [32767, -32768, 32767, -32768, 32767, -32768, 0]


def test_properties():
    rng = np.random.default_rng(7)
    a = np.unique(rng.integers(0, 2 ** 63, 300, dtype=np.uint64))
    b = np.unique(rng.integers(2 ** 63, 2 ** 64 - 1, 200, dtype=np.uint64))
    d = 200
    va, vb = oracle_set(a, d), oracle_set(b, d)
    assert np.array_equal(oracle_set(np.concatenate([a, b]), d), va + vb)        # linearity over disjoint unions
    assert np.array_equal(oracle_set(a, 2048)[:d], va)                           # prefix property
    one = np.array([123456789], dtype=np.uint64)
    assert np.array_equal(oracle_set(one, 256)[64:128], oracle_set(one + np.uint64(64), 64))   # block shift
    assert not oracle_set(np.zeros(0, dtype=np.uint64), 77).any()                # empty set
    assert np.array_equal(oracle_set(a, d, seed=5), oracle_set(a + np.uint64(5), d))           # additive seed


def test_live_reference_if_built():
    if ref_lib() is None:
        pytest.skip("oracle/_ref not built on this host")
    rng = np.random.default_rng(11)
    for n, d in [(0, 64), (1, 1), (17, 63), (500, 2048), (1000, 130), (3, 4160)]:
        h = np.unique(rng.integers(0, 2 ** 64 - 1, n, dtype=np.uint64))
        assert np.array_equal(ref_set(h, d), oracle_set(h, d))


# ---- pairwise comparison: oracle pinned to the reference function ----------------
def _pairwise_cases():
    import importlib.util
    import os
    from oracle_util import GOLDEN
    spec = importlib.util.spec_from_file_location("mgp", os.path.join(GOLDEN, "make_golden_pairwise.py"))
    mgp = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mgp)
    z = np.load(os.path.join(GOLDEN, "pairwise_ref.npz"))
    return mgp, z


def test_pairwise_oracle_matches_reference_goldens():
    from oracle_util import load_toy_ref, oracle_pairwise
    mgp, z = _pairwise_cases()
    ref = load_toy_ref("d2048")
    v = ref["vectors"].reshape(-1, 2048)
    nrm = np.array([float(l.rsplit(" ", 1)[1]) ** 2 for l in ref["norms_txt"].strip().split("\n")])
    r, c, val = oracle_pairwise(v, v, 2048, nrm, nrm)
    assert len(r) == 1291
    assert np.array_equal(r, z["toy_rows"]) and np.array_equal(c, z["toy_cols"]) and np.array_equal(val, z["toy_values"])
    for k, (seed, mi, mj, d, scale, zn) in enumerate(z["specs"].tolist()):
        a, b, ni, nj = mgp.random_case(seed, mi, mj, d, scale, bool(zn))
        r, c, val = oracle_pairwise(a, b, d, ni, nj)
        assert len(r) > 0
        assert np.array_equal(r, z["r%d_rows" % k]) and np.array_equal(c, z["r%d_cols" % k])
        assert np.array_equal(val, z["r%d_values" % k])


def test_pairwise_oracle_matches_live_reference_if_built():
    from oracle_util import oracle_pairwise, ref_pairwise, ref_pairwise_lib, norms_sq_like_reference
    if ref_pairwise_lib() is None:
        pytest.skip("oracle/_ref/*/libref_pairwise.so not built")
    rng = np.random.default_rng(3)
    for mi, mj, d, scale in ((20, 31, 64, 100), (50, 50, 130, 3000), (33, 17, 48, 2 ** 31 - 1)):
        a = rng.integers(-scale, scale + 1, (mi, d), dtype=np.int64).astype(np.int32)
        b = rng.integers(-scale, scale + 1, (mj, d), dtype=np.int64).astype(np.int32)
        b[:10] = a[:10]
        for ni, nj in ((norms_sq_like_reference(a, d), norms_sq_like_reference(b, d)), (np.zeros(mi), np.zeros(mj))):
            want = ref_pairwise(a, b, d, np.nan_to_num(ni), np.nan_to_num(nj))
            got = oracle_pairwise(a, b, d, np.nan_to_num(ni), np.nan_to_num(nj))
            assert all(np.array_equal(x, y) for x, y in zip(got, want[:3]))


def test_jaccard_quantisation_matches_query_goldens():
    """pairwise_comp_optimized.cpp:654-665 on the toy db reproduces the 8-bit values behind KAT-4
    (tests/golden/pairwise_top10.json: DRR000821 vs DRR000822 -> 0.317647 = 81/255)."""
    from oracle_util import load_toy, load_toy_ref, oracle_quantize_jaccard
    names, _, _ = load_toy()
    z = np.load(__import__("os").path.join(__import__("oracle_util").GOLDEN, "pairwise_ref.npz"))
    ref = load_toy_ref("d2048")
    nrm = np.array([float(l.rsplit(" ", 1)[1]) ** 2 for l in ref["norms_txt"].strip().split("\n")])
    i, j = names.index("DRR000821"), names.index("DRR000822")
    k = np.flatnonzero((z["toy_rows"] == i) & (z["toy_cols"] == j))
    assert len(k) == 1
    q = oracle_quantize_jaccard(int(z["toy_values"][k[0]]), 2048, nrm[i], nrm[j])
    assert q == 81 and abs(q / 255.0 - 0.317647) < 1e-6
    assert oracle_quantize_jaccard(int(z["toy_values"][(z["toy_rows"] == i) & (z["toy_cols"] == i)][0]), 2048, nrm[i], nrm[i]) == 255
