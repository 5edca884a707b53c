"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI,
against the CPU oracle (bit-exact: integer work), the committed golden vectors from the
reference, and size-independent properties at BASELINE sizes."""
import hashlib

import numpy as np
import pytest

import metagenome_vector_sketches_b200 as mvs
from oracle_util import (oracle, oracle_set, oracle_csr, ref_lib, ref_set, load_toy, load_toy_ref, load_kat,
                         parse_line_like_reference, oracle_reference_norms, ref_norms_live, load_norms_golden,
                         golden_host_has_same_rsqrt, norms_text)

pytestmark = pytest.mark.gpu

CAP = 16320          # hashes per lane per warp task (kMaxRounds in csrc/mvs_kernels.cuh; fewer for d < 512, see max_rounds)


@pytest.fixture(scope="module")
def sk():
    s = mvs.Sketcher(0)
    yield s
    s.close()


def test_kats_standalone(sk):
    for case in load_kat()["standalone"]:
        d = case["d"]
        for line, ref_out in zip(case["lines"], case["stdout"].split("\n")[:-1]):
            vec = sk.transform_set(parse_line_like_reference(line), d)
            assert " ".join("%g" % float(v) for v in vec) == ref_out


@pytest.mark.parametrize("tag,d", [("d2048", 2048), ("d100", 100)])
def test_toy_golden_bit_exact(sk, tag, d):
    names, offsets, hashes = load_toy()
    ref = load_toy_ref(tag)
    got = sk.project_csr(hashes, offsets, d)
    assert np.array_equal(got, ref["vectors"].astype(np.int32))
    assert hashlib.sha256(got.tobytes()).hexdigest() == ref["vectors_sha256"]


def test_toy_epilogue_int16_and_norms(sk):
    """Fused epilogue (finalize_kernel): int16 copy and exact sum of squares against the oracle, and the
    reference-order float32 norm against the oracle bit for bit and against the reference binary's own
    vector_norms.txt byte for byte (61/61 lines)."""
    names, offsets, hashes = load_toy()
    ref16 = load_toy_ref("d2048_i16")
    got = sk.project_csr(hashes, offsets, 2048)
    sumsq, out16, rn = sk.finalize(got, want_int16=True, want_ref_norms=True)
    assert np.array_equal(out16, ref16["vectors"])
    assert hashlib.sha256(out16.tobytes()).hexdigest() == ref16["vectors_sha256"]
    for row, ss in zip(got, sumsq):
        assert int(ss) == int(oracle().oracle_sumsq(row.ctypes.data, 2048))
    assert np.array_equal(rn.view(np.uint32), oracle_reference_norms(got).view(np.uint32))
    if golden_host_has_same_rsqrt(2048):
        assert norms_text(names, rn) == ref16["norms_txt"]
    # the caller's other choice, sqrt(exact sum / d): tolerance parity only (print precision + float32 evaluation)
    for ss, line in zip(sumsq, ref16["norms_txt"].split("\n")[:-1]):
        printed = float(line.rsplit(" ", 1)[1])
        exact = mvs.norm_from_sumsq(ss, 2048)
        ulp6 = 10.0 ** (np.floor(np.log10(exact)) - 5) if exact > 0 else 0.0
        assert abs(printed - exact) <= 0.5 * ulp6 + 2e-6 * exact


def test_reference_norms_on_device_all_packet_cases(sk):
    """finalize_kernel's float32 norm against the committed rows the reference's own expression was run on
    (every Eigen packet-count case, magnitudes to 2^30), the oracle and -- when built -- the live reference."""
    g = load_norms_golden()
    for d in sorted(int(k[5:]) for k in g if k.startswith("rows_")):
        rows = g["rows_%d" % d]
        _, _, rn = sk.finalize(rows, want_ref_norms=True)
        assert np.array_equal(rn.view(np.uint32), oracle_reference_norms(rows).view(np.uint32)), d
        assert np.array_equal(rn.view(np.uint32), mvs.reference_norms(rows).view(np.uint32)), d      # host twin
        if golden_host_has_same_rsqrt(d, g):
            assert np.array_equal(rn.view(np.uint32), g["norms_%d" % d].view(np.uint32)), d
        if ref_lib() is not None:
            assert np.array_equal(rn.astype(np.float64), ref_norms_live(rows)), d


def test_sketch_batch_epilogue_outputs_and_saturated_norms(sk):
    """mvs_sketch_batch: vectors, int16 copy, exact sum and reference norm in one call.  In --int16 mode the norm
    still comes from the int32 values (reference :328 runs before the saturation :332-347): a sample whose
    columns exceed 32767 shows it."""
    # hashes whose column-0 entry is +1 (bit 0 of the mixer output clear): v[0] = |H| = ~40 000 > 32767
    rng = np.random.default_rng(5)
    # 63 488 of them = exactly one warp task at d = 64 (32 hash slots x 1984 rounds): the flush adds the slots' counts
    # as packed 16-bit pairs, and this column reaches the largest value a task can produce
    cand = np.unique(rng.integers(0, 2 ** 63, 140_000, dtype=np.uint64))
    big = cand[[(oracle().oracle_splitmix64(int(h), 0, 0) & 1) == 0 for h in cand]][:63488]
    assert len(big) == 63488
    small = np.unique(rng.integers(0, 2 ** 63, 700, dtype=np.uint64))
    offsets = np.array([0, len(big), len(big), len(big) + len(small)], dtype=np.uint64)
    hashes = np.concatenate([big, small])
    for d in (64, 100):
        want = oracle_csr(hashes, offsets, d)
        r32 = sk.sketch_csr(hashes, offsets, d)
        r16 = sk.sketch_csr(hashes, offsets, d, int16=True)
        assert np.array_equal(r32["vec"], want) and want[0, 0] == len(big)
        assert np.array_equal(r16["vec16"], np.clip(want, -32768, 32767).astype(np.int16)) and r16["vec16"][0, 0] == 32767
        for r in (r32, r16):
            assert np.array_equal(r["sumsq"], (want.astype(np.int64) ** 2).sum(axis=1).astype(np.uint64))
            assert np.array_equal(r["ref_norms"].view(np.uint32), oracle_reference_norms(want).view(np.uint32))
        assert r32["ref_norms"][1] == 0.0                                 # empty sample: zero vector, norm 0


def test_sketch_batch_multi_two_contexts_on_one_device(sk):
    """The multi-GPU call of the C ABI (mvs_sketch_batch_multi) where the 1-GPU test box can see it: two and three
    contexts on device 0 -- heavy-tailed sizes, empty samples, a sample larger than a whole shard, more contexts
    than samples -- bit-exact against the oracle and identical to the single-context call."""
    extra = [mvs.Sketcher(0), mvs.Sketcher(0)]
    try:
        sizes = [int(x) for x in mvs.synthetic.lognormal_sizes(160, seed=11, median=2000, lo=1, hi=20000)]
        sizes[3] = 0
        sizes[4] = 0
        sizes[50] = 400_000                                               # larger than a whole mean shard of the 3-context split
        sizes[159] = 0
        h, off = mvs.synthetic.ragged_batch(sizes, seed=4)
        for d in (2048, 300):
            want = oracle_csr(h, off, d)
            one = sk.sketch_csr(h, off, d)
            assert np.array_equal(one["vec"], want)
            for ctxs in ([sk, extra[0]], [sk, extra[0], extra[1]]):
                got = mvs.sketch_csr_multi(ctxs, h, off, d)
                assert np.array_equal(got["vec"], want)
                assert np.array_equal(got["sumsq"], one["sumsq"])
                assert np.array_equal(got["ref_norms"].view(np.uint32), one["ref_norms"].view(np.uint32))
            got16 = mvs.sketch_csr_multi([sk, extra[0]], h, off, d, int16=True)
            assert np.array_equal(got16["vec16"], np.clip(want, -32768, 32767).astype(np.int16))
        # more contexts than samples; an all-empty batch
        h2, off2 = mvs.synthetic.ragged_batch([5, 0], seed=2)
        got = mvs.sketch_csr_multi([sk] + extra, h2, off2, 128)
        assert np.array_equal(got["vec"], oracle_csr(h2, off2, 128))
        z = mvs.sketch_csr_multi([sk, extra[0]], np.zeros(0, np.uint64), np.zeros(4, np.uint64), 64)
        assert not z["vec"].any() and z["vec"].shape == (3, 64)
    finally:
        for e in extra:
            e.close()


def test_int16_saturation_edges(sk):
    v = np.array([[40000, -40000, 32767, -32768, 32768, -32769, 0, 1]], dtype=np.int32)
    sumsq, o16, rn = sk.finalize(v, want_int16=True, want_ref_norms=True)
    assert o16.tolist() == [[32767, -32768, 32767, -32768, 32767, -32768, 0, 1]]
    assert np.array_equal(rn.view(np.uint32), oracle_reference_norms(v).view(np.uint32))   # from the unsaturated values
    assert int(sumsq[0]) == int((v.astype(np.int64) ** 2).sum())


@pytest.mark.parametrize("d", [1, 7, 63, 64, 65, 100, 128, 192, 256, 320, 512, 1000, 1024, 1536, 2048,
                               2049, 2112, 4096, 4160, 5000])
def test_random_sets_all_dimensions(sk, d):
    rng = np.random.default_rng(1000 + d)
    sizes = [0, 1, 2, 15, 16, 17, 31, 33, 100, 511, 512, 513, 1025, 4079, 4080, 4081, 9000, CAP - 1, CAP, CAP + 1, 40000]
    if d <= 256:
        sizes += [CAP * (32 // max(1, 1 << int(np.ceil(np.log2(max(1, (d + 63) // 64)))))) + 5, 70000]
    sets = [np.unique(rng.integers(0, 2 ** 64 - 1, n, dtype=np.uint64)) for n in sizes]
    offsets = np.zeros(len(sets) + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum([len(s) for s in sets])
    hashes = np.concatenate(sets)
    got = sk.project_csr(hashes, offsets, d)
    want = oracle_csr(hashes, offsets, d)
    assert np.array_equal(got, want)


def test_seed_is_additive_and_zero_is_reference(sk):
    rng = np.random.default_rng(5)
    h = np.unique(rng.integers(0, 2 ** 63, 3000, dtype=np.uint64))
    for seed in (0, 1, 12345, 2 ** 64 - 1):
        assert np.array_equal(sk.transform_set(h, 2048, seed=seed), oracle_set(h, 2048, seed=seed))
    with np.errstate(over="ignore"):
        assert np.array_equal(sk.transform_set(h, 512, seed=77), sk.transform_set(h + np.uint64(77), 512))


def test_carry_window_of_the_hoisted_add(sk):
    """The kernel adds base = golden + seed + 2048*pass once per hash and 64*block per lane on the
    low word only; hashes whose low word lands in [k*2^30 - 2048, k*2^30) after the base add must take
    the exact path.  Build samples that sit in and around that window for every pass / seed."""
    golden = 0x9E3779B97F4A7C15
    rng = np.random.default_rng(77)
    # d = 512 / 1024: a tile is 4 / 8 Harley-Seal groups, so deferred weight-64 / weight-128 carries are pending when
    # an exact tile follows a fast one (and the other way round)
    for d, seed in ((2048, 0), (2048, 12345), (4096, 0), (256, 0), (100, 2 ** 64 - 5), (5000, 7), (512, 0), (1024, 3),
                    (768, 11)):
        passes = (((d + 63) // 64) + 31) // 32
        sets = []
        for p in range(passes):
            base = (golden + seed + 2048 * p) % 2 ** 64
            # low words of hash + base: the whole window, its two edges, and plain wrap-around values
            # ... and, since the cooked tile pass also fixes bits 30..31 of the low word per hash, the
            # windows below every multiple of 2^30 (where + 64*block reaches bit 30) with both edges
            lows = np.concatenate([np.arange(2 ** 32 - 2112, 2 ** 32, dtype=np.uint64),
                                   np.arange(0, 64, dtype=np.uint64)] +
                                  [np.arange(k * 2 ** 30 - 2112, k * 2 ** 30 + 64, 3, dtype=np.uint64) for k in (1, 2, 3)])
            highs = rng.integers(0, 2 ** 32, len(lows), dtype=np.uint64)
            target = (highs << np.uint64(32)) | lows
            with np.errstate(over="ignore"):
                sets.append(np.unique(target - np.uint64(base)))
            # a sparse mix: one risky hash hidden among ordinary ones (exercises exact and fast tiles)
            plain = rng.integers(0, 2 ** 63, 6000, dtype=np.uint64)
            plain[3333] = sets[-1][-7]
            sets.append(np.unique(plain))
            # several risky hashes spread over a long sample: exact tiles at many positions of the carry cycle
            plain = np.sort(rng.integers(0, 2 ** 63, 40000, dtype=np.uint64))
            risky = sets[-2][rng.integers(0, len(sets[-2]), 9)]
            plain[[300, 700, 1100, 5000, 5300, 12345, 20000, 20300, 39000]] = risky
            sets.append(np.unique(plain))
        offsets = np.zeros(len(sets) + 1, dtype=np.uint64)
        offsets[1:] = np.cumsum([len(x) for x in sets])
        hashes = np.concatenate(sets)
        got = sk.project_csr(hashes, offsets, d, seed=seed)
        assert np.array_equal(got, oracle_csr(hashes, offsets, d, seed=seed)), (d, seed)


def test_small_batches_are_split_finer_and_stay_exact(sk):
    # few samples: the planner shrinks the task cap to fill the GPU; the result must not change
    h, off = mvs.synthetic.uniform_batch(3, 40000, seed=11)
    for d in (256, 2048):
        assert np.array_equal(sk.project_csr(h, off, d), oracle_csr(h, off, d))
    assert sk.timing()["warp_tasks"] >= 3 * (40000 // 2048)   # cap shrunk from CAP to <= 2048 rounds


def test_large_sample_split_across_warp_tasks(sk):
    # one sample >> task capacity: partial rows + fixed-order reduce
    h, off = mvs.synthetic.uniform_batch(1, 300000, seed=3)
    got = sk.project_csr(h, off, 2048)
    assert np.array_equal(got, oracle_csr(h, off, 2048))
    t = sk.timing()
    assert t["partial_rows"] >= 300000 // CAP


def test_many_slices_and_empty_samples(sk):
    sk.set_option("slice_hashes", 50000)
    try:
        sizes = [0, 7000, 0, 0, 120000, 3, 49999, 50001, 0, 16, 80000, 1]
        h, off = mvs.synthetic.ragged_batch(sizes, seed=9)
        got = sk.project_csr(h, off, 1024)
        assert np.array_equal(got, oracle_csr(h, off, 1024))
        assert not got[0].any() and not got[2].any()
    finally:
        sk.set_option("slice_hashes", 16 << 20)


def test_properties_at_scale(sk):
    """Size-independent properties on a batch too big for the scalar oracle (1.6e7 hashes)."""
    n, per, d = 320, 50000, 2048
    h, off = mvs.synthetic.uniform_batch(n, per, seed=0)
    full = sk.project_csr(h, off, d)
    # (1) linearity: halves of every sample projected separately add up
    off2 = np.arange(2 * n + 1, dtype=np.uint64) * np.uint64(per // 2)
    halves = sk.project_csr(h, off2, d)
    assert np.array_equal(halves[0::2] + halves[1::2], full)
    # (2) prefix property: d' = 512 equals the first 512 columns
    assert np.array_equal(sk.project_csr(h, off, 512), full[:, :512])
    # (3) parity of every entry equals parity of |H| (sum of +-1), and |v| <= |H|
    assert np.all((full & 1) == (per & 1)) and np.abs(full).max() <= per
    # (4) spot-check rows against the oracle
    for s in (0, 151, n - 1):
        assert np.array_equal(full[s], oracle_set(h[int(off[s]):int(off[s + 1])], d))
    # (5) one merged sample = sum of all rows (split across many tasks + reduce)
    merged = sk.project_csr(h, np.array([0, len(h)], dtype=np.uint64), d)
    assert np.array_equal(merged[0], full.sum(axis=0, dtype=np.int64).astype(np.int32))


def _project_resident(sk, dh, off, d, ts):
    import torch
    out = torch.empty((len(off) - 1, d), dtype=torch.int32, device="cuda")
    sk.project_device(dh.data_ptr(), off, d, out.data_ptr(), stream=ts.cuda_stream)
    ts.synchronize()
    return out


def test_properties_at_full_config2_size(sk):
    """BASELINE configs[1] at its full size (10 000 x 50 000 hashes = 5e8, d = 2048), device-resident like bench.py's
    `value`: linearity over half-samples, the prefix property, the parity of every entry, a merged sample that is
    split over hundreds of warp tasks, and rows spot-checked against the oracle."""
    import torch
    n, per, d = 10000, 50000, 2048
    ts = torch.cuda.Stream()
    dh, off = mvs.synthetic.uniform_batch_torch(n, per, "cuda", seed=0)
    torch.cuda.synchronize()
    full = _project_resident(sk, dh, off, d, ts)
    off2 = np.arange(2 * n + 1, dtype=np.uint64) * np.uint64(per // 2)
    halves = _project_resident(sk, dh, off2, d, ts)
    assert torch.equal(halves[0::2] + halves[1::2], full)                               # linearity
    del halves
    assert torch.equal(_project_resident(sk, dh, off, 512, ts), full[:, :512])          # prefix property
    assert bool(((full & 1) == (per & 1)).all()) and int(full.abs().max()) <= per       # parity, bound
    for s in (0, 4999, n - 1):                                                          # oracle rows
        hs = dh[s * per:(s + 1) * per].cpu().numpy().view(np.uint64)
        assert np.array_equal(full[s].cpu().numpy(), oracle_set(hs, d)), s
    # 1000 samples merged into one of 5e7 hashes: > 3000 partial rows and the fixed-order reduce
    merged = _project_resident(sk, dh[: 1000 * per], np.array([0, 1000 * per], dtype=np.uint64), d, ts)
    assert torch.equal(merged[0].to(torch.int64), full[:1000].sum(dim=0, dtype=torch.int64))
    # the checksum bench.py prints for this workload (profiles/r02p_bench_1gpu_final.json)
    assert int(full.sum(dtype=torch.int64).item()) == -168614


def test_properties_at_full_config3_size(sk):
    """BASELINE configs[2] (1 024 x 100 000 hashes) over its dimension sweep: every d is a prefix of d = 4096, so the five
    kernel instantiations (4, 8, 16, 32 lanes per hash, two column passes) must agree column for column; d = 256 rows
    against the oracle."""
    import torch
    n, per = 1024, 100000
    ts = torch.cuda.Stream()
    dh, off = mvs.synthetic.uniform_batch_torch(n, per, "cuda", seed=0)
    torch.cuda.synchronize()
    widest = _project_resident(sk, dh, off, 4096, ts)
    for d in (256, 512, 1024, 2048, 100, 4000):
        assert torch.equal(_project_resident(sk, dh, off, d, ts), widest[:, :d]), d
    assert bool(((widest & 1) == (per & 1)).all())
    for s in (0, 777):
        hs = dh[s * per:(s + 1) * per].cpu().numpy().view(np.uint64)
        assert np.array_equal(widest[s, :256].cpu().numpy(), oracle_set(hs, 256)), s


def test_properties_on_a_heavy_tailed_slice_of_config4(sk):
    """One eighth of BASELINE configs[3] (12 500 samples, log-normal sizes 1e3..1e7, about 7.7e8 hashes), i.e. the shard
    one of 8 GPUs owns: parity of every row equals the parity of its size, the device path equals the batch path on a
    sub-range, and the largest sample (split over hundreds of tasks) equals the sum of its halves."""
    import torch
    ts = torch.cuda.Stream()
    sizes = mvs.synthetic.lognormal_sizes(12500, seed=0)
    dh, off = mvs.synthetic.ragged_batch_torch(sizes, "cuda", seed=0)
    torch.cuda.synchronize()
    out = _project_resident(sk, dh, off, 2048, ts)
    par = torch.from_numpy((sizes & 1).astype(np.int32)).cuda()
    assert bool(((out & 1) == par[:, None]).all())
    assert bool((out.abs().max(dim=1).values <= torch.from_numpy(sizes.astype(np.int32)).cuda()).all())
    lo, hi = 100, 164                                                                    # 64 samples through host buffers
    h_host = dh[int(off[lo]):int(off[hi])].cpu().numpy().view(np.uint64)
    assert np.array_equal(sk.project_csr(h_host, off[lo:hi + 1] - off[lo], 2048), out[lo:hi].cpu().numpy())
    big = int(np.argmax(sizes))
    b0, b1 = int(off[big]), int(off[big + 1])
    mid = b0 + (b1 - b0) // 2
    two = _project_resident(sk, dh[b0:b1], np.array([0, mid - b0, b1 - b0], dtype=np.uint64), 2048, ts)
    assert torch.equal(two[0] + two[1], out[big])
    for s in (int(np.argmin(sizes)), 5):
        hs = dh[int(off[s]):int(off[s + 1])].cpu().numpy().view(np.uint64)
        if len(hs) <= 200000:
            assert np.array_equal(out[s].cpu().numpy(), oracle_set(hs, 2048)), s


def test_sketch_batch_multi_on_all_visible_devices(sk):
    """On a box with several B200s: one context per device (what `--gpus N` does), same rows as one context."""
    import torch
    ndev = torch.cuda.device_count()
    if ndev < 2:
        pytest.skip("one visible GPU: the fan-out is covered by the contexts-on-one-device test")
    ctxs = [sk] + [mvs.Sketcher(g) for g in range(1, ndev)]
    try:
        sizes = [int(x) for x in mvs.synthetic.lognormal_sizes(300, seed=5, median=3000, lo=1, hi=40000)]
        sizes[0] = 0
        sizes[100] = 500_000
        h, off = mvs.synthetic.ragged_batch(sizes, seed=8)
        want = oracle_csr(h, off, 2048)
        got = mvs.sketch_csr_multi(ctxs, h, off, 2048)
        assert np.array_equal(got["vec"], want)
        assert np.array_equal(got["ref_norms"].view(np.uint32), oracle_reference_norms(want).view(np.uint32))
    finally:
        for c in ctxs[1:]:
            c.close()


def test_upload_relay_through_a_peer_device(sk):
    """Option relay_device: hash slices go host -> peer GPU -> this GPU (NVLink).  Same rows, u64 and packed; on a
    1-GPU box the option must be refused (no peer), never silently ignored."""
    import torch
    from metagenome_vector_sketches_b200.projection import pinned_empty
    if torch.cuda.device_count() < 2:
        with pytest.raises(mvs.MvsError):
            sk.set_option("relay_device", 1)
        with pytest.raises(mvs.MvsError):
            sk.set_option("relay_device", 0)                              # itself
        sk.set_option("relay_device", -1)
        return
    sizes = [200000, 0, 7, 1500000, 33, 900001]
    h, off = mvs.synthetic.ragged_batch(sizes, seed=77)
    want = oracle_csr(h, off, 2048)
    pin = pinned_empty(len(h), np.uint64)
    pin.a[:] = h
    pk = pinned_empty(7 * len(h), np.uint8)
    mvs.pack56(h, out=pk.a)
    sk.set_option("slice_hashes", 1 << 19)
    try:
        sk.set_option("relay_device", 1)
        assert np.array_equal(sk.project_csr(pin.a, off, 2048), want)
        assert np.array_equal(sk.project_csr_packed56(pk.a, off, 2048), want)
        assert np.array_equal(sk.project_csr(h, off, 2048), want)         # pageable source: direct path
        sk.set_option("relay_device", -1)
        assert np.array_equal(sk.project_csr(pin.a, off, 2048), want)
    finally:
        sk.set_option("relay_device", -1)
        sk.set_option("slice_hashes", 16 << 20)


def test_device_resident_entry_point(sk):
    import torch
    h, off = mvs.synthetic.uniform_batch(64, 20000, seed=2)
    dh = torch.from_numpy(h.view(np.int64)).cuda()
    dout = torch.empty((64, 2048), dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    sk.project_device(dh.data_ptr(), off, 2048, dout.data_ptr(), stream=st)
    torch.cuda.synchronize()
    assert np.array_equal(dout.cpu().numpy(), oracle_csr(h, off, 2048))
    t = sk.timing()
    assert t["kernel_ms"] > 0 and t["launches"] > 0


def test_device_resident_calls_back_to_back_on_a_user_stream(sk):
    """Asynchronous use of mvs_project_device: many calls of different shapes queued on one caller stream without a
    host synchronisation in between.  The plans rotate over the context's slots and are copied on the slots' own
    streams while the previous kernel runs; every output must still be the oracle's (no plan buffer reused early)."""
    import torch
    ts = torch.cuda.Stream()
    shapes = [(40, 9000, 2048), (3, 150000, 256), (500, 37, 100), (1, 700000, 2048), (64, 20000, 512), (7, 65000, 64),
              (200, 3000, 4096), (2, 5, 7), (90, 12000, 1024)]
    work = []
    with torch.cuda.stream(ts):
        for i, (n, per, d) in enumerate(shapes * 2):
            h, off = mvs.synthetic.uniform_batch(n, per, seed=100 + i)
            dh = torch.from_numpy(h.view(np.int64)).to("cuda", non_blocking=False)
            dout = torch.full((n, d), -7, dtype=torch.int32, device="cuda")
            work.append((h, off, d, dh, dout))
        ts.synchronize()
        for h, off, d, dh, dout in work:                                  # 18 launches queued back to back
            sk.project_device(dh.data_ptr(), off, d, dout.data_ptr(), stream=ts.cuda_stream)
        ts.synchronize()
    for h, off, d, dh, dout in work:
        assert np.array_equal(dout.cpu().numpy(), oracle_csr(h, off, d)), (len(off) - 1, d)


def test_packed56_input_gives_the_same_vectors(sk):
    """7-byte packed hashes (mvs_sketch_batch_packed56): ragged samples, empty samples, slices that start at every
    byte phase, values up to 2^56 - 1 -- the same rows as the u64 path and the oracle; a hash >= 2^56 is refused."""
    rng = np.random.default_rng(56)
    sizes = [0, 1, 2, 7, 8, 9, 1000, 0, 33333, 5, 70001, 0]
    sets = [np.unique(rng.integers(0, 2 ** 56, n, dtype=np.uint64)) for n in sizes]
    sets[3][-1] = np.uint64(2 ** 56 - 1)
    sets[4][0] = np.uint64(0)
    offsets = np.zeros(len(sets) + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum([len(x) for x in sets])
    h = np.concatenate(sets)
    packed = mvs.pack56(h)
    assert packed.tobytes() == h.astype("<u8").view(np.uint8).reshape(-1, 8)[:, :7].tobytes()    # the format, restated
    sk.set_option("slice_hashes", 4099)                                   # many slices, odd starts
    try:
        for d in (2048, 200):
            want = oracle_csr(h, offsets, d)
            assert np.array_equal(sk.project_csr_packed56(packed, offsets, d), want)
            assert np.array_equal(sk.project_csr(h, offsets, d), want)
    finally:
        sk.set_option("slice_hashes", 16 << 20)
    assert np.array_equal(sk.project_csr_packed56(packed, offsets, 64), oracle_csr(h, offsets, 64))
    with pytest.raises(mvs.MvsError):
        mvs.pack56(np.array([1, 2 ** 56], dtype=np.uint64))


def test_live_reference_if_built(sk):
    if ref_lib() is None:
        pytest.skip("oracle/_ref not present on this host")
    rng = np.random.default_rng(21)
    for n, d in [(1, 64), (999, 2048), (5000, 300)]:
        h = np.unique(rng.integers(0, 2 ** 64 - 1, n, dtype=np.uint64))
        assert np.array_equal(sk.transform_set(h, d), ref_set(h, d))


def test_error_paths(sk):
    lib = mvs.load_library()
    out = np.zeros(64, dtype=np.int32)
    bad_off = np.array([5, 3], dtype=np.uint64)
    h = np.arange(8, dtype=np.uint64)
    assert lib.mvs_project_batch(sk._h, h.ctypes.data, bad_off.ctypes.data, 1, 64, 0, out.ctypes.data) == -1
    assert lib.mvs_project_batch(sk._h, h.ctypes.data, bad_off.ctypes.data, 1, 0, 0, out.ctypes.data) == -1
    assert lib.mvs_project_batch(sk._h, None, None, 1, 64, 0, out.ctypes.data) == -1
    assert lib.mvs_project_batch(sk._h, h.ctypes.data, bad_off.ctypes.data, 0, 64, 0, None) == 0


def test_host_side_helpers_binding_write_combined_and_staging_modes(sk):
    """mvs_bind_host_to_device never fails on a box without NUMA information; a write-combined input buffer and
    every way of filling the staging ring (pageable source >= 2 MB, odd offsets) give the oracle's vectors."""
    from metagenome_vector_sketches_b200.projection import pinned_empty
    node, applied = mvs.bind_host_to_device(0)
    assert node >= -1 and 0 <= applied <= 3
    sizes = [300001, 7, 0, 123457]
    h, off = mvs.synthetic.ragged_batch(sizes, seed=21)
    want = oracle_csr(h, off, 192)
    wc = pinned_empty(len(h), np.uint64, write_combined=True)
    wc.a[:] = h
    assert np.array_equal(sk.project_csr(wc.a, off, 192), want)
    big = np.empty(len(h) + 3, dtype=np.uint64)[3:]                       # pageable, 24 bytes off the allocation
    big[:] = h
    try:
        for threads in (1, 3, 8, 64):
            sk.set_option("stage_threads", threads)
            assert np.array_equal(sk.project_csr(big, off, 192), want)
    finally:
        sk.set_option("stage_threads", 8)


def test_set_option_rejects_unknown_keys_and_bad_values(sk):
    """A mistyped knob must not silently test the default path: unknown keys, out-of-range values and the round-1
    A/B knobs (which only exist in the experiment build) are MVS_ERR_INVALID_ARG in the shipped library."""
    assert mvs.load_library().mvs_has_experiments() == 0
    for key, value in (("no_such_option", 1), ("stage_thread", 4), ("stage_threads", 0), ("stage_threads", 65),
                       ("slice_hashes", 0), ("slice_rows_bytes", -5), ("pairwise_path", 1), ("pairwise_path", 0),
                       ("stage_copy", 0), ("stage_uploads", 2), ("task_cap", 1024), ("pairwise_limbs", 4), ("", 0)):
        with pytest.raises(mvs.MvsError) as e:
            sk.set_option(key, value)
        assert e.value.status == -1, (key, value)
    sk.set_option("stage_threads", 8)
    sk.set_option("slice_hashes", 16 << 20)
