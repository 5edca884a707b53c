"""mvs_project_device_static (collected last on purpose): the allocation-free variant of the device entry point, whose
scratch -- warp tasks, reduce items, partial rows -- is a static __device__ array inside the library.

Called through the C ABI and compared with the existing entry point mvs_project_device on the same device-resident
inputs.  Integer work: the tolerance is zero (np.array_equal), as in every other parity test here.  Calls come in pairs
and in different shapes one after the other, so that a call always finds the previous call's contents in the array.
"""
import numpy as np
import pytest

import metagenome_vector_sketches_b200 as mvs

pytestmark = pytest.mark.gpu

SCRATCH_MAX = 16 << 20          # MVS_STATIC_SCRATCH_BYTES in include/mvs_b200.h

# (samples, hashes per sample, d): one-task samples, samples split over several warps (partial rows + reduce), two column
# passes (d = 4096), odd d, S > 1 geometries, an empty batch row
SHAPES = [(1, 700000, 2048), (64, 20000, 2048), (3, 150000, 256), (500, 37, 100), (7, 65000, 64), (20, 40000, 4096),
          (2, 5, 7), (90, 12000, 1024)]


@pytest.fixture(scope="module")
def sk():
    s = mvs.Sketcher(0)
    yield s
    s.close()


def _resident(n, per, seed):
    import torch
    h, off = mvs.synthetic.uniform_batch(n, per, seed=seed)
    return h, off, torch.from_numpy(h.view(np.int64)).cuda()


def test_static_scratch_matches_the_allocating_entry_point_twice_in_a_row(sk):
    import torch
    ts = torch.cuda.Stream()
    for i, (n, per, d) in enumerate(SHAPES):
        h, off, dh = _resident(n, per, 900 + i)
        want = torch.full((n, d), -7, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()                 # the fills run on torch's stream, the projections on `ts`
        sk.project_device(dh.data_ptr(), off, d, want.data_ptr(), stream=ts.cuda_stream)      # existing entry point
        ts.synchronize()
        want = want.cpu().numpy()
        for rep in range(2):                                                                  # two calls in a row
            got = torch.full((n, d), -7, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            needed = sk.project_device_static(dh.data_ptr(), off, d, got.data_ptr(), stream=ts.cuda_stream)
            ts.synchronize()
            assert 0 < needed <= SCRATCH_MAX
            assert np.array_equal(got.cpu().numpy(), want), (n, per, d, rep)
    # ragged batch with empty samples
    rng = np.random.default_rng(17)
    sets = [np.unique(rng.integers(0, 2 ** 64 - 1, k, dtype=np.uint64)) for k in (0, 1, 70001, 5, 200000, 0, 33)]
    off = np.zeros(len(sets) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(x) for x in sets])
    dh = torch.from_numpy(np.concatenate(sets).view(np.int64)).cuda()
    for d in (2048, 200):
        want = torch.empty((len(sets), d), dtype=torch.int32, device="cuda")
        got = torch.full((len(sets), d), -7, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        sk.project_device(dh.data_ptr(), off, d, want.data_ptr(), stream=ts.cuda_stream)
        sk.project_device_static(dh.data_ptr(), off, d, got.data_ptr(), stream=ts.cuda_stream)
        ts.synchronize()
        assert torch.equal(got, want), d
    # synchronous form (stream = NULL: the context's own stream), other seed
    h, off, dh = _resident(5, 100000, 77)
    want = torch.empty((5, 512), dtype=torch.int32, device="cuda")
    got = torch.empty((5, 512), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    sk.project_device(dh.data_ptr(), off, 512, want.data_ptr(), seed=3)
    sk.project_device_static(dh.data_ptr(), off, 512, got.data_ptr(), seed=3)
    assert torch.equal(got, want)


def test_static_scratch_calls_queued_back_to_back_and_from_two_contexts(sk):
    """Different shapes queued on ONE stream without a host synchronisation in between (stream order is what keeps
    them from racing on the shared array), then the same through a second context on the same device: the array is
    per GPU, not per context."""
    import torch
    ts = torch.cuda.Stream()
    work = []
    for i, (n, per, d) in enumerate(SHAPES):
        h, off, dh = _resident(n, per, 950 + i)
        want = torch.empty((n, d), dtype=torch.int32, device="cuda")
        sk.project_device(dh.data_ptr(), off, d, want.data_ptr(), stream=ts.cuda_stream)
        work.append((off, d, dh, want, torch.full((n, d), -7, dtype=torch.int32, device="cuda")))
    torch.cuda.synchronize()
    other = mvs.Sketcher(0)
    try:
        for ctx in (sk, other):
            for off, d, dh, want, got in work:
                got.fill_(-7)
            torch.cuda.synchronize()
            for off, d, dh, want, got in work:
                ctx.project_device_static(dh.data_ptr(), off, d, got.data_ptr(), stream=ts.cuda_stream)
            ts.synchronize()
            for off, d, dh, want, got in work:
                assert torch.equal(got, want), (len(off) - 1, d)
    finally:
        other.close()


def test_static_scratch_allocates_nothing(sk):
    """Free device memory as the driver reports it does not move over a run of static calls (after one warm-up call,
    which resolves the symbol and may load the module), whereas the shapes used would make the allocating path grow."""
    import torch
    ts = torch.cuda.Stream()
    work = []
    for i, (n, per, d) in enumerate(SHAPES):
        h, off, dh = _resident(n, per, 990 + i)
        work.append((off, d, dh, torch.empty((n, d), dtype=torch.int32, device="cuda")))
    for off, d, dh, got in work:                                          # warm-up pass
        sk.project_device_static(dh.data_ptr(), off, d, got.data_ptr(), stream=ts.cuda_stream)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for off, d, dh, got in work:
        sk.project_device_static(dh.data_ptr(), off, d, got.data_ptr(), stream=ts.cuda_stream)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0


def test_static_scratch_rejects_what_does_not_fit(sk):
    """64 samples x 600 000 hashes at d = 2048: 37 partial rows of 8 KB per sample = 19 MB > 16 MiB.  The call reports
    the size, returns MVS_ERR_SCRATCH_LIMIT and leaves the output alone; the allocating entry point takes the batch."""
    import ctypes as C
    import torch
    lib = mvs.load_library()
    h, off, dh = _resident(64, 600000, 5)
    got = torch.full((64, 2048), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    with pytest.raises(mvs.MvsError) as e:
        sk.project_device_static(dh.data_ptr(), off, 2048, got.data_ptr())
    assert e.value.status == -7 and "MVS_STATIC_SCRATCH_BYTES" in str(e.value)
    needed = C.c_uint64(0)
    rc = lib.mvs_project_device_static(sk._h, C.c_void_p(dh.data_ptr()), off.ctypes.data_as(C.c_void_p), 64, 2048,
                                       C.c_uint64(0), C.c_void_p(got.data_ptr()), None, C.byref(needed))
    assert rc == -7 and needed.value > SCRATCH_MAX
    torch.cuda.synchronize()
    assert bool((got == -7).all())
    # a following call that fits still works, and the large batch goes through the allocating entry point
    h2, off2, dh2 = _resident(8, 30000, 6)
    a = torch.empty((8, 2048), dtype=torch.int32, device="cuda")
    b = torch.empty((8, 2048), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    sk.project_device_static(dh2.data_ptr(), off2, 2048, a.data_ptr())
    sk.project_device(dh2.data_ptr(), off2, 2048, b.data_ptr())
    assert torch.equal(a, b)
    sk.project_device(dh.data_ptr(), off, 2048, got.data_ptr())
    assert not bool((got == -7).all())
