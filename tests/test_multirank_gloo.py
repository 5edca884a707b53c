"""World-size-2 gloo test (CPU) of the N>1 host logic: rank sharding is a disjoint cover, every
rank's shard projected independently (here by the oracle, the GPU being absent) reassembles to
the full-batch result with no data exchange, and the timing reduction is the max over ranks."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import metagenome_vector_sketches_b200 as mvs
        from metagenome_vector_sketches_b200 import sharding
        from oracle_util import oracle_csr
        sizes = [5, 0, 700, 3, 0, 4000, 12, 900, 1, 64]
        h, off = mvs.synthetic.ragged_batch(sizes, seed=4)
        first, hs, loc = sharding.shard_for_rank(h, off, rank, world)
        part = oracle_csr(hs, loc, 128)
        gathered = [None] * world
        dist.all_gather_object(gathered, (first, part))          # test-only gather; the product writes rows in place
        slowest = sharding.max_over_ranks(10.0 + rank)
        if rank == 0:
            full = oracle_csr(h, off, 128)
            rows = np.concatenate([p for _, p in sorted(gathered, key=lambda t: t[0])])
            firsts = sorted(f for f, _ in gathered)
            q.put((bool(np.array_equal(rows, full)), firsts, slowest, len(rows)))
    finally:
        dist.destroy_process_group()


def test_two_rank_sharding_reassembles():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, firsts, slowest, nrows = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok and nrows == 10 and firsts[0] == 0 and 0 < firsts[1] < 10
    assert slowest == 11.0


def test_bench_relay_share_root():
    """bench.py's secant step for the relay share (8-GPU boxes with unequal PCIe roots): the measured gaps of
    profiles/r02t_bench_8gpu_default.json give the share the run used; flat or degenerate inputs keep the probe's share;
    the result stays inside [0, 100]."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_module", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    f = bench.balanced_relay_share
    assert f(21, 14.33, 36, -9.38) == 30
    assert f(21, 14.84, 36, -6.99) == 31
    assert f(21, 5.0, 36, 5.0) == 21 and f(21, 5.0, 21, -5.0) == 21
    assert f(90, 40.0, 75, 60.0) == 100                      # gap grows as the share falls: everything relayed
    assert f(10, -30.0, 25, -50.0) == 0                      # relaying ranks already faster: nothing relayed
