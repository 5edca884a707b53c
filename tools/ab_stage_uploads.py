"""A/B of the pinned staging ring for pageable caller buffers (option stage_uploads): run on the GPU box."""
import time, numpy as np, sys
sys.path.insert(0, '/root/repo')
import metagenome_vector_sketches_b200 as mvs
n, per, d = 2000, 50000, 2048
rng = np.random.default_rng(1)
h = rng.integers(0, 2**54, n*per, dtype=np.uint64)
off = np.arange(n+1, dtype=np.uint64)*per
out = np.empty((n, d), dtype=np.int32)
sk = mvs.Sketcher(0, experiments=True)       # knobs of the experiment build: make -C metagenome_vector_sketches_b200/csrc experiments
for mode in (0, 2, 0, 2):
    sk.set_option("stage_uploads", mode)
    sk.project_csr(h, off, d, out=out)
    t = time.perf_counter()
    for _ in range(3): sk.project_csr(h, off, d, out=out)
    dt = (time.perf_counter()-t)/3
    print("stage_uploads=%d: %.1f ms per 0.8 GB batch = %.1f GB/s, %.2f G hashes/s" % (mode, dt*1e3, n*per*8/dt/1e9, n*per/dt/1e9))

# how the ring is filled: memcpy / streaming stores x host threads
sk.set_option("stage_uploads", 0)
for copy in (0, 1):
    for threads in (2, 4, 8, 12, 16):
        sk.set_option("stage_copy", copy)
        sk.set_option("stage_threads", threads)
        sk.project_csr(h, off, d, out=out)
        t = time.perf_counter()
        for _ in range(3): sk.project_csr(h, off, d, out=out)
        dt = (time.perf_counter()-t)/3
        print("stage_copy=%d stage_threads=%2d: %.1f ms per 0.8 GB batch = %.1f GB/s" % (copy, threads, dt*1e3, n*per*8/dt/1e9))

# the bench.py shape (4 GB in, 82 MB out, pageable both ways), modes alternated
n, per = 10000, 50000
h = rng.integers(0, 2**54, n*per, dtype=np.uint64)
off = np.arange(n+1, dtype=np.uint64)*per
out = np.empty((n, d), dtype=np.int32)
for rep in range(3):
    for copy, threads in ((0, 4), (1, 8), (0, 8), (1, 4)):
        sk.set_option("stage_copy", copy)
        sk.set_option("stage_threads", threads)
        sk.project_csr(h, off, d, out=out)
        t = time.perf_counter()
        for _ in range(2): sk.project_csr(h, off, d, out=out)
        dt = (time.perf_counter()-t)/2
        print("4 GB batch: stage_copy=%d stage_threads=%d: %.1f ms = %.1f GB/s" % (copy, threads, dt*1e3, n*per*8/dt/1e9))
