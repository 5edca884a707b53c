"""The round-1 A/B variants of the all-pairs comparison, which live in the EXPERIMENT build only
(make -C metagenome_vector_sketches_b200/csrc experiments -> libmvs_b200_exp.so, -DMVS_EXPERIMENTS): every case of
tests/test_gpu_pairwise.py on fifteen runtime-selected paths -- forced 3 / 4 limbs (streamed stages), clusters of
1 x 2 and 2 x 2 CTAs with TMA multicast, forced CTA pairs / single CTAs, the separate count pass, the mma.sync limb
GEMM and the IMAD GEMM -- bit-exact against the oracle.  None of this ships; the module runs only when asked for:

    make -C metagenome_vector_sketches_b200/csrc experiments && MVS_RUN_EXPERIMENTS=1 python -m pytest tests/test_gpu_experiments.py -m gpu -q
"""
import os

import numpy as np
import pytest

import metagenome_vector_sketches_b200 as mvs
import test_gpu_pairwise as cases

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("MVS_RUN_EXPERIMENTS") != "1", reason="experiment build: set MVS_RUN_EXPERIMENTS=1")]

# options pairwise_path, pairwise_limbs, pairwise_cluster, pairwise_pair (1 = always CTA pairs, 2 = single CTAs),
# pairwise_fuse (2 = separate count pass instead of the keep-test in the GEMM epilogue)
PATHS = {"default": (0, 0, 0, 0, 0), "umma": (0, 0, 0, 2, 0), "umma3": (0, 3, 0, 2, 0), "umma4": (0, 4, 0, 2, 0), "umma_1x2": (0, 0, 2, 2, 0),
         "umma_2x2": (0, 0, 4, 2, 0), "umma4_2x2": (0, 4, 4, 2, 0), "pair": (0, 0, 0, 1, 0), "pair3": (0, 3, 0, 1, 0), "pair4": (0, 4, 0, 1, 0),
         "umma_nofuse": (0, 0, 0, 2, 2), "pair_nofuse": (0, 0, 0, 1, 2), "mma": (2, 0, 0, 0, 0), "mma4": (2, 4, 0, 0, 0), "imad": (1, 0, 0, 0, 0)}


@pytest.fixture(scope="module", params=list(PATHS))
def sk(request):
    s = mvs.Sketcher(0, experiments=True)
    for key, value in zip(("pairwise_path", "pairwise_limbs", "pairwise_cluster", "pairwise_pair", "pairwise_fuse"), PATHS[request.param]):
        s.set_option(key, value)
    yield s
    s.close()


# re-export every test function of the shipped-path module; pytest resolves `sk` to the fixture above
for _name in dir(cases):
    if _name.startswith("test_"):
        globals()[_name] = getattr(cases, _name)


def test_upload_staging_variants_and_task_cap():
    """stage_copy / stage_uploads / task_cap: the other A/B knobs of round 1, projection results unchanged."""
    from oracle_util import oracle_csr
    s = mvs.Sketcher(0, experiments=True)
    try:
        h, off = mvs.synthetic.ragged_batch([300001, 7, 0, 123457], seed=21)
        want = oracle_csr(h, off, 192)
        big = np.empty(len(h) + 3, dtype=np.uint64)[3:]
        big[:] = h
        for copy, uploads, cap in ((0, 0, 0), (1, 0, 1024), (1, 2, 0), (0, 2, 4096)):
            s.set_option("stage_copy", copy)
            s.set_option("stage_uploads", uploads)
            s.set_option("task_cap", cap)
            assert np.array_equal(s.project_csr(big, off, 192), want)
    finally:
        s.close()
