"""B200-native random-projection sketching path of metagenome_vector_sketches.

Only the hot path named by BASELINE.json lives here: hash sets -> d-dimensional
int32 sketches (reference src/random_projection.cpp:9-26 driven by
src/project_everything.cpp:238-362 and src/standalone_projection.cpp:11-46).

The compute is hand-written sm_100a CUDA behind a C ABI (include/mvs_b200.h,
csrc/).  This package is the thin host-side mirror used by tests and bench.py;
it has no CPU implementation of the projection and raises if the native library
or a B200 is missing.
"""
from ._lib import MvsError, load_library, library_path          # noqa: F401
from .projection import (Sketcher, transform_set_into_vector,   # noqa: F401
                         norm_from_sumsq, format_norm, bind_host_to_device,
                         sketch_csr_multi, reference_norms, pack56)
from . import synthetic                                           # noqa: F401
from . import pairwise                                            # noqa: F401

__all__ = ["MvsError", "load_library", "library_path", "Sketcher",
           "transform_set_into_vector", "norm_from_sumsq", "format_norm", "bind_host_to_device", "sketch_csr_multi",
           "reference_norms", "pack56", "synthetic", "pairwise"]
