#!/usr/bin/env python
"""Key of the projection kernels as built: sha256 of the sm_100a cubin of csrc/mvs_capi.cu extracted from
libmvs_b200.so (cuobjdump -xelf).  bench.py compares it with the key stored beside the committed ncu summary
(profiles/project_kernel_profile.json) so that instruction counts of a stale capture are never used.

    python tools/cubin_key.py            # prints the key of the library in the tree
"""
import glob
import hashlib
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "metagenome_vector_sketches_b200", "libmvs_b200.so")


def projection_cubin_key(lib=LIB):
    """sha256 hex digest, or None when cuobjdump is unavailable / the library is not built."""
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not (os.path.exists(exe) and os.path.exists(lib)):
        return None
    with tempfile.TemporaryDirectory() as td:
        r = subprocess.run([exe, "-xelf", "all", lib], cwd=td, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        if r.returncode != 0:
            return None
        for f in sorted(glob.glob(os.path.join(td, "*.cubin"))):
            data = open(f, "rb").read()
            if b"project_kernel" in data:
                return hashlib.sha256(data).hexdigest()
    return None


if __name__ == "__main__":
    k = projection_cubin_key(sys.argv[1] if len(sys.argv) > 1 else LIB)
    print(k or "unavailable")
    sys.exit(0 if k else 1)
