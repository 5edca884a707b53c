#!/usr/bin/env python
"""Key of the projection kernels as built: sha256 over the SASS of every project_kernel<..> instantiation in
libmvs_b200.so (cuobjdump -sass; instruction text only, so that unrelated kernels in the same cubin do not change the
key).  bench.py compares it with the key stored beside the committed ncu summary (profiles/project_kernel_profile.json)
so that instruction counts of a stale capture are never used.

    python tools/cubin_key.py            # prints the key of the library in the tree
"""
import hashlib
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "metagenome_vector_sketches_b200", "libmvs_b200.so")


def projection_cubin_key(lib=LIB):
    """sha256 hex digest, or None when cuobjdump is unavailable / the library is not built."""
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not (os.path.exists(exe) and os.path.exists(lib)):
        return None
    r = subprocess.run([exe, "-sass", lib], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    if r.returncode != 0:
        return None
    h = hashlib.sha256()
    found = 0
    for part in re.split(r"\n\s*Function : ", r.stdout)[1:]:
        name = part.split("\n", 1)[0].strip()
        if "project_kernel" not in name:
            continue
        found += 1
        h.update(name.encode())
        for line in part.split("\n"):
            m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
            if m:
                h.update(m.group(1).strip().encode())
                h.update(b"\n")
    return h.hexdigest() if found else None


if __name__ == "__main__":
    k = projection_cubin_key(sys.argv[1] if len(sys.argv) > 1 else LIB)
    print(k or "unavailable")
    sys.exit(0 if k else 1)
