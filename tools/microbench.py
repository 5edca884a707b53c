#!/usr/bin/env python
"""Print the integer issue-rate microbenchmarks of mvs_int_issue_peak (GPU box only)."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import metagenome_vector_sketches_b200 as mvs

NAMES = ["lop3", "imad", "lop3_imad_1to1", "iadd", "shf", "imad_wide(+lop3)", "imad_hi", "shf_funnel",
         "lop3_imad_2to1", "lop3_shf_1to1"]
with mvs.Sketcher(0) as sk:
    out = {n: sk.int_issue_peak(i) / 1e12 for i, n in enumerate(NAMES)}
print(json.dumps(out, indent=1))
