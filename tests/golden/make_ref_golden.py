"""Turns the JSON dump of oracle/_ref/ref_harness (the REFERENCE's own code, compiled by
oracle/build_ref.sh and run on a B200 with
    gpurun -- './oracle/_ref/ref_harness gpurun_out/ref_outputs.json')
into the committed fixture tests/golden/ref_outputs.npz.  Floats are kept as uint32 bit patterns.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, "..", "..", "gpurun_out", "ref_outputs.json")
d = json.load(open(src))
out = {}
for k, v in d.items():
    a = np.array(v, np.int64)
    out[k] = a if k in ("matmul_2x2_int", "kernel_shape_kmnd", "parts_host_grid_kernel") else a.astype(np.uint32)
assert list(out["parts_host_grid_kernel"]) == [1, 1, 1], "harness must run on a GPU box"
np.savez_compressed(os.path.join(HERE, "ref_outputs.npz"), **out)
print({k: v.shape for k, v in out.items()})
