"""Turns the JSON dump of oracle/_ref/ref_fk_harness (the REFERENCE's own device code of stage 2:
generate_trajectories.cu:303-387 rotation builders and :391-466 num_collisions with the stub line
:394 dropped, compiled by oracle/build_ref.sh and run on a B200 with
    gpurun -- './oracle/_ref/ref_fk_harness gpurun_out/ref_fk_outputs.json')
into the committed fixture tests/golden/ref_fk_outputs.npz.  Floats are kept as uint32 bit
patterns."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, "..", "..", "gpurun_out", "ref_fk_outputs.json")
d = json.load(open(src))
out = {}
for k, v in d.items():
    a = np.array(v, np.int64)
    out[k] = a if k in ("coll_counts", "coll_grid_n_seed") else a.astype(np.uint32)
np.savez_compressed(os.path.join(HERE, "ref_fk_outputs.npz"), **out)
print({k: v.shape for k, v in out.items()})
