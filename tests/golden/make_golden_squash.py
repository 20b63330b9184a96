"""Golden fixture for squash clustering from the COMPILED REFERENCE (genesis SquashClustering::run).
Run in the build container only:   python tests/golden/make_golden_squash.py
Writes squash.npz: synthetic 40 samples x 99 edges x 150 placements; reference mergers + tree string for
(bins, p) in {(0,1), (0,2), (4,1), (0,0.5)}."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle  # noqa: E402
from gappa_b200 import synth  # noqa: E402

tree = synth.tree_with_edges(99, seed=17)
mass = synth.sample_masses(tree, 40, 150, seed=17)
out = dict(parent=tree["parent"], child=tree["child"], bl=tree["bl"], root=np.int64(tree["root"]),
           off=mass["off"], edge=mass["edge"], pos=mass["pos"], mass=mass["mass"])
for bins, p in ((0, 1.0), (0, 2.0), (4, 1.0), (0, 0.5)):
    idx, dst, ts, sec = oracle.ref_squash(tree, mass, p=p, bins=bins)
    out[f"idx_b{bins}_p{p}"] = idx
    out[f"dst_b{bins}_p{p}"] = dst
    out[f"tree_b{bins}_p{p}"] = np.array(ts)
    print(bins, p, idx[:3].tolist(), dst[:3].tolist(), ts[:60], f"{sec:.2f}s")
np.savez_compressed(os.path.join(HERE, "squash.npz"), **out)
