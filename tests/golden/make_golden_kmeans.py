"""Golden fixture for phylogenetic k-means from the COMPILED REFERENCE (genesis MassTreeKmeans via Kmeans::run).
Run in the build container only:   python tests/golden/make_golden_kmeans.py
Writes kmeans.npz: synthetic 60 samples x 99 edges x 120 placements; reference assignments / distances /
variances / iteration counts for several (bins, k, initial centroids, max_iterations), among them a start with
duplicate centroids (forces the empty-centroid treatment, utils/math/kmeans.hpp:426-500), and bins = 0 cases: there
gappa leaves accumulate_centroid_masses at 1 and the centroids get ONE mass per branch at the mass centre
(kmeans.cpp:75-78, functions.cpp:181-220)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle  # noqa: E402
from gappa_b200 import synth  # noqa: E402

tree = synth.tree_with_edges(99, seed=23)
mass = synth.sample_masses(tree, 60, 120, seed=23, concentration=0.15)
out = dict(parent=tree["parent"], child=tree["child"], bl=tree["bl"], root=np.int64(tree["root"]),
           off=mass["off"], edge=mass["edge"], pos=mass["pos"], mass=mass["mass"])
rng = np.random.default_rng(5)
cases = []
for bins, k, max_it in ((5, 3, 100), (5, 7, 100), (1, 4, 100), (20, 10, 100), (5, 5, 2), (5, 1, 100)):
    init = rng.choice(60, k, replace=False).astype(np.int32)
    cases.append((bins, k, max_it, init))
cases.append((5, 4, 100, np.array([7, 7, 30, 7], np.int32)))      # duplicates: centroids 1 and 3 start empty
cases.append((12, 3, 100, np.array([11, 11, 11], np.int32)))
# bins = 0 (gappa's default): accumulate_centroid_masses stays 1 -> centroids with one mass per branch at the mass centre
cases.append((0, 3, 100, rng.choice(60, 3, replace=False).astype(np.int32)))
cases.append((0, 6, 100, rng.choice(60, 6, replace=False).astype(np.int32)))
cases.append((0, 4, 100, np.array([20, 20, 5, 20], np.int32)))
cases.append((50, 4, 100, rng.choice(60, 4, replace=False).astype(np.int32)))   # gappa's recommended bin count (general F builder)
names = []
for i, (bins, k, max_it, init) in enumerate(cases):
    a, d, v, it, sec = oracle.ref_kmeans(tree, mass, k, init, bins=bins, max_iterations=max_it)
    names.append(f"c{i}")
    out[f"c{i}_cfg"] = np.array([bins, k, max_it], np.int64)
    out[f"c{i}_init"] = init
    out[f"c{i}_assign"] = a; out[f"c{i}_dist"] = d; out[f"c{i}_var"] = v; out[f"c{i}_iter"] = np.int64(it)
    print(i, bins, k, max_it, init.tolist(), "iter", it, "counts", np.bincount(a, minlength=k).tolist(), f"{sec:.2f}s")
out["cases"] = np.array(names)
np.savez_compressed(os.path.join(HERE, "kmeans.npz"), **out)
