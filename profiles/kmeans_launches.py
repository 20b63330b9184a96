"""One k-means run at the cfg2 shape for the ncu launch list (profiles/r1n_kmeans_launches.csv)."""
import sys, numpy as np
sys.path.insert(0, '.')
import gappa_b200
from gappa_b200 import workloads
spec, tree, masses = workloads.build("cfg2", samples=1000)
plan = gappa_b200.KrdPlan(bins=spec["bins"])
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
print(plan.kmeans(10, np.random.default_rng(3).choice(1000, 10, replace=False))[3])
