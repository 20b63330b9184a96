"""Phylogenetic k-means at the cfg2 shape on one B200 vs genesis MassTreeKmeans on the host cores (bounded sample).
python profiles/kmeans_bench.py [n_gpu_samples] [n_ref_samples] [k]"""
import sys, time, numpy as np
sys.path.insert(0, '.')
import gappa_b200, oracle
from gappa_b200 import workloads
n_gpu = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
n_ref = int(sys.argv[2]) if len(sys.argv) > 2 else 128
k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
spec, tree, masses = workloads.build("cfg2", samples=n_gpu)
bins = spec["bins"]
plan = gappa_b200.KrdPlan(bins=bins)
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
init = np.random.default_rng(3).choice(n_gpu, k, replace=False)
for rep in range(3):
    t0 = time.perf_counter(); a, d, v, it = plan.kmeans(k, init); t1 = time.perf_counter()
    print(f"GPU k-means N={n_gpu} k={k} bins={bins}: {1e3*(t1-t0):.1f} ms wall, {it} iterations -> {1e3*(t1-t0)/it:.2f} ms/iteration "
          f"({n_gpu*k*it/(t1-t0):.3g} sample-centroid distances/s)", flush=True)
sub = workloads.head(masses, n_ref)
init_r = np.random.default_rng(3).choice(n_ref, k, replace=False)
ra, rd, rv, rit, sec = oracle.ref_kmeans(tree, sub, k, init_r, bins=bins)
print(f"reference MassTreeKmeans N={n_ref} k={k}: {sec:.2f} s on {oracle.ref().ref_max_threads()} threads, {rit} iterations "
      f"-> {n_ref*k*rit/sec:.0f} sample-centroid distances/s")
p2 = gappa_b200.KrdPlan(bins=bins)
p2.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
p2.set_masses(sub["off"], sub["edge"], sub["pos"], sub["mass"])
ga, gd, gv, git = p2.kmeans(k, init_r)
print(f"parity at N={n_ref}: same assignments {np.array_equal(ga, ra)}, iterations {git} vs {rit}, "
      f"max rel err {np.max(np.abs(gd - rd) / np.maximum(rd, 1e-12)):.2e}")
