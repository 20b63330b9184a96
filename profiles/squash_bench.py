"""Squash clustering at the cfg2 shape on one B200 vs genesis SquashClustering on the host cores (bounded sample).
python profiles/squash_bench.py [n_gpu_samples] [n_ref_samples]"""
import sys, time, numpy as np
sys.path.insert(0, '.')
import gappa_b200, oracle
from gappa_b200 import workloads
n_gpu = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
n_ref = int(sys.argv[2]) if len(sys.argv) > 2 else 96
spec, tree, masses = workloads.build("cfg2", samples=n_gpu)
plan = gappa_b200.KrdPlan(bins=spec["bins"])
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
for it in range(3):
    t0 = time.perf_counter(); m = plan.squash(p=1.0); t1 = time.perf_counter()
    print(f"GPU squash N={n_gpu}: {1e3*(t1-t0):.1f} ms wall ({n_gpu-1} merges, {n_gpu*(n_gpu-1)//2} initial pairs + ~{(n_gpu-1)*n_gpu*3//4} cluster distances)", flush=True)
sub = workloads.head(masses, n_ref)
idx, dst, ts, sec = oracle.ref_squash(tree, sub, p=1.0, bins=0)
evals = n_ref * (n_ref - 1) // 2 + sum(n_ref - k for k in range(n_ref - 1))
print(f"reference SquashClustering N={n_ref}: {sec:.2f} s on {oracle.ref().ref_max_threads()} threads ({evals} EMD evaluations -> {evals/sec:.0f}/s)")
p2 = gappa_b200.KrdPlan(bins=spec["bins"])
p2.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
p2.set_masses(sub["off"], sub["edge"], sub["pos"], sub["mass"])
g = p2.squash(p=1.0)
same = np.array_equal(g["index_a"], idx[:, 0]) and np.array_equal(g["index_b"], idx[:, 1])
err = np.max(np.abs(np.stack([g["distance_a"], g["distance_b"]], 1) - dst) / dst)
print(f"parity at N={n_ref}: same merge sequence {same}, max rel err {err:.2e}")
