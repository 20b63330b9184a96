"""Pair-kernel time per exponent at the cfg2 shape (1,000 samples x 49,995 segments). Run on a B200."""
import sys
sys.path.insert(0, '.')
import gappa_b200
from gappa_b200 import workloads
spec, tree, masses = workloads.build("cfg2")
plan = gappa_b200.KrdPlan(bins=spec["bins"])
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
ps = [float(x) for x in sys.argv[1:]] or [1.0, 2.0, 0.5, 3.0, 4.0, 7.0, 1.7, 0.3]
for p in ps:
    for _ in range(2):
        plan.compute(p=p)
    st = plan.stats()
    N, S = st["n_samples"], st["n_segments"]
    print(f"p={p:4}: pair kernel {st['ms_pairs']:8.3f} ms  {N*(N-1)/2*S/st['ms_pairs']*1e3/1e12:6.2f} T pair-segments/s")
