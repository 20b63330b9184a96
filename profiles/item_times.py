"""Per-work-item duration of the persistent pair kernel (tuning aid). python profiles/item_times.py p"""
import sys, ctypes as C, numpy as np
sys.path.insert(0, '.')
import gappa_b200
from gappa_b200 import workloads
p = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
spec, tree, masses = workloads.build("cfg2")
plan = gappa_b200.KrdPlan(bins=spec["bins"])
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
for _ in range(3):
    plan.compute(p=p)
L = gappa_b200.lib()
cb = np.zeros(4 * 148, np.uint64); L.krd_debug_cta_times(C.c_void_p(cb.ctypes.data), 148)
ct = cb.reshape(148, 4).astype(np.int64)
n_items = int(ct[:, 3].max())
ib = np.zeros(2 * n_items, np.uint64); L.krd_debug_item_times(C.c_void_p(ib.ctypes.data), n_items)
it = ib.reshape(n_items, 2).astype(np.int64)
t0 = ct[:, 0].min()
for c in range(148):
    parts = []
    for k in range(ct[c, 2], ct[c, 3]):
        parts.append("item%d %d..%d (%d us)" % (k, (it[k, 0] - t0) // 1000, (it[k, 1] - t0) // 1000, (it[k, 1] - it[k, 0]) // 1000))
    print("cta %3d dur %5d us: %s" % (c, (ct[c, 1] - ct[c, 0]) // 1000, "; ".join(parts)))
