"""Per-CTA duration of the persistent pair kernel (stream-K balance). Run on a B200."""
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
buf = np.zeros(4 * 148, np.uint64)
L.krd_debug_cta_times(C.c_void_p(buf.ctypes.data), 148)
t = buf.reshape(148, 4).astype(np.int64)
dur = (t[:, 1] - t[:, 0]) / 1e3
start = (t[:, 0] - t[:, 0].min()) / 1e3
end = (t[:, 1] - t[:, 0].min()) / 1e3
nitems = t[:, 3] - t[:, 2]
print("kernel ms", plan.stats()["ms_pairs"], "span of ends (us): min %.0f mean %.0f max %.0f" % (end.min(), end.mean(), end.max()))
print("duration us: min %.0f mean %.0f max %.0f; start skew max %.0f" % (dur.min(), dur.mean(), dur.max(), start.max()))
order = np.argsort(dur)
print("fastest 8 CTAs (cta, dur, items):", [(int(c), round(float(dur[c])), int(nitems[c])) for c in order[:8]])
print("slowest 8 CTAs (cta, dur, items):", [(int(c), round(float(dur[c])), int(nitems[c])) for c in order[-8:]])
for k in sorted(set(nitems.tolist())):
    print("items=%d: n=%d mean dur %.0f" % (k, (nitems == k).sum(), dur[nitems == k].mean()))
if len(sys.argv) > 2:   # full per-CTA dump (duration in us by CTA index = position along the tile list)
    print("dur by cta:", " ".join(str(int(round(x))) for x in dur))
