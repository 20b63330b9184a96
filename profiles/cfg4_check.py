"""cfg4 SHAPE on one B200: 10,000 samples x 49,999 edges, bins = 1 -> S = 99,998, F = 8.1 GB, 49,995,000 pairs, p = 1.
Placements per sample are reduced (argv[1], default 2000) so that input generation stays cheap; the dense path's
cost does not depend on them (F is dense).  Checks a 48-sample sub-block against the CPU port and size-independent
properties on the full matrix, and prints device times."""
import sys, time, numpy as np
sys.path.insert(0, '.')
import gappa_b200, oracle
from gappa_b200 import workloads
P = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
N = int(sys.argv[2]) if len(sys.argv) > 2 else 10000
t0 = time.time()
spec, tree, masses = workloads.build("cfg4", placements=P, samples=N)
print(f"generated {len(masses['edge'])} masses in {time.time()-t0:.1f} s", flush=True)
plan = gappa_b200.KrdPlan(bins=spec["bins"])
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
for it in range(2):
    t0 = time.time(); m = plan.compute(p=1.0); t1 = time.time()
    st = plan.stats()
    pairs = N * (N - 1) / 2
    print(f"compute {t1-t0:.3f} s wall | F build {st['ms_build']:.2f} ms, pair kernel {st['ms_pairs']:.2f} ms, epilogue {st['ms_epilogue']:.2f} ms"
          f" | {pairs/st['ms_pairs']*1e3:.3e} pairs/s kernel, useful {3*st['n_segments']*pairs/st['ms_pairs']*1e3/1e12:.2f} TFLOP/s"
          f" | S={st['n_segments']} F={st['f_bytes']/1e9:.2f} GB tiles={st['n_tiles']}", flush=True)
    plan.invalidate_F()
sub = workloads.head(masses, 48)
want = oracle.port_matrix(tree, sub, p=1.0, bins=0)
got = m[:48, :48]
off = ~np.eye(48, dtype=bool)
print("sub-block 48x48 vs CPU port: max rel err", np.max(np.abs(got[off] - want[off]) / want[off]))
idx = np.random.default_rng(0).integers(0, N, (3, 2000))
i, j, k = idx
print("symmetric", np.array_equal(m, m.T), "zero diag", np.all(np.diag(m) == 0), "triangle ok", bool(np.all(m[i, k] <= m[i, j] + m[j, k] + 1e-12)),
      "min offdiag", m[~np.eye(N, dtype=bool)].min() if N <= 4000 else float(np.min(m + np.eye(N) * 1e9)))
# last sample block (ragged: 10,000 = 78*128 + 16) against the port
tail = list(range(N - 5, N)) + [0, 1, 2]
sel = dict(off=np.zeros(len(tail) + 1, np.int64))
e, p_, ms = [], [], []
for q, s in enumerate(tail):
    lo, hi = masses["off"][s], masses["off"][s + 1]
    e.append(masses["edge"][lo:hi]); p_.append(masses["pos"][lo:hi]); ms.append(masses["mass"][lo:hi]); sel["off"][q + 1] = sel["off"][q] + hi - lo
sel.update(edge=np.concatenate(e), pos=np.concatenate(p_), mass=np.concatenate(ms))
w2 = oracle.port_matrix(tree, sel, p=1.0, bins=0)
g2 = m[np.ix_(tail, tail)]
o2 = ~np.eye(len(tail), dtype=bool)
print("ragged tail block vs CPU port: max rel err", np.max(np.abs(g2[o2] - w2[o2]) / w2[o2]))
