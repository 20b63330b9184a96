"""p = 2: direct FP64-core path vs Gram/DMMA path — accuracy against the compiled-reference fixtures and timing at cfg3.
Run on a B200: python profiles/p2_modes.py"""
import sys, numpy as np
sys.path.insert(0, '.')
import gappa_b200
from gappa_b200 import workloads
sys.path.insert(0, 'tests')
from conftest import load_golden, rel_err

def run(tree, mass, bins, mode, p=2.0):
    plan = gappa_b200.KrdPlan(bins=bins, p2_mode=mode)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
    m = plan.compute(p=p)
    st = plan.stats(); plan.close()
    return m, st

for name, bins_list in (("cfg1.npz", [0]), ("synth_small.npz", [0, 1, 4]), ("test_ab.npz", [0])):
    tree, mass, z = load_golden(name)
    for bins in bins_list:
        key = "m_p2.0" if name != "synth_small.npz" else f"m_b{bins}_p2.0"
        want = z[key]
        for mode in (1, 2):
            got, _ = run(tree, mass, bins, mode)
            print(f"{name:16s} bins={bins} mode={mode} max rel err vs genesis = {rel_err(got, want):.3e}")

spec, tree, masses = workloads.build("cfg2", p=2.0)
res = {}
for mode in (1, 2):
    plan = gappa_b200.KrdPlan(bins=spec["bins"], p2_mode=mode)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
    for _ in range(3):
        m = plan.compute(p=2.0)
    st = plan.stats(); res[mode] = m
    S, N = st["n_segments"], st["n_samples"]
    pairs = N * (N - 1) / 2
    print(f"cfg3 mode={mode} pair kernel {st['ms_pairs']:.3f} ms  epilogue {st['ms_epilogue']:.3f} ms  -> {pairs / st['ms_pairs'] * 1e3:.3e} pairs/s, "
          f"{pairs * S / st['ms_pairs'] * 1e3 / 1e12:.2f} T pair-segments/s, gram_fixups {st['gram_fixups']}")
    plan.close()
off = ~np.eye(res[1].shape[0], dtype=bool)
print("cfg3 Gram vs direct: max rel diff", np.max(np.abs(res[1][off] - res[2][off]) / res[1][off]), " min distance", res[1][off].min(), "max", res[1].max())
