"""One launch of every kernel worth a full ncu capture (run under `ncu --set full -k regex:...`, see profiles/README.md):
cfg2 shape: F build (k_tile<TOTALS/FILL>, k_subtree_level/top), k_pairs for p = 1, 2 (Gram + k_gram_prescale), 1/2, 1.7;
cfg4 shape at 1,024 samples: the bins = 1 builder; exact mode at 256 samples x 9,999 edges x 30k positions: k_exact_*."""
import sys
sys.path.insert(0, '.')
import gappa_b200
from gappa_b200 import workloads, synth

spec, tree, comp = workloads.build_compact("cfg2", workers=8)
plan = gappa_b200.KrdPlan(bins=spec["bins"])
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
plan.set_masses_binned(comp["off"], comp["cell"], comp["mass"])
for p in (1.0, 2.0, 0.5, 1.7):
    plan.invalidate_F()
    plan.compute(p=p)
plan.close()

spec, tree, comp = workloads.build_compact("cfg4", n_samples=1024, workers=8)
plan = gappa_b200.KrdPlan(bins=1)
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
plan.set_masses_binned(comp["off"], comp["cell"], comp["mass"])
plan.compute(p=1.0)
plan.close()

tree = synth.tree_with_edges(9999, seed=42)
mass = synth.sample_masses(tree, 256, 30000, seed=42)
plan = gappa_b200.KrdPlan(bins=0, max_f_mb=4096)
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
plan.set_masses(mass["off"], mass["edge"], mass["pos"], mass["mass"])
plan.compute(p=1.0)
print(plan.stats())
plan.close()
