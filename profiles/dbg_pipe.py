import sys; sys.path.insert(0,'.')
import numpy as np, gappa_b200
from gappa_b200 import synth
tree = synth.tree_with_edges(301, seed=5)
mass = synth.sample_masses(tree, 640, 80, seed=5)
print("M", len(mass["edge"]), "off[256]", mass["off"][256], "max edge", mass["edge"].max(), "E", len(tree["bl"]))
bad = mass["edge"].copy(); bad[-1] = 10_000_000
pip = gappa_b200.KrdPlan(bins=2, pipeline=1)
pip.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
try:
    pip.set_masses(mass["off"], bad, mass["pos"], mass["mass"]); print("set ok")
    pip.compute(p=1.0)
except Exception as e:
    print("ERR", e)
