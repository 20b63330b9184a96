"""KRD_PIPE_DEBUG=1 python profiles/dbg_pipe_timeline.py [samples] — timeline of the pipelined ingest at the cfg2 shape."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import gappa_b200
from gappa_b200 import workloads
spec, tree, masses = workloads.build("cfg2", samples=int(sys.argv[1]) if len(sys.argv) > 1 else 1000)
plan = gappa_b200.KrdPlan(bins=spec["bins"], pipeline=1)
plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
h = [torch.from_numpy(masses[k]).pin_memory() for k in ("edge", "pos", "mass")]
out = torch.empty((spec["samples"],) * 2, dtype=torch.float64).pin_memory()
for it in range(int(sys.argv[2]) if len(sys.argv) > 2 else 3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    plan.set_masses(masses["off"], *h)
    t1 = time.perf_counter()
    plan.compute(p=1.0, out=out)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"set_masses {1e3*(t1-t0):.2f} ms  compute {1e3*(t2-t1):.2f} ms  total {1e3*(t2-t0):.2f}", file=sys.stderr)
