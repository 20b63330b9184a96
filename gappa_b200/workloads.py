"""Named synthetic workloads (BASELINE.json configs) as flat host arrays.

The arrays handed to BOTH bench arms are the binified mass trees — what
`earth_movers_distance(mass_trees, p)` receives in the reference after
`mass_tree_binify_masses` (genesis tree/mass_tree/functions.cpp:222-265)."""
import os
from concurrent.futures import ProcessPoolExecutor

import numpy as np

from . import synth

WORKLOADS = {
    # name: samples, edges, placements/sample, bins, p
    "cfg2": dict(samples=1000, edges=9999, placements=100_000, bins=4, p=1.0),
    "cfg3": dict(samples=1000, edges=9999, placements=100_000, bins=4, p=2.0),
    "cfg4": dict(samples=10_000, edges=49_999, placements=100_000, bins=1, p=1.0),
    "cfg5": dict(samples=20_000, edges=19_999, placements=100_000, bins=1, p=0.5),
    "tiny": dict(samples=200, edges=999, placements=2000, bins=4, p=1.0),
}


def _gen_chunk(args):
    tree, first, count, placements, bins, seed = args
    raw = synth.sample_masses(tree, count, placements, seed=seed, first_sample=first)
    return synth.binify_host(tree, raw, bins) if bins > 0 else raw


def build(name=None, seed=42, first_sample=0, n_samples=None, workers=None, **override):
    """Returns (spec, tree, masses) with masses = binified CSR for samples
    [first_sample, first_sample + n_samples)."""
    spec = dict(WORKLOADS[name or "cfg2"])
    spec.update({k: v for k, v in override.items() if v is not None})
    spec["name"] = name or "cfg2"
    tree = synth.tree_with_edges(spec["edges"], seed=seed)
    n = spec["samples"] if n_samples is None else n_samples
    workers = workers or min(os.cpu_count() or 1, 32)
    chunk = max(1, min(64, (n + workers - 1) // workers))
    jobs = [(tree, first_sample + s, min(chunk, n - s), spec["placements"], spec["bins"], seed)
            for s in range(0, n, chunk)]
    if workers > 1 and len(jobs) > 1:
        with ProcessPoolExecutor(max_workers=workers) as ex:
            parts = list(ex.map(_gen_chunk, jobs))
    else:
        parts = [_gen_chunk(j) for j in jobs]
    off = np.zeros(n + 1, np.int64)
    k = 0
    for prt in parts:
        c = len(prt["off"]) - 1
        off[k + 1:k + c + 1] = off[k] + prt["off"][1:]
        k += c
    masses = dict(off=off, edge=np.concatenate([q["edge"] for q in parts]),
                  pos=np.concatenate([q["pos"] for q in parts]),
                  mass=np.concatenate([q["mass"] for q in parts]))
    return spec, tree, masses


def head(masses, n):
    """First n samples of a CSR set."""
    m = int(masses["off"][n])
    return dict(off=masses["off"][: n + 1].copy(), edge=masses["edge"][:m], pos=masses["pos"][:m],
                mass=masses["mass"][:m])
