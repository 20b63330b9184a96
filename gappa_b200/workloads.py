"""Named synthetic workloads (BASELINE.json configs) as flat host arrays.

The arrays handed to BOTH bench arms are the binified mass trees — what
`earth_movers_distance(mass_trees, p)` receives in the reference after
`mass_tree_binify_masses` (genesis tree/mass_tree/functions.cpp:222-265) — generated directly in the
compact form `(cell = edge * bins + bin, mass)` (one entry per occupied (sample, edge, bin)) and expanded to
`(edge, pos = bin centre, mass)` where the full form is needed (reference arm, oracle).

Generation is per sample from a counter-based stream (numpy Philox, key = seed, counter = sample index), so any
rank / process regenerates any sample range; `build_compact(..., cache=True)` shares the result between the ranks of one
job and between consecutive jobs on one box through /dev/shm (every rank generates 1/world of the samples).
Pure host-side input generation; no distance arithmetic happens here."""
import hashlib
import json
import os
import time
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
    # exact positions (what `gappa analyze krd` runs without --bins): SURVEY's CPU probe shape at 1,000 samples; the dense matrix
    # would hold ~3e7 segments x 1,024 columns = 245 GB, so the segment axis is streamed (krd_config.max_f_mb)
    "exact1k": dict(samples=1000, edges=9999, placements=30_000, bins=0, p=1.0),
    "exact_tiny": dict(samples=150, edges=999, placements=3000, bins=0, p=1.0),
}


def spec_of(name=None, **override):
    spec = dict(WORKLOADS[name or "cfg2"])
    spec.update({k: v for k, v in override.items() if v is not None})
    spec["name"] = name or "cfg2"
    return spec


def compact_samples(tree, first, count, placements, bins, seed=42, concentration=0.3, point_mass=False):
    """Samples [first, first + count) of the workload, binified, in compact form.  Per sample: `placements` placements;
    edge ~ categorical with per-sample Dirichlet-like weights (gamma(concentration) normalised, so samples differ),
    proximal position ~ U(0, bl), LWR-like weight ~ U(0,1) + 1e-3, masses normalised to 1 per sample (gappa always
    normalises: placement/function/operators.cpp:177); then the reference's bin index (functions.cpp:231-237) and one
    summed mass per occupied (edge, bin).  The same distribution as synth.sample_masses + synth.binify_host, without
    sorting 100k placements per sample (the uniform draws are sorted BEFORE they select edges: 5x faster look-up).
    point_mass: equal masses 1/placements (BASELINE cfg3's `--point-mass` variant)."""
    E = len(tree["bl"])
    bl = np.asarray(tree["bl"], np.float64)
    nb = float(bins)
    hi = np.nextafter(nb, 0.0)
    P = placements
    off = np.zeros(count + 1, np.int64)
    cells, masses = [], []
    for k in range(count):
        rng = np.random.Generator(np.random.Philox(key=seed, counter=[0, 0, 0, first + k + 1]))
        w = rng.standard_gamma(concentration, E) + 1e-12
        cdf = np.cumsum(w)
        cdf /= cdf[-1]
        u = rng.random(P)
        u.sort()
        e = np.searchsorted(cdf, u, side="right")
        np.minimum(e, E - 1, out=e)
        b = bl[e]
        x = rng.random(P) * b
        m = rng.random(P) + 1e-3
        if point_mass:          # `--point-mass`: every pquery is one unit mass at its best placement (jplace_input.cpp:138-141)
            m = np.ones(P)
        m /= m.sum()
        with np.errstate(divide="ignore", invalid="ignore"):
            v = x / b * nb
        pn = np.where(0.0 < v, v, 0.0)
        pn = np.where(hi < pn, hi, pn)
        key = e * bins + pn.astype(np.int64)          # (pn >= 0: truncation = floor)
        msum = np.bincount(key, weights=m, minlength=E * bins)
        nz = np.flatnonzero(msum)
        cells.append(nz.astype(np.uint32))
        masses.append(msum[nz])
        off[k + 1] = off[k] + len(nz)
    return dict(off=off, cell=np.concatenate(cells) if cells else np.zeros(0, np.uint32),
                mass=np.concatenate(masses) if masses else np.zeros(0, np.float64))


def expand(tree, compact, bins):
    """Compact form -> (edge, pos = bin centre, mass): the MassTrees mass_tree_binify_masses leaves (bin centre
    expression of functions.cpp:240: q * bl / nb + bl / nb / 2)."""
    bl = np.asarray(tree["bl"], np.float64)
    nb = float(bins)
    cell = compact["cell"].astype(np.int64)
    ue = (cell // bins).astype(np.int32)
    uq = (cell % bins).astype(np.float64)
    pos = (uq * bl[ue] / nb) + (bl[ue] / nb / 2.0)
    return dict(off=compact["off"], edge=ue, pos=pos, mass=compact["mass"])


def _gen_chunk(args):
    tree, first, count, placements, bins, seed = args
    return compact_samples(tree, first, count, placements, bins, seed)


def _gen_raw(args):
    tree, first, count, placements, seed = args
    return synth.sample_masses(tree, count, placements, seed=seed, first_sample=first)


def _generate(tree, first, count, spec, seed, workers):
    chunk = max(1, min(64, (count + workers - 1) // max(workers, 1)))
    jobs = [(tree, first + s, min(chunk, count - s), spec["placements"], spec["bins"], seed) for s in range(0, count, chunk)]
    if workers > 1 and len(jobs) > 1:
        with ProcessPoolExecutor(max_workers=workers) as ex:
            parts = list(ex.map(_gen_chunk, jobs))
    else:
        parts = [_gen_chunk(j) for j in jobs]
    return _concat(parts, count)


def _concat(parts, n):
    off = np.zeros(n + 1, np.int64)
    k = 0
    for prt in parts:
        c = len(prt["off"]) - 1
        off[k + 1:k + c + 1] = off[k] + prt["off"][1:]
        k += c
    return dict(off=off, cell=np.concatenate([q["cell"] for q in parts]), mass=np.concatenate([q["mass"] for q in parts]))


CACHE_PARTS = 32   # the cached cohort is cut into this many sample ranges, one file each (independent of the rank count)


def build_compact(name=None, seed=42, n_samples=None, workers=None, rank=0, world=1, cache=False, timeout=840.0, **override):
    """Returns (spec, tree, compact) for samples [0, n).  With cache=True the samples are generated once per box: the cohort
    is cut into CACHE_PARTS sample ranges, rank r of `world` produces the missing parts k = r (mod world) into /dev/shm,
    every rank then reads all parts — later jobs with the same spec (any rank count) reuse them."""
    spec = spec_of(name, **override)
    if spec["bins"] < 1:
        raise ValueError("build_compact needs a binned workload")
    tree = synth.tree_with_edges(spec["edges"], seed=seed)
    n = spec["samples"] if n_samples is None else n_samples
    workers = workers or min(os.cpu_count() or 1, 32)
    if not cache:
        return spec, tree, _generate(tree, 0, n, spec, seed, workers)
    key = hashlib.sha1(json.dumps([spec["edges"], spec["placements"], spec["bins"], seed, n, CACHE_PARTS, "v3"]).encode()).hexdigest()[:16]
    base = os.path.join("/dev/shm", "krd_bench_" + key)
    os.makedirs(base, exist_ok=True)
    bounds = [k * n // CACHE_PARTS for k in range(CACHE_PARTS + 1)]
    todo = [k for k in range(rank, CACHE_PARTS, world) if bounds[k + 1] > bounds[k] and not os.path.exists(os.path.join(base, f"part{k}.npz"))]
    if todo:
        jobs, owner = [], []
        for k in todo:
            cnt = bounds[k + 1] - bounds[k]
            chunk = max(1, min(32, (cnt * len(todo) + workers - 1) // workers))
            for s in range(0, cnt, chunk):
                jobs.append((tree, bounds[k] + s, min(chunk, cnt - s), spec["placements"], spec["bins"], seed))
                owner.append(k)
        if workers > 1 and len(jobs) > 1:
            with ProcessPoolExecutor(max_workers=workers) as ex:
                res = list(ex.map(_gen_chunk, jobs))
        else:
            res = [_gen_chunk(j) for j in jobs]
        for k in todo:
            part = _concat([r for r, o in zip(res, owner) if o == k], bounds[k + 1] - bounds[k])
            tmp = os.path.join(base, f".tmp{k}.{os.getpid()}.npz")
            np.savez(tmp, **part)
            os.replace(tmp, os.path.join(base, f"part{k}.npz"))
        del res
    parts = []
    t0 = time.time()
    for k in range(CACHE_PARTS):
        if bounds[k + 1] == bounds[k]:
            continue
        f = os.path.join(base, f"part{k}.npz")
        while not os.path.exists(f):
            if time.time() - t0 > timeout:
                raise TimeoutError(f"workload part {f} did not appear")
            time.sleep(0.05)
        with np.load(f) as z:
            parts.append(dict(off=z["off"], cell=z["cell"], mass=z["mass"]))
    return spec, tree, _concat(parts, n)


def build(name=None, seed=42, first_sample=0, n_samples=None, workers=None, **override):
    """Returns (spec, tree, masses) with masses = binified CSR (edge, pos, mass) for samples
    [first_sample, first_sample + n_samples)."""
    spec = spec_of(name, **override)
    tree = synth.tree_with_edges(spec["edges"], seed=seed)
    n = spec["samples"] if n_samples is None else n_samples
    workers = workers or min(os.cpu_count() or 1, 32)
    if spec["bins"] > 0:
        return spec, tree, expand(tree, _generate(tree, first_sample, n, spec, seed, workers), spec["bins"])
    chunk = max(1, min(32, (n + workers - 1) // workers))
    jobs = [(tree, first_sample + s0, min(chunk, n - s0), spec["placements"], seed) for s0 in range(0, n, chunk)]
    if workers > 1 and len(jobs) > 1:
        with ProcessPoolExecutor(max_workers=workers) as ex:
            parts = list(ex.map(_gen_raw, jobs))
    else:
        parts = [_gen_raw(j) for j in jobs]
    off = np.zeros(n + 1, np.int64)
    k = 0
    for prt in parts:
        c = len(prt["off"]) - 1
        off[k + 1:k + c + 1] = off[k] + prt["off"][1:]
        k += c
    raw = dict(off=off, edge=np.concatenate([q["edge"] for q in parts]), pos=np.concatenate([q["pos"] for q in parts]),
               mass=np.concatenate([q["mass"] for q in parts]))
    return spec, tree, raw


def head(masses, n):
    """First n samples of a CSR set (either form)."""
    m = int(masses["off"][n])
    return {k: (v[: n + 1].copy() if k == "off" else v[:m]) for k, v in masses.items()}
