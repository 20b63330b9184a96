"""Synthetic inputs of the shapes BASELINE.json names: a random rooted binary
reference tree and per-sample placement masses, produced from counter-based RNG streams
(numpy Philox, key = seed, stream = sample index) so every rank / test regenerates the same data.

Pure host-side input generation; no distance arithmetic happens here.
"""
import numpy as np


def random_tree(n_leaves, seed=42):
    """Random rooted binary tree: n_leaves leaves -> E = 2*n_leaves - 2 edges, nodes 0..E, root 0.
    (gappa's `simulate random-tree --leaf-count n` is unrooted: 2n-3 edges; use drop_root_edge to
    get that count.)  Built by random leaf splitting; branch lengths ~ U(0,1)."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    parent, child = [], []
    leaves = [0]
    next_node = 1
    while len(leaves) < n_leaves:
        k = int(rng.integers(len(leaves)))
        v = leaves[k]
        a, b = next_node, next_node + 1
        next_node += 2
        parent += [v, v]
        child += [a, b]
        leaves[k] = a
        leaves.append(b)
    E = len(parent)
    bl = rng.random(E)
    return dict(parent=np.asarray(parent, np.int32), child=np.asarray(child, np.int32),
                bl=bl.astype(np.float64), root=0)


def tree_with_edges(n_edges, seed=42):
    """Tree with exactly n_edges edges (odd counts get a trifurcating root, like an unrooted
    Newick tree read by genesis: 9,999 edges = 5,001 leaves)."""
    if n_edges % 2 == 0:
        return random_tree(n_edges // 2 + 1, seed)
    t = random_tree((n_edges - 1) // 2 + 1, seed)
    rng = np.random.Generator(np.random.Philox(key=seed + 1))
    new = len(t["bl"]) + 1
    t["parent"] = np.append(t["parent"], np.int32(0))
    t["child"] = np.append(t["child"], np.int32(new))
    t["bl"] = np.append(t["bl"], rng.random())
    return t


def sample_masses(tree, n_samples, placements_per_sample, seed=42, first_sample=0,
                  concentration=0.3):
    """Per sample: `placements_per_sample` placements; edge ~ categorical with per-sample
    Dirichlet-like weights (gamma(concentration) normalised, so samples differ), proximal
    position ~ U(0, bl), LWR-like weight ~ U(0,1); masses normalised to 1 per sample
    (gappa always normalises: placement/function/operators.cpp:177).
    Returns CSR dict(off, edge, pos, mass) in MassTree order (edge asc, pos asc)."""
    E = len(tree["bl"])
    bl = np.asarray(tree["bl"], np.float64)
    P = placements_per_sample
    off = np.arange(n_samples + 1, dtype=np.int64) * P
    edge = np.empty(n_samples * P, np.int32)
    pos = np.empty(n_samples * P, np.float64)
    mass = np.empty(n_samples * P, np.float64)
    for k in range(n_samples):
        i = first_sample + k
        rng = np.random.Generator(np.random.Philox(key=seed, counter=[0, 0, 0, i + 1]))
        w = rng.standard_gamma(concentration, E) + 1e-12
        cdf = np.cumsum(w)
        cdf /= cdf[-1]
        e = np.searchsorted(cdf, rng.random(P), side="right").astype(np.int32)
        np.minimum(e, E - 1, out=e)
        x = rng.random(P) * bl[e]
        m = rng.random(P) + 1e-3
        m /= m.sum()
        order = np.lexsort((x, e))
        s = slice(k * P, (k + 1) * P)
        edge[s] = e[order]
        pos[s] = x[order]
        mass[s] = m[order]
    return dict(off=off, edge=edge, pos=pos, mass=mass)


def binify_host(tree, masses, bins):
    """Host-side input preparation for the binned configs: what mass_tree_binify_masses
    (genesis tree/mass_tree/functions.cpp:222-265) would hand to earth_movers_distance — one
    entry per occupied (sample, edge, bin), at the bin centre.  Used only to SHAPE bench inputs
    (both arms then receive the same binified mass trees); parity of the binning itself is tested
    against the oracle in tests/."""
    bl = np.asarray(tree["bl"], np.float64)
    nb = float(bins)
    off, edge, pos, mass = masses["off"], masses["edge"], masses["pos"], masses["mass"]
    n = len(off) - 1
    out_off = np.zeros(n + 1, np.int64)
    oe, op, om = [], [], []
    hi = np.nextafter(nb, 0.0)
    for i in range(n):
        s = slice(off[i], off[i + 1])
        e = edge[s]
        b = bl[e]
        with np.errstate(divide="ignore", invalid="ignore"):
            v = pos[s] / b * nb
        pn = np.where(0.0 < v, v, 0.0)
        pn = np.where(hi < pn, hi, pn)
        q = np.floor(pn).astype(np.int64)
        key = e.astype(np.int64) * bins + q
        uk, inv = np.unique(key, return_inverse=True)
        msum = np.zeros(len(uk))
        np.add.at(msum, inv, mass[s])
        ue = (uk // bins).astype(np.int32)
        uq = (uk % bins).astype(np.float64)
        centre = (uq * bl[ue] / nb) + (bl[ue] / nb / 2.0)
        oe.append(ue); op.append(centre); om.append(msum)
        out_off[i + 1] = out_off[i] + len(uk)
    return dict(off=out_off, edge=np.concatenate(oe), pos=np.concatenate(op), mass=np.concatenate(om))
