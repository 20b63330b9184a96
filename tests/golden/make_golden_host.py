"""Golden fixtures for the HOST side (jplace ingest, matrix writer, CLI) — generated from the
COMPILED REFERENCE (oracle/_ref/libkrd_ref.so).  Run in the build container only:

    python tests/golden/make_golden_host.py

Writes
  jplace/*.jplace[.gz]     small synthetic jplace documents (ours, seeded) that exercise: gzip,
                           'nm' and 'n'+'m' multiplicities, proximal_length vs distal_length,
                           numbers with exponents / many digits, LWR ties (point mass), key order
                           (fields before/after placements), per-sample branch lengths (averaging),
                           several placements on one position, jplace version 1 ([n] edge numbers).
  host_ingest.npz          reference mass_tree_set() arrays for the option combinations
                           (default / --point-mass / --ignore-multiplicities) + diameter + pendant,
                           reference matrices p = 1, 2, 0.5.
  csv/cluster.newick       `analyze squash`: reference SquashClustering mergers as the Newick text squash.cpp writes
  csv/k_2_assignments.csv, overview.csv   `analyze phylogenetic-kmeans`: reference MassTreeKmeans result in the files' format
  csv/*.csv                reference MatrixWriter output of the reference p = 1 matrix in the three
                           formats (with/without labels), plus the normalised one; and cfg1's
                           krd_matrix.csv (7847 bytes, test/scripts/analyze-krd-1.sh:7).
"""
import glob
import gzip
import json
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402


def random_tree_newick(rng, n_leaves, scale, version):
    """Random rooted binary topology (shared via the rng seed), per-sample branch-length scale."""
    nodes = [f"L{i}" for i in range(n_leaves)]
    topo_rng = np.random.default_rng(1234)   # same topology and base lengths for all samples
    items = [(n, None) for n in nodes]
    while len(items) > 1:
        i, j = sorted(topo_rng.choice(len(items), 2, replace=False))
        b = items.pop(j); a = items.pop(i)
        items.append(((a, b), None))
    counter = [0]
    base_rng = np.random.default_rng(99)

    def emit(item, is_root):
        node = item[0]
        if isinstance(node, tuple):
            s = "(" + emit(node[0], False) + "," + emit(node[1], False) + ")"
        else:
            s = node
        if is_root:
            return s
        bl = float(base_rng.uniform(0.05, 1.0)) * scale
        k = counter[0]; counter[0] += 1
        tag = f"[{k}]" if version == 1 else f"{{{k}}}"
        return f"{s}:{bl:.6f}{tag}"

    return emit(items[0], True) + ";", counter[0]


def make_jplace(path, seed, version=3, fields_first=False, use_prox=False, gz=False, scale=1.0):
    rng = np.random.default_rng(seed)
    newick, n_edges = random_tree_newick(rng, 12, scale, version)
    # branch lengths by edge_num, to draw valid distal lengths
    import re
    bls = {int(m.group(2)): float(m.group(1)) for m in re.finditer(r":([0-9.]+)[\[{](\d+)[\]}]", newick)}
    fields = ["edge_num", "likelihood", "like_weight_ratio", "proximal_length" if use_prox else "distal_length", "pendant_length"]
    if seed % 2:
        fields.insert(2, "classification")   # a non-numeric, unused column
    pqs = []
    for q in range(40):
        k = int(rng.integers(1, 5))
        lw = rng.dirichlet(np.ones(k))
        if q % 7 == 0 and k >= 2:
            lw[:] = 1.0 / k                    # exact LWR ties for --point-mass
        rows = []
        for j in range(k):
            e = int(rng.integers(0, n_edges))
            d = float(rng.uniform(0, bls[e]))
            if q % 9 == 0:
                d = round(d, 2)                # repeated positions on an edge
            cells = {"edge_num": e, "likelihood": -1234.5 - q, "like_weight_ratio": float(lw[j]),
                     "distal_length": d, "proximal_length": d, "pendant_length": float(rng.uniform(0, 0.3)),
                     "classification": "x"}
            rows.append([cells[f] for f in fields])
        pq = {"p": rows}
        if q % 3 == 0:
            pq["nm"] = [[f"q{q}_{t}", float(rng.integers(1, 9)) + (0.5 if t else 0.0)] for t in range(1 + q % 2)]
        elif q % 3 == 1:
            pq["n"] = [f"q{q}"]; pq["m"] = float(rng.integers(1, 5))
        else:
            pq["n"] = [f"q{q}a", f"q{q}b"]
        pqs.append(pq)
    doc = {}
    if fields_first:
        doc["fields"] = fields
    doc["tree"] = newick
    doc["placements"] = pqs
    if not fields_first:
        doc["fields"] = fields
    doc["metadata"] = {"invocation": "make_golden_host.py", "nested": {"a": [1, 2, {"b": None}], "t": True}}
    doc["version"] = version
    text = json.dumps(doc, indent=1)
    # numbers with exponents: rewrite some LWRs in scientific notation (same value text-wise is not
    # required; the reference parses whatever is written)
    text = text.replace("-1234.5", "-1.2345e3", 3)
    if gz:
        with open(path, "wb") as raw, gzip.GzipFile(filename="", mode="wb", fileobj=raw, mtime=0) as f:   # reproducible bytes
            f.write(text.encode())
    else:
        open(path, "w").write(text)


def pack(prefix, out, tree, mass, ex):
    for k in ("parent", "child", "bl"):
        out[f"{prefix}_{k}"] = tree[k]
    out[f"{prefix}_root"] = np.int64(tree["root"])
    for k in ("off", "edge", "pos", "mass"):
        out[f"{prefix}_{k}"] = mass[k]
    out[f"{prefix}_edge_num"] = ex["edge_num"]
    out[f"{prefix}_pendant"] = ex["pendant"]
    out[f"{prefix}_diameter"] = np.float64(ex["diameter"])


def main():
    jd = os.path.join(HERE, "jplace"); cd = os.path.join(HERE, "csv")
    for d in (jd, cd):
        shutil.rmtree(d, ignore_errors=True); os.makedirs(d)
    make_jplace(os.path.join(jd, "s_a.jplace"), 1, scale=1.0)
    make_jplace(os.path.join(jd, "s_b.jplace.gz"), 2, gz=True, fields_first=True, scale=1.1)
    make_jplace(os.path.join(jd, "s_c.jplace"), 3, use_prox=True, scale=0.9)
    make_jplace(os.path.join(jd, "s_d.jplace"), 4, version=1, scale=1.0)
    make_jplace(os.path.join(jd, "s_e.jplace.gz"), 5, gz=True, scale=1.3)
    paths = sorted(glob.glob(jd + "/*.jplace*"))
    labels = [os.path.basename(p).split(".")[0] for p in paths]
    out = {"labels": np.array(labels)}
    for name, kw in (("default", {}), ("pm", dict(point_mass=True)), ("im", dict(ignore_multiplicities=True))):
        tree, mass, ex = oracle.ref_jplace(paths, **kw)
        pack(name, out, tree, mass, ex)
        for p in (1.0, 2.0, 0.5):
            out[f"{name}_m_p{p}"] = oracle.ref_matrix(tree, mass, p)[0]
    np.savez_compressed(os.path.join(HERE, "host_ingest.npz"), **out)
    m = out["default_m_p1.0"]
    for fmt in ("matrix", "list", "triangular"):
        oracle.ref_write_matrix(os.path.join(cd, f"{fmt}.csv"), m, fmt, labels)
        oracle.ref_write_matrix(os.path.join(cd, f"{fmt}_nolabels.csv"), m, fmt, None)
    oracle.ref_write_matrix(os.path.join(cd, "matrix_normalized.csv"), m / out["default_diameter"], "matrix", labels)
    oracle.ref_write_matrix(os.path.join(cd, "matrix_pm_p2.csv"), out["pm_m_p2.0"], "matrix", labels)
    # `analyze squash` on the same documents: reference mergers -> cluster.newick text (squash.cpp:171-174)
    from gappa_b200 import krd
    tree, mass, ex = oracle.ref_jplace(paths)
    idx, dst, _, _ = oracle.ref_squash(tree, mass, p=1.0)
    mg = np.zeros(len(idx), krd.MERGER_DTYPE)
    mg["index_a"], mg["index_b"], mg["distance_a"], mg["distance_b"] = idx[:, 0], idx[:, 1], dst[:, 0], dst[:, 1]
    open(os.path.join(cd, "cluster.newick"), "w").write(krd.squash_tree_string(labels, mg))
    # `analyze phylogenetic-kmeans --bins 5 --k 2` started from samples 0 and 3: reference assignments table
    # (write_assignment_file, src/commands/analyze/kmeans.cpp:163-192) and overview (kmeans.cpp:194-269)
    a, d, v, it, _ = oracle.ref_kmeans(tree, mass, 2, [0, 3], bins=5)
    with open(os.path.join(cd, "k_2_assignments.csv"), "w") as f:
        f.write("Sample\tCluster\tDistance\n")
        for lab, ai, di in zip(labels, a, d):
            f.write("%s\t%d\t%g\n" % (lab, ai, di))
    with open(os.path.join(cd, "overview.csv"), "w") as f:
        f.write("k\tDistance\tVariance\n%d\t%g\t%g\n" % (2, d.sum() / len(d), (d * d).sum() / len(d)))
    g = np.load(os.path.join(HERE, "cfg1.npz"))
    oracle.ref_write_matrix(os.path.join(cd, "cfg1_krd_matrix.csv"), g["m_p1.0"], "matrix", [str(x) for x in g["names"]])
    print("cfg1 csv bytes", os.path.getsize(os.path.join(cd, "cfg1_krd_matrix.csv")))
    print("fixtures:", paths, m[0, 1])


if __name__ == "__main__":
    main()
