"""Generates the golden fixtures in this directory from the COMPILED REFERENCE
(oracle/_ref/libkrd_ref.so = unmodified genesis sources, built by oracle/Makefile).
Run in the build container only (it reads /root/reference test data):

    python tests/golden/make_golden.py

Fixtures (numpy .npz, float64 kept bit-exact):
  test_ab.npz      genesis' own unit-test inputs test_a.jplace / test_b.jplace
                   (libs/genesis/test/src/placement/measures.cpp:68-110) as flat arrays, with the
                   reference matrices for p = 1, 2, 0.5 and the per-sample pendant work.
  cfg1.npz         gappa's 30 test samples (test/data/jplace, test/scripts/analyze-krd-1.sh) after
                   gappa's ingest (mass_tree_set), reference matrices p = 1, 2, 0.5, diameter.
  cfg1_pm.npz      same with --point-mass, p = 1.
  synth_small.npz  synthetic 24 samples x 199 edges x 400 placements; reference matrices for
                   (bins, p) in {0,1,4} x {1, 2, 0.5, 1.7}; reference binify output for bins=4.
"""
import glob
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from gappa_b200 import synth  # noqa: E402

REF = "/root/reference"


def pack(tree, mass):
    return dict(parent=tree["parent"], child=tree["child"], bl=tree["bl"], root=np.int64(tree["root"]),
                off=mass["off"], edge=mass["edge"], pos=mass["pos"], mass=mass["mass"])


def main():
    d = REF + "/libs/genesis/test/data/placement/"
    tree, mass, ex = oracle.ref_jplace([d + "test_a.jplace", d + "test_b.jplace"])
    out = pack(tree, mass)
    out["pendant"] = ex["pendant"]
    for p in (1.0, 2.0, 0.5):
        out[f"m_p{p}"] = oracle.ref_matrix(tree, mass, p)[0]
    np.savez_compressed(os.path.join(HERE, "test_ab.npz"), **out)
    print("test_ab", out["m_p1.0"][0, 1], out["m_p2.0"][0, 1], out["m_p0.5"][0, 1])

    paths = sorted(glob.glob(REF + "/test/data/jplace/*.jplace.gz"))
    tree, mass, ex = oracle.ref_jplace(paths)
    out = pack(tree, mass)
    out["diameter"] = np.float64(ex["diameter"])
    out["edge_num"] = ex["edge_num"]
    out["names"] = np.array([os.path.basename(p).replace(".jplace.gz", "") for p in paths])
    for p in (1.0, 2.0, 0.5):
        out[f"m_p{p}"] = oracle.ref_matrix(tree, mass, p)[0]
    np.savez_compressed(os.path.join(HERE, "cfg1.npz"), **out)
    print("cfg1", out["m_p1.0"][0, 1], len(mass["edge"]))

    tree, mass, ex = oracle.ref_jplace(paths, point_mass=True)
    out = pack(tree, mass)
    out["m_p1.0"] = oracle.ref_matrix(tree, mass, 1.0)[0]
    np.savez_compressed(os.path.join(HERE, "cfg1_pm.npz"), **out)

    tree = synth.tree_with_edges(199, seed=3)
    mass = synth.sample_masses(tree, 24, 400, seed=3)
    out = pack(tree, mass)
    for bins in (0, 1, 4):
        for p in (1.0, 2.0, 0.5, 1.7):
            out[f"m_b{bins}_p{p}"] = oracle.ref_matrix(tree, mass, p, bins=bins)[0]
    b = oracle.ref_binify(tree, mass, 4)
    out["bin4_off"], out["bin4_edge"], out["bin4_pos"], out["bin4_mass"] = b["off"], b["edge"], b["pos"], b["mass"]
    np.savez_compressed(os.path.join(HERE, "synth_small.npz"), **out)
    print("synth_small done")


if __name__ == "__main__":
    main()
