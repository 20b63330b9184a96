import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    tree = dict(parent=z["parent"], child=z["child"], bl=z["bl"], root=int(z["root"]))
    mass = dict(off=z["off"], edge=z["edge"], pos=z["pos"], mass=z["mass"])
    return tree, mass, z


def rel_err(got, want):
    return float(np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300)))


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle
    oracle.port()  # builds the C restatement on demand
    return oracle


def subset_masses(mass, idx):
    """CSR sub-set (samples idx, in that order) of a full-form mass set."""
    off = np.zeros(len(idx) + 1, np.int64)
    e, p, m = [], [], []
    for q, s in enumerate(idx):
        lo, hi = mass["off"][s], mass["off"][s + 1]
        e.append(mass["edge"][lo:hi]); p.append(mass["pos"][lo:hi]); m.append(mass["mass"][lo:hi])
        off[q + 1] = off[q] + hi - lo
    cat = lambda xs, dt: np.concatenate(xs) if xs else np.zeros(0, dt)
    return dict(off=off, edge=cat(e, np.int32), pos=cat(p, np.float64), mass=cat(m, np.float64))


def assert_matches_reference_p_below_one(tree, mass, idx, p, bins, got_block, tol=1e-10):
    """Parity for an exponent p < 1 on the samples idx (got_block = their sub-matrix).

    Pass 1: within `tol` relative of the FP64 port (= genesis), as for every other exponent.
    Pass 2, only where pass 1 fails: against the binary128 ground truth with the bound every FP64 evaluation is subject to.
    For p <= 1, | |a|^p - |b|^p | <= |a - b|^p; a cumulative mass is a sum of at most K masses (K = entries of the sample),
    so it carries a rounding error of at most K eps F_max whatever the summation order, and the distance of a pair is off
    by at most  sum_s len_s (2 K eps F_max)^p  from the exact distance of the same double-valued trees — attained where the
    two samples' cumulative masses tie (tests/test_p_below_one.py shows genesis' own walk sits at that distance from the
    truth).  A result inside that bound AND no further from the truth than 1e-10 relative plus the bound is as correct as
    the reference's."""
    import oracle
    assert p < 1.0
    sub = subset_masses(mass, idx)
    k = len(idx)
    want = oracle.port_matrix(tree, sub, p, bins=bins)
    off = ~np.eye(k, dtype=bool)
    if float(np.max(np.abs(got_block[off] - want[off]) / np.maximum(want[off], 1e-300))) <= tol:
        return
    pairs = [(a, b) for a in range(k) for b in range(a + 1, k)]
    T, _ = oracle.truth_pairs(tree, sub, p, pairs, bins=bins)
    K = int(np.max(np.diff(sub["off"]))) + 2
    f_max = max(1.0, float(np.max(np.add.reduceat(sub["mass"], sub["off"][:-1][np.diff(sub["off"]) > 0]))) if len(sub["mass"]) else 1.0)
    bound = float(np.sum(np.abs(tree["bl"]))) * (2.0 * K * 2.0 ** -53 * f_max) ** p
    got = np.array([got_block[a, b] for a, b in pairs])
    err = np.abs(got - T)
    assert np.all(err <= tol * np.abs(T) + bound), (float(np.max(err - tol * np.abs(T))), bound)
