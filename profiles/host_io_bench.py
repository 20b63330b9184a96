"""f1 / f2 of SURVEY §8(f): the host pieces around the distance call, timed against the compiled reference on the same
host (CPU only; run where /root/reference and oracle/_ref exist):
  f1  jplace ingest -> mass trees: libkrd_host (single-pass scanner, threads over files) vs genesis' JplaceReader + gappa's
      mass_tree_set() (oracle/_ref ref_jplace_load), on gappa's own 30 test samples and on 8 renamed copies of them
  f2  matrix writer: write_matrix_file (rows formatted in parallel) vs genesis MatrixWriter, n x n doubles, "matrix" format
Usage: python profiles/host_io_bench.py [n_matrix] > profiles/r2_host_io.txt"""
import glob
import os
import shutil
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle
from gappa_b200 import host

threads = len(os.sched_getaffinity(0))
src = sorted(glob.glob("/root/reference/test/data/jplace/*.jplace.gz"))
print(f"host threads: {threads}; cfg1 inputs: {len(src)} files, {sum(os.path.getsize(f) for f in src) / 1e6:.2f} MB gzip")


def best(fn, reps=3):
    t = []
    for _ in range(reps):
        t0 = time.perf_counter(); r = fn(); t.append(time.perf_counter() - t0)
    return min(t), r


with tempfile.TemporaryDirectory() as td:
    for rep in (1, 8):
        d = os.path.join(td, f"x{rep}"); os.makedirs(d)
        for c in range(rep):
            for f in src:
                shutil.copy(f, os.path.join(d, f"c{c}_" + os.path.basename(f)))
        fl = sorted(glob.glob(os.path.join(d, "*.jplace.gz")))
        mb = sum(os.path.getsize(f) for f in fl) / 1e6
        t_ours, (tree, m, ex) = best(lambda: host.load_jplace(fl, threads=threads))
        t_ref, (rtree, rm, rex) = best(lambda: oracle.ref_jplace(fl))
        same = all(np.array_equal(m[k], rm[k]) for k in ("off", "edge", "pos", "mass")) and np.array_equal(tree["bl"], rtree["bl"])
        print(f"f1 ingest x{rep}: {len(fl)} files, {int(m['off'][-1])} mass entries | libkrd_host {t_ours * 1e3:8.1f} ms ({mb / t_ours:6.1f} MB/s gzip) "
              f"| genesis mass_tree_set {t_ref * 1e3:8.1f} ms ({mb / t_ref:6.1f} MB/s) | x{t_ref / t_ours:.1f}, bit-identical mass trees: {same}")

    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
    rng = np.random.default_rng(0)
    D = rng.random((n, n)); D = (D + D.T) / 2; np.fill_diagonal(D, 0.0)
    labels = [f"sample_{i}" for i in range(n)]
    a, b = os.path.join(td, "a.csv"), os.path.join(td, "b.csv")
    t_ours, _ = best(lambda: host.write_matrix(a, D, labels=labels, threads=threads), 2)
    t_ref, _ = best(lambda: oracle.ref_write_matrix(b, D, labels=labels), 1)
    same = open(a, "rb").read() == open(b, "rb").read()
    print(f"f2 writer: {n} x {n} cells ({os.path.getsize(a) / 1e6:.0f} MB text) | libkrd_host {t_ours:6.2f} s ({n * n / t_ours / 1e6:6.1f} M cells/s) "
          f"| genesis MatrixWriter {t_ref:6.2f} s ({n * n / t_ref / 1e6:6.1f} M cells/s) | x{t_ref / t_ours:.1f}, byte-identical: {same}")
    print(f"   extrapolated to 10,000 x 10,000: ours {t_ours * (1e4 / n) ** 2:.1f} s, genesis {t_ref * (1e4 / n) ** 2:.1f} s")
