"""TEST INFRASTRUCTURE ONLY — ctypes access to the CPU oracle.

`port`  = oracle/libkrd_oracle.so  (our plain-C restatement, oracle/krd_oracle.c)
`ref`   = oracle/_ref/libkrd_ref.so (the UNMODIFIED genesis sources compiled by oracle/Makefile
          behind our small driver oracle/ref_driver.cpp); present only when it was built in the
          container that has /root/reference — it then travels to the GPU box as a binary.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm import this
package.  gappa_b200/ never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PORT = os.path.join(_HERE, "libkrd_oracle.so")
_TRUTH = os.path.join(_HERE, "libkrd_truth.so")
_REF = os.path.join(_HERE, "_ref", "libkrd_ref.so")

_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


def build(ref=True):
    """Compile the port (always) and the reference driver (when /root/reference exists)."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "port"])
    if ref and os.path.isdir("/root/reference/libs/genesis/lib/genesis"):
        subprocess.check_call(["make", "-s", "-j8", "-C", _HERE, "ref"])


_port = None
_ref = None


def port():
    global _port
    if _port is None:
        if not os.path.exists(_PORT):
            build(ref=False)
        L = C.CDLL(_PORT)
        L.okrd_matrix.restype = C.c_int
        L.okrd_matrix.argtypes = [C.c_int, _i32p, _i32p, _f64p, C.c_int, C.c_int, _i64p, _i32p,
                                  _f64p, _f64p, C.c_double, C.c_int, C.c_int64, _f64p]
        L.okrd_binify.restype = C.c_int64
        L.okrd_binify.argtypes = [C.c_int, _i64p, _i32p, _f64p, _f64p, _f64p, C.c_int,
                                  _i64p, _i32p, _f64p, _f64p]
        L.okrd_canonicalize.restype = C.c_int64
        L.okrd_canonicalize.argtypes = [C.c_int, _i64p, _i32p, _f64p, _f64p, _i64p, _i32p, _f64p, _f64p]
        L.okrd_bin_index.restype = C.c_int
        L.okrd_bin_index.argtypes = [C.c_double, C.c_double, C.c_int]
        L.okrd_bin_center.restype = C.c_double
        L.okrd_bin_center.argtypes = [C.c_int, C.c_double, C.c_int]
        L.okrd_postorder.restype = C.c_int
        L.okrd_postorder.argtypes = [C.c_int, _i32p, _i32p, C.c_int, _i32p]
        L.okrd_max_threads.restype = C.c_int
        _port = L
    return _port


_truth = None


def truth_pairs(tree, mass, p, pairs, bins=0):
    """binary128 ground truth (oracle/krd_truth.c) of the listed (a, b) pairs: canonicalise, optionally binify with the
    port (so the masses are the doubles every implementation receives), exact-arithmetic walk.  Returns (value rounded
    to double, residual value - double(value))."""
    global _truth
    if _truth is None:
        if not os.path.exists(_TRUTH):
            build(ref=False)
        L = C.CDLL(_TRUTH)
        L.okrd_truth_pairs.restype = C.c_int
        L.okrd_truth_pairs.argtypes = [C.c_int, _i32p, _i32p, _f64p, C.c_int, C.c_int, _i64p, _i32p, _f64p, _f64p,
                                       C.c_double, C.c_int64, _i32p, _i32p, _f64p, _f64p]
        _truth = L
    mc = canonicalize(mass)
    if bins > 0:
        mc = binify(tree, mc, bins)
    pr = np.ascontiguousarray(pairs, np.int32).reshape(-1, 2)
    a = np.ascontiguousarray(pr[:, 0]); b = np.ascontiguousarray(pr[:, 1])
    out = np.zeros(len(pr)); lo = np.zeros(len(pr))
    rc = _truth.okrd_truth_pairs(*_t(tree), *_m(mc), float(p), len(pr), a, b, out, lo)
    if rc != 0:
        raise RuntimeError("okrd_truth_pairs failed: %d" % rc)
    return out, lo


def have_ref():
    return os.path.exists(_REF)


def ref():
    global _ref
    if _ref is None:
        L = C.CDLL(_REF)
        L.ref_last_error.restype = C.c_char_p
        L.ref_max_threads.restype = C.c_int
        L.ref_emd_matrix.restype = C.c_int
        L.ref_emd_matrix.argtypes = [C.c_int, _i32p, _i32p, _f64p, C.c_int, C.c_int, _i64p, _i32p,
                                     _f64p, _f64p, C.c_int, C.c_double, C.c_int, _f64p,
                                     C.POINTER(C.c_double)]
        L.ref_binify_flat.restype = C.c_int
        L.ref_binify_flat.argtypes = [C.c_int, _i32p, _i32p, _f64p, C.c_int, C.c_int, _i64p, _i32p,
                                      _f64p, _f64p, C.c_int, _i64p, _i32p, _f64p, _f64p]
        L.ref_jplace_load.restype = C.c_void_p
        L.ref_jplace_load.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.c_int, C.c_int]
        L.ref_dump_sizes.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                     C.POINTER(C.c_int64), C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.ref_dump_copy.argtypes = [C.c_void_p, _i32p, _i32p, _i32p, _f64p, _i64p, _i32p, _f64p,
                                    _f64p, _f64p]
        L.ref_dump_free.argtypes = [C.c_void_p]
        L.ref_kmeans.restype = C.c_int
        L.ref_kmeans.argtypes = [C.c_int, _i32p, _i32p, _f64p, C.c_int, C.c_int, _i64p, _i32p, _f64p, _f64p,
                                 C.c_int, C.c_int, C.c_int, _i32p, _i64p, _f64p, _f64p, _i64p, C.POINTER(C.c_double)]
        L.ref_squash.restype = C.c_int
        L.ref_squash.argtypes = [C.c_int, _i32p, _i32p, _f64p, C.c_int, C.c_int, _i64p, _i32p, _f64p, _f64p,
                                 C.c_int, C.c_double, _i64p, _f64p, C.c_char_p, C.c_int64, C.POINTER(C.c_double)]
        L.ref_write_matrix.restype = C.c_int
        L.ref_write_matrix.argtypes = [C.c_char_p, _f64p, C.c_int, C.c_int, C.POINTER(C.c_char_p), C.c_char_p]
        _ref = L
    return _ref


# ---- convenience wrappers on the flat data model (dict with numpy arrays) ------------------------
# tree  = dict(parent=int32[E], child=int32[E], bl=float64[E], root=int)
# mass  = dict(off=int64[N+1], edge=int32[M], pos=float64[M], mass=float64[M])

def _t(tree):
    return (len(tree["bl"]), np.ascontiguousarray(tree["parent"], np.int32),
            np.ascontiguousarray(tree["child"], np.int32),
            np.ascontiguousarray(tree["bl"], np.float64), int(tree["root"]))


def _m(mass):
    return (len(mass["off"]) - 1, np.ascontiguousarray(mass["off"], np.int64),
            np.ascontiguousarray(mass["edge"], np.int32),
            np.ascontiguousarray(mass["pos"], np.float64),
            np.ascontiguousarray(mass["mass"], np.float64))


def canonicalize(mass):
    n, off, e, p, m = _m(mass)
    oo = np.zeros(n + 1, np.int64); oe = np.zeros_like(e); op = np.zeros_like(p); om = np.zeros_like(m)
    w = port().okrd_canonicalize(n, off, e, p, m, oo, oe, op, om)
    return dict(off=oo, edge=oe[:w].copy(), pos=op[:w].copy(), mass=om[:w].copy())


def binify(tree, mass, bins):
    """mass must be canonical."""
    n, off, e, p, m = _m(mass)
    oo = np.zeros(n + 1, np.int64); oe = np.zeros_like(e); op = np.zeros_like(p); om = np.zeros_like(m)
    w = port().okrd_binify(n, off, e, p, m, np.ascontiguousarray(tree["bl"], np.float64), bins,
                           oo, oe, op, om)
    return dict(off=oo, edge=oe[:w].copy(), pos=op[:w].copy(), mass=om[:w].copy())


def port_matrix(tree, mass, p=1.0, bins=0, threads=0, pairs_limit=0):
    """Port oracle: canonicalise, optionally binify, all-pairs KR matrix."""
    mc = canonicalize(mass)
    if bins > 0:
        mc = binify(tree, mc, bins)
    n = len(mc["off"]) - 1
    out = np.zeros((n, n), np.float64)
    rc = port().okrd_matrix(*_t(tree), *_m(mc), float(p), threads, pairs_limit, out)
    if rc == 1:
        raise RuntimeError("Invalid exponent value p for earth mover's distance calculation. Has to be > 0.0.")
    if rc != 0:
        raise ValueError("Incompatible MassTrees.")
    return out


def ref_matrix(tree, mass, p=1.0, bins=0, threads=0):
    """The compiled reference (genesis) on the same flat inputs. Returns (matrix, seconds)."""
    n = len(mass["off"]) - 1
    out = np.zeros((n, n), np.float64)
    sec = C.c_double(0)
    rc = ref().ref_emd_matrix(*_t(tree), *_m(mass), bins, float(p), threads, out, C.byref(sec))
    if rc != 0:
        raise RuntimeError(ref().ref_last_error().decode())
    return out, sec.value


def ref_binify(tree, mass, bins):
    n, off, e, p, m = _m(mass)
    oo = np.zeros(n + 1, np.int64); oe = np.zeros_like(e); op = np.zeros_like(p); om = np.zeros_like(m)
    rc = ref().ref_binify_flat(*_t(tree), n, off, e, p, m, bins, oo, oe, op, om)
    if rc != 0:
        raise RuntimeError(ref().ref_last_error().decode())
    w = oo[-1]
    return dict(off=oo, edge=oe[:w].copy(), pos=op[:w].copy(), mass=om[:w].copy())


def ref_jplace(paths, point_mass=False, ignore_multiplicities=False):
    """gappa's mass_tree_set() ingest via genesis; returns (tree, mass, extras)."""
    L = ref()
    arr = (C.c_char_p * len(paths))(*[p.encode() for p in paths])
    h = L.ref_jplace_load(len(paths), arr, int(point_mass), int(ignore_multiplicities))
    if not h:
        raise RuntimeError(L.ref_last_error().decode())
    ne, ns, root = C.c_int(), C.c_int(), C.c_int()
    nm = C.c_int64(); dia = C.c_double()
    L.ref_dump_sizes(h, C.byref(ne), C.byref(ns), C.byref(nm), C.byref(root), C.byref(dia))
    E, N, M = ne.value, ns.value, nm.value
    parent = np.zeros(E, np.int32); child = np.zeros(E, np.int32); en = np.zeros(E, np.int32)
    bl = np.zeros(E); off = np.zeros(N + 1, np.int64); edge = np.zeros(M, np.int32)
    pos = np.zeros(M); mass = np.zeros(M); pend = np.zeros(N)
    L.ref_dump_copy(h, parent, child, en, bl, off, edge, pos, mass, pend)
    L.ref_dump_free(h)
    tree = dict(parent=parent, child=child, bl=bl, root=root.value)
    m = dict(off=off, edge=edge, pos=pos, mass=mass)
    return tree, m, dict(edge_num=en, pendant=pend, diameter=dia.value)


def ref_write_matrix(path, matrix, fmt="matrix", labels=None, corner="Sample"):
    """genesis MatrixWriter<double> configured like gappa's write_matrix (matrix_output.cpp:76-111)."""
    m = np.ascontiguousarray(matrix, np.float64)
    n = m.shape[0]
    lab = None if labels is None else (C.c_char_p * n)(*[s.encode() for s in labels])
    rc = ref().ref_write_matrix(os.fspath(path).encode(), m, n, {"matrix": 0, "list": 1, "triangular": 2}[fmt], lab,
                                corner.encode())
    if rc != 0:
        raise RuntimeError(ref().ref_last_error().decode())


def ref_kmeans(tree, mass, k, init_centroids, bins=0, max_iterations=100):
    """genesis MassTreeKmeans::run started from data points as centroids.
    Returns (assignments[n], distances[n], variances[k], iterations, seconds)."""
    n = len(mass["off"]) - 1
    a = np.zeros(n, np.int64); d = np.zeros(n, np.float64); v = np.zeros(k, np.float64); it = np.zeros(1, np.int64)
    init = np.ascontiguousarray(init_centroids, np.int32)
    assert len(init) == k
    sec = C.c_double(0)
    rc = ref().ref_kmeans(*_t(tree), *_m(mass), bins, k, max_iterations, init, a, d, v, it, C.byref(sec))
    if rc != 0:
        raise RuntimeError(ref().ref_last_error().decode())
    return a, d, v, int(it[0]), sec.value


def ref_squash(tree, mass, p=1.0, bins=0):
    """genesis SquashClustering::run on the flat inputs. Returns (index[n-1,2], dist[n-1,2], tree_string, seconds)."""
    n = len(mass["off"]) - 1
    idx = np.zeros(2 * (n - 1), np.int64); dst = np.zeros(2 * (n - 1), np.float64)
    buf = C.create_string_buffer(64 * n + 64)
    sec = C.c_double(0)
    rc = ref().ref_squash(*_t(tree), *_m(mass), bins, float(p), idx, dst, buf, len(buf), C.byref(sec))
    if rc != 0:
        raise RuntimeError(ref().ref_last_error().decode())
    return idx.reshape(-1, 2), dst.reshape(-1, 2), buf.value.decode(), sec.value
