"""ctypes binding of libkrd_host.so (include/krd_host.h): jplace ingest -> flat mass arrays and the
distance-matrix writer — the host-side pieces of `gappa analyze krd` around the distance call
(gappa src/options/jplace_input.cpp:295-332, src/options/matrix_output.cpp:76-111).
No GPU needed; nothing under oracle/ is imported here."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkrd_host.so")
CLI_PATH = os.path.join(_HERE, "gappa-b200")

EXPORTS = ["krd_host_last_error", "krd_host_load_jplace", "krd_host_free_set", "krd_host_set_sizes",
           "krd_host_set_copy", "krd_host_set_label", "krd_host_write_matrix", "krd_host_parse_number",
           "krd_host_add_pendant_work"]

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `make -C gappa_b200/host`")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.krd_host_last_error.restype = C.c_char_p
        L.krd_host_load_jplace.restype = vp
        L.krd_host_load_jplace.argtypes = [C.c_int32, C.POINTER(C.c_char_p), C.c_int32, C.c_int32, C.c_int32]
        L.krd_host_free_set.argtypes = [vp]
        L.krd_host_set_sizes.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int64),
                                         C.POINTER(C.c_int32), C.POINTER(C.c_double)]
        L.krd_host_set_copy.argtypes = [vp] * 10
        L.krd_host_set_label.restype = C.c_char_p
        L.krd_host_set_label.argtypes = [vp, C.c_int32]
        L.krd_host_write_matrix.restype = C.c_int64
        L.krd_host_write_matrix.argtypes = [C.c_char_p, vp, C.c_int32, C.c_char_p, C.POINTER(C.c_char_p), C.c_char_p,
                                            C.c_int32, C.c_int32]
        L.krd_host_add_pendant_work.argtypes = [vp, C.c_int32, vp]
        L.krd_host_parse_number.restype = C.c_double
        L.krd_host_parse_number.argtypes = [C.c_char_p]
        _lib = L
    return _lib


def load_jplace(paths, point_mass=False, ignore_multiplicities=False, threads=4):
    """Files and/or directories -> (tree, masses, extras) on the flat data model of krd_b200.h.
    Raises RuntimeError with the reference's message on malformed input."""
    L = lib()
    arr = (C.c_char_p * len(paths))(*[os.fspath(p).encode() for p in paths])
    h = L.krd_host_load_jplace(len(paths), arr, int(point_mass), int(ignore_multiplicities), threads)
    if not h:
        raise RuntimeError(L.krd_host_last_error().decode())
    try:
        ne, ns, root, nm, dia = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int64(), C.c_double()
        L.krd_host_set_sizes(h, C.byref(ne), C.byref(ns), C.byref(nm), C.byref(root), C.byref(dia))
        E, N, M = ne.value, ns.value, nm.value
        parent = np.zeros(E, np.int32); child = np.zeros(E, np.int32); en = np.zeros(E, np.int64)
        bl = np.zeros(E); off = np.zeros(N + 1, np.int64); edge = np.zeros(M, np.int32)
        pos = np.zeros(M); mass = np.zeros(M); pend = np.zeros(N)
        L.krd_host_set_copy(h, *[a.ctypes.data for a in (parent, child, en, bl, off, edge, pos, mass, pend)])
        labels = [L.krd_host_set_label(h, i).decode() for i in range(N)]
    finally:
        L.krd_host_free_set(h)
    tree = dict(parent=parent, child=child, bl=bl, root=root.value)
    masses = dict(off=off, edge=edge, pos=pos, mass=mass)
    return tree, masses, dict(edge_num=en, pendant=pend, diameter=dia.value, labels=labels)


def write_matrix(path, matrix, fmt="matrix", labels=None, corner="Sample", gzip=False, threads=4):
    """Byte-compatible with the reference's MatrixWriter output. Returns uncompressed bytes written."""
    m = np.ascontiguousarray(matrix, np.float64)
    n = m.shape[0]
    lab = None
    if labels is not None:
        lab = (C.c_char_p * n)(*[s.encode() for s in labels])
    r = lib().krd_host_write_matrix(os.fspath(path).encode(), m.ctypes.data, n, fmt.encode(), lab, corner.encode(),
                                    int(gzip), threads)
    if r < 0:
        raise RuntimeError(lib().krd_host_last_error().decode())
    return r


def parse_number(text):
    return float(lib().krd_host_parse_number(text.encode()))


def add_pendant_work(matrix, pendant):
    """placement::earth_movers_distance(SampleSet, p, with_pendant_length=true) from the mass-tree matrix
    (placement/function/emd.cpp:107-129): returns matrix + pendant[i] + pendant[j] in every cell."""
    m = np.array(matrix, np.float64, order="C")
    pend = np.ascontiguousarray(pendant, np.float64)
    assert m.ndim == 2 and m.shape[0] == m.shape[1] == len(pend)
    lib().krd_host_add_pendant_work(m.ctypes.data, m.shape[0], pend.ctypes.data)
    return m
