"""ctypes binding of libkrd_b200.so (the C ABI in include/krd_b200.h) plus a thin host-side mirror
of the reference operator

    genesis::tree::earth_movers_distance(std::vector<MassTree> const&, double p)
        libs/genesis/lib/genesis/tree/mass_tree/emd.hpp:124, emd.cpp:183-235

There is NO CPU fallback: if the CUDA library is missing or no B200 is visible, every entry point
raises.  Nothing under oracle/ is imported here.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkrd_b200.so")

KRD_OK = 0
KRD_ERR_INVALID_EXPONENT = 1
KRD_ERR_INCOMPATIBLE_TREES = 2
KRD_ERR_TOO_FEW_SAMPLES = 3
KRD_ERR_INVALID_ARGUMENT = 4
KRD_ERR_CUDA = 5
KRD_ERR_STATE = 6
KRD_ERR_NO_DEVICE = 7

EXPORTS = [
    "krd_create", "krd_destroy", "krd_last_error", "krd_set_stream", "krd_set_tree",
    "krd_set_masses", "krd_set_masses_device", "krd_compute", "krd_compute_device",
    "krd_shard_doubles", "krd_compute_shard", "krd_unpack_shards", "krd_get_stats",
    "krd_squash", "krd_kmeans", "krd_kmeans_ex", "krd_matrix_from_F_noalloc", "krd_set_masses_binned", "krd_set_masses_binned_device", "krd_set_masses_binned16", "krd_set_masses_binned16_device", "krd_unpack_shards_rows", "krd_compute_allgather", "krd_compute_allgather_host", "krd_scatter_shard_rows", "krd_scatter_tile_range_rows", "krd_set_late_blocks", "krd_copy_rect_to_host_async", "krd_shard_to_host_async", "krd_compute_tile_range", "krd_tile_range_to_host_async", "krd_p2p_alloc", "krd_p2p_open", "krd_p2p_close",
    "krd_p2p_free", "krd_p2p_enable_peer", "krd_nccl_init_all", "krd_nccl_destroy", "krd_compute_shard_host", "krd_unpack_shards_host", "krd_tile_count", "krd_tile_list", "krd_tile_list_late", "krd_rank_tiles", "krd_invalidate_F", "krd_debug_copy_F", "krd_host_alloc", "krd_host_free", "krd_measure_fp64_peak",
]


class KrdConfig(C.Structure):
    _fields_ = [("device", C.c_int), ("bins", C.c_int), ("split", C.c_int), ("builder", C.c_int),
                ("p2_mode", C.c_int), ("pipeline", C.c_int), ("max_f_mb", C.c_int), ("reserved", C.c_int * 1)]


class KrdStats(C.Structure):
    _fields_ = [("n_samples", C.c_int64), ("n_edges", C.c_int64), ("n_masses", C.c_int64),
                ("n_segments", C.c_int64), ("ld_samples", C.c_int64), ("f_bytes", C.c_int64),
                ("n_tiles", C.c_int64), ("split", C.c_int32), ("pair_ctas", C.c_int32),
                ("levels", C.c_int32), ("launches", C.c_int32), ("ms_build", C.c_float),
                ("ms_pairs", C.c_float), ("ms_epilogue", C.c_float), ("ms_slabs", C.c_float),
                ("gram_fixups", C.c_int32), ("reserved_i", C.c_int32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


MERGER_DTYPE = np.dtype([("index_a", np.int32), ("index_b", np.int32), ("distance_a", np.float64), ("distance_b", np.float64)])


def squash_tree_string(labels, mergers):
    """SquashClustering::tree_string (squash_clustering.cpp:251-274): Newick of the cluster tree; std::to_string(double) = %f."""
    lst = list(labels)
    n = len(lst)
    for i, m in enumerate(mergers):
        a = "%s:%f" % (lst[m["index_a"]], m["distance_a"])
        b = "%s:%f" % (lst[m["index_b"]], m["distance_b"])
        lst[m["index_a"]] = ""; lst[m["index_b"]] = ""
        lst.append("(%s,%s)%d" % (a, b, i + n))
    return lst[-1] + ";"


class KrdError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(message)
        self.status = status


_lib = None


def lib():
    """Load the CUDA library; fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("KRD_B200_LIB", LIB_PATH)   # tuning variants built with `make EXTRA=... LIB=...` (csrc/Makefile)
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). gappa_b200 has no CPU fallback.")
    L = C.CDLL(path)
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    L.krd_create.restype = C.c_int
    L.krd_create.argtypes = [C.POINTER(KrdConfig), C.POINTER(vp)]
    L.krd_destroy.restype = None
    L.krd_destroy.argtypes = [vp]
    L.krd_last_error.restype = C.c_char_p
    L.krd_last_error.argtypes = [vp]
    L.krd_set_stream.restype = C.c_int
    L.krd_set_stream.argtypes = [vp, vp]
    L.krd_set_tree.restype = C.c_int
    L.krd_set_tree.argtypes = [vp, i32, vp, vp, vp, i32]
    L.krd_set_masses.restype = C.c_int
    L.krd_set_masses.argtypes = [vp, i32, vp, vp, vp, vp]
    L.krd_set_masses_device.restype = C.c_int
    L.krd_set_masses_device.argtypes = [vp, i32, vp, vp, vp, vp]
    L.krd_set_masses_binned.restype = C.c_int
    L.krd_set_masses_binned.argtypes = [vp, i32, vp, vp, vp]
    L.krd_set_masses_binned_device.restype = C.c_int
    L.krd_set_masses_binned_device.argtypes = [vp, i32, vp, vp, vp]
    L.krd_set_masses_binned16.restype = C.c_int
    L.krd_set_masses_binned16.argtypes = [vp, i32, vp, vp, vp]
    L.krd_set_masses_binned16_device.restype = C.c_int
    L.krd_set_masses_binned16_device.argtypes = [vp, i32, vp, vp, vp]
    L.krd_unpack_shards_rows.restype = C.c_int
    L.krd_unpack_shards_rows.argtypes = [vp, i32, vp, i32, i32, vp]
    L.krd_compute_allgather.restype = C.c_int
    L.krd_compute_allgather.argtypes = [vp, dbl, dbl, i32, i32, vp, i32, i32, vp]
    L.krd_compute_allgather_host.restype = C.c_int
    L.krd_compute_allgather_host.argtypes = [vp, dbl, dbl, i32, i32, vp, i32, i32, vp]
    L.krd_scatter_shard_rows.restype = C.c_int
    L.krd_scatter_shard_rows.argtypes = [vp, i32, i32, vp, vp, vp]
    L.krd_scatter_tile_range_rows.restype = C.c_int
    L.krd_scatter_tile_range_rows.argtypes = [vp, i64, i64, vp, i32, vp, vp]
    L.krd_set_late_blocks.restype = C.c_int
    L.krd_set_late_blocks.argtypes = [vp, i32, C.POINTER(i64)]
    L.krd_copy_rect_to_host_async.restype = C.c_int
    L.krd_copy_rect_to_host_async.argtypes = [vp, i64, vp, i64, i64, i64, vp]
    L.krd_shard_to_host_async.restype = C.c_int
    L.krd_shard_to_host_async.argtypes = [vp, i32, i32, vp, vp, vp, vp]
    L.krd_compute_tile_range.restype = C.c_int
    L.krd_compute_tile_range.argtypes = [vp, dbl, dbl, i64, i64, vp]
    L.krd_tile_range_to_host_async.restype = C.c_int
    L.krd_tile_range_to_host_async.argtypes = [vp, i64, i64, vp, vp, vp, vp]
    L.krd_p2p_alloc.restype = C.c_int
    L.krd_p2p_alloc.argtypes = [i32, i64, C.POINTER(vp), vp]
    L.krd_p2p_open.restype = C.c_int
    L.krd_p2p_open.argtypes = [i32, vp, C.POINTER(vp)]
    L.krd_p2p_close.restype = None
    L.krd_p2p_close.argtypes = [vp]
    L.krd_p2p_free.restype = None
    L.krd_p2p_free.argtypes = [vp]
    L.krd_p2p_enable_peer.restype = C.c_int
    L.krd_p2p_enable_peer.argtypes = [i32, i32]
    L.krd_nccl_init_all.restype = C.c_int
    L.krd_nccl_init_all.argtypes = [i32, vp, vp]
    L.krd_nccl_destroy.restype = None
    L.krd_nccl_destroy.argtypes = [i32, vp]
    L.krd_compute.restype = C.c_int
    L.krd_compute.argtypes = [vp, dbl, dbl, vp]
    L.krd_compute_device.restype = C.c_int
    L.krd_compute_device.argtypes = [vp, dbl, dbl, vp]
    L.krd_shard_doubles.restype = i64
    L.krd_shard_doubles.argtypes = [vp, i32]
    L.krd_compute_shard.restype = C.c_int
    L.krd_compute_shard.argtypes = [vp, dbl, dbl, i32, i32, vp]
    L.krd_unpack_shards.restype = C.c_int
    L.krd_unpack_shards.argtypes = [vp, i32, vp, vp]
    L.krd_compute_shard_host.restype = C.c_int
    L.krd_compute_shard_host.argtypes = [vp, dbl, dbl, i32, i32, vp]
    L.krd_unpack_shards_host.restype = C.c_int
    L.krd_unpack_shards_host.argtypes = [i32, i32, vp, vp]
    L.krd_kmeans.restype = C.c_int
    L.krd_kmeans.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, vp, C.POINTER(C.c_int32)]
    L.krd_kmeans_ex.restype = C.c_int
    L.krd_kmeans_ex.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, vp, vp, C.POINTER(C.c_int32)]
    L.krd_matrix_from_F_noalloc.restype = C.c_int
    L.krd_matrix_from_F_noalloc.argtypes = [i32, vp, vp, i64, i32, i64, dbl, dbl, vp, vp]
    L.krd_squash.restype = C.c_int
    L.krd_squash.argtypes = [vp, dbl, vp]
    L.krd_get_stats.restype = C.c_int
    L.krd_get_stats.argtypes = [vp, C.POINTER(KrdStats)]
    L.krd_tile_count.restype = i64
    L.krd_tile_count.argtypes = [i32]
    L.krd_tile_list.restype = C.c_int
    L.krd_tile_list.argtypes = [i32, vp, vp]
    L.krd_tile_list_late.restype = C.c_int
    L.krd_tile_list_late.argtypes = [i32, i32, vp, vp, vp]
    L.krd_rank_tiles.restype = None
    L.krd_rank_tiles.argtypes = [i64, i32, i32, C.POINTER(i64), C.POINTER(i64)]
    L.krd_invalidate_F.restype = C.c_int
    L.krd_invalidate_F.argtypes = [vp]
    L.krd_debug_copy_F.restype = C.c_int
    L.krd_debug_copy_F.argtypes = [vp, vp, vp]
    L.krd_host_alloc.restype = vp
    L.krd_host_alloc.argtypes = [i64]
    L.krd_host_free.restype = None
    L.krd_host_free.argtypes = [vp]
    L.krd_measure_fp64_peak.restype = dbl
    L.krd_measure_fp64_peak.argtypes = [C.c_int, C.c_int, dbl]
    _lib = L
    return L


def _ptr(a):
    """Raw address of a numpy array / torch tensor / int."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


def _itemsize(a):
    """Bytes per element of a numpy array / torch tensor."""
    return a.dtype.itemsize if isinstance(a, np.ndarray) else a.element_size()


class KrdPlan:
    """One handle of the C ABI: tree + masses resident in HBM, distance matrices on demand."""

    def __init__(self, device=0, bins=0, split=0, builder=0, p2_mode=0, pipeline=0, max_f_mb=0):
        self._L = lib()
        self._h = C.c_void_p()
        cfg = KrdConfig(device=device, bins=bins, split=split, builder=builder, p2_mode=p2_mode, pipeline=pipeline, max_f_mb=max_f_mb)
        st = self._L.krd_create(C.byref(cfg), C.byref(self._h))
        if st != KRD_OK:
            raise KrdError(st, self._L.krd_last_error(None).decode())
        self.device = device
        self.bins = bins
        self.n_samples = 0
        self._keep = []

    def _check(self, st):
        if st != KRD_OK:
            raise KrdError(st, self._L.krd_last_error(self._h).decode())

    def close(self):
        if self._h:
            self._L.krd_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream):
        self._check(self._L.krd_set_stream(self._h, cuda_stream))

    def set_tree(self, parent, child, branch_length, root):
        parent = np.ascontiguousarray(parent, np.int32)
        child = np.ascontiguousarray(child, np.int32)
        bl = np.ascontiguousarray(branch_length, np.float64)
        self._check(self._L.krd_set_tree(self._h, len(bl), _ptr(parent), _ptr(child), _ptr(bl), int(root)))

    def set_masses(self, off, edge, pos, mass):
        """HOST arrays (numpy, or pinned CPU torch tensors of the right dtype)."""
        if isinstance(off, np.ndarray) or isinstance(off, (list, tuple)):
            off = np.ascontiguousarray(off, np.int64)
        if isinstance(edge, np.ndarray):
            edge = np.ascontiguousarray(edge, np.int32)
            pos = np.ascontiguousarray(pos, np.float64)
            mass = np.ascontiguousarray(mass, np.float64)
        n = len(off) - 1
        self._keep_host = [off, edge, pos, mass]   # krd_config.pipeline: the arrays are read until the next compute returns
        self._check(self._L.krd_set_masses(self._h, n, _ptr(off), _ptr(edge), _ptr(pos), _ptr(mass)))
        self.n_samples = n

    def set_masses_device(self, off, d_edge, d_pos, d_mass):
        """off: host int64; the rest are CUDA tensors that must outlive the plan's use of them."""
        off = np.ascontiguousarray(off, np.int64)
        n = len(off) - 1
        self._keep = [d_edge, d_pos, d_mass]
        self._check(self._L.krd_set_masses_device(self._h, n, _ptr(off), _ptr(d_edge), _ptr(d_pos), _ptr(d_mass)))
        self.n_samples = n

    def set_masses_binned(self, off, cell, mass):
        """Compact binned form: cell = edge * bins + bin, sorted within a sample; HOST arrays.  16-bit cells (numpy uint16 or a
        2-byte torch tensor) go through krd_set_masses_binned16 (10 B per entry), anything else through krd_set_masses_binned
        (uint32, 12 B per entry)."""
        if isinstance(off, (np.ndarray, list, tuple)):
            off = np.ascontiguousarray(off, np.int64)
        if isinstance(cell, np.ndarray):
            cell = np.ascontiguousarray(cell, np.uint16 if cell.dtype == np.uint16 else np.uint32)
            mass = np.ascontiguousarray(mass, np.float64)
        n = len(off) - 1
        self._keep_host = [off, cell, mass]
        fn = self._L.krd_set_masses_binned16 if _itemsize(cell) == 2 else self._L.krd_set_masses_binned
        self._check(fn(self._h, n, _ptr(off), _ptr(cell), _ptr(mass)))
        self.n_samples = n

    def set_masses_binned_device(self, off, d_cell, d_mass):
        """DEVICE arrays; a 2-byte cell tensor selects krd_set_masses_binned16_device."""
        off = np.ascontiguousarray(off, np.int64)
        n = len(off) - 1
        self._keep = [d_cell, d_mass]
        fn = self._L.krd_set_masses_binned16_device if _itemsize(d_cell) == 2 else self._L.krd_set_masses_binned_device
        self._check(fn(self._h, n, _ptr(off), _ptr(d_cell), _ptr(d_mass)))
        self.n_samples = n

    def unpack_shards_rows(self, d_gathered, d_out_rows, world, row0, row1):
        self._check(self._L.krd_unpack_shards_rows(self._h, world, _ptr(d_gathered), int(row0), int(row1), _ptr(d_out_rows)))

    def scatter_shard_rows(self, rank, world, d_packed, row_bounds, row_ptrs):
        """krd_scatter_shard_rows: this rank's packed tiles -> the owners' row blocks through peer memory.
        row_ptrs: world device addresses (ints) valid in this process."""
        rb = np.ascontiguousarray(row_bounds, np.int32)
        ptrs = (C.c_void_p * world)(*[C.c_void_p(int(q)) for q in row_ptrs])
        self._check(self._L.krd_scatter_shard_rows(self._h, rank, world, _ptr(d_packed), _ptr(rb), C.cast(ptrs, C.c_void_p)))

    def shard_to_host_async(self, rank, world, d_packed, d_scratch, host_matrix, copy_stream):
        """krd_shard_to_host_async: packed tiles of (rank, world) -> pinned n x n host matrix with the copy engines."""
        self._check(self._L.krd_shard_to_host_async(self._h, rank, world, _ptr(d_packed), _ptr(d_scratch), _ptr(host_matrix), copy_stream))

    def scatter_tile_range_rows(self, t0, t1, d_packed, row_bounds, row_ptrs):
        rb = np.ascontiguousarray(row_bounds, np.int32)
        n = len(rb) - 1
        ptrs = (C.c_void_p * n)(*[C.c_void_p(int(q)) for q in row_ptrs])
        self._check(self._L.krd_scatter_tile_range_rows(self._h, int(t0), int(t1), _ptr(d_packed), n, _ptr(rb), C.cast(ptrs, C.c_void_p)))

    def set_late_blocks(self, n_late_blocks):
        """krd_set_late_blocks: reorder the tile list (tiles touching the last blocks come last); returns the split index."""
        sp = C.c_int64(0)
        self._check(self._L.krd_set_late_blocks(self._h, int(n_late_blocks), C.byref(sp)))
        return int(sp.value)

    def compute_tile_range(self, d_packed, t0, t1, p=1.0, divisor=1.0):
        self._check(self._L.krd_compute_tile_range(self._h, float(p), float(divisor), int(t0), int(t1), _ptr(d_packed)))

    def tile_range_to_host_async(self, t0, t1, d_packed, d_scratch, host_matrix, copy_stream):
        self._check(self._L.krd_tile_range_to_host_async(self._h, int(t0), int(t1), _ptr(d_packed), _ptr(d_scratch), _ptr(host_matrix), copy_stream))

    def compute_allgather_host(self, rank, world, comm, row0, row1, p=1.0, divisor=1.0, out_rows=None):
        """krd_compute_allgather_host: this rank's tile share, NCCL all-gather (comm = ncclComm_t as int), rows [row0, row1) on the host."""
        if out_rows is None:
            out_rows = np.zeros((row1 - row0, self.n_samples), np.float64)
        self._check(self._L.krd_compute_allgather_host(self._h, float(p), float(divisor), rank, world, C.c_void_p(comm), row0, row1, _ptr(out_rows)))
        return out_rows

    def compute(self, p=1.0, divisor=1.0, out=None):
        """Host result (numpy n x n)."""
        n = self.n_samples
        if out is None:
            out = np.zeros((n, n), np.float64)
        self._check(self._L.krd_compute(self._h, float(p), float(divisor), _ptr(out)))
        return out

    def compute_device(self, d_out, p=1.0, divisor=1.0):
        self._check(self._L.krd_compute_device(self._h, float(p), float(divisor), _ptr(d_out)))
        return d_out

    def shard_doubles(self, world):
        return int(self._L.krd_shard_doubles(self._h, world))

    def compute_shard(self, d_packed, rank, world, p=1.0, divisor=1.0):
        self._check(self._L.krd_compute_shard(self._h, float(p), float(divisor), rank, world, _ptr(d_packed)))

    def unpack_shards(self, d_gathered, d_out, world):
        self._check(self._L.krd_unpack_shards(self._h, world, _ptr(d_gathered), _ptr(d_out)))

    def compute_shard_host(self, rank, world, p=1.0, divisor=1.0):
        """This rank's packed tiles in a HOST array (krd_compute_shard_host)."""
        out = np.zeros(self.shard_doubles(world), np.float64)
        self._check(self._L.krd_compute_shard_host(self._h, float(p), float(divisor), rank, world, _ptr(out)))
        return out

    def squash(self, p=1.0):
        """genesis SquashClustering::run on the resident samples: returns the n-1 mergers as a structured array
        (index_a, index_b, distance_a, distance_b), squash_clustering.hpp:98-119."""
        out = np.zeros(max(self.n_samples - 1, 0), MERGER_DTYPE)
        self._check(self._L.krd_squash(self._h, float(p), _ptr(out)))
        return out

    def kmeans(self, k, init_centroids, max_iterations=100, init_assignments=None):
        """genesis MassTreeKmeans (tree/mass_tree/kmeans.cpp:59-208) via Kmeans::run (utils/math/kmeans.hpp:118-168) on
        the resident binned samples, started from the given samples as centroids.  Returns (assignments[n] int32,
        distances[n] of every sample to its centroid, cluster variances[k], iterations); the variances are
        KmeansClusteringInfo::variances (kmeans.hpp:384-424)."""
        init = np.ascontiguousarray(init_centroids, np.int32)
        if init.shape != (k,):
            raise ValueError("init_centroids must hold k sample indices")
        a = np.zeros(self.n_samples, np.int32); d = np.zeros(self.n_samples, np.float64)
        it = C.c_int32(0)
        ia = None if init_assignments is None else np.ascontiguousarray(init_assignments, np.int32)
        self._check(self._L.krd_kmeans_ex(self._h, int(k), int(max_iterations), _ptr(init), _ptr(ia), _ptr(a), _ptr(d), C.byref(it)))
        cnt = np.bincount(a, minlength=k)
        var = np.zeros(k, np.float64)
        np.add.at(var, a, d * d)     # sample order, like the reference's serial accumulation
        var = np.where(cnt > 0, var / np.maximum(cnt, 1), var)
        return a, d, var, int(it.value)

    def invalidate_F(self):
        self._check(self._L.krd_invalidate_F(self._h))

    def stats(self):
        s = KrdStats()
        self._check(self._L.krd_get_stats(self._h, C.byref(s)))
        return s.as_dict()

    def debug_F(self):
        s = self.stats()
        F = np.zeros((s["n_segments"], s["ld_samples"]), np.float64)
        ln = np.zeros(s["n_segments"], np.float64)
        self._check(self._L.krd_debug_copy_F(self._h, _ptr(F), _ptr(ln)))
        return F, ln


def unpack_shards_host(n_samples, packed_per_rank):
    """Host-only: the ranks' packed tile buffers (list, rank order) -> symmetric n x n matrix (krd_unpack_shards_host)."""
    allp = np.ascontiguousarray(np.concatenate(packed_per_rank), np.float64)
    out = np.zeros((n_samples, n_samples), np.float64)
    st = lib().krd_unpack_shards_host(n_samples, len(packed_per_rank), _ptr(allp), _ptr(out))
    if st != 0:
        raise KrdError(st, "krd_unpack_shards_host failed")
    return out


def binned_cells(tree, masses, bins):
    """Host-side encoding of BINIFIED mass trees (positions = bin centres, e.g. workloads.build output or the result of
    mass_tree_binify_masses) as the compact form of krd_set_masses_binned: cell = edge * bins + bin, with the reference's bin
    index expression (functions.cpp:231-237).  Pure re-encoding; no arithmetic on masses."""
    bl = np.asarray(tree["bl"], np.float64)[masses["edge"]]
    nb = float(bins)
    with np.errstate(divide="ignore", invalid="ignore"):
        v = masses["pos"] / bl * nb
    pn = np.where(0.0 < v, v, 0.0)
    pn = np.where(np.nextafter(nb, 0.0) < pn, np.nextafter(nb, 0.0), pn)
    return (masses["edge"].astype(np.int64) * bins + np.floor(pn).astype(np.int64)).astype(np.uint32)


def earth_movers_distance(tree, masses, p=1.0, bins=0, device=0, divisor=1.0):
    """Mirror of genesis `earth_movers_distance(std::vector<MassTree> const&, double p)`
    (tree/mass_tree/emd.cpp:183-235) on the flat data model:
    tree = dict(parent, child, bl, root); masses = dict(off, edge, pos, mass).
    Raises KrdError with the reference's message for p <= 0 (emd.cpp:186-190)."""
    plan = KrdPlan(device=device, bins=bins)
    try:
        plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
        plan.set_masses(masses["off"], masses["edge"], masses["pos"], masses["mass"])
        return plan.compute(p=p, divisor=divisor)
    finally:
        plan.close()


def copy_rect_to_host_async(host_dst_ptr, host_pitch, d_src_ptr, src_pitch, cols, rows, stream):
    """krd_copy_rect_to_host_async (addresses as ints, pitches / cols in doubles)."""
    st = lib().krd_copy_rect_to_host_async(C.c_void_p(int(host_dst_ptr)), int(host_pitch), C.c_void_p(int(d_src_ptr)), int(src_pitch), int(cols), int(rows), stream)
    if st != KRD_OK:
        raise KrdError(st, lib().krd_last_error(None).decode())


def p2p_alloc(device, nbytes):
    """krd_p2p_alloc: (device address, 64-byte CUDA IPC handle as bytes)."""
    ptr = C.c_void_p()
    hd = (C.c_ubyte * 64)()
    st = lib().krd_p2p_alloc(device, int(nbytes), C.byref(ptr), C.cast(hd, C.c_void_p))
    if st != KRD_OK:
        raise KrdError(st, lib().krd_last_error(None).decode())
    return int(ptr.value), bytes(hd)


def p2p_open(device, handle):
    ptr = C.c_void_p()
    buf = (C.c_ubyte * 64)(*handle)
    st = lib().krd_p2p_open(device, C.cast(buf, C.c_void_p), C.byref(ptr))
    if st != KRD_OK:
        raise KrdError(st, lib().krd_last_error(None).decode())
    return int(ptr.value)


def p2p_close(ptr):
    lib().krd_p2p_close(C.c_void_p(ptr))


def p2p_free(ptr):
    lib().krd_p2p_free(C.c_void_p(ptr))


def p2p_enable_peer(device, peer):
    st = lib().krd_p2p_enable_peer(device, peer)
    if st != KRD_OK:
        raise KrdError(st, lib().krd_last_error(None).decode())


def nccl_init_all(devices):
    """ncclCommInitAll for devices of this process (krd_nccl_init_all): list of communicators (ints)."""
    devs = np.ascontiguousarray(devices, np.int32)
    comms = (C.c_void_p * len(devs))()
    st = lib().krd_nccl_init_all(len(devs), _ptr(devs), C.cast(comms, C.c_void_p))
    if st != KRD_OK:
        raise KrdError(st, lib().krd_last_error(None).decode())
    return [int(c) for c in comms]


def nccl_destroy(comms):
    arr = (C.c_void_p * len(comms))(*comms)
    lib().krd_nccl_destroy(len(comms), C.cast(arr, C.c_void_p))


def matrix_from_F_noalloc(d_F, d_len, n_segments, n_samples, ld, d_out, p=1.0, divisor=1.0, device=0, stream=None):
    """krd_matrix_from_F_noalloc: allocation-free pair reduction on DEVICE tensors (n_samples <= 128)."""
    st = lib().krd_matrix_from_F_noalloc(device, _ptr(d_F), _ptr(d_len), int(n_segments), int(n_samples), int(ld), float(p),
                                         float(divisor), _ptr(d_out), stream)
    if st != KRD_OK:
        raise KrdError(st, lib().krd_last_error(None).decode())
    return d_out


def measure_fp64_peak(device=0, mode=0, ms=200.0):
    return float(lib().krd_measure_fp64_peak(device, mode, ms))


TILE = 128


def tile_list(n_samples):
    """Upper-triangular 128x128 tile coordinates in the library's order (diagonal tiles interleaved, see krd_b200.h)."""
    n = int(lib().krd_tile_count(n_samples))
    ti = np.zeros(n, np.int32); tj = np.zeros(n, np.int32)
    lib().krd_tile_list(n_samples, _ptr(ti), _ptr(tj))
    return ti, tj


def tile_list_late(n_samples, n_late_blocks):
    """The two-part order of krd_set_late_blocks (host-only): (tile_i, tile_j, t_split)."""
    n = int(lib().krd_tile_count(n_samples))
    ti = np.zeros(max(n, 1), np.int32); tj = np.zeros(max(n, 1), np.int32)
    split = C.c_int64()
    st = lib().krd_tile_list_late(n_samples, n_late_blocks, _ptr(ti), _ptr(tj), C.byref(split))
    if st != 0:
        raise KrdError(st, "krd_tile_list_late failed")
    return ti[:n], tj[:n], split.value


def rank_tiles(n_tiles, rank, world):
    t0, t1 = C.c_int64(), C.c_int64()
    lib().krd_rank_tiles(n_tiles, rank, world, C.byref(t0), C.byref(t1))
    return t0.value, t1.value
