/* krd_b200.h — C ABI of the B200-native Kantorovich-Rubinstein (KR / earth mover's) distance path.
 *
 * This is the drop-in boundary for ONE call of the reference:
 *
 *     auto krd_matrix = earth_movers_distance( mass_trees, options.exponent );
 *         gappa   src/commands/analyze/krd.cpp:140
 *         genesis libs/genesis/lib/genesis/tree/mass_tree/emd.hpp:124  (declaration)
 *                 libs/genesis/lib/genesis/tree/mass_tree/emd.cpp:183-235 (all pairs)
 *                 libs/genesis/lib/genesis/tree/mass_tree/emd.cpp:60-181  (one pair)
 *
 * plus the optional steps around it: mass_tree_binify_masses
 * (tree/mass_tree/functions.cpp:222-265, selected with krd_config.bins) and the
 * "--normalize" division (src/commands/analyze/krd.cpp:144-150, the `divisor` argument).
 *
 * Plain C, POD only, no exceptions cross this boundary.  Every function returns a krd_status;
 * krd_last_error() gives the message (same wording as the reference's exceptions where one exists).
 *
 * Flat data model
 *   tree    n_edges edges, nodes 0..n_edges.  Edge e joins parent_node[e] (root side; genesis
 *           "primary_node") and child_node[e] (leaf side; "secondary_node").
 *   masses  CSR over samples (= one genesis MassTree each): sample_off[n_samples+1]; entry m has
 *           edge[m], pos[m] = distance from the edge's parent node ("proximal_length",
 *           tree/mass_tree/tree.hpp:186-199) and mass[m].  Any order is accepted; the iteration
 *           order of a MassTree (edge ascending, pos ascending) is the fast path.
 *   result  n_samples x n_samples row-major doubles, symmetric, zero diagonal
 *           (utils/containers/matrix.hpp:85-94,214).
 */
#ifndef KRD_B200_H
#define KRD_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum krd_status {
    KRD_OK = 0,
    KRD_ERR_INVALID_EXPONENT = 1, /* emd.cpp:186-190  p <= 0 */
    KRD_ERR_INCOMPATIBLE_TREES = 2, /* emd.cpp:78,109,123,172 */
    KRD_ERR_TOO_FEW_SAMPLES = 3,  /* krd.cpp:131-133 (raised by the host command, not here) */
    KRD_ERR_INVALID_ARGUMENT = 4,
    KRD_ERR_CUDA = 5,
    KRD_ERR_STATE = 6,            /* call order violated (e.g. masses before tree) */
    KRD_ERR_NO_DEVICE = 7         /* no CUDA device: this library has NO CPU fallback */
} krd_status;

typedef struct krd_handle krd_handle;

typedef struct krd_config {
    int device;        /* CUDA device ordinal */
    int bins;          /* 0 = exact positions (what `gappa analyze krd` does);
                          B>0 = mass_tree_binify_masses(tree, B) first (functions.cpp:222) */
    int split;         /* 0 = auto (one persistent CTA per SM); >0 forces the CTA count of the pair kernel */
    int builder;       /* F builder: 0 = auto (tiled single-write builder in binned mode), 1 = force the
                          general scatter + level-scan builder (the only one in exact mode) */
    int p2_mode;       /* exponent 2: 0 = auto (= 2), 1 = direct (a-b)^2 on the FP64 CUDA cores,
                          2 = Gram contraction |a|^2+|b|^2-2ab on the FP64 tensor cores (DMMA m8n8k4) with
                          direct recomputation of the cells where the subtraction cancels */
    int pipeline;      /* 1 = pipelined ingest for krd_set_masses + krd_compute[_device] from HOST arrays (binned mode,
                          1 <= bins <= 31, >= 512 samples; ignored otherwise): the samples are uploaded in up to eight
                          sample-range blocks and block q's F columns and the pair tiles that need only blocks <= q are
                          computed while block q+1 is still crossing PCIe.  krd_set_masses returns once the first block
                          is resident; the rest of the ingest completes inside the first krd_compute*, so (a) the input
                          arrays must stay valid until that call returns, (b) input errors of the later blocks are
                          reported by it (their checks are read back once, after the last phase; a later block whose
                          entries are not in MassTree order makes the call start over on the ordinary, sorting path).  Results agree with the ordinary path to rounding (different partial-sum
                          splits).  Sharded compute is not available on such a handle; krd_squash / krd_kmeans /
                          introspection first turn it into an ordinary one. */
    int max_f_mb;      /* exact mode (bins = 0): cap, in MiB, on the resident dense matrix F (0 = auto: 60 % of the free device
                          memory).  F has one row per segment, S = n_edges + #distinct (edge, position) over all samples, so
                          it outgrows HBM quickly (1,000 samples x 30k positions: 245 GB).  Beyond the cap the segment axis is
                          STREAMED: only the n_edges x n subtree-mass matrices stay resident, F is rebuilt slab after slab
                          inside every krd_compute* and the pair kernel accumulates over the slabs.  krd_squash / krd_kmeans /
                          krd_debug_copy_F need the resident form and fail with a sized message otherwise. */
    int reserved[1];   /* must be 0 */
} krd_config;

typedef struct krd_stats {
    int64_t n_samples, n_edges, n_masses;
    int64_t n_segments;     /* S: union segments (exact) or n_edges*(bins+1) (binned) */
    int64_t ld_samples;     /* padded sample stride of the dense matrix F[S][ld] */
    int64_t f_bytes;        /* bytes of F resident in HBM */
    int64_t n_tiles;        /* 128x128 upper-triangular tiles (incl. diagonal) */
    int32_t split;          /* work items per tile (rounded up) in the last pair launch */
    int32_t pair_ctas;      /* CTAs of the last pair launch */
    int32_t levels;         /* tree levels = launches of the F builder */
    int32_t launches;       /* kernels launched by the last krd_compute* call */
    float   ms_build;       /* device time of the F builder in the last compute (CUDA events) */
    float   ms_pairs;       /* device time of the pair kernel alone in the last compute */
    float   ms_epilogue;    /* device time of the split-sum / root / mirror epilogue */
    float   ms_slabs;       /* streamed exact mode: device time of the slab builds inside the last compute (ms_pairs excludes it) */
    int32_t gram_fixups;    /* exponent 2, Gram path: pairs recomputed directly because |a|^2+|b|^2-2ab cancelled
                               (-1: more than the list holds; the launch was repeated on the direct path by the library) */
    int32_t reserved_i;
} krd_stats;

krd_status  krd_create(const krd_config* cfg, krd_handle** out);
void        krd_destroy(krd_handle* h);
const char* krd_last_error(const krd_handle* h); /* h may be NULL: error of a failed krd_create */

/* Run all later work of this handle on the given cudaStream_t (NULL = legacy default stream). */
krd_status krd_set_stream(krd_handle* h, void* cuda_stream);

/* Topology + branch lengths shared by all samples (after
 * mass_trees_make_average_branch_lengths, functions.cpp:271-335).  HOST pointers. */
krd_status krd_set_tree(krd_handle* h, int32_t n_edges, const int32_t* parent_node,
                        const int32_t* child_node, const double* branch_length, int32_t root_node);

/* The N mass trees.  HOST pointers (pinned memory makes the copy faster; see krd_host_alloc).
 * Copies to HBM, assigns every mass to its segment slot and builds the segment table.
 * Asynchronous w.r.t. the host when the buffers are pinned. */
krd_status krd_set_masses(krd_handle* h, int32_t n_samples, const int64_t* sample_off,
                          const int32_t* edge, const double* pos, const double* mass);

/* Same, but the four arrays are already DEVICE pointers on cfg.device (no copy is made of
 * edge/pos/mass: they must stay valid until the next krd_set_masses*).  sample_off is host. */
krd_status krd_set_masses_device(krd_handle* h, int32_t n_samples, const int64_t* sample_off,
                                 const int32_t* d_edge, const double* d_pos, const double* d_mass);

/* The same N mass trees in COMPACT BINNED form (krd_config.bins = B >= 1 only): what mass_tree_binify_masses
 * (tree/mass_tree/functions.cpp:222-265) leaves is one mass per occupied (edge, bin), sitting at the bin centre, so an
 * entry is fully described by  cell = edge * B + bin  (bin = floor(min(max(0, pos / bl * B), nextafter(B, 0))),
 * functions.cpp:231-237) and its mass: 12 bytes instead of 20 across PCIe.  The command `gappa-b200 analyze krd | phylogenetic-kmeans
 * --bins B` bins on the host and ships this form (host/krd_command.cpp, set_masses).  Within a sample the entries must be sorted by cell (= MassTree order); equal cells are
 * summed.  n_edges * B must be < 2^32.  HOST pointers; honours krd_config.pipeline like krd_set_masses. */
krd_status krd_set_masses_binned(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const uint32_t* cell,
                                 const double* mass);
/* Same with DEVICE arrays (no copy is made: they must stay valid until the next krd_set_masses*). sample_off is host. */
krd_status krd_set_masses_binned_device(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const uint32_t* d_cell,
                                        const double* d_mass);
/* The compact form with 16-bit cells, for trees with n_edges * B <= 65536 (every BASELINE config: 9,999 x 4, 49,999 x 1,
 * 19,999 x 1): 10 bytes per entry across PCIe — half of what the MassTree form (edge, pos, mass) ships.  Same rules, same
 * results (the cells are widened on the device by the pass that decodes them).  Replaces nothing in the reference: it is the
 * transport form of the MassTrees that earth_movers_distance (emd.hpp:124) receives after mass_tree_binify_masses. */
krd_status krd_set_masses_binned16(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const uint16_t* cell,
                                   const double* mass);
krd_status krd_set_masses_binned16_device(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const uint16_t* d_cell,
                                          const double* d_mass);

/* earth_movers_distance(trees, p) [/ divisor].  Builds F (level-synchronous subtree scan), runs
 * the all-pairs kernel, writes the n x n matrix.  `out` is a HOST pointer; returns after the
 * matrix is complete in host memory. divisor = 1.0 for no normalisation.  p <= 0 (or NaN) fails with
 * KRD_ERR_INVALID_EXPONENT and genesis' message (emd.cpp:186-190); every p > 0 is accepted (p > 30000 runs a slower
 * library-pow operator, the table-driven one needs 32 p log2 x inside 31 bits). */
krd_status krd_compute(krd_handle* h, double p, double divisor, double* out);

/* Same with a DEVICE output pointer (n*n doubles); asynchronous on the handle's stream (except for p = 2 on the Gram
 * path, which reads its 4-byte fix-up count back before returning). */
krd_status krd_compute_device(krd_handle* h, double p, double divisor, double* d_out);

/* Multi-GPU sharding (one process per GPU, F replicated): rank r of `world` computes its share
 * of the upper-triangular tiles into a packed DEVICE buffer of krd_shard_doubles() doubles
 * (equal on every rank, so it can feed an NCCL all-gather directly). */
int64_t    krd_shard_doubles(const krd_handle* h, int32_t world);
krd_status krd_compute_shard(krd_handle* h, double p, double divisor, int32_t rank, int32_t world,
                             double* d_packed);
/* Scatter the all-gathered packed tiles (world * krd_shard_doubles doubles, DEVICE) into the
 * full symmetric n x n DEVICE matrix. */
krd_status krd_unpack_shards(krd_handle* h, int32_t world, const double* d_gathered, double* d_out);
/* Same for the rows [row0, row1) of the matrix only: d_out_rows is (row1 - row0) x n, DEVICE.  With one process per GPU
 * every rank unpacks just the row block it then copies to the host over ITS PCIe link (bench.py, N > 1). */
krd_status krd_unpack_shards_rows(krd_handle* h, int32_t world, const double* d_gathered, int32_t row0, int32_t row1,
                                  double* d_out_rows);

/* The three steps above in ONE call for a caller that owns an NCCL communicator (ncclComm_t passed as void*; one process per
 * GPU, or one thread per GPU with communicators from krd_nccl_init_all): this rank's tile share, ncclAllGather of the packed
 * tiles on the handle's stream, unpack of the rows [row0, row1) of the matrix into d_out ((row1 - row0) x n doubles, DEVICE;
 * row0 = 0, row1 = n for the whole matrix).  Every rank of the communicator must make the call.  NCCL is loaded at run
 * time (libnccl.so.2, or $KRD_NCCL_LIB); without it the call fails with KRD_ERR_STATE and nothing else is affected.
 * This is the north-star's "NCCL all-gather of result tiles over NVLink"  behind the C ABI. */
krd_status krd_compute_allgather(krd_handle* h, double p, double divisor, int32_t rank, int32_t world, void* nccl_comm,
                                 int32_t row0, int32_t row1, double* d_out);
/* Same with a HOST output: the rows [row0, row1) of the matrix ((row1 - row0) x n doubles) are copied to out_rows over this
 * GPU's own PCIe link; returns when they are complete in host memory. */
krd_status krd_compute_allgather_host(krd_handle* h, double p, double divisor, int32_t rank, int32_t world, void* nccl_comm,
                                      int32_t row0, int32_t row1, double* out_rows);
/* Communicators for n devices of THIS process (ncclCommInitAll), and their release. */
krd_status krd_nccl_init_all(int32_t n, const int32_t* devices, void** comms);
void       krd_nccl_destroy(int32_t n, void** comms);

/* Exchange over PEER MEMORY instead of a collective (NVLink / NVSwitch stores; csrc/krd_p2p.cu): rank `rank` scatters the
 * packed tiles it computed (krd_compute_shard's output) straight into the row blocks of the ranks that OWN those rows —
 * the tile into the block of its rows, its mirror into the block of its columns.  row_bounds (HOST, world + 1 entries,
 * 0 ... n_samples) says which rows each rank owns; d_rows (HOST array of world DEVICE pointers valid in THIS process: the
 * own block and peer mappings) are the blocks, (row_bounds[r+1] - row_bounds[r]) x n_samples doubles each.  Every cell
 * crosses NVLink once, to the one GPU that needs it (the all-gather moves every tile to every rank).  Asynchronous on the
 * handle's stream; a rank's block is complete once EVERY rank's call has finished — order that with any stream-ordered
 * barrier (a one-element all-reduce) before reading it, and do not rewrite a block a peer may still be reading (bench.py
 * alternates two sets of blocks).  world <= 16. */
krd_status krd_scatter_shard_rows(krd_handle* h, int32_t rank, int32_t world, const double* d_packed,
                                  const int32_t* row_bounds, double* const* d_rows);
/* This rank's packed tiles (krd_compute_shard's output) straight into the caller's n x n HOST matrix (pinned / registered
 * memory) with the COPY ENGINES: the tiles are transposed on the handle's stream into d_scratch (DEVICE, as large as
 * d_packed), then one 2-D copy per tile and one per mirror are queued on copy_stream (a cudaStream_t).  Every cell of the
 * matrix is written by exactly one rank, so no collective and no barrier is needed, and because the copies do not occupy
 * SMs they run UNDER the next pair launch: split a rank's share into sub-shares (rank * k + q of world * k, q = 0..k-1:
 * contiguous parts of the same tile list) and only the last sub-share's copies are exposed (bench.py --exchange host_tiles).
 * d_packed and d_scratch must stay untouched until copy_stream has passed the copies. */
krd_status krd_shard_to_host_async(krd_handle* h, int32_t rank, int32_t world, const double* d_packed, double* d_scratch,
                                   double* host_matrix, void* copy_stream);
/* The same two steps for an arbitrary range [t0, t1) of the tile list (krd_tile_list order; krd_rank_tiles gives a rank's
 * range): d_packed receives / holds (t1 - t0) tiles.  Lets a caller cut its share UNEVENLY — a long first launch whose
 * copies are hidden under a short last one — which equal sub-shares cannot (every extra pair launch costs ~1 ms of stream-K
 * balance at the cfg4 shape). */
krd_status krd_compute_tile_range(krd_handle* h, double p, double divisor, int64_t t0, int64_t t1, double* d_packed);
krd_status krd_tile_range_to_host_async(krd_handle* h, int64_t t0, int64_t t1, const double* d_packed, double* d_scratch,
                                        double* host_matrix, void* copy_stream);

/* The same scatter for an arbitrary range [t0, t1) of the handle's tile list; d_packed holds (t1 - t0) tiles; n_owners row
 * blocks (<= 16). */
krd_status krd_scatter_tile_range_rows(krd_handle* h, int64_t t0, int64_t t1, const double* d_packed, int32_t n_owners,
                                       const int32_t* row_bounds, double* const* d_rows);
/* Tile ORDER for overlapping the device-to-host copy with compute.  n_late_blocks = L > 0 reorders the handle's tile list:
 * first every tile whose sample blocks both lie before the last L blocks (in the default order of that smaller matrix),
 * then the tiles that touch the last L blocks; *t_split = the index where the second part starts.  Once the first part is
 * computed (and, across GPUs, scattered and a barrier passed) all cells (i, j) with i, j < n_samples' = (T - L) * 128 are
 * final, so that rectangle can be copied to the host (krd_copy_rect_to_host_async, copy engine) UNDER the pair launch of the
 * short second part; afterwards only the thin strip of late columns / rows is left to copy.  The order is a property of the
 * handle (it survives krd_set_masses* of the same size); L = 0 restores the default.  The host-only helpers krd_tile_list /
 * krd_unpack_shards_host describe the DEFAULT order only: use the handle-based calls with a reordered handle. */
krd_status krd_set_late_blocks(krd_handle* h, int32_t n_late_blocks, int64_t* t_split);
/* cudaMemcpy2DAsync device -> (pinned) host of a cols x rows rectangle of doubles, pitches in doubles; no handle needed. */
krd_status krd_copy_rect_to_host_async(double* host_dst, int64_t host_pitch_doubles, const double* d_src,
                                       int64_t src_pitch_doubles, int64_t cols, int64_t rows, void* cuda_stream);

/* Row blocks that other PROCESSES can map: cudaMalloc + CUDA IPC handle (64 bytes, to be sent to the peers by any means),
 * mapping of a peer's handle on `device`, unmapping, release.  One process driving several GPUs can use ordinary device
 * pointers after krd_p2p_enable_peer(device, peer). */
krd_status krd_p2p_alloc(int32_t device, int64_t bytes, void** d_ptr, unsigned char* ipc_handle_64);
krd_status krd_p2p_open(int32_t device, const unsigned char* ipc_handle_64, void** d_ptr);
void       krd_p2p_close(void* d_ptr);
void       krd_p2p_free(void* d_ptr);
krd_status krd_p2p_enable_peer(int32_t device, int32_t peer);

/* The same for a caller without an NCCL communicator (e.g. one host thread per GPU in one process, as
 * `gappa-b200 --gpus G` does): the rank's packed tiles go to a HOST buffer of krd_shard_doubles() doubles
 * (D2H over that GPU's own PCIe link), and the concatenation of all ranks' buffers, in rank order, is
 * scattered into the symmetric n x n HOST matrix on the host (no GPU, no handle needed). */
krd_status krd_compute_shard_host(krd_handle* h, double p, double divisor, int32_t rank, int32_t world,
                                  double* packed);
krd_status krd_unpack_shards_host(int32_t n_samples, int32_t world, const double* packed_all, double* out);

/* Host-only helpers (no GPU needed) describing the sharding, so that callers and tests can reason
 * about it: the upper triangle of 128x128 tiles (off-diagonal tiles in diagonal-major order (i,i+1), (i,i+2), ...
 * with the cheaper diagonal tiles (i,i) spread evenly through the list, so that equal-count shares are
 * equal-cost shares), and the contiguous share [t0,t1) of that list owned by `rank`. */
int64_t    krd_tile_count(int32_t n_samples);
krd_status krd_tile_list(int32_t n_samples, int32_t* tile_i, int32_t* tile_j);
void       krd_rank_tiles(int64_t n_tiles, int32_t rank, int32_t world, int64_t* t0, int64_t* t1);
/* The tile list a handle uses after krd_set_late_blocks(h, n_late_blocks, &t_split) (n_late_blocks is clamped to the number
 * of sample blocks - 1; 0 gives krd_tile_list's order and t_split = the tile count). */
krd_status krd_tile_list_late(int32_t n_samples, int32_t n_late_blocks, int32_t* tile_i, int32_t* tile_j, int64_t* t_split);

/* Squash clustering on the resident dense matrix:
 *     genesis::tree::SquashClustering::run()   libs/genesis/lib/genesis/tree/mass_tree/squash_clustering.cpp:57-249
 * as called by `gappa analyze squash` (src/commands/analyze/squash.cpp:94-130).  The n samples are the initial
 * clusters 0..n-1; merge m (m = 0..n-2) joins the closest pair of active clusters into cluster n+m, whose mass
 * tree is the count-weighted, re-normalised average (mass_tree_merge_trees + mass_tree_normalize_masses,
 * tree/mass_tree/functions.cpp:60-140) — in the dense representation the same weighted average of two columns
 * of F — and whose distances to all active clusters are computed by a one-vs-all kernel.  The whole loop runs
 * on the device without host round trips.  `out` (HOST, n-1 entries) = SquashClustering::mergers()
 * (squash_clustering.hpp:98-119).  Needs n >= 2. */
typedef struct krd_merger {
    int32_t index_a, index_b;     /* a < b: the two clusters that were merged */
    double  distance_a, distance_b; /* their KR distances to the new cluster node */
} krd_merger;
krd_status krd_squash(krd_handle* h, double p, krd_merger* out);

/* Phylogenetic k-means (`gappa analyze phylogenetic-kmeans`, src/commands/analyze/kmeans_phylogenetic.cpp:134-175)
 * = genesis::tree::MassTreeKmeans (tree/mass_tree/kmeans.cpp:59-208) on the Lloyd loop of
 * utils/math/kmeans.hpp:118-168, run on the resident dense matrix.  The assignment step is the p = 1 KR reduction of
 * every sample against the k centroids, first minimum wins; empty centroids are treated as in kmeans.hpp:426-500.
 * krd_config.bins >= 1 (gappa --bins B: accumulate_centroid_masses = B): centroids are extra columns of F, the
 * re-normalised sum of their members' columns.  bins == 0 (gappa's default: accumulate_centroid_masses stays 1):
 * centroids carry one mass per branch at the members' mass centre (functions.cpp:181-220), kept as (mass below, mass,
 * centre) per edge.
 * `init_centroids` (HOST, k sample indices) are the initial centroids, i.e. the state after the reference's
 * initialisation (kmeans.hpp:254-306: random samples or k-means++, both pick data points); all assignments
 * start at 0.  Outputs (HOST): assignments[n], distances[n] = KR distance of every sample to its centroid
 * (KmeansClusteringInfo::distances, kmeans.hpp:384-424), *iterations = Kmeans::run()'s return value.
 * 1 <= k <= n (any k: the assignment step runs 32 centroids per pass over F; any n); max_iterations = the reference's
 * default is 100 (kmeans.hpp:699). */
krd_status krd_kmeans(krd_handle* h, int32_t k, int32_t max_iterations, const int32_t* init_centroids,
                      int32_t* assignments, double* distances, int32_t* iterations);
/* Same with an initial assignment vector (HOST, n cluster indices < k, or NULL = all 0): `gappa analyze phylogenetic-kmeans`
 * runs several k on ONE Kmeans object, whose assignments survive from the previous k (kmeans.hpp:296-306) — which decides
 * whether the first Lloyd step counts as "changed" and hence the iteration count. */
krd_status krd_kmeans_ex(krd_handle* h, int32_t k, int32_t max_iterations, const int32_t* init_centroids,
                         const int32_t* init_assignments, int32_t* assignments, double* distances, int32_t* iterations);

/* Allocation-free pair reduction.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Computes the matrix krd_compute_device computes — KR_p(i,j) = (sum_s len[s] |F[s][i] - F[s][j]|^p)^(min(1, 1/p)) / divisor,
 * symmetric, zero diagonal — from a dense cumulative-mass matrix the CALLER holds on `device`: d_F is n_segments x ld doubles
 * (segment-major, ld >= n_samples; the layout krd_debug_copy_F returns), d_len the n_segments segment lengths, d_out
 * n_samples x n_samples doubles; all DEVICE pointers; asynchronous on cuda_stream (a cudaStream_t, NULL = default stream).
 * It calls no cudaMalloc / cudaFree / allocating routine and needs no handle: its only scratch (per-chunk partial sums) is
 * a file-scope static __device__ array of KRD_NOALLOC_SCRATCH_BYTES that every GPU loading this library holds until the
 * process exits.  Limit: n_samples <= KRD_NOALLOC_MAX_SAMPLES (more fails with KRD_ERR_INVALID_ARGUMENT and a message saying
 * so; krd_last_error(NULL)).  All calls share the array and it keeps the last call's contents (every cell a call reads it
 * has written first, so nothing needs clearing) — calls that OVERLAP on the GPU (different streams / host threads) race on it;
 * calls on one stream are ordered.  p <= 0 fails like krd_compute. */
#define KRD_NOALLOC_MAX_SAMPLES 128
#define KRD_NOALLOC_SCRATCH_BYTES (64 * 128 * 128 * 8)   /* 8 MiB */
krd_status krd_matrix_from_F_noalloc(int32_t device, const double* d_F, const double* d_len, int64_t n_segments,
                                     int32_t n_samples, int64_t ld, double p, double divisor, double* d_out,
                                     void* cuda_stream);

/* Drop the dense matrix F so that the next krd_compute* rebuilds it from the resident masses
 * (bench: one "step" = F build + pair kernel + epilogue). */
krd_status krd_invalidate_F(krd_handle* h);

/* Introspection for parity tests: copy the dense cumulative-mass matrix F (n_segments x
 * ld_samples doubles) and the segment lengths (n_segments doubles) to HOST buffers. */
krd_status krd_get_stats(const krd_handle* h, krd_stats* out);
krd_status krd_debug_copy_F(krd_handle* h, double* out_F, double* out_len);

/* Pinned host memory helpers for callers that want full-speed copies. */
void* krd_host_alloc(int64_t bytes);
void  krd_host_free(void* p);

/* Roofline denominator: measured FP64 FMA throughput (TFLOP/s, FMA = 2 flop) of `device`, from a
 * register-resident DFMA loop run for about `ms` milliseconds.  mode 0 = DFMA, 1 = DADD+DFMA mix
 * (the p = 1 instruction mix), 2 = DMMA m8n8k4 tensor-core path. */
double krd_measure_fp64_peak(int device, int mode, double ms);

#ifdef __cplusplus
}
#endif
#endif /* KRD_B200_H */
