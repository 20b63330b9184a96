// Shared internals of libkrd_b200.so (sm_100a only).  See include/krd_b200.h for the C ABI and
// DESIGN.md for the data layout.  Reference semantics: genesis tree/mass_tree/emd.cpp:60-235.
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/krd_b200.h"

namespace krd {

constexpr int TILE = 128;  // samples per tile edge (pair kernel CTA tile is TILE x TILE)
constexpr int KC   = 16;   // segments per pipeline stage
constexpr int NUM_SMS_FALLBACK = 148;

enum PairMode : int {
    MODE_P1 = 0,      // p == 1      : acc += len * |a-b|
    MODE_P2 = 1,      // p == 2      : acc += (sqrt(len)*(a-b))^2   (direct, FP64 cores)
    MODE_PHALF = 2,   // p == 0.5    : acc += len * sqrt(|a-b|)
    MODE_GENERIC = 3, // any p > 0   : acc += len * |a-b|^p
    MODE_INT = 5,     // p = 5..16   : |a-b|^p by repeated multiplication
    MODE_P3 = 6,      // p == 3      : acc += len * |d| * d^2
    MODE_P4 = 7,      // p == 4      : acc += len * (d^2)^2
    MODE_LIBPOW = 8,
    MODE_P2S = 9,     // p == 2      : acc += len * (a-b)^2 without the sqrt(len) pre-scaling (some len < 0: emd.cpp:156,166 keep the sign)  // p > 30000  : acc += len * pow(|a-b|, p) with the library pow (outside kr_pow's exponent range)
    MODE_P2G = 4      // p == 2      : Gram form on the FP64 tensor cores (DMMA): acc += len * (a-c)(b-c),
                      //               W = |a-c|^2 + |b-c|^2 - 2 acc   (c = a reference column, cancels exactly in a-b)
};

struct Tile { int32_t i, j; };  // tile-block coordinates, i <= j

}  // namespace krd

struct krd_handle {
    krd_config cfg{};
    int device = 0;
    int num_sms = krd::NUM_SMS_FALLBACK;
    cudaStream_t stream = nullptr;
    std::string err;

    // ---- tree (host copies + device arrays) ----
    int32_t E = 0, root = 0;
    bool have_tree = false;
    std::vector<int32_t> parent, child;
    std::vector<double> bl;
    std::vector<int32_t> lvl_off;     // host: edges of level L are lvl_edges[lvl_off[L]..lvl_off[L+1])
    int32_t* d_lvl_edges = nullptr;   // edges sorted by level (level = height of the child node)
    int32_t* d_kid_off = nullptr;     // E+1: child edges of edge e's child node
    int32_t* d_kid_edge = nullptr;
    double*  d_bl = nullptr;
    // top of the tree (the last levels, <= TOP_MAX edges together): finished by ONE launch in shared memory
    int32_t  top_level = 0, top_edges = 0, top_kids = 0, top_levels = 0;
    int32_t* d_top = nullptr;         // [edge ids][kid offsets][kids: >=0 edge below the top, <0 ~index in the top][level offsets]

    // ---- masses ----
    int32_t N = 0;
    int64_t M = 0;
    bool have_masses = false;
    bool owns_masses = false;
    int64_t* d_off = nullptr;   // N+1
    int32_t* d_edge = nullptr;  // M
    double*  d_pos = nullptr;   // M
    double*  d_mass = nullptr;  // M
    uint32_t* d_sid = nullptr;  // M: sample of each mass entry
    int32_t* d_slot = nullptr;  // M: segment index within the edge that receives this mass
    void*    d_owned[3] = {nullptr, nullptr, nullptr};  // edge/pos/mass buffers we allocated
    // compact binned input (krd_set_masses_binned): entries arrive as (cell = edge * bins + bin, mass); d_edge / d_slot are
    // decoded from it by K0 and d_pos stays null (the bin centre is implied)
    bool     binned_input = false;
    const void* d_cell32 = nullptr;       // M cells of cell_bytes each (uint32_t, or uint16_t from krd_set_masses_binned16)
    int      cell_bytes = 4;
    int32_t* d_edge_dec = nullptr;        // M: decoded edge ids (owned, grow-only)

    // ---- segments + dense matrix ----
    int64_t S = 0, S_pad = 0;
    int32_t ld = 0;              // column stride of F: N (+ extra_cols) rounded up to TILE
    int32_t extra_cols = 0;      // columns reserved after the samples (squash clustering: one per merged cluster)
    int32_t ncols = 0;           // columns this handle's F builder writes (= ld, except for the children of a pipelined ingest)
    bool     ld_fixed = false;   // child of a pipelined ingest: ld (the parent's), d_F (inside the parent's F) are set by the parent
    int32_t* d_seg_off = nullptr;  // E+1
    double*  d_len = nullptr;      // S_pad (+ KC slack), zero padded
    double*  d_sqrt_len = nullptr; // same, sqrt(len) (direct p=2 path)
    bool     len_negative = false; // exact mode: a mass position outside its branch made a segment length negative
    double*  d_Fg = nullptr;       // S_pad x ld: Gram path operands, sqrt(len) (F - F[.][0]) (k_gram_prescale)
    double*  d_norm = nullptr;     // ld: sum_s len (F[s][i] - F[s][0])^2
    double*  d_norm_part = nullptr;
    bool     norms_valid = false;
    void*    d_fix = nullptr;      // int4[1 + cap]: Gram-path cells to recompute directly
    bool     fix_pending = false;  // a Gram compute was issued; its fix-up count has not been read back yet
    double*  d_F = nullptr;        // S_pad x ld, row = segment, N contiguous
    size_t   F_capacity = 0;
    bool F_valid = false;

    // ---- tiled F builder (binned mode): see krd_build.cu "K1 tiled" ----
    int32_t  build_launches = 0;   // kernels launched by the last build_F
    bool     tiled = false;        // fast path usable for the resident masses
    int32_t  EB = 0, n_eblocks = 0;   // edges per tile, number of edge blocks
    uint16_t* d_cell = nullptr;    // M, tile-major order: (row << 6 | sample-in-tile) of every mass entry
    double*  d_tmass = nullptr;    // M, tile-major order: the masses
    int64_t* d_tile_ptr = nullptr; // n_eblocks x N (+1): first tile-major entry of (edge block b, sample i)
    double*  d_M = nullptr;        // E x ncols: subtree mass at the proximal end of every edge
    void*    d_scratch[3] = {nullptr, nullptr, nullptr};   // K0 scratch, grow-only
    void*    d_flags = nullptr;    // K0: canonical-order / range check flags
    double*  d_D = nullptr;        // E x ncols: mass strictly below every edge (sum of its child edges' M)

    // ---- exact mode (bins = 0): single-write builder and segment-axis streaming, krd_build.cu "K1 exact" ----
    bool     exact_fast = false;   // the resident masses can use the single-write exact builder
    int32_t* d_grow = nullptr;     // M: global segment (row of F) of every entry = seg_off[edge] + slot
    double*  d_cum = nullptr;      // M: value of F at that row for the entry's sample (mass below the edge + own masses at or above)
    bool     cum_valid = false;    // d_M / d_D / d_cum describe the resident masses
    bool     streamed = false;     // F does not fit: d_F is ONE slab of slab_rows rows, rebuilt slab after slab inside every compute
    int64_t  slab_rows = 0;
    double*  d_acc = nullptr;      // streamed: per-tile sums carried from slab to slab
    std::vector<cudaEvent_t> slab_ev;   // (begin, end) of the first slab builds of the last streamed compute (timing only)
    int32_t  slabs_timed = 0, slabs_total = 0;

    // ---- pair kernel workspace ----
    double* d_ws = nullptr;
    size_t  ws_capacity = 0;
    krd::Tile* d_tiles = nullptr;
    std::vector<krd::Tile> h_tiles;
    void* d_sched = nullptr;      // work items + per-CTA offsets + per-tile offsets of the pair kernel (current schedule)
    int64_t sched_key = -1, sched_items = 0, sched_cta_bytes = 0;
    struct Sched { void* d; int64_t items, cta_bytes; };
    std::unordered_map<int64_t, Sched> scheds;   // schedules by key (tile range, chunks, CTAs, diagonal cost); d_sched points into one
    int64_t force_t0 = -1, force_t1 = -1;        // pipelined ingest: explicit tile range of the next run_pairs
    bool custom_tiles = false;                   // tile list set by the pipelined ingest (phase-major), not krd_tile_list
    int32_t late_blocks = 0;                     // krd_set_late_blocks: tiles touching the last `late_blocks` sample blocks come last
    int64_t late_split = 0;                      // ... and start at this index of the list

    // ---- pipelined ingest (krd_config.pipeline, krd_pipe.cu): sample-range children whose uploads overlap the pair kernel ----
    std::vector<krd_handle*> kids;
    std::vector<int32_t> kid_first;              // first sample of every child (+ N at the end)
    std::vector<int64_t> phase_off;              // tile ranges of the phases in h_tiles
    std::vector<std::vector<int64_t>> kid_off;   // rebased host offsets of every child
    std::vector<cudaEvent_t> kid_ev;             // upload of child q complete (copy stream)
    cudaStream_t copy_stream = nullptr;
    cudaStream_t cap_stream = nullptr;           // capture-only stream of the phase graphs (never executes anything)
    int32_t kids_ready = 0;                      // children whose masses are prepared
    int64_t tree_version = 0, kids_tree_version = -1;
    bool piped = false;
    const int64_t* d_off_from = nullptr;         // child: its sample offsets are this device array (the parent's), rebased
    bool defer_check = false;                    // child q >= 1: K0 leaves its check flags on the device (no host round trip)
    bool graph_ok = false;                       // a phase has run launch by launch once (buffers, schedules exist)
    std::vector<cudaGraphExec_t> phase_exec;     // executable graph of every phase, updated in place call after call
    bool capturing = false;                      // inside a stream capture (a phase of the pipelined ingest): no timing events
    int64_t n_tiles_alloc = 0;
    double* d_shard = nullptr;    // krd_compute_allgather: this rank's packed tiles ...
    double* d_gathered = nullptr; // ... and all ranks' after the NCCL all-gather
    double* d_out_tmp = nullptr;  // device N*N for krd_compute (host out)
    double* d_out_tmp_sq = nullptr;   // squash clustering: initial N*N distances
    void*   d_squash = nullptr;       // squash clustering state (distance matrix of 2N-1 clusters, counts, mergers, ...)
    void*   d_kmeans = nullptr;       // k-means state (distance partials, assignments, member lists)
    size_t  out_tmp_capacity = 0;

    std::unordered_map<void**, size_t> caps;   // capacities of buffers managed by krd::reserve

    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_copy = nullptr;   // krd_shard_to_host_async: the tile transposes are done (compute stream -> copy stream)
    krd_stats stats{};
};

namespace krd {

krd_status fail(krd_handle* h, krd_status st, const std::string& msg);
krd_status cuda_fail(krd_handle* h, cudaError_t e, const char* what);

#define KRD_CUDA(h, call)                                              \
    do {                                                               \
        cudaError_t _e = (call);                                       \
        if (_e != cudaSuccess) return krd::cuda_fail((h), _e, #call);  \
    } while (0)

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE setting: remember it per device (one process may hold
// handles on several GPUs, e.g. `gappa-b200 --gpus G`), not per process.
template <typename Kernel>
inline cudaError_t allow_dynamic_smem(Kernel kernel, int bytes, std::atomic<unsigned long long>& done_mask)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 64 && ((done_mask.load(std::memory_order_acquire) >> dev) & 1ull)) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess && dev < 64) done_mask.fetch_or(1ull << dev, std::memory_order_release);
    return e;
}

krd_status alloc_sized(krd_handle* h, void** p, size_t bytes, const char* what);   // cudaMalloc with a sized error message

// grow-only device buffer: reallocates *p only when `bytes` exceeds its recorded capacity
krd_status reserve(krd_handle* h, void** p, size_t bytes);

// krd_build.cu
void launch_rebase_off(const int64_t* d_src, int n, int64_t* d_dst, cudaStream_t stream);
krd_status build_tree(krd_handle* h);                 // host preprocessing + upload
krd_status prepare_masses(krd_handle* h);             // canonical order, slots, segment table
krd_status build_F(krd_handle* h);                    // K1
krd_status needs_resident_F(krd_handle* h);           // the sized error of a streamed handle asked for something that needs all of F

// krd_pairs.cu
krd_status run_pairs(krd_handle* h, double p, double divisor, int rank, int world,
                     double* d_out_matrix, double* d_out_packed);
// One slab of a streamed compute: rows [row0, row0 + rows) of F are at `F`; the per-tile sums are carried in h->d_acc.
struct SlabArgs { const double* F; int64_t row0, rows; bool first, last; };
krd_status run_pairs_slab(krd_handle* h, double p, double divisor, int rank, int world, double* d_out_matrix, double* d_out_packed,
                          const SlabArgs& slab);
krd_status compute_streamed(krd_handle* h, double p, double divisor, int rank, int world, double* d_out, double* d_packed);   // krd_build.cu
krd_status unpack_shards(krd_handle* h, int world, const double* d_gathered, double* d_out);
krd_status unpack_shard_rows(krd_handle* h, const double* d_gathered, int row0, int row1, double* d_out_rows);
krd_status set_late_blocks(krd_handle* h, int n_late, int64_t* split);
krd_status rank_tile_range(krd_handle* h, int rank, int world, const Tile** d_tiles, int64_t* t0, int64_t* t1);   // device tile list + the rank's share

// krd_pipe.cu — pipelined ingest (krd_config.pipeline)
bool pipe_wanted(const krd_handle* h, int32_t n_samples);
krd_status pipe_set_masses(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const int32_t* edge, const double* pos, const double* mass,
                           const void* cell32, int cell_bytes);   // cell_bytes 4 / 2: compact input (cells, mass), edge / pos unused; 0: (edge, pos, mass)
// krd_api.cu: masses already on the device; `binned` selects the compact form (d_cell32, d_mass)
krd_status set_masses_device_impl(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const int32_t* d_edge, const double* d_pos,
                                  const double* d_mass, const void* d_cell32, int cell_bytes);
krd_status pipe_compute(krd_handle* h, double p, double divisor, double* d_out);   // all phases, or everything at once when the children are resident
krd_status pipe_flatten(krd_handle* h);     // finish the uploads and turn the handle into an ordinary one (squash, k-means, sharding, introspection)
void pipe_destroy(krd_handle* h);
krd_status prepare_len_binned(krd_handle* h);   // krd_build.cu: segment lengths / offsets of binned mode (no mass data needed)
krd_status set_custom_tiles(krd_handle* h, const std::vector<krd::Tile>& tiles);   // krd_pairs.cu

// krd_squash.cu
krd_status run_squash(krd_handle* h, double p, krd_merger* out);
int64_t tiles_per_rank(int64_t n_tiles, int world);

// krd_kmeans.cu
krd_status run_kmeans(krd_handle* h, int k, int max_iterations, const int32_t* init_centroids, const int32_t* init_assignments,
                      int32_t* assignments, double* distances, int32_t* iterations_out);

}  // namespace krd
