// Pipelined ingest (krd_config.pipeline): the end-to-end path from HOST mass arrays is PCIe-bound (cfg2: 391 MB at 54 GB/s
// = 7.3 ms), and everything that follows an upload — K0 (slots, tile order), K1 (F columns) and the pair tiles — depends
// only on the samples involved.  The samples are therefore cut into up to eight sample-range CHILDREN (whole 128-sample
// blocks; when the block count is not a multiple of eight the larger children come first, because what is left when the
// upload ends is the LAST child's K0 + K1 and the tiles of its columns).  Child q's entries cross PCIe on a copy stream
// while the compute stream runs "phase q-1": K0 and K1 of child q-1 and the pair tiles whose blocks all lie in children
// <= q-1.
//
// What made the phases keep up with the upload (profiles/r1r_pipe_timeline.txt):
//  * no host round trip per child: only child 0 is checked by krd_set_masses itself (input errors surface there); the later
//    children leave their check flags on the device (k_clamp_edges keeps a bad edge id inside the buffers, the guards in
//    k_slice_len / k_tile_order make a mis-ordered child a no-op) and ONE read-back follows the last phase: a bad edge fails
//    the call, entries out of MassTree order restart the cohort on the ordinary (sorting) path;
//  * no copy-engine work on the compute stream while the bulk upload runs (a 1 KB offsets copy queued 0.3 ms behind it):
//    the child's offsets are rebased on the device from the parent's, zero fills are kernels;
//  * every phase (~30 launches) is captured into a CUDA graph and submitted as one launch — launch by launch, each
//    submission is slow to reach the GPU while the H2D stream saturates PCIe (phases 3.5x slower than after the upload).
//    The executable graph of a phase is kept and updated in place (cudaGraphExecUpdate) call after call; a phase that
//    cannot be captured (first call: buffers grow, schedules are not cached yet) is issued launch by launch instead.
//
// The parent owns F, the segment table, the tile list (phase-major) and the device copy of the entries; a child is an
// ordinary handle whose masses are a sub-range of the parent's device arrays (krd_set_masses_device path) and whose F
// builder writes ncols columns of the parent's F (ld_fixed).  krd_set_masses returns after child 0 is prepared; the first
// compute finishes the rest.  Anything else that needs all masses in one handle (squash, k-means, sharding, introspection)
// calls pipe_flatten, which waits for the uploads and prepares the parent the ordinary way.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "krd_common.cuh"

extern "C" krd_status krd_create(const krd_config* cfg, krd_handle** out);
extern "C" void krd_destroy(krd_handle* h);
extern "C" krd_status krd_set_tree(krd_handle* h, int32_t n_edges, const int32_t* parent_node, const int32_t* child_node,
                                   const double* branch_length, int32_t root_node);

namespace krd {

constexpr int PIPE_KIDS = 8;
constexpr int PIPE_MIN_SAMPLES = 4 * TILE;

bool pipe_wanted(const krd_handle* h, int32_t n_samples)
{
    return h->cfg.pipeline == 1 && !h->ld_fixed && h->cfg.builder != 1 && h->cfg.bins >= 1 && h->cfg.bins + 1 <= 32 &&
           h->E > 0 && n_samples >= PIPE_MIN_SAMPLES;
}

// KRD_PIPE_DEBUG=1: timeline of the uploads and phases in ms from the first upload (tuning aid)
static const bool g_dbg = std::getenv("KRD_PIPE_DEBUG") != nullptr;
static cudaEvent_t g_wait[PIPE_KIDS], g_t0, g_up[PIPE_KIDS], g_pair[PIPE_KIDS];
static void dbg_init()
{
    static bool done = false;
    if (done) return;
    done = true;
    cudaEventCreate(&g_t0);
    for (int q = 0; q < PIPE_KIDS; ++q) { cudaEventCreate(&g_up[q]); cudaEventCreate(&g_wait[q]); cudaEventCreate(&g_pair[q]); }
}

static krd_status kid_error(krd_handle* h, krd_handle* kid, krd_status st)
{
    return fail(h, st, kid ? kid->err : std::string("pipelined ingest: child handle failed"));
}

static krd_status kid_prepare(krd_handle* h, int q, bool wait = true);

krd_status pipe_set_masses(krd_handle* h, int32_t N, const int64_t* off, const int32_t* edge, const double* pos, const double* mass,
                           const void* cell32, int cell_bytes)
{
    const bool binned = cell_bytes != 0;
    const int64_t M = off[N];
    const int T = (N + TILE - 1) / TILE;
    int K = std::min(T, PIPE_KIDS);
    std::vector<int> blk0((size_t)K + 1, 0);                  // first 128-sample block of every child; T % K children get one more, first
    for (int q = 0; q < K; ++q) blk0[(size_t)q + 1] = blk0[(size_t)q] + T / K + (q < T % K ? 1 : 0);
    if (const char* ov = std::getenv("KRD_PIPE_SIZES")) {     // tuning aid: "3,2,1,1,1" = blocks per child (must add up to T)
        std::vector<int> b{0};
        for (const char* c = ov; *c;) { b.push_back(b.back() + std::max(1, std::atoi(c))); while (*c && *c != ',') ++c; if (*c) ++c; }
        if (b.back() == T && (int)b.size() - 1 <= PIPE_KIDS) { blk0 = b; K = (int)b.size() - 1; }
    }
    h->piped = false; h->have_masses = false; h->F_valid = false; h->kids_ready = 0;
    h->N = N; h->M = M;

    // ---- children (created once, re-used; the tree follows the parent's) ----
    while ((int)h->kids.size() < K) {
        krd_config c = h->cfg;
        c.pipeline = 0;
        krd_handle* kid = nullptr;
        if (krd_create(&c, &kid) != KRD_OK) return fail(h, KRD_ERR_CUDA, "pipelined ingest: cannot create a child handle");
        kid->ld_fixed = true;
        h->kids.push_back(kid);
        h->kids_tree_version = -1;
    }
    if (h->kids_tree_version != h->tree_version) {
        for (krd_handle* kid : h->kids) {
            kid->stream = h->stream;
            krd_status st = krd_set_tree(kid, h->E, h->parent.data(), h->child.data(), h->bl.data(), h->root);
            if (st != KRD_OK) return kid_error(h, kid, st);
        }
        h->kids_tree_version = h->tree_version;
    }
    if (!h->copy_stream) KRD_CUDA(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    while ((int)h->kid_ev.size() < K) {
        cudaEvent_t ev;
        KRD_CUDA(h, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        h->kid_ev.push_back(ev);
    }

    // ---- parent: segment table, F, device copy of the entries ----
    krd_status st = prepare_len_binned(h);
    if (st != KRD_OK) return st;
    h->ld = ((N + h->extra_cols + TILE - 1) / TILE) * TILE; h->ncols = h->ld;
    const size_t needF = (size_t)h->S_pad * h->ld * sizeof(double);
    if (needF > h->F_capacity) {
        cudaFree(h->d_F); h->d_F = nullptr; h->F_capacity = 0;
        if ((st = alloc_sized(h, (void**)&h->d_F, needF, "the dense cumulative-mass matrix F (segments x samples x 8 B)")) != KRD_OK) return st;
        h->F_capacity = needF;
    }
    if (h->ld > T * TILE)   // reserved columns after the samples (nothing writes them here)
        KRD_CUDA(h, cudaMemset2DAsync(h->d_F + (size_t)T * TILE, (size_t)h->ld * 8, 0, (size_t)(h->ld - T * TILE) * 8, (size_t)h->S_pad, h->stream));
    const size_t Mx = (size_t)std::max<int64_t>(M, 1);
    if ((st = reserve(h, (void**)&h->d_off, (size_t)(N + 1) * 8)) != KRD_OK) return st;
    if ((st = reserve(h, &h->d_owned[0], Mx * 4)) != KRD_OK) return st;
    if (!binned && (st = reserve(h, &h->d_owned[1], Mx * 8)) != KRD_OK) return st;
    if ((st = reserve(h, &h->d_owned[2], Mx * 8)) != KRD_OK) return st;
    h->d_mass = (double*)h->d_owned[2];
    h->binned_input = binned;
    if (binned) { h->d_cell32 = h->d_owned[0]; h->cell_bytes = cell_bytes; h->d_edge = nullptr; h->d_pos = nullptr; }
    else { h->d_cell32 = nullptr; h->cell_bytes = 4; h->d_edge = (int32_t*)h->d_owned[0]; h->d_pos = (double*)h->d_owned[1]; }

    // ---- uploads, child by child, on the copy stream (after whatever the compute stream still does with the buffers) ----
    KRD_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
    KRD_CUDA(h, cudaStreamWaitEvent(h->copy_stream, h->ev[0], 0));
    if (g_dbg) { dbg_init(); cudaEventRecord(g_t0, h->copy_stream); }
    KRD_CUDA(h, cudaMemcpyAsync(h->d_off, off, (size_t)(N + 1) * 8, cudaMemcpyHostToDevice, h->copy_stream));
    h->kid_first.assign((size_t)K + 1, N);
    h->kid_off.assign((size_t)K, {});
    for (int q = 0; q < K; ++q) {
        const int s0 = blk0[(size_t)q] * TILE, s1 = std::min(N, blk0[(size_t)q + 1] * TILE);
        h->kid_first[(size_t)q] = s0;
        const int64_t m0 = off[s0], m1 = off[s1];
        auto& ko = h->kid_off[(size_t)q];
        ko.resize((size_t)(s1 - s0) + 1);
        for (int i = s0; i <= s1; ++i) ko[(size_t)(i - s0)] = off[i] - m0;
        if (m1 > m0) {
            if (binned) KRD_CUDA(h, cudaMemcpyAsync((char*)h->d_owned[0] + m0 * cell_bytes, (const char*)cell32 + m0 * cell_bytes, (size_t)(m1 - m0) * cell_bytes, cudaMemcpyHostToDevice, h->copy_stream));
            else {
                KRD_CUDA(h, cudaMemcpyAsync(h->d_edge + m0, edge + m0, (size_t)(m1 - m0) * 4, cudaMemcpyHostToDevice, h->copy_stream));
                KRD_CUDA(h, cudaMemcpyAsync(h->d_pos + m0, pos + m0, (size_t)(m1 - m0) * 8, cudaMemcpyHostToDevice, h->copy_stream));
            }
            KRD_CUDA(h, cudaMemcpyAsync(h->d_mass + m0, mass + m0, (size_t)(m1 - m0) * 8, cudaMemcpyHostToDevice, h->copy_stream));
        }
        KRD_CUDA(h, cudaEventRecord(h->kid_ev[(size_t)q], h->copy_stream));
        if (g_dbg) cudaEventRecord(g_up[q], h->copy_stream);
    }
    h->kid_first[(size_t)K] = N;

    // ---- phase-major tile list: phase q = tiles whose larger block lies in child q ----
    std::vector<Tile> tiles;
    h->phase_off.assign((size_t)K + 1, 0);
    for (int q = 0; q < K; ++q) {
        const int b0 = blk0[(size_t)q], b1 = blk0[(size_t)q + 1];
        for (int j = b0; j < b1; ++j) {
            tiles.push_back(Tile{j, j});
            for (int i = 0; i < j; ++i) tiles.push_back(Tile{i, j});
        }
        h->phase_off[(size_t)q + 1] = (int64_t)tiles.size();
    }
    const bool same_tiles = h->custom_tiles && h->h_tiles.size() == tiles.size() &&
                            std::equal(tiles.begin(), tiles.end(), h->h_tiles.begin(), [](const Tile& a, const Tile& b) { return a.i == b.i && a.j == b.j; });
    if (!same_tiles && (st = set_custom_tiles(h, tiles)) != KRD_OK) return st;   // (keeps the cached schedules when nothing changed)

    // ---- child 0 now (the caller gets its input errors here); the others inside the first compute ----
    h->piped = true;
    h->kids_ready = 0;
    if ((st = kid_prepare(h, 0)) != KRD_OK) { h->piped = false; return st; }
    h->kids_ready = 1;
    h->have_masses = true;
    h->stats.n_samples = N; h->stats.n_edges = h->E; h->stats.n_masses = M;
    h->stats.n_segments = h->S; h->stats.ld_samples = h->ld; h->stats.f_bytes = (int64_t)h->S_pad * h->ld * 8;
    return KRD_OK;
}

// masses of child q: wait for its upload, then the ordinary K0 on a sub-range of the parent's device arrays
static krd_status kid_prepare(krd_handle* h, int q, bool wait)
{
    krd_handle* kid = h->kids[(size_t)q];
    const int s0 = h->kid_first[(size_t)q];
    if (wait) {
        KRD_CUDA(h, cudaStreamWaitEvent(h->stream, h->kid_ev[(size_t)q], 0));
        if (g_dbg) { dbg_init(); cudaEventRecord(g_wait[q], h->stream); }
    }
    kid->stream = h->stream;
    kid->d_off_from = h->d_off + s0;   // (uploaded first on the copy stream, before child 0's entries)
    kid->defer_check = q > 0;      // child 0 is checked by krd_set_masses itself; the others by check_deferred below
    kid->ld = h->ld;
    kid->d_F = h->d_F + s0;
    const auto& ko = h->kid_off[(size_t)q];
    // first entry of the child inside the parent's arrays = the parent's offset of its first sample
    int64_t base = 0;
    for (int k = 0; k < q; ++k) base += h->kid_off[(size_t)k].back();
    if (g_dbg) std::fprintf(stderr, "[pipe] child %d: samples %d..%d, entries %lld..%lld of %lld\n", q, s0, s0 + (int)ko.size() - 1, (long long)base, (long long)(base + ko.back()), (long long)h->M);
    krd_status st = h->binned_input
        ? set_masses_device_impl(kid, (int32_t)ko.size() - 1, ko.data(), nullptr, nullptr, h->d_mass + base, (const char*)h->d_cell32 + base * h->cell_bytes, h->cell_bytes)
        : set_masses_device_impl(kid, (int32_t)ko.size() - 1, ko.data(), h->d_edge + base, h->d_pos + base, h->d_mass + base, nullptr, 0);
    if (st != KRD_OK) return kid_error(h, kid, st);
    if (!kid->tiled) return fail(h, KRD_ERR_STATE, "pipelined ingest: child without a tiled builder");
    return KRD_OK;
}

// Flags of the children prepared without a host round trip: one read-back when everything has been issued.
// *unsorted = some child's entries were not in MassTree order (the caller re-runs the cohort the ordinary way, which sorts).
static krd_status check_deferred(krd_handle* h, int K, bool* unsorted)
{
    int flags[PIPE_KIDS] = {0};
    for (int q = 1; q < K; ++q)
        KRD_CUDA(h, cudaMemcpyAsync(&flags[q], h->kids[(size_t)q]->d_flags, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    int all = 0;
    for (int q = 1; q < K; ++q) all |= flags[q];
    if (all & 2) {
        h->have_masses = false; h->piped = false; h->custom_tiles = false; h->n_tiles_alloc = -1; h->F_valid = false;
        return fail(h, KRD_ERR_INVALID_ARGUMENT, "mass entry refers to an edge outside the tree");
    }
    *unsorted = (all & 1) != 0;
    return KRD_OK;
}

static void drop_graphs(krd_handle* h)
{
    if (h->phase_exec.empty()) return;
    cudaStreamSynchronize(h->stream);
    for (cudaGraphExec_t ge : h->phase_exec) if (ge) cudaGraphExecDestroy(ge);
    h->phase_exec.clear();
}

krd_status pipe_compute(krd_handle* h, double p, double divisor, double* d_out)
{

    const int K = (int)h->kid_first.size() - 1;
    const int mode_gram = (p == 2.0 && h->cfg.p2_mode != 1);
    h->stats.launches = 0;
    cudaEventRecord(h->ev[0], h->stream);
    const bool phased = !mode_gram && h->kids_ready < K;   // uploads still in flight: overlap them, phase by phase
    if (!phased) {
        if (h->kids_ready < K) {
            for (int q = h->kids_ready; q < K; ++q) { krd_status st = kid_prepare(h, q); if (st != KRD_OK) return st; h->kids_ready = q + 1; }
            bool unsorted = false;
            krd_status st = check_deferred(h, K, &unsorted);
            if (st != KRD_OK) return st;
            if (unsorted) return pipe_flatten(h);      // (h->piped is false now: the caller continues on the ordinary path)
        }
        if (!h->F_valid) {
            for (int q = 0; q < K; ++q) {
                krd_status st = build_F(h->kids[(size_t)q]);
                if (st != KRD_OK) return kid_error(h, h->kids[(size_t)q], st);
                h->stats.launches += h->kids[(size_t)q]->build_launches;
            }
            h->F_valid = true; h->norms_valid = false;
        }
        cudaEventRecord(h->ev[1], h->stream);
        h->force_t0 = 0; h->force_t1 = h->phase_off[(size_t)K];
        krd_status st = run_pairs(h, p, divisor, 0, 1, d_out, nullptr);
        h->force_t0 = h->force_t1 = -1;
        if (st != KRD_OK) return st;
        cudaEventRecord(h->ev[3], h->stream);
        return KRD_OK;
    }
    // One phase = the child's F columns (K1: ~14 launches) + the pair tiles that are complete with it + their epilogue.
    // While uploads saturate PCIe every launch is slow to reach the GPU, so a phase is captured into a CUDA graph and
    // submitted as ONE launch; anything that cannot be captured (a first call that still allocates, a schedule not yet
    // cached) ends the capture and the phase is issued launch by launch instead — the work is idempotent.
    static const bool use_graph = std::getenv("KRD_PIPE_NO_GRAPH") == nullptr;
    auto phase = [&](int q, bool prep) -> krd_status {
        krd_status st = prep ? kid_prepare(h, q, false) : KRD_OK;     // K0 of the child (its upload was waited for outside)
        if (st != KRD_OK) return st;
        st = build_F(h->kids[(size_t)q]);
        if (st != KRD_OK) return kid_error(h, h->kids[(size_t)q], st);
        h->force_t0 = h->phase_off[(size_t)q]; h->force_t1 = h->phase_off[(size_t)q + 1];
        st = run_pairs(h, p, divisor, 0, 1, d_out, nullptr);
        h->force_t0 = h->force_t1 = -1;
        return st;
    };
    for (int q = 0; q < K; ++q) {
        const bool prep = q >= h->kids_ready;
        if (prep) {
            KRD_CUDA(h, cudaStreamWaitEvent(h->stream, h->kid_ev[(size_t)q], 0));
        }
        if (g_dbg) cudaEventRecord(g_wait[q], h->stream);
        if (q == K - 1) cudaEventRecord(h->ev[1], h->stream);   // ms_build / ms_pairs of the stats describe the LAST phase
        bool launched = false;
        // (captured on a stream of our own: the caller's stream may be the legacy default stream, which cannot capture)
        cudaStream_t const real = h->stream;
        if (use_graph && h->graph_ok && !h->cap_stream && cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking) != cudaSuccess) { h->cap_stream = nullptr; cudaGetLastError(); }
        if (use_graph && h->graph_ok && h->cap_stream && cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeRelaxed) != cudaSuccess) cudaGetLastError();
        else if (use_graph && h->graph_ok && h->cap_stream) {
            h->capturing = true; h->kids[(size_t)q]->capturing = true;
            h->stream = h->cap_stream; h->kids[(size_t)q]->stream = h->cap_stream;
            const std::string err_before = h->err;
            const int64_t launches_before = h->stats.launches;
            const krd_status st = phase(q, prep);
            h->capturing = false; h->kids[(size_t)q]->capturing = false;
            h->stream = real; h->kids[(size_t)q]->stream = real;
            cudaGraph_t g = nullptr;
            const cudaError_t ce = cudaStreamEndCapture(h->cap_stream, &g);
            // the executable graph of phase q is kept and UPDATED in place with this call's capture (same topology, new
            // arguments): instantiating and destroying eight graphs per call cost a 2.5-4 ms stall every fifth call
            if (st == KRD_OK && ce == cudaSuccess && g) {
                if ((int)h->phase_exec.size() <= q) h->phase_exec.resize((size_t)q + 1, nullptr);
                cudaGraphExec_t& ge = h->phase_exec[(size_t)q];
                if (ge) {
                    cudaGraphExecUpdateResultInfo info;
                    if (cudaGraphExecUpdate(ge, g, &info) != cudaSuccess) { cudaGetLastError(); cudaGraphExecDestroy(ge); ge = nullptr; }
                }
                if (!ge && cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) { cudaGetLastError(); ge = nullptr; }
                launched = ge && cudaGraphLaunch(ge, h->stream) == cudaSuccess;
            }
            if (g) cudaGraphDestroy(g);
            if (!launched) { cudaGetLastError(); h->err = err_before; h->stats.launches = launches_before; }
        }
        if (!launched) {
            krd_status st = phase(q, prep);
            if (st != KRD_OK) return st;
            h->graph_ok = true;                               // buffers and schedules exist now: the next call can be captured
        }
        if (prep) h->kids_ready = q + 1;
        h->stats.launches += h->kids[(size_t)q]->build_launches;
        if (q == K - 1) cudaEventRecord(h->ev[2], h->stream);
        if (g_dbg) cudaEventRecord(g_pair[q], h->stream);
    }
    h->F_valid = true; h->norms_valid = false;
    cudaEventRecord(h->ev[3], h->stream);
    {
        bool unsorted = false;
        krd_status st = check_deferred(h, K, &unsorted);
        if (st != KRD_OK) return st;
        if (unsorted) return pipe_flatten(h);          // results so far are discarded; the caller recomputes the ordinary way
    }
    if (g_dbg) {
        cudaStreamSynchronize(h->stream);
        for (int q = 0; q < K; ++q) {
            float a = 0, w = 0, d = 0;
            cudaEventElapsedTime(&a, g_t0, g_up[q]); cudaEventElapsedTime(&w, g_t0, g_wait[q]); cudaEventElapsedTime(&d, g_t0, g_pair[q]);
            std::fprintf(stderr, "[pipe] child %d: upload done %.3f  phase (K0 + K1 + pair tiles) %.3f .. %.3f ms\n", q, a, w, d);
        }
    }
    return KRD_OK;
}

krd_status pipe_flatten(krd_handle* h)
{
    if (!h->piped) return KRD_OK;
    KRD_CUDA(h, cudaStreamSynchronize(h->copy_stream));   // all entries are on the device
    h->piped = false; h->custom_tiles = false; h->n_tiles_alloc = -1; h->kids_ready = 0;
    h->F_valid = false;
    krd_status st = prepare_masses(h);                    // ordinary K0 on the parent's complete device arrays
    if (st != KRD_OK) { h->have_masses = false; return st; }
    return KRD_OK;
}

void pipe_destroy(krd_handle* h)
{
    if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    drop_graphs(h);
    if (h->cap_stream) { cudaStreamDestroy(h->cap_stream); h->cap_stream = nullptr; }
    for (krd_handle* kid : h->kids) { kid->d_F = nullptr; kid->d_edge = nullptr; kid->d_pos = nullptr; kid->d_mass = nullptr; kid->d_cell32 = nullptr; krd_destroy(kid); }
    h->kids.clear();
    for (cudaEvent_t ev : h->kid_ev) cudaEventDestroy(ev);
    h->kid_ev.clear();
    if (h->copy_stream) { cudaStreamDestroy(h->copy_stream); h->copy_stream = nullptr; }
}

}  // namespace krd
