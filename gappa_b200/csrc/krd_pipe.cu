// Pipelined ingest (krd_config.pipeline): the end-to-end path from HOST mass arrays is PCIe-bound (cfg2: 391 MB at 51 GB/s
// = 7.6 ms of an 11.7 ms call), and everything that follows an upload — K0 (slots, tile order), K1 (F columns) and the
// pair tiles — depends only on the samples involved.  The samples are therefore cut into up to four sample-range CHILDREN
// (whole 128-sample blocks); child q's entries are uploaded on a copy stream while the compute stream prepares child q-1,
// builds its columns of the parent's F and runs the pair tiles whose blocks all lie in children <= q-1 ("phase q-1").
// Only the last child's K0 + K1 and the last phase are left when the upload ends.
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
extern "C" krd_status krd_set_masses_device(krd_handle* h, int32_t n_samples, const int64_t* sample_off, const int32_t* d_edge,
                                            const double* d_pos, const double* d_mass);

namespace krd {

constexpr int PIPE_KIDS = 4;
constexpr int PIPE_MIN_SAMPLES = 4 * TILE;

bool pipe_wanted(const krd_handle* h, int32_t n_samples)
{
    return h->cfg.pipeline == 1 && !h->ld_fixed && h->cfg.builder != 1 && h->cfg.bins >= 1 && h->cfg.bins + 1 <= 32 &&
           h->E > 0 && n_samples >= PIPE_MIN_SAMPLES;
}

static krd_status kid_error(krd_handle* h, krd_handle* kid, krd_status st)
{
    return fail(h, st, kid ? kid->err : std::string("pipelined ingest: child handle failed"));
}

static krd_status kid_prepare(krd_handle* h, int q);

krd_status pipe_set_masses(krd_handle* h, int32_t N, const int64_t* off, const int32_t* edge, const double* pos, const double* mass)
{
    const int64_t M = off[N];
    const int T = (N + TILE - 1) / TILE;
    const int bpk = (T + PIPE_KIDS - 1) / PIPE_KIDS;          // blocks per child
    const int K = (T + bpk - 1) / bpk;
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
        KRD_CUDA(h, cudaMalloc(&h->d_F, needF));
        h->F_capacity = needF;
    }
    if (h->ld > T * TILE)   // reserved columns after the samples (nothing writes them here)
        KRD_CUDA(h, cudaMemset2DAsync(h->d_F + (size_t)T * TILE, (size_t)h->ld * 8, 0, (size_t)(h->ld - T * TILE) * 8, (size_t)h->S_pad, h->stream));
    const size_t Mx = (size_t)std::max<int64_t>(M, 1);
    if ((st = reserve(h, (void**)&h->d_off, (size_t)(N + 1) * 8)) != KRD_OK) return st;
    if ((st = reserve(h, &h->d_owned[0], Mx * 4)) != KRD_OK) return st;
    if ((st = reserve(h, &h->d_owned[1], Mx * 8)) != KRD_OK) return st;
    if ((st = reserve(h, &h->d_owned[2], Mx * 8)) != KRD_OK) return st;
    h->d_edge = (int32_t*)h->d_owned[0]; h->d_pos = (double*)h->d_owned[1]; h->d_mass = (double*)h->d_owned[2];

    // ---- uploads, child by child, on the copy stream (after whatever the compute stream still does with the buffers) ----
    KRD_CUDA(h, cudaEventRecord(h->ev[0], h->stream));
    KRD_CUDA(h, cudaStreamWaitEvent(h->copy_stream, h->ev[0], 0));
    KRD_CUDA(h, cudaMemcpyAsync(h->d_off, off, (size_t)(N + 1) * 8, cudaMemcpyHostToDevice, h->copy_stream));
    h->kid_first.assign((size_t)K + 1, N);
    h->kid_off.assign((size_t)K, {});
    for (int q = 0; q < K; ++q) {
        const int s0 = q * bpk * TILE, s1 = std::min(N, (q + 1) * bpk * TILE);
        h->kid_first[(size_t)q] = s0;
        const int64_t m0 = off[s0], m1 = off[s1];
        auto& ko = h->kid_off[(size_t)q];
        ko.resize((size_t)(s1 - s0) + 1);
        for (int i = s0; i <= s1; ++i) ko[(size_t)(i - s0)] = off[i] - m0;
        if (m1 > m0) {
            KRD_CUDA(h, cudaMemcpyAsync(h->d_edge + m0, edge + m0, (size_t)(m1 - m0) * 4, cudaMemcpyHostToDevice, h->copy_stream));
            KRD_CUDA(h, cudaMemcpyAsync(h->d_pos + m0, pos + m0, (size_t)(m1 - m0) * 8, cudaMemcpyHostToDevice, h->copy_stream));
            KRD_CUDA(h, cudaMemcpyAsync(h->d_mass + m0, mass + m0, (size_t)(m1 - m0) * 8, cudaMemcpyHostToDevice, h->copy_stream));
        }
        KRD_CUDA(h, cudaEventRecord(h->kid_ev[(size_t)q], h->copy_stream));
    }
    h->kid_first[(size_t)K] = N;

    // ---- phase-major tile list: phase q = tiles whose larger block lies in child q ----
    std::vector<Tile> tiles;
    h->phase_off.assign((size_t)K + 1, 0);
    for (int q = 0; q < K; ++q) {
        const int b0 = q * bpk, b1 = std::min(T, (q + 1) * bpk);
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
static krd_status kid_prepare(krd_handle* h, int q)
{
    krd_handle* kid = h->kids[(size_t)q];
    const int s0 = h->kid_first[(size_t)q];
    KRD_CUDA(h, cudaStreamWaitEvent(h->stream, h->kid_ev[(size_t)q], 0));
    kid->stream = h->stream;
    kid->ld = h->ld;
    kid->d_F = h->d_F + s0;
    const auto& ko = h->kid_off[(size_t)q];
    // first entry of the child inside the parent's arrays = the parent's offset of its first sample
    int64_t base = 0;
    for (int k = 0; k < q; ++k) base += h->kid_off[(size_t)k].back();
    if (std::getenv("KRD_PIPE_DEBUG")) std::fprintf(stderr, "[pipe] child %d: samples %d..%d, entries %lld..%lld of %lld\n", q, s0, s0 + (int)ko.size() - 1, (long long)base, (long long)(base + ko.back()), (long long)h->M);
    krd_status st = krd_set_masses_device(kid, (int32_t)ko.size() - 1, ko.data(), h->d_edge + base, h->d_pos + base, h->d_mass + base);
    if (st != KRD_OK) return kid_error(h, kid, st);
    if (!kid->tiled) return fail(h, KRD_ERR_STATE, "pipelined ingest: child without a tiled builder");
    return KRD_OK;
}

krd_status pipe_compute(krd_handle* h, double p, double divisor, double* d_out)
{
    const int K = (int)h->kid_first.size() - 1;
    const int mode_gram = (p == 2.0 && h->cfg.p2_mode != 1);
    h->stats.launches = 0;
    cudaEventRecord(h->ev[0], h->stream);
    const bool phased = !mode_gram && h->kids_ready < K;   // uploads still in flight: overlap them, phase by phase
    if (!phased) {
        for (int q = h->kids_ready; q < K; ++q) { krd_status st = kid_prepare(h, q); if (st != KRD_OK) return st; h->kids_ready = q + 1; }
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
    for (int q = 0; q < K; ++q) {
        if (q >= h->kids_ready) { krd_status st = kid_prepare(h, q); if (st != KRD_OK) return st; h->kids_ready = q + 1; }
        krd_status st = build_F(h->kids[(size_t)q]);
        if (st != KRD_OK) return kid_error(h, h->kids[(size_t)q], st);
        h->stats.launches += h->kids[(size_t)q]->build_launches;
        if (q == K - 1) cudaEventRecord(h->ev[1], h->stream);   // ms_build / ms_pairs of the stats describe the LAST phase
        h->force_t0 = h->phase_off[(size_t)q]; h->force_t1 = h->phase_off[(size_t)q + 1];
        st = run_pairs(h, p, divisor, 0, 1, d_out, nullptr);
        h->force_t0 = h->force_t1 = -1;
        if (st != KRD_OK) return st;
    }
    h->F_valid = true; h->norms_valid = false;
    cudaEventRecord(h->ev[3], h->stream);
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
    for (krd_handle* kid : h->kids) { kid->d_F = nullptr; kid->d_edge = nullptr; kid->d_pos = nullptr; kid->d_mass = nullptr; krd_destroy(kid); }
    h->kids.clear();
    for (cudaEvent_t ev : h->kid_ev) cudaEventDestroy(ev);
    h->kid_ev.clear();
    if (h->copy_stream) { cudaStreamDestroy(h->copy_stream); h->copy_stream = nullptr; }
}

}  // namespace krd
