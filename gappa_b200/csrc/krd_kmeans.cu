// Phylogenetic k-means on the dense matrix F — SURVEY §8f rank 4.
//
// Reference: genesis::tree::MassTreeKmeans (tree/mass_tree/kmeans.cpp:59-208) on the generic Lloyd loop of
// utils/math/kmeans.hpp:118-168, as run by `gappa analyze phylogenetic-kmeans --bins B`
// (src/commands/analyze/kmeans_phylogenetic.cpp:134-175).  With binned masses a centroid is the
// re-normalised SUM of its members' mass trees (update_centroids, kmeans.cpp:106-150) — in the dense
// representation the sum of the members' columns of F divided by the members' total mass — and the
// assignment step is a rectangular N x k version of the KR reduction (distance = earth_movers_distance with
// p = 1, kmeans.cpp:204-207).  Both steps are one pass over F each (HBM-bound).  Centroids live in the k
// columns after the samples.
#include <algorithm>
#include <vector>

#include "krd_common.cuh"

namespace krd {

constexpr int KM_CHUNKS = 128;   // segment chunks of the assignment reduction (partials summed in fixed order)

__global__ void k_km_copy_cols(double* __restrict__ F, int64_t ld, int64_t S_pad, const int32_t* __restrict__ src, int N, int k)
{
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= S_pad * k) return;
    const int64_t s = g / k; const int c = (int)(g % k);
    F[s * ld + N + c] = F[s * ld + src[c]];
}

__global__ void k_km_totals(const double* __restrict__ F, int64_t ld, const int32_t* __restrict__ root_seg, int n_root, int N,
                            double* __restrict__ total)
{
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= N) return;
    double t = 0.0;
    for (int r = 0; r < n_root; ++r) t += F[(int64_t)root_seg[r] * ld + d];
    total[d] = t;
}

// part[chunk][c][d] = sum over the chunk's segments of len * |F[s][d] - C[s][c]|
// The k centroid values and the length of KM_ROWS segments are staged in shared memory (read by every thread);
// the sample values stream from HBM with KM_UNROLL independent row loads in flight per thread.
constexpr int KM_ROWS = 64, KM_UNROLL = 8;
template <int KMAX>
__global__ void __launch_bounds__(128)
k_km_assign_part(const double* __restrict__ F, const double* __restrict__ len, int64_t ld, int64_t S_pad, int64_t rows_per_chunk,
                 int N, int k_all, int c0, double* __restrict__ part)
{
    const int k = min(KMAX, k_all - c0);     // this launch: centroids c0 .. c0 + k - 1 of k_all (more than 32 centroids: several launches)
    __shared__ double sc[KM_ROWS][KMAX];
    __shared__ double sl[KM_ROWS];
    const int64_t d = (int64_t)blockIdx.x * 128 + threadIdx.x;
    const int64_t s0 = (int64_t)blockIdx.y * rows_per_chunk, s1 = min(S_pad, s0 + rows_per_chunk);
    double acc[KMAX];
#pragma unroll
    for (int c = 0; c < KMAX; ++c) acc[c] = 0.0;
    for (int64_t r0 = s0; r0 < s1; r0 += KM_ROWS) {
        const int nr = (int)min((int64_t)KM_ROWS, s1 - r0);
        __syncthreads();
        for (int i = threadIdx.x; i < KM_ROWS * KMAX; i += 128) {
            const int r = i / KMAX, c = i % KMAX;
            sc[r][c] = (r < nr && c < k) ? F[(r0 + r) * ld + N + c0 + c] : 0.0;
        }
        if (threadIdx.x < KM_ROWS) sl[threadIdx.x] = (threadIdx.x < nr) ? len[r0 + threadIdx.x] : 0.0;   // rows past the chunk: length 0
        __syncthreads();
        for (int r = 0; r < nr; r += KM_UNROLL) {
            double x[KM_UNROLL];
#pragma unroll
            for (int u = 0; u < KM_UNROLL; ++u) x[u] = F[(r0 + min(r + u, nr - 1)) * ld + d];
#pragma unroll
            for (int u = 0; u < KM_UNROLL; ++u) {
                const double l = sl[r + u];
#pragma unroll
                for (int c = 0; c < KMAX; ++c) if (c < k) acc[c] = fma(l, fabs(x[u] - sc[r + u][c]), acc[c]);
            }
        }
    }
#pragma unroll
    for (int c = 0; c < KMAX; ++c)
        if (c < k) part[((int64_t)blockIdx.y * k_all + c0 + c) * ld + d] = acc[c];
}

// nearest centroid per sample, first minimum wins (find_nearest_cluster, kmeans.hpp:365-382);
// reassign = 0: only the distance to the CURRENT assignment (cluster_info, kmeans.hpp:384-424).
// Block = 32 samples x 32 centroid lanes: thread (x, y) sums the chunk partials of (sample x, centroids y, y + 32, ...) in
// fixed order and keeps its first strict minimum; lane y = 0 then takes the minimum over the lanes, ties to the smaller
// centroid index — the same "first minimum" for any k (the reference has no limit on k, kmeans.hpp:118-168).
__global__ void k_km_assign_final(const double* __restrict__ part, int64_t ld, int N, int k, int reassign,
                                  int32_t* __restrict__ assign, double* __restrict__ dist, int* __restrict__ changed)
{
    __shared__ double bv[32][33], curv[32];
    __shared__ int bc[32][33];
    const int x = threadIdx.x, y = threadIdx.y;
    const int d = blockIdx.x * 32 + x;
    const int cur = (d < N) ? assign[d] : 0;
    double best_v = 1.7976931348623157e308; int best_c = 0x7fffffff;
    for (int c = y; c < k; c += 32) {
        double v = 0.0;
        if (d < N) {
#pragma unroll 8
            for (int ch = 0; ch < KM_CHUNKS; ++ch) v += part[((int64_t)ch * k + c) * ld + d];
        }
        if (c == cur) curv[x] = v;
        if (v < best_v) { best_v = v; best_c = c; }
    }
    bv[y][x] = best_v; bc[y][x] = best_c;
    __syncthreads();
    if (y != 0 || d >= N) return;
    if (!reassign) { dist[d] = curv[x]; return; }
    for (int q = 1; q < 32; ++q)
        if (bv[q][x] < best_v || (bv[q][x] == best_v && bc[q][x] < best_c)) { best_v = bv[q][x]; best_c = bc[q][x]; }
    dist[d] = best_v;
    if (best_c != cur) { assign[d] = best_c; atomicOr(changed, 1); }
}

// centroid c = (sum of its members' columns, in sample order) / (sum of the members' total masses)
// (mass_tree_merge_trees_inplace per member, then mass_tree_normalize_masses: kmeans.cpp:124-149)
// STAGED: the segment row of the samples goes through shared memory (coalesced read once; N <= 25,600); otherwise the
// members are gathered from the row directly — every sample belongs to exactly one centroid, so the row is still read once.
template <bool STAGED>
__global__ void __launch_bounds__(256)
k_km_update(double* __restrict__ F, int64_t ld, int64_t S_pad, int N, int k, const int32_t* __restrict__ mem_off,
            const int32_t* __restrict__ mem, const double* __restrict__ ctot)
{
    extern __shared__ double row_sm[];   // one segment row of the samples
    for (int64_t s = blockIdx.x; s < S_pad; s += gridDim.x) {
        const double* row = STAGED ? row_sm : F + s * ld;
        if (STAGED) {
            for (int d = threadIdx.x; d < N; d += 256) row_sm[d] = F[s * ld + d];
            __syncthreads();
        }
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        for (int c = warp; c < k; c += 8) {
            double acc = 0.0;
            for (int q = mem_off[c] + lane; q < mem_off[c + 1]; q += 32) acc += row[mem[q]];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
            if (lane == 0) F[s * ld + N + c] = (ctot[c] == 0.0) ? acc : acc / ctot[c];
        }
        if (STAGED) __syncthreads();
    }
}

// ---- exact mode (bins = 0): branch-centred centroids -------------------------------------------------------------
// gappa leaves MassTreeKmeans::accumulate_centroid_masses at 1 when --bins is 0 (kmeans_phylogenetic.cpp:160-167), so a
// centroid carries ONE mass per branch, at the mass centre of everything its members have there
// (mass_tree_center_masses_on_branches_averaged, functions.cpp:181-220; the initial centroids too: pre_loop_hook,
// kmeans.cpp:67-84).  Its cumulative function on edge e is D (mass strictly below e) plus m_e for x < c_e: a centroid is
// three numbers per edge (D, m_e, c_e), and a segment [lo, hi] of e contributes
//   (clamp(c_e) - lo) |x - (D + m_e)| + (hi - clamp(c_e)) |x - D|.
// D comes from the members' F at the edge's last (distal) segment, m_e and c_e from per-sample per-edge sums of mass and
// position x mass taken once from the mass entries.
__global__ void k_km_seg_geom(const int32_t* __restrict__ seg_off, const double* __restrict__ len, int E, double* __restrict__ seg_lo,
                              int32_t* __restrict__ seg_edge)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    double lo = 0.0;
    for (int s = seg_off[e]; s < seg_off[e + 1]; ++s) { seg_lo[s] = lo; seg_edge[s] = e; lo += len[s]; }
}

// Q[e][i] = sum of sample i's masses on edge e, P[e][i] = sum of position * mass (entries are in (sample, edge, pos) order:
// the first entry of every run adds its run up sequentially)
__global__ void k_km_moments(const uint32_t* __restrict__ sid, const int32_t* __restrict__ edge, const double* __restrict__ pos,
                             const double* __restrict__ mass, int64_t M, int64_t ld, double* __restrict__ Q, double* __restrict__ P)
{
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= M) return;
    const uint32_t s = sid[g]; const int32_t e = edge[g];
    if (g > 0 && sid[g - 1] == s && edge[g - 1] == e) return;
    double q = 0.0, p = 0.0;
    for (int64_t j = g; j < M && sid[j] == s && edge[j] == e; ++j) { q += mass[j]; p += pos[j] * mass[j]; }
    Q[(int64_t)e * ld + s] = q; P[(int64_t)e * ld + s] = p;
}

// cent[e][c] = (D, m, centre) of centroid c on edge e from its members (fixed order: lane-strided partial sums + shuffle tree)
__global__ void __launch_bounds__(128)
k_km_update_exact(const double* __restrict__ F, const double* __restrict__ Q, const double* __restrict__ P, const int32_t* __restrict__ seg_off,
                  int64_t ld, int E, int k, const int32_t* __restrict__ mem_off, const int32_t* __restrict__ mem, const double* __restrict__ ctot,
                  double* __restrict__ cent)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int e = blockIdx.x; e < E; e += gridDim.x) {
        const int64_t last = (int64_t)(seg_off[e + 1] - 1) * ld, row = (int64_t)e * ld;
        for (int c = warp; c < k; c += 4) {
            double sD = 0.0, sQ = 0.0, sP = 0.0;
            for (int q = mem_off[c] + lane; q < mem_off[c + 1]; q += 32) { const int m = mem[q]; sD += F[last + m]; sQ += Q[row + m]; sP += P[row + m]; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                sD += __shfl_down_sync(0xffffffffu, sD, o); sQ += __shfl_down_sync(0xffffffffu, sQ, o); sP += __shfl_down_sync(0xffffffffu, sP, o);
            }
            if (lane == 0) {
                const double t = ctot[c] == 0.0 ? 1.0 : ctot[c];
                double* o3 = cent + ((int64_t)e * k + c) * 3;
                o3[0] = sD / t; o3[1] = sQ / t; o3[2] = sQ > 0.0 ? sP / sQ : 0.0;
            }
        }
    }
}

constexpr int KME_ROWS = 16;
template <int KMAX>
__global__ void __launch_bounds__(128)
k_km_assign_part_exact(const double* __restrict__ F, const double* __restrict__ len, const double* __restrict__ seg_lo,
                       const int32_t* __restrict__ seg_edge, const double* __restrict__ cent, int64_t ld, int64_t S, int64_t rows_per_chunk,
                       int N, int k_all, int c0, double* __restrict__ part)
{
    const int k = min(KMAX, k_all - c0);
    __shared__ double sv1[KME_ROWS][KMAX], sv2[KME_ROWS][KMAX], sw1[KME_ROWS][KMAX], sw2[KME_ROWS][KMAX];
    const int64_t d = (int64_t)blockIdx.x * 128 + threadIdx.x;
    const int64_t s0 = (int64_t)blockIdx.y * rows_per_chunk, s1 = min(S, s0 + rows_per_chunk);
    double acc[KMAX];
#pragma unroll
    for (int c = 0; c < KMAX; ++c) acc[c] = 0.0;
    for (int64_t r0 = s0; r0 < s1; r0 += KME_ROWS) {
        const int nr = (int)min((int64_t)KME_ROWS, s1 - r0);
        __syncthreads();
        for (int i = threadIdx.x; i < KME_ROWS * KMAX; i += 128) {
            const int r = i / KMAX, c = i % KMAX;
            double v1 = 0.0, v2 = 0.0, w1 = 0.0, w2 = 0.0;
            if (r < nr && c < k) {
                const int64_t s = r0 + r;
                const double* c3 = cent + ((int64_t)seg_edge[s] * k_all + c0 + c) * 3;
                const double lo = seg_lo[s], l = len[s];
                const double cc = fmin(fmax(c3[2], lo), lo + l);
                w1 = fmin(cc - lo, l); w2 = l - w1;
                v2 = c3[0]; v1 = v2 + c3[1];
            }
            sv1[r][c] = v1; sv2[r][c] = v2; sw1[r][c] = w1; sw2[r][c] = w2;
        }
        __syncthreads();
        for (int r = 0; r < nr; r += 4) {
            double x[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) x[u] = F[(r0 + min(r + u, nr - 1)) * ld + d];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (r + u < nr) {
#pragma unroll
                    for (int c = 0; c < KMAX; ++c)
                        if (c < k) acc[c] = fma(sw1[r + u][c], fabs(x[u] - sv1[r + u][c]), fma(sw2[r + u][c], fabs(x[u] - sv2[r + u][c]), acc[c]));
                }
            }
        }
    }
#pragma unroll
    for (int c = 0; c < KMAX; ++c)
        if (c < k) part[((int64_t)blockIdx.y * k_all + c0 + c) * ld + d] = acc[c];
}

template <int KMAX>
static void launch_assign_exact(krd_handle* h, int k, int c0, int64_t rpc, const double* seg_lo, const int32_t* seg_edge, const double* cent, double* part)
{
    k_km_assign_part_exact<KMAX><<<dim3((unsigned)((h->N + 127) / 128), KM_CHUNKS), 128, 0, h->stream>>>(
        h->d_F, h->d_len, seg_lo, seg_edge, cent, h->ld, h->S, rpc, h->N, k, c0, part);
}

template <int KMAX>
static void launch_assign(krd_handle* h, int k, int c0, int64_t rpc, double* part)
{
    k_km_assign_part<KMAX><<<dim3((unsigned)((h->N + 127) / 128), KM_CHUNKS), 128, 0, h->stream>>>(h->d_F, h->d_len, h->ld, h->S_pad, rpc, h->N, k, c0, part);
}

krd_status run_kmeans(krd_handle* h, int k, int max_iterations, const int32_t* init_centroids, const int32_t* init_assignments,
                      int32_t* assignments, double* distances, int32_t* iterations_out)
{
    const int N = h->N;
    const bool exact = h->cfg.bins == 0;   // branch-centred centroids (three numbers per edge) instead of centroid columns
    if (!exact && h->ld < N + k) {     // room for the centroid columns
        h->extra_cols = std::max(h->extra_cols, k);
        krd_status st = prepare_masses(h);
        if (st != KRD_OK) return st;
    }
    krd_status st = build_F(h);   // always rebuild: earlier clusterings leave their columns behind
    if (st != KRD_OK) return st;

    std::vector<int32_t> seg_off((size_t)h->E + 1), root_seg;
    KRD_CUDA(h, cudaMemcpyAsync(seg_off.data(), h->d_seg_off, seg_off.size() * 4, cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    for (int e = 0; e < h->E; ++e) if (h->parent[(size_t)e] == h->root) root_seg.push_back(seg_off[(size_t)e]);

    auto al = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t b_part = (size_t)KM_CHUNKS * k * h->ld * 8, b_n8 = (size_t)N * 8, b_n4 = (size_t)N * 4, b_k = (size_t)(k + 1) * 8;
    const size_t b_root = std::max<size_t>(root_seg.size(), 1) * 4;
    const size_t b_seg8 = exact ? (size_t)h->S_pad * 8 : 0, b_seg4 = exact ? (size_t)h->S_pad * 4 : 0;
    const size_t b_mom = exact ? (size_t)std::max(h->E, 1) * h->ld * 8 : 0, b_cent = exact ? (size_t)std::max(h->E, 1) * k * 3 * 8 : 0;
    if ((st = reserve(h, &h->d_kmeans, al(b_part) + 2 * al(b_n8) + 2 * al(b_n4) + 3 * al(b_k) + al(b_root) + 256 +
                                       al(b_seg8) + al(b_seg4) + 2 * al(b_mom) + al(b_cent))) != KRD_OK) return st;
    char* q = static_cast<char*>(h->d_kmeans);
    double* part = (double*)q; q += al(b_part);
    double* d_dist = (double*)q; q += al(b_n8);
    double* d_total = (double*)q; q += al(b_n8);
    int32_t* d_assign = (int32_t*)q; q += al(b_n4);
    int32_t* d_mem = (int32_t*)q; q += al(b_n4);
    int32_t* d_mem_off = (int32_t*)q; q += al(b_k);
    double* d_ctot = (double*)q; q += al(b_k);
    int32_t* d_init = (int32_t*)q; q += al(b_k);
    int32_t* d_root = (int32_t*)q; q += al(b_root);
    int* d_changed = (int*)q; q += 256;
    double* seg_lo = (double*)q; q += al(b_seg8);
    int32_t* seg_edge = (int32_t*)q; q += al(b_seg4);
    double* d_Q = (double*)q; q += al(b_mom);
    double* d_P = (double*)q; q += al(b_mom);
    double* d_cent = (double*)q;

    KRD_CUDA(h, cudaMemcpyAsync(d_root, root_seg.data(), root_seg.size() * 4, cudaMemcpyHostToDevice, h->stream));
    KRD_CUDA(h, cudaMemcpyAsync(d_init, init_centroids, (size_t)k * 4, cudaMemcpyHostToDevice, h->stream));
    std::vector<int32_t> a((size_t)N, 0);   // kmeans.hpp:303-305: assignments start at 0 when only centroids are given;
    // gappa runs several k on ONE Kmeans object, whose assignment vector survives from the previous k (kmeans.hpp:296-306)
    if (init_assignments) std::copy(init_assignments, init_assignments + N, a.begin());
    KRD_CUDA(h, cudaMemcpyAsync(d_assign, a.data(), b_n4, cudaMemcpyHostToDevice, h->stream));
    k_km_totals<<<(unsigned)((N + 255) / 256), 256, 0, h->stream>>>(h->d_F, h->ld, d_root, (int)root_seg.size(), N, d_total);
    if (exact) {
        KRD_CUDA(h, cudaMemsetAsync(d_Q, 0, 2 * al(b_mom), h->stream));   // Q and P are adjacent
        k_km_seg_geom<<<(unsigned)((h->E + 127) / 128), 128, 0, h->stream>>>(h->d_seg_off, h->d_len, h->E, seg_lo, seg_edge);
        if (h->M > 0) k_km_moments<<<(unsigned)((h->M + 255) / 256), 256, 0, h->stream>>>(h->d_sid, h->d_edge, h->d_pos, h->d_mass, h->M, h->ld, d_Q, d_P);
    } else {
        k_km_copy_cols<<<(unsigned)((h->S_pad * k + 255) / 256), 256, 0, h->stream>>>(h->d_F, h->ld, h->S_pad, d_init, N, k);
    }
    std::vector<double> total((size_t)N);
    KRD_CUDA(h, cudaMemcpyAsync(total.data(), d_total, b_n8, cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));

    const int64_t rpc = (h->S_pad + KM_CHUNKS - 1) / KM_CHUNKS;
    auto assign_pass = [&](int reassign) {
        // 32 centroids per launch (one pass over F each); the last launch takes the narrowest kernel that fits
        for (int c0 = 0; c0 < k; c0 += 32) {
            const int kk = std::min(32, k - c0);
            if (exact) {
                if (kk <= 4) launch_assign_exact<4>(h, k, c0, rpc, seg_lo, seg_edge, d_cent, part);
                else if (kk <= 8) launch_assign_exact<8>(h, k, c0, rpc, seg_lo, seg_edge, d_cent, part);
                else if (kk <= 16) launch_assign_exact<16>(h, k, c0, rpc, seg_lo, seg_edge, d_cent, part);
                else launch_assign_exact<32>(h, k, c0, rpc, seg_lo, seg_edge, d_cent, part);
            }
            else if (kk <= 4) launch_assign<4>(h, k, c0, rpc, part);
            else if (kk <= 8) launch_assign<8>(h, k, c0, rpc, part);
            else if (kk <= 16) launch_assign<16>(h, k, c0, rpc, part);
            else launch_assign<32>(h, k, c0, rpc, part);
        }
        k_km_assign_final<<<(unsigned)((N + 31) / 32), dim3(32, 32), 0, h->stream>>>(part, h->ld, N, k, reassign, d_assign, d_dist, d_changed);
    };
    static std::atomic<unsigned long long> configured{0};
    const size_t smem_row = (size_t)N * 8;
    const bool staged = smem_row <= 200 * 1024;      // beyond 25,600 samples the members are gathered from the row directly
    if (!exact && staged) KRD_CUDA(h, allow_dynamic_smem(k_km_update<true>, 200 * 1024, configured));

    int iteration = 0, changed = 0, launches = 2;
    std::vector<int32_t> mem((size_t)N), mem_off((size_t)k + 1);
    std::vector<double> ctot((size_t)k), dist((size_t)N);
    // update_centroids from the host copy of the assignments: member lists in sample order + the members'
    // total mass per centroid (host: N integers), then one pass over F.  A centroid without members becomes
    // all-zero (mass_tree_normalize_masses returns early on an empty tree, functions.cpp:126-132).
    auto push_members = [&]() -> krd_status {
        KRD_CUDA(h, cudaMemcpyAsync(d_mem, mem.data(), b_n4, cudaMemcpyHostToDevice, h->stream));
        KRD_CUDA(h, cudaMemcpyAsync(d_mem_off, mem_off.data(), (size_t)(k + 1) * 4, cudaMemcpyHostToDevice, h->stream));
        KRD_CUDA(h, cudaMemcpyAsync(d_ctot, ctot.data(), (size_t)k * 8, cudaMemcpyHostToDevice, h->stream));
        if (exact)
            k_km_update_exact<<<(unsigned)std::min(std::max(h->E, 1), 16 * h->num_sms), 128, 0, h->stream>>>(h->d_F, d_Q, d_P, h->d_seg_off, h->ld, h->E, k,
                                                                                                           d_mem_off, d_mem, d_ctot, d_cent);
        else if (staged)
            k_km_update<true><<<(unsigned)std::min<int64_t>(h->S_pad, 16 * h->num_sms), 256, smem_row, h->stream>>>(h->d_F, h->ld, h->S_pad, N, k, d_mem_off, d_mem, d_ctot);
        else
            k_km_update<false><<<(unsigned)std::min<int64_t>(h->S_pad, 16 * h->num_sms), 256, 0, h->stream>>>(h->d_F, h->ld, h->S_pad, N, k, d_mem_off, d_mem, d_ctot);
        KRD_CUDA(h, cudaGetLastError());
        return KRD_OK;
    };
    if (exact) {   // pre_loop_hook: the initial centroids are the start samples with their masses centred on the branches
        for (int c = 0; c < k; ++c) { mem[(size_t)c] = init_centroids[c]; mem_off[(size_t)c] = c; ctot[(size_t)c] = total[(size_t)init_centroids[c]]; }
        mem_off[(size_t)k] = k;
        if ((st = push_members()) != KRD_OK) return st;
        KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    }
    auto update = [&]() -> krd_status {
        std::fill(mem_off.begin(), mem_off.end(), 0);
        for (int d = 0; d < N; ++d) mem_off[(size_t)a[(size_t)d] + 1]++;
        for (int c = 0; c < k; ++c) mem_off[(size_t)c + 1] += mem_off[(size_t)c];
        std::vector<int32_t> fill(mem_off.begin(), mem_off.end() - 1);
        std::fill(ctot.begin(), ctot.end(), 0.0);
        for (int d = 0; d < N; ++d) { mem[(size_t)fill[(size_t)a[(size_t)d]]++] = d; ctot[(size_t)a[(size_t)d]] += total[(size_t)d]; }
        { krd_status ps = push_members(); if (ps != KRD_OK) return ps; }
        KRD_CUDA(h, cudaStreamSynchronize(h->stream));   // the host vectors are reused
        launches += 1;
        return KRD_OK;
    };
    do {
        // lloyd_step (kmeans.hpp:324-336): assign, then update
        KRD_CUDA(h, cudaMemsetAsync(d_changed, 0, sizeof(int), h->stream));
        assign_pass(1);
        KRD_CUDA(h, cudaMemcpyAsync(a.data(), d_assign, b_n4, cudaMemcpyDeviceToHost, h->stream));
        KRD_CUDA(h, cudaMemcpyAsync(&changed, d_changed, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        KRD_CUDA(h, cudaStreamSynchronize(h->stream));
        launches += 2;
        if ((st = update()) != KRD_OK) return st;

        // treat_empty_centroids (kmeans.hpp:426-500): give each empty centroid the sample that is furthest from
        // the centroid of the cluster with the largest variance.  The reference walks a std::unordered_set of
        // the empty indices; libstdc++ yields small sets of integers in reverse insertion order, so highest first.
        std::vector<int> empties;
        for (int c = k - 1; c >= 0; --c) if (mem_off[(size_t)c + 1] == mem_off[(size_t)c]) empties.push_back(c);
        bool treated = false;
        for (size_t ei = 0; ei < empties.size(); ++ei) {
            assign_pass(0);
            KRD_CUDA(h, cudaMemcpyAsync(dist.data(), d_dist, b_n8, cudaMemcpyDeviceToHost, h->stream));
            KRD_CUDA(h, cudaStreamSynchronize(h->stream));
            launches += 2;
            std::vector<double> var((size_t)k, 0.0); std::vector<int64_t> cnt((size_t)k, 0);
            for (int d = 0; d < N; ++d) { var[(size_t)a[(size_t)d]] += dist[(size_t)d] * dist[(size_t)d]; cnt[(size_t)a[(size_t)d]]++; }
            for (int c = 0; c < k; ++c) if (cnt[(size_t)c] > 0) var[(size_t)c] /= (double)cnt[(size_t)c];
            const int mv = (int)(std::max_element(var.begin(), var.end()) - var.begin());
            if (var[(size_t)mv] == 0.0) { treated = false; break; }   // the reference returns false here
            int far = -1; double far_d = -1.7976931348623157e308;
            for (int d = 0; d < N; ++d) if (a[(size_t)d] == mv && dist[(size_t)d] > far_d) { far = d; far_d = dist[(size_t)d]; }
            a[(size_t)far] = empties[ei];
            KRD_CUDA(h, cudaMemcpyAsync(d_assign, a.data(), b_n4, cudaMemcpyHostToDevice, h->stream));
            if ((st = update()) != KRD_OK) return st;
            treated = true;
        }
        if (treated) changed = 1;
        ++iteration;
    } while (changed && iteration < max_iterations);

    assign_pass(0);   // cluster_info: distance of every sample to its (final) centroid
    KRD_CUDA(h, cudaGetLastError());
    KRD_CUDA(h, cudaMemcpyAsync(distances, d_dist, b_n8, cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    std::copy(a.begin(), a.end(), assignments);
    if (iterations_out) *iterations_out = iteration;
    h->F_valid = false;      // the centroid columns are scratch
    h->stats.launches = launches + 2;
    return KRD_OK;
}

}  // namespace krd
