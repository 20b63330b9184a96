// K0 (mass -> segment slot, segment table) and K1 (dense cumulative subtree-mass matrix F).
//
// What the reference does once PER PAIR — a postorder walk that carries subtree mass through
// node_masses[] and sweeps every edge distal -> proximal (genesis tree/mass_tree/emd.cpp:89-167) —
// is done here once PER SAMPLE, level by level, and stored densely:
//     F[s][i] = mass of sample i strictly below segment s
//             = M_i(child_node(e)) + sum of sample i's masses on e at positions above segment s.
// The pair kernel (krd_pairs.cu) then only needs  sum_s len[s] * |F[s][i] - F[s][j]|^p.
#include <cub/cub.cuh>

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>

#include "krd_common.cuh"

namespace krd {

// ------------------------------------------------------------------------------------------------
//  Tree preprocessing (host): validate, levels, child-edge lists
// ------------------------------------------------------------------------------------------------

// Top-of-tree kernel shape, measured at cfg2 (K1 per build, `bench.py`): 32 columns x 256 threads (round 1) 0.404 ms;
// 32 x 1,024 threads 0.372; 16 x 1,024 (512 edges) 0.358; 8 x 1,024 (1,024 edges: two more levels leave the per-level
// launches) 0.354.  The kernel's time is (edges per thread group) x (latency of three dependent loads), not bandwidth.
#ifndef KRD_TOP_COLS
#define KRD_TOP_COLS 8
#endif
constexpr int TOP_COLS = KRD_TOP_COLS;       // samples per CTA of k_subtree_top
constexpr int TOP_MAX = 256 * 32 / TOP_COLS; // edges finished by k_subtree_top (2 x TOP_MAX x TOP_COLS doubles = 128 KB of shared memory)
#ifndef KRD_TOP_THREADS
#define KRD_TOP_THREADS 1024
#endif
constexpr int TOP_THREADS = KRD_TOP_THREADS;   // the kernel is a chain of dependent loads per (edge, column): its time is (edges per thread group) x latency —
                                              // 1,024 threads = 32 groups of 32 columns: 8 edges per group instead of 32 (46 -> ~15 us at cfg2)

krd_status build_tree(krd_handle* h)
{
    const int E = h->E, n_nodes = E + 1;
    if (h->root < 0 || h->root >= n_nodes) return fail(h, KRD_ERR_INVALID_ARGUMENT, "root_node out of range");
    std::vector<int> up(n_nodes, -1);        // edge above each node
    std::vector<int> kcnt(n_nodes + 1, 0);
    for (int e = 0; e < E; ++e) {
        int p = h->parent[e], c = h->child[e];
        if (p < 0 || p >= n_nodes || c < 0 || c >= n_nodes || p == c)
            return fail(h, KRD_ERR_INVALID_ARGUMENT, "edge endpoints out of range");
        if (up[c] != -1 || c == h->root)
            return fail(h, KRD_ERR_INCOMPATIBLE_TREES, "Incompatible MassTrees.");
        up[c] = e;
        kcnt[p + 1]++;
    }
    for (int n = 0; n < n_nodes; ++n) kcnt[n + 1] += kcnt[n];
    std::vector<int> kids(E > 0 ? E : 1), fill(kcnt.begin(), kcnt.end() - 1);
    for (int e = 0; e < E; ++e) kids[fill[h->parent[e]]++] = e;   // listing order = child order

    // BFS from the root gives a top-down node order; reversed it is bottom-up.
    std::vector<int> order; order.reserve(n_nodes);
    order.push_back(h->root);
    for (size_t q = 0; q < order.size(); ++q) {
        int n = order[q];
        for (int k = kcnt[n]; k < kcnt[n + 1]; ++k) order.push_back(h->child[kids[k]]);
    }
    if ((int)order.size() != n_nodes)  // unreachable nodes <=> cycle / forest
        return fail(h, KRD_ERR_INCOMPATIBLE_TREES, "Incompatible MassTrees.");
    std::vector<int> height(n_nodes, 0);
    int max_h = 0;
    for (int q = n_nodes - 1; q > 0; --q) {
        int n = order[q], pn = h->parent[up[n]];
        height[pn] = std::max(height[pn], height[n] + 1);
    }
    for (int e = 0; e < E; ++e) max_h = std::max(max_h, height[h->child[e]]);
    const int L = (E > 0) ? max_h + 1 : 0;
    h->lvl_off.assign(L + 1, 0);
    for (int e = 0; e < E; ++e) h->lvl_off[height[h->child[e]] + 1]++;
    for (int l = 0; l < L; ++l) h->lvl_off[l + 1] += h->lvl_off[l];
    std::vector<int> lvl_edges(E > 0 ? E : 1), lf(h->lvl_off.begin(), h->lvl_off.end() - (L > 0 ? 1 : 0));
    for (int e = 0; e < E; ++e) lvl_edges[lf[height[h->child[e]]]++] = e;

    std::vector<int> kid_off(E + 1, 0), kid_edge(E > 0 ? E : 1);
    for (int e = 0; e < E; ++e) {
        int c = h->child[e];
        kid_off[e + 1] = kid_off[e] + (kcnt[c + 1] - kcnt[c]);
        std::copy(kids.begin() + kcnt[c], kids.begin() + kcnt[c + 1], kid_edge.begin() + kid_off[e]);
    }

    cudaFree(h->d_lvl_edges); cudaFree(h->d_kid_off); cudaFree(h->d_kid_edge); cudaFree(h->d_bl);
    h->d_lvl_edges = h->d_kid_off = h->d_kid_edge = nullptr; h->d_bl = nullptr;
    size_t ne = (size_t)std::max(E, 1);
    KRD_CUDA(h, cudaMalloc(&h->d_lvl_edges, ne * sizeof(int)));
    KRD_CUDA(h, cudaMalloc(&h->d_kid_off, (ne + 1) * sizeof(int)));
    KRD_CUDA(h, cudaMalloc(&h->d_kid_edge, ne * sizeof(int)));
    KRD_CUDA(h, cudaMalloc(&h->d_bl, ne * sizeof(double)));
    KRD_CUDA(h, cudaMemcpyAsync(h->d_lvl_edges, lvl_edges.data(), (size_t)E * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    KRD_CUDA(h, cudaMemcpyAsync(h->d_kid_off, kid_off.data(), (size_t)(E + 1) * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    KRD_CUDA(h, cudaMemcpyAsync(h->d_kid_edge, kid_edge.data(), (size_t)E * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    KRD_CUDA(h, cudaMemcpyAsync(h->d_bl, h->bl.data(), (size_t)E * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    // ---- top of the tree: the trailing levels that together hold <= TOP_MAX edges ----
    {
        int lt = L;
        while (lt > 1 && E - h->lvl_off[lt - 1] <= TOP_MAX) --lt;
        h->top_level = lt;
        h->top_edges = (L > 0) ? E - h->lvl_off[lt] : 0;
        h->top_levels = L - lt;
        std::vector<int32_t> pack;
        if (h->top_edges > 0) {
            const int base = h->lvl_off[lt], Kt = h->top_edges;
            std::vector<int> top_index(E, -1);
            for (int j = 0; j < Kt; ++j) top_index[lvl_edges[base + j]] = j;
            std::vector<int32_t> koff(Kt + 1, 0), kids2;
            for (int j = 0; j < Kt; ++j) {
                const int e = lvl_edges[base + j];
                for (int k = kid_off[e]; k < kid_off[e + 1]; ++k) if (top_index[kid_edge[k]] < 0) kids2.push_back(kid_edge[k]);
                for (int k = kid_off[e]; k < kid_off[e + 1]; ++k) if (top_index[kid_edge[k]] >= 0) kids2.push_back(~top_index[kid_edge[k]]);
                koff[j + 1] = (int32_t)kids2.size();
            }
            h->top_kids = (int32_t)kids2.size();
            for (int j = 0; j < Kt; ++j) pack.push_back(lvl_edges[base + j]);
            pack.insert(pack.end(), koff.begin(), koff.end());
            pack.insert(pack.end(), kids2.begin(), kids2.end());
            for (int l = lt; l <= L; ++l) pack.push_back(h->lvl_off[l] - base);
        }
        cudaFree(h->d_top); h->d_top = nullptr;
        KRD_CUDA(h, cudaMalloc(&h->d_top, std::max<size_t>(pack.size(), 1) * sizeof(int32_t)));
        KRD_CUDA(h, cudaMemcpyAsync(h->d_top, pack.data(), pack.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
        KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    }
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));  // host vectors above go out of scope
    h->stats.levels = L;
    return KRD_OK;
}


// ------------------------------------------------------------------------------------------------
//  K0: canonical order, slots, segment table
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ int sample_of(const int64_t* __restrict__ off, int n, int64_t m)
{
    int lo = 0, hi = n;  // largest i with off[i] <= m
    while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (off[mid] <= m) lo = mid; else hi = mid; }
    return lo;
}

__global__ void k_sample_ids(const int64_t* __restrict__ off, int n, int64_t M, uint32_t* __restrict__ sid)
{
    int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m < M) sid[m] = (uint32_t)sample_of(off, n, m);
}

// flags[0] |= 1 if some sample is not sorted by (edge asc, pos asc); |= 2 if an edge id is invalid
__global__ void k_check_canonical(const uint32_t* __restrict__ sid, const int32_t* __restrict__ edge,
                                  const double* __restrict__ pos, int64_t M, int E, int* __restrict__ flags)
{
    int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    int e = edge[m];
    if (e < 0 || e >= E) { atomicOr(flags, 2); return; }
    if (m > 0 && sid[m - 1] == sid[m]) {
        int ep = edge[m - 1];
        if (ep > e || (ep == e && !(pos[m - 1] <= pos[m]))) atomicOr(flags, 1);
    }
}

// Deferred check (children of a pipelined ingest, whose entries are the parent's OWN device copy): an edge id outside
// the tree is recorded (flags |= 2) and replaced by 0 so that everything downstream stays inside its buffers; the host
// reads the flags once, when the whole pipeline has been issued, and fails the call then.
__global__ void k_clamp_edges(int32_t* __restrict__ edge, int64_t M, int E, int* __restrict__ flags)
{
    int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    const int e = edge[m];
    if (e < 0 || e >= E) { atomicOr(flags, 2); edge[m] = 0; }
}

// Sample offsets of a pipelined child = the parent's (already on the device), rebased to the child's first entry.  (A host
// copy here would queue behind the cohort's bulk upload on the copy engine: measured 0.3 ms per child at cfg2.)
__global__ void k_rebase_off(const int64_t* __restrict__ src, int n, int64_t* __restrict__ dst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i] - src[0];
}
void launch_rebase_off(const int64_t* d_src, int n, int64_t* d_dst, cudaStream_t stream)
{
    k_rebase_off<<<(n + 255) / 256, 256, 0, stream>>>(d_src, n, d_dst);
}

__global__ void k_iota(uint32_t* a, int64_t M)
{
    int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m < M) a[m] = (uint32_t)m;
}

template <typename T, typename U>
__global__ void k_gather(const T* __restrict__ src, const uint32_t* __restrict__ idx, int64_t M, U* __restrict__ dst)
{
    int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m < M) dst[m] = (U)src[idx[m]];
}

// mass_tree_binify_masses' bin index (genesis tree/mass_tree/functions.cpp:231-237), same
// operation order: pn = min(max(0, pos / bl * nb), nextafter(nb, 0)); q = floor(pn).
// std::max(0.0, v) returns 0.0 unless 0.0 < v (so NaN -> 0.0); std::min(pn, hi) returns pn unless hi < pn.
__global__ void k_slots_binned(const int32_t* __restrict__ edge, const double* __restrict__ pos,
                               const double* __restrict__ bl, int64_t M, double nb, double nb_below,
                               int32_t* __restrict__ slot)
{
    int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    double v = __dmul_rn(__ddiv_rn(pos[m], bl[edge[m]]), nb);
    double pn = (0.0 < v) ? v : 0.0;
    pn = (nb_below < pn) ? nb_below : pn;
    slot[m] = (int32_t)floor(pn);
}

// Compact binned input (krd_set_masses_binned): cell = edge * B + bin.  One pass decodes edge / slot / sample id and checks
// the range and the order (ascending cell within a sample = MassTree order of the binified tree); flags as in
// k_check_canonical.  `clamp`: deferred check of a pipelined child — a bad edge is recorded and replaced by edge 0.
template <typename CT>
__global__ void k_decode_binned(const int64_t* __restrict__ off, int n, const CT* __restrict__ cell, int64_t M, uint32_t B, int E,
                                int clamp, uint32_t* __restrict__ sid, int32_t* __restrict__ edge, int32_t* __restrict__ slot,
                                int* __restrict__ flags)
{
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    const int i = sample_of(off, n, m);
    const uint32_t c = cell[m];
    uint32_t e = c / B;
    const uint32_t q = c - e * B;
    if (e >= (uint32_t)E) { atomicOr(flags, 2); if (!clamp) return; e = 0; }
    if (m > off[i] && (uint32_t)cell[m - 1] > c) atomicOr(flags, 1);
    sid[m] = (uint32_t)i; edge[m] = (int32_t)e; slot[m] = (int32_t)q;
}

// Segment table of the binned mode: edge e owns segments e*(B+1)+t, t = 0..B, bounded by the bin
// centres c_q = q*bl/nb + bl/nb/2 (functions.cpp:240): [0,c_0], (c_0,c_1], ..., (c_{B-1}, bl].
__global__ void k_len_binned(const double* __restrict__ bl, int E, int B, double* __restrict__ len,
                             int32_t* __restrict__ seg_off)
{
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g > (int64_t)E * (B + 1)) return;
    if (g <= E) seg_off[g] = (int32_t)(g * (B + 1));
    if (g == (int64_t)E * (B + 1)) return;
    int e = (int)(g / (B + 1)), t = (int)(g % (B + 1));
    double b = bl[e], nb = (double)B;
    double hi = (t == B) ? b : __dadd_rn(__ddiv_rn(__dmul_rn((double)t, b), nb), __ddiv_rn(__ddiv_rn(b, nb), 2.0));
    double lo = (t == 0) ? 0.0 : __dadd_rn(__ddiv_rn(__dmul_rn((double)(t - 1), b), nb), __ddiv_rn(__ddiv_rn(b, nb), 2.0));
    len[g] = hi - lo;
}

// ---- exact mode -------------------------------------------------------------------------------

__global__ void k_heads(const uint32_t* __restrict__ sedge, const uint32_t* __restrict__ perm,
                        const double* __restrict__ pos, int64_t M, uint32_t* __restrict__ head)
{
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= M) return;
    uint32_t hd = 1;
    if (k > 0 && sedge[k] == sedge[k - 1] && pos[perm[k]] == pos[perm[k - 1]]) hd = 0;
    head[k] = hd;
}

__global__ void k_unique_scatter(const uint32_t* __restrict__ sedge, const uint32_t* __restrict__ perm,
                                 const double* __restrict__ pos, const uint32_t* __restrict__ head,
                                 const uint32_t* __restrict__ rank, int64_t M,
                                 uint32_t* __restrict__ uidx, double* __restrict__ upos,
                                 uint32_t* __restrict__ uedge, uint32_t* __restrict__ cnt)
{
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= M) return;
    uint32_t u = rank[k] - 1, m = perm[k];
    uidx[m] = u;
    if (head[k]) { upos[u] = pos[m]; uedge[u] = sedge[k]; atomicAdd(&cnt[sedge[k]], 1u); }
}

__global__ void k_exact_slots(const int32_t* __restrict__ edge, const uint32_t* __restrict__ uidx,
                              const uint32_t* __restrict__ ubase, int64_t M, int32_t* __restrict__ slot)
{
    int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m < M) slot[m] = (int32_t)(uidx[m] - ubase[edge[m]]);
}

// Segment t of edge e (k_e distinct positions b_1<..<b_k): t<k: (b_t, b_{t+1}] (b_0 = 0); t=k: (b_k, bl].
// Lengths are the same differences genesis forms while walking (emd.cpp:154-166).
__global__ void k_exact_len_unique(const double* __restrict__ upos, const uint32_t* __restrict__ uedge,
                                   const uint32_t* __restrict__ ubase, int64_t U, double* __restrict__ len, int* __restrict__ neg)
{
    int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= U) return;
    uint32_t e = uedge[u], r = (uint32_t)u - ubase[e];
    double lo = (r > 0) ? upos[u - 1] : 0.0;
    const double l = upos[u] - lo;
    len[(int64_t)ubase[e] + e + r] = l;
    if (l < 0.0) *neg = 1;     // a position below 0: the reference keeps the signed length (emd.cpp:156)
}

__global__ void k_exact_len_last(const double* __restrict__ upos, const uint32_t* __restrict__ ubase,
                                 const double* __restrict__ bl, int E, double* __restrict__ len,
                                 int32_t* __restrict__ seg_off, int* __restrict__ neg)
{
    int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e > E) return;
    seg_off[e] = (int32_t)(ubase[e] + e);
    if (e == E) return;
    uint32_t k = ubase[e + 1] - ubase[e];
    const double l = (k > 0) ? bl[e] - upos[ubase[e] + k - 1] : bl[e];
    len[(int64_t)ubase[e] + e + k] = l;
    if (l < 0.0) *neg = 1;     // a position beyond the branch (or a negative branch length): emd.cpp:166 keeps the sign
}

__global__ void k_sqrt(const double* __restrict__ a, int64_t n, double* __restrict__ b)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) b[i] = sqrt(a[i]);
}

static inline unsigned nblk(int64_t n, int bs = 256) { return (unsigned)((n + bs - 1) / bs); }

// Zero fill by a KERNEL (rows x width 8-byte words, row pitch in words).  cudaMemset*Async may run on a copy engine, where
// it queues behind the bulk upload of a pipelined ingest (measured: +0.25 ms per child at cfg2).
__global__ void k_zero_words(unsigned long long* __restrict__ p, int64_t pitch, int64_t width, int64_t rows)
{
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g < width * rows) p[(g / width) * pitch + g % width] = 0ull;
}
static inline void zero_words(void* p, int64_t pitch, int64_t width, int64_t rows, cudaStream_t st)
{
    if (width * rows > 0) k_zero_words<<<nblk(width * rows), 256, 0, st>>>(static_cast<unsigned long long*>(p), pitch, width, rows);
}

struct DevBuf {  // scoped device scratch
    void* p = nullptr;
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 16); }
    ~DevBuf() { if (p) cudaFree(p); }
    template <typename T> T* as() { return static_cast<T*>(p); }
};

static int bits_for(uint32_t max_value) { int b = 1; while (b < 32 && (max_value >> b)) ++b; return b; }

// stable sort of `perm` by 32-bit keys gathered through perm: perm_out[k] ordered by key[perm[..]]
template <typename SrcT>
static krd_status stable_pass_u32(krd_handle* h, const SrcT* d_src, uint32_t max_key, int64_t M,
                                  uint32_t* perm_in, uint32_t* perm_out, uint32_t* keys_sorted_out)
{
    DevBuf kin, tmp;
    KRD_CUDA(h, kin.alloc((size_t)M * 4));
    k_gather<SrcT, uint32_t><<<nblk(M), 256, 0, h->stream>>>(d_src, perm_in, M, kin.as<uint32_t>());
    size_t bytes = 0;
    KRD_CUDA(h, cub::DeviceRadixSort::SortPairs(nullptr, bytes, kin.as<uint32_t>(), keys_sorted_out, perm_in, perm_out, M, 0, bits_for(max_key), h->stream));
    KRD_CUDA(h, tmp.alloc(bytes));
    KRD_CUDA(h, cub::DeviceRadixSort::SortPairs(tmp.p, bytes, kin.as<uint32_t>(), keys_sorted_out, perm_in, perm_out, M, 0, bits_for(max_key), h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    return KRD_OK;
}

// perm sorted by (edge asc, pos asc), stable in the input index; sedge = edge keys in that order
static krd_status sort_edge_pos(krd_handle* h, uint32_t* perm /*M*/, uint32_t* sedge /*M*/)
{
    const int64_t M = h->M;
    DevBuf iota, pos_sorted, p1, tmp;
    KRD_CUDA(h, iota.alloc((size_t)M * 4));
    KRD_CUDA(h, pos_sorted.alloc((size_t)M * 8));
    KRD_CUDA(h, p1.alloc((size_t)M * 4));
    k_iota<<<nblk(M), 256, 0, h->stream>>>(iota.as<uint32_t>(), M);
    size_t bytes = 0;
    KRD_CUDA(h, cub::DeviceRadixSort::SortPairs(nullptr, bytes, h->d_pos, pos_sorted.as<double>(), iota.as<uint32_t>(), p1.as<uint32_t>(), M, 0, 64, h->stream));
    KRD_CUDA(h, tmp.alloc(bytes));
    KRD_CUDA(h, cub::DeviceRadixSort::SortPairs(tmp.p, bytes, h->d_pos, pos_sorted.as<double>(), iota.as<uint32_t>(), p1.as<uint32_t>(), M, 0, 64, h->stream));
    return stable_pass_u32<int32_t>(h, h->d_edge, (uint32_t)std::max(h->E - 1, 1), M, p1.as<uint32_t>(), perm, sedge);
}

static krd_status prepare_tiled(krd_handle* h);
static krd_status prepare_exact(krd_handle* h);

// Segment table of binned mode without any mass data (the parent of a pipelined ingest needs len / seg_off for the pair
// kernel while its children hold the masses): same kernels as prepare_masses.
krd_status prepare_len_binned(krd_handle* h)
{
    const int E = h->E, B = h->cfg.bins;
    if (B <= 0) return fail(h, KRD_ERR_STATE, "prepare_len_binned needs bins > 0");
    if ((int64_t)E * (B + 1) > 0x7FFFFFF0) return fail(h, KRD_ERR_INVALID_ARGUMENT, "too many segments");
    h->S = (int64_t)E * (B + 1);
    h->S_pad = ((h->S + KC - 1) / KC) * KC;
    if (h->S_pad == 0) h->S_pad = KC;
    krd_status rs;
    if ((rs = reserve(h, (void**)&h->d_seg_off, (size_t)(E + 1) * 4)) != KRD_OK) return rs;
    if ((rs = reserve(h, (void**)&h->d_len, (size_t)h->S_pad * 8)) != KRD_OK) return rs;
    if ((rs = reserve(h, (void**)&h->d_sqrt_len, (size_t)h->S_pad * 8)) != KRD_OK) return rs;
    zero_words(h->d_len, h->S_pad, h->S_pad, 1, h->stream);
    k_len_binned<<<nblk((int64_t)E * (B + 1) + 1), 256, 0, h->stream>>>(h->d_bl, E, B, h->d_len, h->d_seg_off);
    k_sqrt<<<nblk(h->S_pad), 256, 0, h->stream>>>(h->d_len, h->S_pad, h->d_sqrt_len);
    KRD_CUDA(h, cudaGetLastError());
    return KRD_OK;
}


krd_status prepare_masses(krd_handle* h)
{
    const int64_t M = h->M;
    const int E = h->E, N = h->N;
    h->F_valid = false;
    if (M >= (int64_t)0xFFFFFFF0u) return fail(h, KRD_ERR_INVALID_ARGUMENT, "more than 2^32 masses are not supported");

    // --- canonical order check (edge asc, pos asc per sample = MassTree iteration order) ---
    // (the flag word lives in the handle: a scoped cudaMalloc/cudaFree here would synchronise the whole device on every
    //  call, including the copy stream of a pipelined ingest)
    { krd_status rs = reserve(h, &h->d_flags, 16); if (rs != KRD_OK) return rs; }
    int* const d_flags = static_cast<int*>(h->d_flags);
    { krd_status rs = reserve(h, (void**)&h->d_sid, (size_t)std::max<int64_t>(M, 1) * 4); if (rs != KRD_OK) return rs; }
    zero_words(d_flags, 2, 2, 1, h->stream);   // (16 bytes reserved: int [0] order / range check, int [1] negative segment length)
    if (h->binned_input) {
        if (h->cfg.bins < 1) return fail(h, KRD_ERR_STATE, "binned input without bins");
        krd_status rs;
        if ((rs = reserve(h, (void**)&h->d_slot, (size_t)std::max<int64_t>(M, 1) * 4)) != KRD_OK) return rs;
        if ((rs = reserve(h, (void**)&h->d_edge_dec, (size_t)std::max<int64_t>(M, 1) * 4)) != KRD_OK) return rs;
        h->d_edge = h->d_edge_dec;
        if (M > 0 && h->cell_bytes == 2)
            k_decode_binned<uint16_t><<<nblk(M), 256, 0, h->stream>>>(h->d_off, N, static_cast<const uint16_t*>(h->d_cell32), M, (uint32_t)h->cfg.bins, E,
                                                                      h->defer_check ? 1 : 0, h->d_sid, h->d_edge, h->d_slot, d_flags);
        else if (M > 0)
            k_decode_binned<uint32_t><<<nblk(M), 256, 0, h->stream>>>(h->d_off, N, static_cast<const uint32_t*>(h->d_cell32), M, (uint32_t)h->cfg.bins, E,
                                                                      h->defer_check ? 1 : 0, h->d_sid, h->d_edge, h->d_slot, d_flags);
    } else if (M > 0) {
        k_sample_ids<<<nblk(M), 256, 0, h->stream>>>(h->d_off, N, M, h->d_sid);
        if (h->defer_check) k_clamp_edges<<<nblk(M), 256, 0, h->stream>>>(h->d_edge, M, E, d_flags);
        k_check_canonical<<<nblk(M), 256, 0, h->stream>>>(h->d_sid, h->d_edge, h->d_pos, M, E, d_flags);
    }
    int hflags = 0;
    if (!h->defer_check) {   // (deferred: no host round trip here; krd_pipe.cu reads d_flags after the last phase is issued)
        KRD_CUDA(h, cudaMemcpyAsync(&hflags, d_flags, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    }
    if (hflags & 2) return fail(h, KRD_ERR_INVALID_ARGUMENT, "mass entry refers to an edge outside the tree");
    if ((hflags & 1) && h->binned_input)
        return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_set_masses_binned: the entries of a sample must be sorted by cell (MassTree order)");
    if (hflags & 1) {
        // sort by (sample, edge, pos), stable: three LSD passes, then gather into owned buffers
        DevBuf perm2, sedge, perm3, skeys;
        KRD_CUDA(h, perm2.alloc((size_t)M * 4)); KRD_CUDA(h, sedge.alloc((size_t)M * 4));
        KRD_CUDA(h, perm3.alloc((size_t)M * 4)); KRD_CUDA(h, skeys.alloc((size_t)M * 4));
        krd_status st = sort_edge_pos(h, perm2.as<uint32_t>(), sedge.as<uint32_t>());
        if (st != KRD_OK) return st;
        st = stable_pass_u32<uint32_t>(h, h->d_sid, (uint32_t)std::max(N - 1, 1), M, perm2.as<uint32_t>(), perm3.as<uint32_t>(), skeys.as<uint32_t>());
        if (st != KRD_OK) return st;
        int32_t* ne; double *np, *nm;
        KRD_CUDA(h, cudaMalloc(&ne, (size_t)M * 4)); KRD_CUDA(h, cudaMalloc(&np, (size_t)M * 8)); KRD_CUDA(h, cudaMalloc(&nm, (size_t)M * 8));
        k_gather<int32_t, int32_t><<<nblk(M), 256, 0, h->stream>>>(h->d_edge, perm3.as<uint32_t>(), M, ne);
        k_gather<double, double><<<nblk(M), 256, 0, h->stream>>>(h->d_pos, perm3.as<uint32_t>(), M, np);
        k_gather<double, double><<<nblk(M), 256, 0, h->stream>>>(h->d_mass, perm3.as<uint32_t>(), M, nm);
        KRD_CUDA(h, cudaStreamSynchronize(h->stream));
        for (void*& q : h->d_owned) { if (q) cudaFree(q); q = nullptr; }
        h->d_owned[0] = ne; h->d_owned[1] = np; h->d_owned[2] = nm;
        h->caps[&h->d_owned[0]] = (size_t)M * 4; h->caps[&h->d_owned[1]] = (size_t)M * 8; h->caps[&h->d_owned[2]] = (size_t)M * 8;
        h->d_edge = ne; h->d_pos = np; h->d_mass = nm;
    }

    { krd_status rs = reserve(h, (void**)&h->d_slot, (size_t)std::max<int64_t>(M, 1) * 4); if (rs != KRD_OK) return rs; }
    { krd_status rs = reserve(h, (void**)&h->d_seg_off, (size_t)(E + 1) * 4); if (rs != KRD_OK) return rs; }

    auto alloc_len = [&](int64_t S) -> krd_status {
        h->S = S;
        h->S_pad = ((S + KC - 1) / KC) * KC;
        if (h->S_pad == 0) h->S_pad = KC;
        krd_status rs = reserve(h, (void**)&h->d_len, (size_t)h->S_pad * 8);
        if (rs != KRD_OK) return rs;
        if ((rs = reserve(h, (void**)&h->d_sqrt_len, (size_t)h->S_pad * 8)) != KRD_OK) return rs;
        zero_words(h->d_len, h->S_pad, h->S_pad, 1, h->stream);
        return KRD_OK;
    };

    h->len_negative = false;
    if (h->cfg.bins > 0) {
        const int B = h->cfg.bins;
        if ((int64_t)E * (B + 1) > 0x7FFFFFF0) return fail(h, KRD_ERR_INVALID_ARGUMENT, "too many segments");
        for (double b : h->bl) if (b < 0.0) h->len_negative = true;
        krd_status st = alloc_len((int64_t)E * (B + 1));
        if (st != KRD_OK) return st;
        double nb = (double)B;
        if (M > 0 && !h->binned_input) k_slots_binned<<<nblk(M), 256, 0, h->stream>>>(h->d_edge, h->d_pos, h->d_bl, M, nb, std::nextafter(nb, 0.0), h->d_slot);
        k_len_binned<<<nblk((int64_t)E * (B + 1) + 1), 256, 0, h->stream>>>(h->d_bl, E, B, h->d_len, h->d_seg_off);
    } else {
        DevBuf perm, sedge, head, rank, uidx, upos, uedge, cnt, ubase, tmp;
        const int64_t Mx = std::max<int64_t>(M, 1);
        KRD_CUDA(h, perm.alloc(Mx * 4)); KRD_CUDA(h, sedge.alloc(Mx * 4)); KRD_CUDA(h, head.alloc(Mx * 4));
        KRD_CUDA(h, rank.alloc(Mx * 4)); KRD_CUDA(h, uidx.alloc(Mx * 4)); KRD_CUDA(h, upos.alloc(Mx * 8));
        KRD_CUDA(h, uedge.alloc(Mx * 4)); KRD_CUDA(h, cnt.alloc((size_t)(E + 1) * 4)); KRD_CUDA(h, ubase.alloc((size_t)(E + 1) * 4));
        KRD_CUDA(h, cudaMemsetAsync(cnt.p, 0, (size_t)(E + 1) * 4, h->stream));
        uint32_t U = 0;
        if (M > 0) {
            krd_status st = sort_edge_pos(h, perm.as<uint32_t>(), sedge.as<uint32_t>());
            if (st != KRD_OK) return st;
            k_heads<<<nblk(M), 256, 0, h->stream>>>(sedge.as<uint32_t>(), perm.as<uint32_t>(), h->d_pos, M, head.as<uint32_t>());
            size_t bytes = 0;
            KRD_CUDA(h, cub::DeviceScan::InclusiveSum(nullptr, bytes, head.as<uint32_t>(), rank.as<uint32_t>(), M, h->stream));
            KRD_CUDA(h, tmp.alloc(bytes));
            KRD_CUDA(h, cub::DeviceScan::InclusiveSum(tmp.p, bytes, head.as<uint32_t>(), rank.as<uint32_t>(), M, h->stream));
            k_unique_scatter<<<nblk(M), 256, 0, h->stream>>>(sedge.as<uint32_t>(), perm.as<uint32_t>(), h->d_pos, head.as<uint32_t>(), rank.as<uint32_t>(), M,
                                                             uidx.as<uint32_t>(), upos.as<double>(), uedge.as<uint32_t>(), cnt.as<uint32_t>());
            KRD_CUDA(h, cudaMemcpyAsync(&U, rank.as<uint32_t>() + (M - 1), 4, cudaMemcpyDeviceToHost, h->stream));
        }
        {
            DevBuf tmp2; size_t bytes = 0;
            KRD_CUDA(h, cub::DeviceScan::ExclusiveSum(nullptr, bytes, cnt.as<uint32_t>(), ubase.as<uint32_t>(), E + 1, h->stream));
            KRD_CUDA(h, tmp2.alloc(bytes));
            KRD_CUDA(h, cub::DeviceScan::ExclusiveSum(tmp2.p, bytes, cnt.as<uint32_t>(), ubase.as<uint32_t>(), E + 1, h->stream));
            KRD_CUDA(h, cudaStreamSynchronize(h->stream));
        }
        if ((int64_t)U + E > 0x7FFFFFF0) return fail(h, KRD_ERR_INVALID_ARGUMENT, "too many segments");
        krd_status st = alloc_len((int64_t)U + E);
        if (st != KRD_OK) return st;
        if (M > 0) {
            k_exact_slots<<<nblk(M), 256, 0, h->stream>>>(h->d_edge, uidx.as<uint32_t>(), ubase.as<uint32_t>(), M, h->d_slot);
            k_exact_len_unique<<<nblk(U), 256, 0, h->stream>>>(upos.as<double>(), uedge.as<uint32_t>(), ubase.as<uint32_t>(), U, h->d_len, d_flags + 1);
        }
        k_exact_len_last<<<nblk(E + 1), 256, 0, h->stream>>>(upos.as<double>(), ubase.as<uint32_t>(), h->d_bl, E, h->d_len, h->d_seg_off, d_flags + 1);
        int neg = 0;
        KRD_CUDA(h, cudaMemcpyAsync(&neg, d_flags + 1, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        KRD_CUDA(h, cudaStreamSynchronize(h->stream));  // scratch buffers die here
        h->len_negative = neg != 0;
    }
    k_sqrt<<<nblk(h->S_pad), 256, 0, h->stream>>>(h->d_len, h->S_pad, h->d_sqrt_len);
    KRD_CUDA(h, cudaGetLastError());
    if (!h->ld_fixed) KRD_CUDA(h, cudaStreamSynchronize(h->stream));   // (a child of a pipelined ingest keeps the stream running)

    if (!h->ld_fixed) { h->ld = ((N + h->extra_cols + TILE - 1) / TILE) * TILE; h->ncols = h->ld; }
    else h->ncols = ((N + TILE - 1) / TILE) * TILE;     // child of a pipelined ingest: own columns inside the parent's F
    h->stats.n_segments = h->S;
    h->stats.ld_samples = h->ld;
    h->stats.f_bytes = (int64_t)h->S_pad * h->ld * 8;
    h->exact_fast = false; h->streamed = false; h->cum_valid = false;
    if (h->cfg.builder == 1) { h->tiled = false; return KRD_OK; }   // forced legacy builder (A/B tests)
    if (h->cfg.bins == 0) { h->tiled = false; return prepare_exact(h); }
    return prepare_tiled(h);
}


// ------------------------------------------------------------------------------------------------
//  K1: scatter + level-synchronous postorder scan -> F[S_pad][ld]
// ------------------------------------------------------------------------------------------------
//
// Phase A (k_scatter, one thread per mass entry, coalesced reads): entries of one sample that fall
// on the same segment slot are adjacent (canonical order), so the first entry of each run sums the
// run and stores it — one writer per cell, no atomics, bit-reproducible.
// Phase B (k_scan_level, one thread per (edge of the level, sample), all traffic coalesced over
// the sample axis): levels are heights of the edge's child node, so every child edge was finished
// by an earlier launch.  The thread starts from the subtree mass below the edge
//     sum over child edges c of F[seg_off[c]][i]        (emd.cpp:145,167 without node_masses[])
// and turns the per-slot masses into the running distal -> proximal sum (emd.cpp:149-166).
__global__ void __launch_bounds__(256)
k_scatter(const uint32_t* __restrict__ sid, const int64_t* __restrict__ off, const int32_t* __restrict__ edge,
          const int32_t* __restrict__ slot, const double* __restrict__ mass, const int32_t* __restrict__ seg_off,
          int64_t M, int64_t ld, double* __restrict__ F)
{
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    const uint32_t i = sid[m];
    const int e = edge[m], s = slot[m];
    if (m > off[i] && edge[m - 1] == e && slot[m - 1] == s) return;   // not the head of its run
    const int64_t end = off[i + 1];
    double sum = mass[m];
    for (int64_t q = m + 1; q < end && edge[q] == e && slot[q] == s; ++q) sum += mass[q];
    F[(int64_t)(seg_off[e] + s) * ld + i] = sum;
}

__global__ void __launch_bounds__(128)
k_scan_level(const int32_t* __restrict__ lvl_edges, int lvl_begin, int lvl_count,
             const int32_t* __restrict__ kid_off, const int32_t* __restrict__ kid_edge,
             const int32_t* __restrict__ seg_off, int64_t ld, double* __restrict__ F)
{
    const int64_t i = (int64_t)blockIdx.y * blockDim.x + threadIdx.x;
    for (int le = blockIdx.x; le < lvl_count; le += gridDim.x) {
        const int e = lvl_edges[lvl_begin + le];
        const int k0 = kid_off[e], k1 = kid_off[e + 1];
        const int s0 = seg_off[e], s1 = seg_off[e + 1];
        double acc = 0.0;
        for (int k = k0; k < k1; ++k) acc += F[(int64_t)seg_off[kid_edge[k]] * ld + i];
        double* col = F + i;
        int t = s1 - 1;
        for (; t - 3 >= s0; t -= 4) {   // four loads in flight before the dependent adds
            const double v0 = col[(int64_t)t * ld], v1 = col[(int64_t)(t - 1) * ld];
            const double v2 = col[(int64_t)(t - 2) * ld], v3 = col[(int64_t)(t - 3) * ld];
            acc += v0; col[(int64_t)t * ld] = acc;
            acc += v1; col[(int64_t)(t - 1) * ld] = acc;
            acc += v2; col[(int64_t)(t - 2) * ld] = acc;
            acc += v3; col[(int64_t)(t - 3) * ld] = acc;
        }
        for (; t >= s0; --t) { acc += col[(int64_t)t * ld]; col[(int64_t)t * ld] = acc; }
    }
}

// ------------------------------------------------------------------------------------------------
//  K1 tiled (binned mode): F is written exactly once, straight from the CSR
// ------------------------------------------------------------------------------------------------
//
// Every edge owns B+1 consecutive segments, so the (edge block) x (64 samples) rectangle of F is a
// shared-memory tile.  Three steps, all HBM traffic coalesced:
//   k_tile<TOTALS>   T[e][i] = mass of sample i on edge e (warp per sample reads its contiguous CSR
//                    slice of the edge block; one writer per cell, sums in fixed order)
//   k_subtree_level  M[e][i] = T[e][i] + sum_kids M[kid][i], leaves first — the postorder
//                    accumulation of emd.cpp:145,167 on the small E x N matrix only
//   k_tile<FILL>     slot masses -> tile, running distal -> proximal sum started from the mass
//                    below the edge (emd.cpp:141-166), tile rows streamed to F.
// Algorithmic bytes: 8*S*ld (F, written once) + 2 * 12*M (CSR read twice) + ~5 * 8*E*ld (T/M).
constexpr int FT = 64;           // samples per tile
#ifndef KRD_FT_ROWS_BINNED
#define KRD_FT_ROWS_BINNED 48
#endif
#ifndef KRD_FT_MAX_ROWS
#define KRD_FT_MAX_ROWS 96
#endif
constexpr int FT_MAX_ROWS = KRD_FT_MAX_ROWS;  // tile rows (segments) -> 48 KB of shared memory, 4 CTAs per SM

// blk_ptr[b][i] = first entry of sample i whose edge is >= b*EB (b = n_eblocks: end of the sample)
__global__ void k_block_ptr(const int64_t* __restrict__ off, const int32_t* __restrict__ edge, int N, int n_eblocks, int EB,
                            int64_t* __restrict__ blk_ptr)
{
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= (int64_t)(n_eblocks + 1) * N) return;
    const int b = (int)(g / N), i = (int)(g % N);
    int64_t lo = off[i], hi = off[i + 1];
    const int e0 = b * EB;
    while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (edge[mid] < e0) lo = mid + 1; else hi = mid; }
    blk_ptr[g] = lo;
}

// `guard` (deferred check only, else nullptr): a non-zero flag word means the entries are not in canonical order or were
// clamped — the block pointers are then meaningless, so every slice is made empty (k_tile sees no entries) and
// k_tile_order does nothing; the host discards the results when it reads the flags.
__global__ void k_slice_len(const int64_t* __restrict__ blk_ptr, int64_t n, int N, int64_t* __restrict__ len, const int* __restrict__ guard)
{
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (guard && *guard) { if (g <= n) len[g] = 0; return; }
    if (g < n) len[g] = blk_ptr[g + N] - blk_ptr[g];
    else if (g == n) len[g] = 0;
}

// Tile-major copy of the CSR payload: entries ordered by (edge block, sample, segment), so that the
// entries of one (edge block) x (64 samples) tile are ONE contiguous range.  Per entry only what K1
// needs: where it lands in its tile, 16 bits (edge-in-block << 11 | slot << 6 | sample-in-tile), and its mass.
__global__ void k_tile_order(const uint32_t* __restrict__ sid, const int32_t* __restrict__ edge, const int32_t* __restrict__ slot,
                             const double* __restrict__ mass, const int64_t* __restrict__ blk_ptr,
                             const int64_t* __restrict__ tile_ptr, int64_t M, int N, int EB, int B1,
                             uint16_t* __restrict__ cell, double* __restrict__ tmass, const int* __restrict__ guard)
{
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M || (guard && *guard)) return;
    const int i = (int)sid[m], e = edge[m], b = e / EB;
    const int64_t g = (int64_t)b * N + i;
    const int64_t dst = tile_ptr[g] + (m - blk_ptr[g]);
    cell[dst] = (uint16_t)(((e - b * EB) << 11) | (slot[m] << 6) | (i & (FT - 1)));
    tmass[dst] = mass[m];
}

constexpr int FTS = FT + 1;   // shared-memory row stride (doubles): odd, so that one sample's rows spread over the banks

template <bool FILL>
__global__ void __launch_bounds__(256)
k_tile(const uint16_t* __restrict__ cell, const double* __restrict__ tmass, const int64_t* __restrict__ tile_ptr,
       int N, int64_t ld, int64_t ldm, int E, int EB, int B1, const double* __restrict__ Dsub, double* __restrict__ Msub,
       double* __restrict__ F, const int32_t* __restrict__ kid_off, int64_t zero_row_off)
{
    extern __shared__ double tile[];   // [rows][FTS]
    const int b = blockIdx.x, i0 = blockIdx.y * FT;
    const int e0 = b * EB, ne = min(EB, E - e0);
    const int rows = FILL ? ne * B1 : ne;
    const int base = FILL ? e0 * B1 : e0;      // first global segment / edge of the tile
    const int tid = threadIdx.x;
    int64_t r0 = 0, r1 = 0;
    if (i0 < N) { r0 = tile_ptr[(int64_t)b * N + i0]; r1 = tile_ptr[(int64_t)b * N + min(i0 + FT, N)]; }
    for (int k = tid; k < rows * FTS; k += 256) tile[k] = 0.0;
    __syncthreads();

    // The tile's entries are one contiguous range, sorted by (sample, segment).  An entry's key is its
    // cell (FILL: segment row, sample) or the cell without the slot bits (TOTALS: edge row, sample);
    // entries of equal key are adjacent and the first (FILL) / last (TOTALS) of a run owns the sum.
    const uint16_t* cp = cell + r0;
    const double* mp = tmass + r0;
    const int n = (int)(r1 - r0);
    constexpr uint32_t KMASK = FILL ? 0xFFFFu : 0xF83Fu;
    for (int k0 = tid; k0 < n; k0 += 256 * 4) {
        uint32_t c[4], cn[4];
        double v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int k = k0 + u * 256;
            const bool ok = k < n;
            c[u] = ok ? (uint32_t)cp[k] : 0x10000u;
            v[u] = ok ? mp[k] : 0.0;
            if (FILL) cn[u] = (ok && k > 0) ? (uint32_t)cp[k - 1] : 0x20000u;        // previous entry
            else cn[u] = (ok && k + 1 < n) ? (uint32_t)cp[k + 1] : 0x20000u;          // next entry
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t key = c[u] & KMASK;
            if (c[u] == 0x10000u || (cn[u] & (KMASK | 0x30000u)) == key) continue;   // padding / not the owner of its run
            const int k = k0 + u * 256;
            double sum = v[u];
            if (FILL) {   // sums forward (ascending position)
                for (int q = k + 1; q < n && (uint32_t)cp[q] == key; ++q) sum += mp[q];
                tile[((key >> 11) * B1 + ((key >> 6) & 31u)) * FTS + (key & 63u)] = sum;
            } else {      // sums backward (distal -> proximal)
                for (int q = k - 1; q >= 0 && ((uint32_t)cp[q] & KMASK) == key; --q) sum += mp[q];
                tile[(key >> 11) * FTS + (key & 63u)] = sum;
            }
        }
    }
    __syncthreads();

    if (FILL) {
        // thread = (edge group, sample): running sum down each edge, started from the mass below it
        // (emd.cpp:141 current_mass = node_masses[secondary]; :149-166 distal -> proximal)
        const int il = tid & (FT - 1), eg = tid / FT;
        const double* dcol = Dsub + (int64_t)e0 * ldm + i0 + il;
        for (int el = eg; el < ne; el += 2 * (256 / FT)) {
            const int el2 = el + 256 / FT;
            // (leaf edges have nothing below them: their D row is never read)
            double acc = (kid_off[e0 + el + 1] > kid_off[e0 + el]) ? dcol[(int64_t)el * ldm] : 0.0;
            double acc2 = (el2 < ne && kid_off[e0 + el2 + 1] > kid_off[e0 + el2]) ? dcol[(int64_t)el2 * ldm] : 0.0;
            double* t = tile + (size_t)el * B1 * FTS + il;
            for (int s = B1 - 1; s >= 0; --s) { acc += t[s * FTS]; t[s * FTS] = acc; }
            if (el2 < ne) {
                double* t2 = tile + (size_t)el2 * B1 * FTS + il;
                for (int s = B1 - 1; s >= 0; --s) { acc2 += t2[s * FTS]; t2[s * FTS] = acc2; }
            }
        }
        __syncthreads();
    }
    const int64_t ldo = FILL ? ld : ldm;       // F has the (parent's) sample stride, M/D the handle's own column count
    double* dst = (FILL ? F : Msub) + (int64_t)base * ldo + i0;
    for (int k = tid; k < rows * FT; k += 256) {
        const int r = k / FT, cc = k % FT;
        dst[(int64_t)r * ldo + cc] = tile[r * FTS + cc];
        if (!FILL && zero_row_off) dst[(int64_t)r * ldo + zero_row_off + cc] = 0.0;   // bins = 1: the D row of the edge, inside F
    }
}

// D[e][i] = sum over child edges of M[kid][i]; M[e][i] = T[e][i] + D[e][i].  One launch per tree
// level, leaves first (their D stays 0 from prepare_tiled's memset and their M is T).
__global__ void __launch_bounds__(128)
k_subtree_level(const int32_t* __restrict__ lvl_edges, int lvl_begin, int lvl_count, const int32_t* __restrict__ kid_off,
                const int32_t* __restrict__ kid_edge, int64_t ld, double* __restrict__ Msub, double* __restrict__ Dsub)
{
    const int64_t i = (int64_t)blockIdx.y * blockDim.x + threadIdx.x;
    for (int le = blockIdx.x; le < lvl_count; le += gridDim.x) {
        const int e = lvl_edges[lvl_begin + le];
        const int k0 = kid_off[e], k1 = kid_off[e + 1];
        if (k0 == k1) continue;
        double acc = 0.0;
        for (int k = k0; k < k1; ++k) acc += Msub[(int64_t)kid_edge[k] * ld + i];
        Dsub[(int64_t)e * ld + i] = acc;
        Msub[(int64_t)e * ld + i] += acc;
    }
}

// The last `n_lvl` levels of the tree (Kt <= TOP_MAX edges) in one launch: a CTA owns 32 sample
// columns and keeps M and D of those edges in shared memory, so the level-by-level dependency costs a
// CTA barrier instead of a kernel launch plus a chain of DRAM round trips.
__global__ void __launch_bounds__(TOP_THREADS)
k_subtree_top(const int32_t* __restrict__ top, int Kt, int nk, int n_lvl, int64_t ld, double* __restrict__ Msub,
              double* __restrict__ Dsub)
{
    extern __shared__ double sm[];
    double* Msm = sm;                              // [Kt][TOP_COLS]
    double* Dsm = sm + (size_t)Kt * TOP_COLS;      // [Kt][TOP_COLS]
    int32_t* meta = reinterpret_cast<int32_t*>(Dsm + (size_t)Kt * TOP_COLS);
    const int n_meta = Kt + (Kt + 1) + nk + (n_lvl + 1);
    for (int k = threadIdx.x; k < n_meta; k += TOP_THREADS) meta[k] = top[k];
    __syncthreads();
    const int32_t* edge_id = meta;
    const int32_t* koff = meta + Kt;
    const int32_t* kid = koff + Kt + 1;
    const int32_t* loff = kid + nk;
    const int col = threadIdx.x & (TOP_COLS - 1), grp = threadIdx.x / TOP_COLS;
    const int64_t gcol = (int64_t)blockIdx.x * TOP_COLS + col;

    for (int j = grp; j < Kt; j += TOP_THREADS / TOP_COLS) {      // everything that comes from below the top: independent loads
        double d = 0.0;
        for (int k = koff[j]; k < koff[j + 1]; ++k) if (kid[k] >= 0) d += Msub[(int64_t)kid[k] * ld + gcol];
        Dsm[j * TOP_COLS + col] = d;
        Msm[j * TOP_COLS + col] = Msub[(int64_t)edge_id[j] * ld + gcol];   // T
    }
    __syncthreads();
    for (int l = 0; l < n_lvl; ++l) {
        for (int j = loff[l] + grp; j < loff[l + 1]; j += TOP_THREADS / TOP_COLS) {
            double d = Dsm[j * TOP_COLS + col];
            for (int k = koff[j]; k < koff[j + 1]; ++k) if (kid[k] < 0) d += Msm[(~kid[k]) * TOP_COLS + col];
            Dsm[j * TOP_COLS + col] = d;
            Msm[j * TOP_COLS + col] += d;
        }
        __syncthreads();
    }
    for (int j = grp; j < Kt; j += TOP_THREADS / TOP_COLS) {
        Msub[(int64_t)edge_id[j] * ld + gcol] = Msm[j * TOP_COLS + col];
        Dsub[(int64_t)edge_id[j] * ld + gcol] = Dsm[j * TOP_COLS + col];
    }
}

// K0 part of the tiled path: per-entry global segment ids and per-(edge block, sample) CSR pointers
static krd_status prepare_tiled(krd_handle* h)
{
    h->tiled = false;
    const int B1 = h->cfg.bins + 1;
    if (h->cfg.bins <= 0 || B1 > 32 || h->E == 0 || h->N == 0) return KRD_OK;   // 5 slot bits in the entry code
    // Tile height: FT_MAX_ROWS rows (32 edges) for bins = 1, whose builder is a pure stream; about half of that for bins > 1,
    // where a CTA goes through its phases (zero, scatter, running sums, store) in sequence and more, smaller CTAs hide them
    // better: cfg2 K1 0.352 -> 0.333 ms with 48 rows (9 edges); 24 rows: 0.376 (profiles/r3g_k1_top_kernel_shapes.txt).
    const int rows_max = (B1 == 2) ? FT_MAX_ROWS : std::min(FT_MAX_ROWS, KRD_FT_ROWS_BINNED);
    h->EB = std::max(1, std::min(32, rows_max / B1));                           // 5 edge bits
    h->n_eblocks = (h->E + h->EB - 1) / h->EB;
    const int64_t n_slices = (int64_t)h->n_eblocks * h->N;
    krd_status st;
    if ((st = reserve(h, (void**)&h->d_cell, (size_t)std::max<int64_t>(h->M, 1) * 2)) != KRD_OK) return st;
    if ((st = reserve(h, (void**)&h->d_tmass, (size_t)std::max<int64_t>(h->M, 1) * 8)) != KRD_OK) return st;
    if ((st = reserve(h, (void**)&h->d_tile_ptr, (size_t)(n_slices + 1) * 8)) != KRD_OK) return st;
    if (B1 != 2) {   // (bins = 1 keeps M and D inside F: build_F_tiled)
        if ((st = reserve(h, (void**)&h->d_M, (size_t)h->E * h->ncols * 8)) != KRD_OK) return st;   // E x ncols (stride ncols)
        if ((st = reserve(h, (void**)&h->d_D, (size_t)h->E * h->ncols * 8)) != KRD_OK) return st;   // (leaf rows are never read)
    }
    // scratch lives in the handle (grow-only): no cudaMalloc/cudaFree, no host sync on this path
    if ((st = reserve(h, &h->d_scratch[0], (size_t)(n_slices + h->N) * 8)) != KRD_OK) return st;
    if ((st = reserve(h, &h->d_scratch[1], (size_t)(n_slices + 1) * 8)) != KRD_OK) return st;
    int64_t* blk = static_cast<int64_t*>(h->d_scratch[0]);
    int64_t* len = static_cast<int64_t*>(h->d_scratch[1]);
    k_block_ptr<<<nblk(n_slices + h->N), 256, 0, h->stream>>>(h->d_off, h->d_edge, h->N, h->n_eblocks, h->EB, blk);
    const int* guard = h->defer_check ? static_cast<const int*>(h->d_flags) : nullptr;
    k_slice_len<<<nblk(n_slices + 1), 256, 0, h->stream>>>(blk, n_slices, h->N, len, guard);
    size_t bytes = 0;
    KRD_CUDA(h, cub::DeviceScan::ExclusiveSum(nullptr, bytes, len, h->d_tile_ptr, n_slices + 1, h->stream));
    if ((st = reserve(h, &h->d_scratch[2], bytes)) != KRD_OK) return st;
    KRD_CUDA(h, cub::DeviceScan::ExclusiveSum(h->d_scratch[2], bytes, len, h->d_tile_ptr, n_slices + 1, h->stream));
    if (h->M > 0)
        k_tile_order<<<nblk(h->M), 256, 0, h->stream>>>(h->d_sid, h->d_edge, h->d_slot, h->d_mass, blk, h->d_tile_ptr,
                                                        h->M, h->N, h->EB, B1, h->d_cell, h->d_tmass, guard);
    KRD_CUDA(h, cudaGetLastError());
    h->tiled = true;
    return KRD_OK;
}

// M[e] = T[e] + D[e], D[e] = sum over child edges of M[kid] on the E x ldm matrices (T is in d_M on entry): one launch per
// wide level, the narrow top of the tree in one launch.
static krd_status subtree_passes(krd_handle* h, int64_t ldm, int* n_launch, double* Msub = nullptr, double* Dsub = nullptr)
{
    if (!Msub) { Msub = h->d_M; Dsub = h->d_D; }
    const int L = (int)h->lvl_off.size() - 1;
    const size_t smem_top = (size_t)2 * h->top_edges * TOP_COLS * 8 +
                            (size_t)(2 * h->top_edges + 1 + h->top_kids + h->top_levels + 1) * 4;
    const bool use_top = h->top_edges > 0 && smem_top <= 190 * 1024;   // (a huge multifurcation would not fit)
    const int l_end = use_top ? h->top_level : L;
    for (int l = 1; l < l_end; ++l, ++*n_launch) {   // level 0 = leaf edges: nothing below them
        const int cnt = h->lvl_off[l + 1] - h->lvl_off[l];
        dim3 g2((unsigned)std::min(cnt, 1 << 20), (unsigned)(h->ncols / 128));
        k_subtree_level<<<g2, 128, 0, h->stream>>>(h->d_lvl_edges, h->lvl_off[l], cnt, h->d_kid_off, h->d_kid_edge, ldm, Msub, Dsub);
    }
    if (use_top) {
        static std::atomic<unsigned long long> top_configured{0};
        KRD_CUDA(h, allow_dynamic_smem(k_subtree_top, 200 * 1024, top_configured));
        k_subtree_top<<<(unsigned)(h->ncols / TOP_COLS), TOP_THREADS, smem_top, h->stream>>>(h->d_top, h->top_edges, h->top_kids, h->top_levels,
                                                                                 ldm, Msub, Dsub);
        ++*n_launch;
    }
    return KRD_OK;
}

static krd_status build_F_tiled(krd_handle* h)
{
    const int B1 = h->cfg.bins + 1;
    const size_t smem_fill = (size_t)h->EB * B1 * FTS * sizeof(double), smem_tot = (size_t)h->EB * FTS * sizeof(double);
    static std::atomic<unsigned long long> configured{0};
    KRD_CUDA(h, allow_dynamic_smem(k_tile<true>, FT_MAX_ROWS * FTS * 8, configured));
    if (h->S_pad > h->S)   // rows past the last segment feed the pair kernel's last chunk: keep them zero (own columns only)
        zero_words(h->d_F + (size_t)h->S * h->ld, h->ld, h->ncols, h->S_pad - h->S, h->stream);
    dim3 grid((unsigned)h->n_eblocks, (unsigned)(h->ncols / FT));
    if (B1 == 2) {
        // bins = 1: edge e owns the rows 2e = M[e] (mass at its proximal end) and 2e+1 = D[e] (mass below it) — F IS the pair
        // of subtree matrices, interleaved.  The totals go straight into the even rows (zeros into the odd ones) and the
        // subtree passes run on F itself with a row stride of two: no separate M / D, no fill pass (4.5 instead of ~7.5
        // passes over E x ld at cfg4 / cfg5, the multi-GPU configs).
        const int64_t ld2 = 2 * h->ld;
        k_tile<false><<<grid, 256, smem_tot, h->stream>>>(h->d_cell, h->d_tmass, h->d_tile_ptr, h->N, h->ld, ld2, h->E, h->EB, B1,
                                                          nullptr, h->d_F, nullptr, h->d_kid_off, h->ld);
        int n_launch = 1;
        { krd_status sp = subtree_passes(h, ld2, &n_launch, h->d_F, h->d_F + h->ld); if (sp != KRD_OK) return sp; }
        KRD_CUDA(h, cudaGetLastError());
        h->build_launches = n_launch;
        return KRD_OK;
    }
    const int64_t ldm = h->ncols;   // stride of M / D: the handle's own columns (a pipelined child owns a slice of the parent's F)
    k_tile<false><<<grid, 256, smem_tot, h->stream>>>(h->d_cell, h->d_tmass, h->d_tile_ptr, h->N, h->ld, ldm, h->E, h->EB, B1,
                                                      nullptr, h->d_M, nullptr, h->d_kid_off, 0);
    int n_launch = 2;
    { krd_status sp = subtree_passes(h, ldm, &n_launch); if (sp != KRD_OK) return sp; }
    k_tile<true><<<grid, 256, smem_fill, h->stream>>>(h->d_cell, h->d_tmass, h->d_tile_ptr, h->N, h->ld, ldm, h->E, h->EB, B1,
                                                     h->d_D, nullptr, h->d_F, h->d_kid_off, 0);
    KRD_CUDA(h, cudaGetLastError());
    h->build_launches = n_launch;   // totals + level passes + top-of-tree + fill
    return KRD_OK;
}


// ------------------------------------------------------------------------------------------------
//  K1 exact (bins = 0): F written exactly once, whole or slab by slab
// ------------------------------------------------------------------------------------------------
//
// In exact mode every distinct (edge, position) of ANY sample cuts a segment, so F has S ~ (number of mass entries) rows
// while a column changes only where ITS sample has a mass: column i over edge e is piecewise constant with
// (entries of i on e) + 1 pieces.  The builder therefore never scatters and scans; per entry it prepares
//     grow[m] = seg_off[edge] + slot        the row of F at which the entry's mass starts to count (going proximal)
//     cum[m]  = D[e][i] + masses of (i, e) at rows >= grow[m], summed distal -> proximal  (emd.cpp:141-166)
// from the E x N subtree matrices (T -> M, D: the same level passes as the binned builder), and a row of F is then a pure
// look-up:  F[r][i] = cum[m*] if the first entry m* of sample i with grow >= r lies on the edge of row r, else D[e(r)][i].
// k_exact_slab sweeps a row range in order, one thread per sample column carrying its entry pointer in registers: one
// binary search per (column, row range), one coalesced 512 B store per row and 64 columns, no read of F at all.
// The same kernel fills the whole matrix (resident) or one slab of it (streamed, krd_config.max_f_mb).

__global__ void k_exact_grow(const int32_t* __restrict__ edge, const int32_t* __restrict__ slot, const int32_t* __restrict__ seg_off,
                             int64_t M, int32_t* __restrict__ grow)
{
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m < M) grow[m] = seg_off[edge[m]] + slot[m];
}

// The LAST entry of every (sample, edge) run walks its run distal -> proximal.  Entries of equal slot (duplicate
// positions in the input) are one mass: summed in input order first, like the std::map's += (operators.cpp:169).
//   CUM = false: T[e][i] = mass of sample i on edge e            (-> subtree passes)
//   CUM = true : cum[m]  = D[e][i] + running sum                 (needs the finished D)
template <bool CUM>
__global__ void __launch_bounds__(256)
k_exact_runs(const uint32_t* __restrict__ sid, const int64_t* __restrict__ off, const int32_t* __restrict__ edge,
             const int32_t* __restrict__ slot, const double* __restrict__ mass, int64_t M, int64_t ldm,
             const double* __restrict__ D, double* __restrict__ T, double* __restrict__ cum)
{
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    const uint32_t i = sid[m];
    const int e = edge[m];
    if (m + 1 < off[i + 1] && edge[m + 1] == e) return;      // not the last entry of its run
    const int64_t lo = off[i];
    double acc = CUM ? D[(int64_t)e * ldm + i] : 0.0;
    int64_t q = m;
    while (q >= lo && edge[q] == e) {
        const int s = slot[q];
        int64_t g0 = q;
        while (g0 - 1 >= lo && edge[g0 - 1] == e && slot[g0 - 1] == s) --g0;
        double sum = mass[g0];
        for (int64_t k = g0 + 1; k <= q; ++k) sum += mass[k];
        acc += sum;
        if (CUM) for (int64_t k = g0; k <= q; ++k) cum[k] = acc;
        q = g0 - 1;
    }
    if (!CUM) T[(int64_t)e * ldm + i] = acc;
}

constexpr int XS_COLS = 64;   // sample columns per CTA of k_exact_slab (one thread each)

__global__ void __launch_bounds__(XS_COLS)
k_exact_slab(const int64_t* __restrict__ off, const int32_t* __restrict__ grow, const int32_t* __restrict__ edge,
             const double* __restrict__ cum, const int32_t* __restrict__ seg_off, int E, const double* __restrict__ D,
             int64_t ldm, int N, int64_t S, int64_t row0, int64_t row1, int64_t rows_per_cta, double* __restrict__ F, int64_t ld)
{
    const int i = blockIdx.y * XS_COLS + threadIdx.x;
    int64_t r = row0 + (int64_t)blockIdx.x * rows_per_cta;
    const int64_t r_end = min(row1, r + rows_per_cta);
    if (r >= r_end) return;
    // edge of row r: the largest e with seg_off[e] <= r (rows >= S are the zero padding of the last chunk)
    int e = E;
    if (r < S) {
        int lo = 0, hi = E;      // seg_off[lo] <= r < seg_off[hi]
        while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (seg_off[mid] <= r) lo = mid; else hi = mid; }
        e = lo;
    }
    int64_t e_end = (e < E) ? (int64_t)seg_off[e + 1] : INT64_MAX;
    const bool col = i < N;
    double d = (col && e < E) ? D[(int64_t)e * ldm + i] : 0.0;
    // first entry of this sample with grow >= r
    int64_t m = 0, m_end = 0;
    if (col) {
        m = off[i]; m_end = off[i + 1];
        int64_t lo = m, hi = m_end;
        while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (grow[mid] < r) lo = mid + 1; else hi = mid; }
        m = lo;
    }
    int64_t g_next = INT64_MAX; int e_next = -1; double c_next = 0.0;
    if (m < m_end) { g_next = grow[m]; e_next = edge[m]; c_next = cum[m]; }
    double* out = F + (r - row0) * ld + i;
    for (; r < r_end; ++r, out += ld) {
        if (r >= e_end) {        // (every edge owns at least one row)
            ++e;
            e_end = (e < E) ? (int64_t)seg_off[e + 1] : INT64_MAX;
            d = (col && e < E) ? D[(int64_t)e * ldm + i] : 0.0;
        }
        while (g_next < r) {
            ++m;
            if (m < m_end) { g_next = grow[m]; e_next = edge[m]; c_next = cum[m]; }
            else { g_next = INT64_MAX; e_next = -1; }
        }
        *out = (r < S && col) ? ((e_next == e) ? c_next : d) : 0.0;
    }
}

krd_status needs_resident_F(krd_handle* h)
{
    char msg[320];
    std::snprintf(msg, sizeof msg, "this operation needs the dense matrix F resident (%.2f GB for %lld segments x %d samples); "
                  "it is being streamed in slabs - use --bins, fewer samples, or raise krd_config.max_f_mb",
                  (double)h->S_pad * h->ld * 8 / 1e9, (long long)h->S, h->N);
    return fail(h, KRD_ERR_STATE, msg);
}

// K0 part: per-entry rows, buffers, and the resident / streamed decision
static krd_status prepare_exact(krd_handle* h)
{
    h->exact_fast = false; h->streamed = false; h->cum_valid = false; h->slab_rows = 0;
    if (h->cfg.bins != 0 || h->cfg.builder == 1 || h->E == 0 || h->N == 0 || h->ld_fixed) return KRD_OK;
    const size_t Mx = (size_t)std::max<int64_t>(h->M, 1);
    krd_status st;
    if ((st = reserve(h, (void**)&h->d_grow, Mx * 4)) != KRD_OK) return st;
    if ((st = reserve(h, (void**)&h->d_cum, Mx * 8)) != KRD_OK) return st;
    if ((st = reserve(h, (void**)&h->d_M, (size_t)h->E * h->ncols * 8)) != KRD_OK) return st;
    if ((st = reserve(h, (void**)&h->d_D, (size_t)h->E * h->ncols * 8)) != KRD_OK) return st;
    if (h->M > 0) k_exact_grow<<<nblk(h->M), 256, 0, h->stream>>>(h->d_edge, h->d_slot, h->d_seg_off, h->M, h->d_grow);
    KRD_CUDA(h, cudaGetLastError());
    // resident or streamed?
    size_t free_b = 0, total_b = 0;
    KRD_CUDA(h, cudaMemGetInfo(&free_b, &total_b));
    free_b += h->F_capacity;                                       // (our own old F is reusable)
    const size_t need = (size_t)h->S_pad * h->ld * sizeof(double);
    const size_t cap = h->cfg.max_f_mb > 0 ? (size_t)h->cfg.max_f_mb << 20 : (size_t)(0.6 * (double)free_b);
    if (need > cap) {
        if (h->extra_cols > 0) return needs_resident_F(h);
        const size_t row_bytes = (size_t)h->ld * sizeof(double);
        size_t slab = h->cfg.max_f_mb > 0 ? cap : std::min(cap, std::max<size_t>((size_t)256 << 20, std::min<size_t>((size_t)2 << 30, free_b / 4)));
        int64_t rows = (int64_t)(slab / row_bytes) / KC * KC;
        if (rows < KC) rows = KC;
        h->streamed = true;
        h->slab_rows = std::min<int64_t>(rows, h->S_pad);
        h->stats.f_bytes = (int64_t)h->slab_rows * h->ld * 8;
    }
    h->exact_fast = true;
    return KRD_OK;
}

// T -> M, D -> cum: everything of the exact builder that does not depend on the row range
static krd_status exact_prefix(krd_handle* h, int* n_launch)
{
    const int64_t ldm = h->ncols;
    zero_words(h->d_M, ldm, ldm, h->E, h->stream);
    zero_words(h->d_D, ldm, ldm, h->E, h->stream);
    *n_launch = 2;
    if (h->M > 0) {
        k_exact_runs<false><<<nblk(h->M), 256, 0, h->stream>>>(h->d_sid, h->d_off, h->d_edge, h->d_slot, h->d_mass, h->M, ldm, nullptr, h->d_M, nullptr);
        ++*n_launch;
    }
    krd_status st = subtree_passes(h, ldm, n_launch);
    if (st != KRD_OK) return st;
    if (h->M > 0) {
        k_exact_runs<true><<<nblk(h->M), 256, 0, h->stream>>>(h->d_sid, h->d_off, h->d_edge, h->d_slot, h->d_mass, h->M, ldm, h->d_D, nullptr, h->d_cum);
        ++*n_launch;
    }
    KRD_CUDA(h, cudaGetLastError());
    h->cum_valid = true;
    return KRD_OK;
}

static krd_status exact_rows(krd_handle* h, int64_t row0, int64_t row1, double* F)
{
    const int64_t rows = row1 - row0;
    const int col_groups = (int)(h->ld / XS_COLS);
    const int64_t want_ctas = (int64_t)h->num_sms * 32;            // 32 CTAs of 64 threads per SM
    int64_t parts = std::max<int64_t>(1, want_ctas / col_groups);
    int64_t rpc = std::max<int64_t>(32, (rows + parts - 1) / parts);
    parts = (rows + rpc - 1) / rpc;
    dim3 grid((unsigned)parts, (unsigned)col_groups);
    k_exact_slab<<<grid, XS_COLS, 0, h->stream>>>(h->d_off, h->d_grow, h->d_edge, h->d_cum, h->d_seg_off, h->E, h->d_D, h->ncols, h->N,
                                                  h->S, row0, row1, rpc, F, h->ld);
    KRD_CUDA(h, cudaGetLastError());
    return KRD_OK;
}

static krd_status ensure_F_buffer(krd_handle* h, size_t need)
{
    if (need <= h->F_capacity) return KRD_OK;
    cudaFree(h->d_F); h->d_F = nullptr; h->F_capacity = 0;
    krd_status sa = alloc_sized(h, (void**)&h->d_F, need, "the dense cumulative-mass matrix F (segments x samples x 8 B)");
    if (sa != KRD_OK) return sa;
    h->F_capacity = need;
    return KRD_OK;
}

// Streamed exact mode: slab after slab, the pair kernel accumulates per-tile sums across the slabs (krd_pairs.cu).
krd_status compute_streamed(krd_handle* h, double p, double divisor, int rank, int world, double* d_out, double* d_packed)
{
    krd_status st = ensure_F_buffer(h, (size_t)h->slab_rows * h->ld * sizeof(double));
    if (st != KRD_OK) return st;
    h->stats.launches = 0;
    cudaEventRecord(h->ev[0], h->stream);
    if (!h->cum_valid) {
        int nl = 0;
        if ((st = exact_prefix(h, &nl)) != KRD_OK) return st;
        h->stats.launches += nl;
    }
    cudaEventRecord(h->ev[1], h->stream);
    const int64_t n_slabs = (h->S_pad + h->slab_rows - 1) / h->slab_rows;
    constexpr int TIMED = 32;
    while ((int)h->slab_ev.size() < 2 * TIMED) { cudaEvent_t ev; KRD_CUDA(h, cudaEventCreate(&ev)); h->slab_ev.push_back(ev); }
    h->slabs_timed = (int32_t)std::min<int64_t>(n_slabs, TIMED); h->slabs_total = (int32_t)n_slabs;
    for (int64_t b = 0; b < n_slabs; ++b) {
        const int64_t row0 = b * h->slab_rows, row1 = std::min(h->S_pad, row0 + h->slab_rows);
        if (b < TIMED) cudaEventRecord(h->slab_ev[(size_t)2 * b], h->stream);
        if ((st = exact_rows(h, row0, row1, h->d_F)) != KRD_OK) return st;
        if (b < TIMED) cudaEventRecord(h->slab_ev[(size_t)2 * b + 1], h->stream);
        SlabArgs slab{h->d_F, row0, row1 - row0, b == 0, b == n_slabs - 1};
        if ((st = run_pairs_slab(h, p, divisor, rank, world, d_out, d_packed, slab)) != KRD_OK) return st;
        h->stats.launches += 1;
    }
    cudaEventRecord(h->ev[3], h->stream);
    return KRD_OK;
}

krd_status build_F(krd_handle* h)
{
    h->norms_valid = false;
    if (h->streamed) return needs_resident_F(h);
    const size_t need = (size_t)h->S_pad * h->ld * sizeof(double);
    if (h->ld_fixed) {   // child of a pipelined ingest: d_F points into the parent's matrix
        if (!h->d_F || !h->tiled) return fail(h, KRD_ERR_STATE, "pipelined child without a tiled builder");
    } else {
        krd_status sa = ensure_F_buffer(h, need);
        if (sa != KRD_OK) return sa;
    }
    if (h->exact_fast) {
        int nl = 0;
        krd_status st = exact_prefix(h, &nl);
        if (st == KRD_OK) st = exact_rows(h, 0, h->S_pad, h->d_F);
        if (st != KRD_OK) return st;
        h->build_launches = nl + 1;
        h->F_valid = true;
        return KRD_OK;
    }
    if (h->tiled) {
        krd_status st = build_F_tiled(h);
        if (st == KRD_OK) h->F_valid = true;
        return st;
    }
    KRD_CUDA(h, cudaMemsetAsync(h->d_F, 0, need, h->stream));
    if (h->M > 0)
        k_scatter<<<nblk(h->M), 256, 0, h->stream>>>(h->d_sid, h->d_off, h->d_edge, h->d_slot, h->d_mass, h->d_seg_off, h->M, h->ld, h->d_F);
    const int L = (int)h->lvl_off.size() - 1;
    dim3 block(128);
    for (int l = 0; l < L; ++l) {
        int cnt = h->lvl_off[l + 1] - h->lvl_off[l];
        dim3 grid((unsigned)std::min(cnt, 1 << 20), (unsigned)(h->ld / 128));
        k_scan_level<<<grid, block, 0, h->stream>>>(h->d_lvl_edges, h->lvl_off[l], cnt, h->d_kid_off, h->d_kid_edge,
                                                    h->d_seg_off, h->ld, h->d_F);
    }
    KRD_CUDA(h, cudaGetLastError());
    h->build_launches = L + 2;   // memset + scatter + one scan per level
    h->F_valid = true;
    return KRD_OK;
}

}  // namespace krd
