// Squash clustering on the dense matrix F — SURVEY §8f rank 3.
//
// Reference: genesis::tree::SquashClustering (tree/mass_tree/squash_clustering.cpp:57-249), the engine of
// `gappa analyze squash`.  There every merge builds a new MassTree (std::map merge of two trees, :146-150) and
// runs N sparse pair walks against it (:181-189).  Here a cluster is a COLUMN of F: F is linear in the masses,
// so the merged, re-normalised tree is the count-weighted average of two columns (k_merge_cols), and the N
// distances of the new cluster are one pass over F (k_one_vs_all, HBM-bound: 8*S*(N+m) bytes per merge).
// The n-1 merges are enqueued back to back; the pair to merge travels through device memory, the host never
// waits inside the loop.
#include <algorithm>
#include <vector>

#include "krd_common.cuh"
#include "krd_ops.cuh"

namespace krd {

constexpr int OVA_CHUNKS = 64;     // segment chunks of the one-vs-all reduction (partials summed in fixed order)
constexpr int ARGMIN_CTAS = 148;

struct Best { double v; int i, j; };

__device__ __forceinline__ bool better(const Best& a, const Best& b)   // strict order: value, then row, then column
{                                                                      // (= the first minimum of the serial scan, :103-131)
    if (a.v != b.v) return a.v < b.v;
    if (a.i != b.i) return a.i < b.i;
    return a.j < b.j;
}

__device__ __forceinline__ Best block_best(Best b, Best* sh)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Best c;
        c.v = __shfl_down_sync(0xffffffffu, b.v, o); c.i = __shfl_down_sync(0xffffffffu, b.i, o); c.j = __shfl_down_sync(0xffffffffu, b.j, o);
        if (better(c, b)) b = c;
    }
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = b;
    __syncthreads();
    if (threadIdx.x == 0) for (int w = 1; w < (int)(blockDim.x >> 5); ++w) if (better(sh[w], b)) b = sh[w];
    return b;   // valid in thread 0
}

__global__ void k_sq_init(int N, int Nt, const double* __restrict__ M0, const double* __restrict__ F, int64_t ld,
                          const int32_t* __restrict__ root_seg, int n_root, double* __restrict__ D, int* __restrict__ active,
                          double* __restrict__ count, double* __restrict__ total)
{
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g < (int64_t)N * N) { const int i = (int)(g / N), j = (int)(g % N); D[(int64_t)i * Nt + j] = M0[g]; }
    if (g < Nt) {
        active[g] = g < N; count[g] = 1.0;
        double t = 0.0;   // total mass of a sample = what arrives at the root: sum over the root's child edges of F at their first segment
        if (g < N) for (int r = 0; r < n_root; ++r) t += F[(int64_t)root_seg[r] * ld + g];
        total[g] = t;
    }
}

__global__ void __launch_bounds__(256)
k_sq_argmin(const double* __restrict__ D, const int* __restrict__ active, int n_cur, int Nt, Best* __restrict__ part)
{
    __shared__ Best sh[8];
    Best b{1.7976931348623157e308, 0x7fffffff, 0x7fffffff};
    for (int i = blockIdx.x; i < n_cur; i += gridDim.x) {
        if (!active[i]) continue;
        for (int j = threadIdx.x; j < i; j += 256) {
            if (!active[j]) continue;
            const Best c{D[(int64_t)i * Nt + j], i, j};
            if (better(c, b)) b = c;
        }
    }
    b = block_best(b, sh);
    if (threadIdx.x == 0) part[blockIdx.x] = b;
}

__global__ void __launch_bounds__(256) k_sq_argmin_final(const Best* __restrict__ part, int n_part, int* __restrict__ pair)
{
    __shared__ Best sh[8];
    Best b{1.7976931348623157e308, 0x7fffffff, 0x7fffffff};
    for (int k = threadIdx.x; k < n_part; k += 256) if (better(part[k], b)) b = part[k];
    b = block_best(b, sh);
    if (threadIdx.x == 0) { pair[0] = b.j; pair[1] = b.i; }   // returned as (smaller, larger) index, :133-135
}

// new column c = (w_a F_a + w_b F_b) / (w_a T_a + w_b T_b)      (functions.cpp:71-94 then :126-140)
__global__ void k_sq_merge_cols(double* __restrict__ F, int64_t ld, int64_t S_pad, const int* __restrict__ pair,
                                const double* __restrict__ count, const double* __restrict__ total, int c)
{
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S_pad) return;
    const int a = pair[0], b = pair[1];
    const double wa = count[a], wb = count[b];
    const double tot = wa * total[a] + wb * total[b];
    const double v = wa * F[s * ld + a] + wb * F[s * ld + b];
    F[s * ld + c] = (tot == 0.0) ? v : v / tot;
}

template <int MODE>
__global__ void __launch_bounds__(128)
k_sq_one_vs_all(const double* __restrict__ F, const double* __restrict__ len, int64_t ld, int64_t S_pad, int64_t rows_per_chunk,
                int c, double p, double* __restrict__ part)
{
    __shared__ double s_tab[MODE == MODE_GENERIC ? POW_TAB : 1];
    if (MODE == MODE_GENERIC) { for (int k = threadIdx.x; k < POW_TAB; k += 128) s_tab[k] = g_pow_tab[k]; __syncthreads(); }
    const int64_t k = (int64_t)blockIdx.x * 128 + threadIdx.x;     // column (cluster) k < ld; columns >= c are not used
    const int64_t s0 = (int64_t)blockIdx.y * rows_per_chunk, s1 = min(S_pad, s0 + rows_per_chunk);
    double acc = 0.0;
#pragma unroll 4
    for (int64_t s = s0; s < s1; ++s) {
        const double x = F[s * ld + c], y = F[s * ld + k], l = len[s];
        if (MODE == MODE_P2) { const double d = x - y; acc = fma(l * d, d, acc); }
        else acc = pair_op<MODE>(x, y, l, p, s_tab, acc);
    }
    part[(int64_t)blockIdx.y * ld + k] = acc;
}

__global__ void k_sq_ova_final(const double* __restrict__ part, int64_t ld, int c, int Nt, double p, const int* __restrict__ active,
                               const int* __restrict__ pair, double* __restrict__ D, krd_merger* __restrict__ merger)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= c || !active[k]) return;     // the two clusters being merged are still active here (:176-189)
    double v = 0.0;
    for (int ch = 0; ch < OVA_CHUNKS; ++ch) v += part[(int64_t)ch * ld + k];
    if (p > 1.0) v = (p == 2.0) ? sqrt(v) : pow(v, 1.0 / p);
    D[(int64_t)c * Nt + k] = v;
    D[(int64_t)k * Nt + c] = v;
    if (k == pair[0]) { merger->index_a = k; merger->distance_a = v; }
    if (k == pair[1]) { merger->index_b = k; merger->distance_b = v; }
}

__global__ void k_sq_merge_state(const int* __restrict__ pair, int* __restrict__ active, double* __restrict__ count,
                                 double* __restrict__ total, int c)
{
    const int a = pair[0], b = pair[1];
    count[c] = count[a] + count[b];
    total[c] = 1.0;                 // the merged tree was re-normalised
    active[a] = 0; active[b] = 0; active[c] = 1;
}

template <int MODE>
static void launch_ova(krd_handle* h, int c, double p, int64_t rpc, double* part)
{
    k_sq_one_vs_all<MODE><<<dim3((unsigned)((c + 127) / 128), OVA_CHUNKS), 128, 0, h->stream>>>(h->d_F, h->d_len, h->ld, h->S_pad, rpc, c, p, part);
}

krd_status run_squash(krd_handle* h, double p, krd_merger* out)
{
    const int N = h->N, Nt = 2 * N - 1;
    if (h->streamed) return needs_resident_F(h);
    if (h->ld < Nt) {     // widen F: one extra column per merged cluster, then redo the (resident) masses' preparation
        h->extra_cols = N;
        krd_status st = prepare_masses(h);
        if (st != KRD_OK) {      // (e.g. the widened matrix no longer fits: back to the plain layout, the handle stays usable)
            const std::string msg = h->err;
            h->extra_cols = 0;
            if (prepare_masses(h) != KRD_OK) h->have_masses = false;
            h->err = msg;
            return st;
        }
    }
    if (!h->F_valid) { krd_status st = build_F(h); if (st != KRD_OK) return st; }

    // initial clusters = samples: all-pairs matrix from the tile kernel (init_, squash_clustering.cpp:57-95)
    krd_status st = reserve(h, (void**)&h->d_out_tmp_sq, (size_t)N * N * 8);
    if (st != KRD_OK) return st;
    if ((st = run_pairs(h, p, 1.0, 0, 1, h->d_out_tmp_sq, nullptr)) != KRD_OK) return st;

    // state: D | count | total | part | active | pair | argmin partials | mergers | root segments
    std::vector<int32_t> root_seg;
    std::vector<int32_t> seg_off((size_t)h->E + 1);
    KRD_CUDA(h, cudaMemcpyAsync(seg_off.data(), h->d_seg_off, seg_off.size() * 4, cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    for (int e = 0; e < h->E; ++e) if (h->parent[(size_t)e] == h->root) root_seg.push_back(seg_off[(size_t)e]);
    const size_t b_D = (size_t)Nt * Nt * 8, b_vec = (size_t)Nt * 8, b_part = (size_t)OVA_CHUNKS * h->ld * 8;
    const size_t b_act = (size_t)Nt * 4, b_best = ARGMIN_CTAS * sizeof(Best), b_mer = (size_t)(N - 1) * sizeof(krd_merger);
    const size_t b_root = std::max<size_t>(root_seg.size(), 1) * 4;
    auto al = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t total_bytes = al(b_D) + 2 * al(b_vec) + al(b_part) + al(b_act) + 256 + al(b_best) + al(b_mer) + al(b_root);
    if ((st = reserve(h, &h->d_squash, total_bytes)) != KRD_OK) return st;
    char* q = static_cast<char*>(h->d_squash);
    double* D = (double*)q; q += al(b_D);
    double* count = (double*)q; q += al(b_vec);
    double* total = (double*)q; q += al(b_vec);
    double* part = (double*)q; q += al(b_part);
    int* active = (int*)q; q += al(b_act);
    int* pair = (int*)q; q += 256;
    Best* best = (Best*)q; q += al(b_best);
    krd_merger* mergers = (krd_merger*)q; q += al(b_mer);
    int32_t* d_root = (int32_t*)q;
    KRD_CUDA(h, cudaMemcpyAsync(d_root, root_seg.data(), root_seg.size() * 4, cudaMemcpyHostToDevice, h->stream));
    KRD_CUDA(h, cudaMemsetAsync(D, 0, b_D, h->stream));
    const int64_t n_init = std::max<int64_t>((int64_t)N * N, Nt);
    k_sq_init<<<(unsigned)((n_init + 255) / 256), 256, 0, h->stream>>>(N, Nt, h->d_out_tmp_sq, h->d_F, h->ld, d_root, (int)root_seg.size(),
                                                                      D, active, count, total);
    const int mode = mode_for_exponent(p, false);
    if (mode == MODE_GENERIC) { cudaError_t e = make_pow_tables(); if (e != cudaSuccess) return cuda_fail(h, e, "pow tables"); }
    const int64_t rpc = (h->S_pad + OVA_CHUNKS - 1) / OVA_CHUNKS;
    for (int m = 0; m < N - 1; ++m) {
        const int c = N + m;
        k_sq_argmin<<<ARGMIN_CTAS, 256, 0, h->stream>>>(D, active, c, Nt, best);
        k_sq_argmin_final<<<1, 256, 0, h->stream>>>(best, ARGMIN_CTAS, pair);
        k_sq_merge_cols<<<(unsigned)((h->S_pad + 255) / 256), 256, 0, h->stream>>>(h->d_F, h->ld, h->S_pad, pair, count, total, c);
        switch (mode) {
            case MODE_P1: launch_ova<MODE_P1>(h, c, p, rpc, part); break;
            case MODE_P2: launch_ova<MODE_P2>(h, c, p, rpc, part); break;
            case MODE_PHALF: launch_ova<MODE_PHALF>(h, c, p, rpc, part); break;
            case MODE_P3: launch_ova<MODE_P3>(h, c, p, rpc, part); break;
            case MODE_P4: launch_ova<MODE_P4>(h, c, p, rpc, part); break;
            case MODE_INT: launch_ova<MODE_INT>(h, c, p, rpc, part); break;
            case MODE_LIBPOW: launch_ova<MODE_LIBPOW>(h, c, p, rpc, part); break;
            default: launch_ova<MODE_GENERIC>(h, c, p, rpc, part); break;
        }
        k_sq_ova_final<<<(unsigned)((c + 255) / 256), 256, 0, h->stream>>>(part, h->ld, c, Nt, p, active, pair, D, mergers + m);
        k_sq_merge_state<<<1, 1, 0, h->stream>>>(pair, active, count, total, c);
    }
    KRD_CUDA(h, cudaGetLastError());
    KRD_CUDA(h, cudaMemcpyAsync(out, mergers, b_mer, cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    h->F_valid = false;     // the cluster columns are scratch: the next compute rebuilds F
    h->stats.launches += 1 + 6 * (N - 1);
    return KRD_OK;
}

}  // namespace krd
