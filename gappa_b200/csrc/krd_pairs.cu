// K2/K3: all-pairs weighted-Lp reduction over the dense matrix F[S_pad][ld] (segment-major):
//     W_p(i,j) = sum_s len[s] * |F[s][i] - F[s][j]|^p          (genesis emd.cpp:154,166)
//     KR_p     = W_p            if p <= 1
//              = W_p^(1/p)      if p  > 1                       (genesis emd.cpp:176-178)
// for every i < j, written symmetrically with a zero diagonal (emd.cpp:193,214-215).
//
// Shape: a "GEMM with a different inner operator".  CTA tile 128 x 128 pairs, 256 threads, each
// thread an 8 x 8 register tile (64 FP64 accumulators), K (= segment) axis streamed through a
// 4-stage shared-memory ring filled by the TMA engine (cp.async.bulk, 1 KB rows, completion on an
// mbarrier).  The upper triangle of tiles is crossed with a split of the segment axis so that the
// grid fills 148 SMs in whole waves; partial tiles go to a workspace and a small deterministic
// epilogue kernel sums the splits in fixed order, applies the outer root / divisor and mirrors.
// FP64 CUDA cores only: per (pair, segment) p = 1 costs DADD + DFMA, the floor for this operator.
#include <algorithm>
#include <cmath>

#include "krd_common.cuh"

namespace krd {

constexpr int ST = 4;          // pipeline stages
constexpr int NT = 256;        // threads per CTA
constexpr int STAGE_DOUBLES = 2 * KC * TILE + KC;  // A rows, B rows, len
constexpr size_t SMEM_BYTES = (size_t)ST * STAGE_DOUBLES * sizeof(double) + ST * sizeof(uint64_t);

struct PairParams {
    const double* F;
    const double* len;     // len or sqrt(len) (MODE_P2)
    const Tile* tiles;
    double* ws;            // [split][n_tiles][TILE*TILE]
    int64_t ld;
    int n_chunks;          // S_pad / KC
    int chunks_per_split;
    int n_tiles;
    double p;
};

// ---- PTX wrappers -------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// TMA engine, 1-D bulk copy global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- inner operators ----------------------------------------------------------------------------

template <int MODE>
__device__ __forceinline__ double pair_op(double a, double b, double l, double p, double acc)
{
    if (MODE == MODE_P1) return fma(l, fabs(a - b), acc);
    if (MODE == MODE_P2) { double d = a - b; return fma(d, d, acc); }   // operands pre-scaled by sqrt(len)
    if (MODE == MODE_PHALF) return fma(l, sqrt(fabs(a - b)), acc);
    return fma(l, pow(fabs(a - b), p), acc);
}

// ---- the tile kernel ------------------------------------------------------------------------------

template <int MODE>
__global__ void __launch_bounds__(NT, 1) k_pairs(const PairParams P)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double* stage_base = reinterpret_cast<double*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)ST * STAGE_DOUBLES * sizeof(double));

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tx = tid & 15, ty = tid >> 4;
    const Tile tile = P.tiles[blockIdx.x];
    const bool diag = (tile.i == tile.j);
    const int c0 = blockIdx.y * P.chunks_per_split;
    const int c1 = min(c0 + P.chunks_per_split, P.n_chunks);
    const int nc = max(c1 - c0, 0);

    if (tid == 0) {
        for (int s = 0; s < ST; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    const double* gA = P.F + (int64_t)tile.i * TILE;
    const double* gB = P.F + (int64_t)tile.j * TILE;
    const uint32_t stage_bytes = (uint32_t)((diag ? 1 : 2) * KC * TILE * 8 + KC * 8);

    // producer: warp 0; lane r < 16 fetches A row r, lane 16 + r fetches B row r, lane 0 also len
    auto issue = [&](int chunk_local) {
        const int s = chunk_local % ST;
        double* sA = stage_base + (size_t)s * STAGE_DOUBLES;
        double* sB = sA + KC * TILE;
        double* sL = sB + KC * TILE;
        const int64_t k0 = (int64_t)(c0 + chunk_local) * KC;
        if (lane == 0) mbar_expect_tx(&full[s], stage_bytes);
        __syncwarp();
        const int r = lane & 15;
        if (lane < 16) tma_load_1d(sA + r * TILE, gA + (k0 + r) * P.ld, TILE * 8, &full[s]);
        else if (!diag) tma_load_1d(sB + r * TILE, gB + (k0 + r) * P.ld, TILE * 8, &full[s]);
        if (lane == 0) tma_load_1d(sL, P.len + k0, KC * 8, &full[s]);
    };

    if (warp == 0) {
        for (int c = 0; c < min(ST - 1, nc); ++c) issue(c);
    }

    double acc[8][8];
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[r][c] = 0.0;

    for (int c = 0; c < nc; ++c) {
        // every warp has finished chunk c-1 -> its stage may be refilled with chunk c+ST-1
        __syncthreads();
        if (warp == 0 && c + ST - 1 < nc) issue(c + ST - 1);

        const int s = c % ST;
        mbar_wait(&full[s], (uint32_t)((c / ST) & 1));
        const double* sA = stage_base + (size_t)s * STAGE_DOUBLES;
        const double* sB = diag ? sA : sA + KC * TILE;
        const double* sL = sA + 2 * KC * TILE;
        const double* pa = sA + 8 * ty;
        const double* pb = sB + 2 * tx;

#pragma unroll 4
        for (int kk = 0; kk < KC; ++kk) {
            double a[8], b[8];
            const double l = sL[kk];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double2 va = *reinterpret_cast<const double2*>(pa + kk * TILE + 2 * q);
                const double2 vb = *reinterpret_cast<const double2*>(pb + kk * TILE + 32 * q);
                a[2 * q] = va.x; a[2 * q + 1] = va.y;
                b[2 * q] = vb.x; b[2 * q + 1] = vb.y;
            }
            if (MODE == MODE_P2) {
#pragma unroll
                for (int q = 0; q < 8; ++q) { a[q] *= l; b[q] *= l; }
            }
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int cc = 0; cc < 8; ++cc) acc[r][cc] = pair_op<MODE>(a[r], b[cc], l, P.p, acc[r][cc]);
        }
    }

    // partial tile -> workspace; row = 8*ty + r, col = 32*q + 2*tx + {0,1}
    double* w = P.ws + ((size_t)blockIdx.y * P.n_tiles + blockIdx.x) * (size_t)(TILE * TILE);
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q)
            *reinterpret_cast<double2*>(w + (8 * ty + r) * TILE + 32 * q + 2 * tx) = make_double2(acc[r][2 * q], acc[r][2 * q + 1]);
}

// ---- epilogue: sum splits in fixed order, outer root, divisor, mirror ---------------------------

struct EpiParams {
    const double* ws;
    const Tile* tiles;
    int n_tiles, split, N;
    double p, divisor;
    double* out;        // N x N row-major or nullptr
    double* packed;     // [n_tiles][TILE*TILE] or nullptr
};

__global__ void __launch_bounds__(256) k_epilogue(const EpiParams P)
{
    const Tile tile = P.tiles[blockIdx.x];
    const double inv_p = 1.0 / P.p;
    for (int idx = threadIdx.x; idx < TILE * TILE; idx += blockDim.x) {
        double v = 0.0;
        for (int s = 0; s < P.split; ++s) v += P.ws[((size_t)s * P.n_tiles + blockIdx.x) * (size_t)(TILE * TILE) + idx];
        if (P.p > 1.0) v = (P.p == 2.0) ? sqrt(v) : pow(v, inv_p);
        v = v / P.divisor;
        const int r = idx / TILE, c = idx % TILE;
        const int64_t i = (int64_t)tile.i * TILE + r, j = (int64_t)tile.j * TILE + c;
        if (P.packed) P.packed[(size_t)blockIdx.x * (TILE * TILE) + idx] = (i < j) ? v : 0.0;
        if (P.out && i < P.N && j < P.N) {
            if (i < j) { P.out[i * P.N + j] = v; P.out[j * P.N + i] = v; }
            else if (i == j) P.out[i * P.N + i] = 0.0;
        }
    }
}

// gathered packed tiles of all ranks -> full matrix
__global__ void __launch_bounds__(256) k_unpack(const double* __restrict__ gathered, const Tile* __restrict__ tiles,
                                                int64_t n_tiles, int64_t per_rank, int N, double* __restrict__ out)
{
    // blockIdx.x = global tile number t; it lives on rank t / per_rank at slot t % per_rank
    const int64_t t = blockIdx.x;
    if (t >= n_tiles) return;
    const Tile tile = tiles[t];
    const double* src = gathered + (size_t)t * (TILE * TILE);  // ranks are contiguous: rank*per_rank + slot == t
    for (int idx = threadIdx.x; idx < TILE * TILE; idx += blockDim.x) {
        const int r = idx / TILE, c = idx % TILE;
        const int64_t i = (int64_t)tile.i * TILE + r, j = (int64_t)tile.j * TILE + c;
        if (i < N && j < N) {
            if (i < j) { const double v = src[idx]; out[i * N + j] = v; out[j * N + i] = v; }
            else if (i == j) out[i * N + i] = 0.0;
        }
    }
    (void)per_rank;
}

// ---- host side ------------------------------------------------------------------------------------

int64_t tiles_per_rank(int64_t n_tiles, int world) { return (n_tiles + world - 1) / world; }

static krd_status ensure_tiles(krd_handle* h)
{
    const int T = h->ld / TILE;
    const int64_t n_tiles = (int64_t)T * (T + 1) / 2;
    if (h->n_tiles_alloc == n_tiles && h->d_tiles) return KRD_OK;
    std::vector<Tile> tiles; tiles.reserve((size_t)n_tiles);
    // diagonal-major order: every contiguous range of tiles mixes rows and columns evenly, so a
    // contiguous per-rank share has equal cost and concurrently running CTAs share F row blocks.
    for (int d = 0; d < T; ++d)
        for (int i = 0; i + d < T; ++i) tiles.push_back(Tile{i, i + d});
    cudaFree(h->d_tiles); h->d_tiles = nullptr;
    KRD_CUDA(h, cudaMalloc(&h->d_tiles, (size_t)std::max<int64_t>(n_tiles, 1) * sizeof(Tile)));
    KRD_CUDA(h, cudaMemcpyAsync(h->d_tiles, tiles.data(), (size_t)n_tiles * sizeof(Tile), cudaMemcpyHostToDevice, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    h->n_tiles_alloc = n_tiles;
    h->stats.n_tiles = n_tiles;
    return KRD_OK;
}

// Split of the segment axis: minimise waves * chunks_per_split (time in chunk units).
static int choose_split(int64_t n_tiles, int n_chunks, int sms)
{
    int best = 1; int64_t best_cost = INT64_MAX;
    const int max_split = std::max(1, std::min(n_chunks / 8, 512));
    for (int s = 1; s <= max_split; ++s) {
        int64_t cps = (n_chunks + s - 1) / s;
        int64_t waves = (n_tiles * s + sms - 1) / sms;
        int64_t cost = waves * (cps + 6);   // +6 chunks ~ prologue/epilogue of a CTA
        if (cost < best_cost) { best_cost = cost; best = s; }
    }
    return best;
}

template <int MODE>
static cudaError_t launch_pairs(const PairParams& P, dim3 grid, cudaStream_t stream)
{
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(k_pairs<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    k_pairs<MODE><<<grid, NT, SMEM_BYTES, stream>>>(P);
    return cudaGetLastError();
}

krd_status run_pairs(krd_handle* h, double p, double divisor, int rank, int world,
                     double* d_out_matrix, double* d_out_packed)
{
    krd_status st = ensure_tiles(h);
    if (st != KRD_OK) return st;
    const int64_t n_tiles_all = h->n_tiles_alloc;
    const int64_t per_rank = tiles_per_rank(n_tiles_all, world);
    const int64_t t0 = std::min<int64_t>((int64_t)rank * per_rank, n_tiles_all);
    const int64_t t1 = std::min<int64_t>(t0 + per_rank, n_tiles_all);
    const int n_tiles = (int)(t1 - t0);
    if (d_out_packed && n_tiles < per_rank)   // ragged tail: keep the all-gather payload defined
        KRD_CUDA(h, cudaMemsetAsync(d_out_packed + (size_t)n_tiles * TILE * TILE, 0, (size_t)(per_rank - n_tiles) * TILE * TILE * 8, h->stream));
    h->stats.pair_ctas = 0;
    if (n_tiles == 0) return KRD_OK;

    const int n_chunks = (int)(h->S_pad / KC);
    int split = h->cfg.split > 0 ? std::min(h->cfg.split, n_chunks) : choose_split(n_tiles, n_chunks, h->num_sms);
    const int cps = (n_chunks + split - 1) / split;
    split = (n_chunks + cps - 1) / cps;   // drop empty splits
    const size_t ws_need = (size_t)split * n_tiles * TILE * TILE * sizeof(double);
    if (ws_need > h->ws_capacity) {
        cudaFree(h->d_ws); h->d_ws = nullptr; h->ws_capacity = 0;
        KRD_CUDA(h, cudaMalloc(&h->d_ws, ws_need));
        h->ws_capacity = ws_need;
    }

    int mode = MODE_GENERIC;
    if (p == 1.0) mode = MODE_P1; else if (p == 2.0) mode = MODE_P2; else if (p == 0.5) mode = MODE_PHALF;

    PairParams P;
    P.F = h->d_F; P.len = (mode == MODE_P2) ? h->d_sqrt_len : h->d_len;
    P.tiles = h->d_tiles + t0; P.ws = h->d_ws; P.ld = h->ld;
    P.n_chunks = n_chunks; P.chunks_per_split = cps; P.n_tiles = n_tiles; P.p = p;
    dim3 grid((unsigned)n_tiles, (unsigned)split);
    cudaError_t e;
    switch (mode) {
        case MODE_P1: e = launch_pairs<MODE_P1>(P, grid, h->stream); break;
        case MODE_P2: e = launch_pairs<MODE_P2>(P, grid, h->stream); break;
        case MODE_PHALF: e = launch_pairs<MODE_PHALF>(P, grid, h->stream); break;
        default: e = launch_pairs<MODE_GENERIC>(P, grid, h->stream); break;
    }
    if (e != cudaSuccess) return cuda_fail(h, e, "k_pairs launch");
    cudaEventRecord(h->ev[2], h->stream);

    EpiParams Q;
    Q.ws = h->d_ws; Q.tiles = h->d_tiles + t0; Q.n_tiles = n_tiles; Q.split = split; Q.N = h->N;
    Q.p = p; Q.divisor = divisor; Q.out = d_out_matrix; Q.packed = d_out_packed;
    k_epilogue<<<(unsigned)n_tiles, 256, 0, h->stream>>>(Q);
    KRD_CUDA(h, cudaGetLastError());
    h->stats.split = split;
    h->stats.pair_ctas = n_tiles * split;
    h->stats.launches += 2;
    return KRD_OK;
}

krd_status unpack_shards(krd_handle* h, int world, const double* d_gathered, double* d_out)
{
    krd_status st = ensure_tiles(h);
    if (st != KRD_OK) return st;
    const int64_t n_tiles = h->n_tiles_alloc;
    if (n_tiles == 0) return KRD_OK;
    k_unpack<<<(unsigned)n_tiles, 256, 0, h->stream>>>(d_gathered, h->d_tiles, n_tiles, tiles_per_rank(n_tiles, world), h->N, d_out);
    KRD_CUDA(h, cudaGetLastError());
    return KRD_OK;
}

}  // namespace krd
