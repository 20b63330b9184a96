// Result exchange over peer memory (NVLink / NVSwitch), no collective: every rank scatters the tiles it computed
// STRAIGHT into the row blocks of the ranks that own those rows — the tile itself into the block of its rows, its mirror
// into the block of its columns — with ordinary stores through peer mappings.  What the NCCL variant does in three steps
// (all-gather of every rank's packed tiles to every rank: world x the matrix in flight; unpack of a row range; D2H), this
// does in one kernel that moves each cell exactly once to the one GPU that needs it (2 x 1/world of the matrix per rank).
// A stream-ordered barrier (any tiny collective, or krd_* callers' own) then tells a rank that its block is complete.
//
// One process per GPU: the row blocks are allocated here (cudaMalloc: IPC handles need whole allocations) and shared
// through CUDA IPC handles (krd_p2p_alloc / krd_p2p_open / krd_p2p_close); one process with several GPUs can pass plain
// device pointers after cudaDeviceEnablePeerAccess (krd_p2p_enable_peer).
#include <algorithm>

#include "krd_common.cuh"

namespace krd {

constexpr int P2P_MAX_WORLD = 16;
struct ScatterParams {
    const double* packed;          // this rank's tiles, [n_tiles][TILE*TILE]
    const Tile* tiles;             // this rank's share of the tile list
    int n_tiles, N, world;
    int32_t bounds[P2P_MAX_WORLD + 1];   // rank r owns matrix rows [bounds[r], bounds[r+1])
    double* rows[P2P_MAX_WORLD];         // rank r's row block ((bounds[r+1]-bounds[r]) x N), mapped into THIS process
};

__device__ __forceinline__ int owner_of(const ScatterParams& P, int row)
{
    int r = 0;
    while (r + 1 < P.world && row >= P.bounds[r + 1]) ++r;
    return r;
}

// One CTA per (tile, 16-row band).  Direct part: rows of block i, consecutive threads along j (1 KB contiguous per row).
// Mirror part: rows of block j, consecutive threads along i (128 B contiguous per row and band) — the band is staged in
// shared memory so that both the local reads and the remote writes are contiguous.
__global__ void __launch_bounds__(256) k_scatter_rows(const ScatterParams P)
{
    __shared__ double band[16][TILE + 1];
    const int t = blockIdx.x >> 3, b = blockIdx.x & 7;
    const Tile tile = P.tiles[t];
    const int i0 = tile.i * TILE + 16 * b, j0 = tile.j * TILE;
    const double* src = P.packed + (size_t)t * (TILE * TILE) + (size_t)16 * b * TILE;
    for (int k = threadIdx.x; k < 16 * TILE; k += 256) band[k / TILE][k % TILE] = src[k];
    __syncthreads();
    // direct: cell (i, j), i < j (the packed tile holds 0 elsewhere); the diagonal cell is written as 0
    for (int k = threadIdx.x; k < 16 * TILE; k += 256) {
        const int r = k / TILE, c = k % TILE;
        const int i = i0 + r, j = j0 + c;
        if (i >= P.N || j >= P.N || i > j) continue;
        const int o = owner_of(P, i);
        P.rows[o][(size_t)(i - P.bounds[o]) * P.N + j] = (i == j) ? 0.0 : band[r][c];
    }
    // mirror: cell (j, i) for i < j; thread -> (c = column of the tile = matrix row j, r = band row = matrix column i)
    for (int k = threadIdx.x; k < 16 * TILE; k += 256) {
        const int c = k / 16, r = k % 16;
        const int i = i0 + r, j = j0 + c;
        if (i >= P.N || j >= P.N || i >= j) continue;
        const int o = owner_of(P, j);
        P.rows[o][(size_t)(j - P.bounds[o]) * P.N + i] = band[r][c];
    }
}

// packedT[t] = the transpose of packed tile t (off-diagonal tiles), or the full symmetric tile with a zero diagonal
// (diagonal tiles: packed holds their strict upper triangle).  32 x 32 blocks through shared memory, both sides coalesced.
__global__ void __launch_bounds__(256) k_tile_transpose(const double* __restrict__ packed, const Tile* __restrict__ tiles,
                                                        double* __restrict__ packedT)
{
    __shared__ double a[32][33], b[32][33];
    const int t = blockIdx.x, br = blockIdx.y >> 2, bc = blockIdx.y & 3;     // 4 x 4 blocks of 32 x 32
    const bool diag = tiles[t].i == tiles[t].j;
    const double* src = packed + (size_t)t * (TILE * TILE);
    double* dst = packedT + (size_t)t * (TILE * TILE);
    const int x = threadIdx.x & 31, y0 = threadIdx.x >> 5;
    for (int y = y0; y < 32; y += 8) {
        a[y][x] = src[(32 * br + y) * TILE + 32 * bc + x];                    // block (br, bc)
        if (diag) b[y][x] = src[(32 * bc + y) * TILE + 32 * br + x];          // block (bc, br)
    }
    __syncthreads();
    for (int y = y0; y < 32; y += 8) {
        if (!diag) dst[(32 * bc + y) * TILE + 32 * br + x] = a[x][y];         // transposed block goes to (bc, br)
        else {
            const int r = 32 * br + y, c = 32 * bc + x;                       // symmetric: upper from (r, c), lower from (c, r)
            dst[r * TILE + c] = (r < c) ? a[y][x] : (r > c) ? b[x][y] : 0.0;
        }
    }
}

}  // namespace krd

using namespace krd;

extern "C" {

krd_status krd_scatter_shard_rows(krd_handle* h, int32_t rank, int32_t world, const double* d_packed, const int32_t* row_bounds,
                                  double* const* d_rows)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!h->have_masses) return fail(h, KRD_ERR_STATE, "no masses set");
    if (world < 1 || rank < 0 || rank >= world) return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_scatter_shard_rows: bad rank/world");
    int64_t t0 = 0, t1 = 0;
    krd_rank_tiles(krd_tile_count(h->N), rank, world, &t0, &t1);
    return krd_scatter_tile_range_rows(h, t0, t1, d_packed, world, row_bounds, d_rows);
}

krd_status krd_scatter_tile_range_rows(krd_handle* h, int64_t t0, int64_t t1, const double* d_packed, int32_t world,
                                       const int32_t* row_bounds, double* const* d_rows)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!h->have_masses) return fail(h, KRD_ERR_STATE, "no masses set");
    if (world < 1 || world > P2P_MAX_WORLD || !d_packed || !row_bounds || !d_rows || t0 < 0 || t1 < t0 || t1 > krd_tile_count(h->N))
        return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_scatter_tile_range_rows: bad arguments (at most 16 owners)");
    if (row_bounds[0] != 0 || row_bounds[world] != h->N) return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_scatter_shard_rows: row_bounds must run from 0 to n_samples");
    for (int r = 0; r < world; ++r)
        if (row_bounds[r + 1] < row_bounds[r] || (row_bounds[r + 1] > row_bounds[r] && !d_rows[r]))
            return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_scatter_shard_rows: row_bounds must be non-decreasing and every non-empty block needs a pointer");
    KRD_CUDA(h, cudaSetDevice(h->device));
    if (h->piped) { krd_status fs = pipe_flatten(h); if (fs != KRD_OK) return fs; }
    const krd::Tile* d_tiles = nullptr; int64_t u0 = 0, u1 = 0;
    krd_status st = rank_tile_range(h, 0, 1, &d_tiles, &u0, &u1);
    if (st != KRD_OK) return st;
    if (t1 == t0) return KRD_OK;
    ScatterParams P;
    P.packed = d_packed; P.tiles = d_tiles + t0; P.n_tiles = (int)(t1 - t0); P.N = h->N; P.world = world;
    for (int r = 0; r <= world; ++r) P.bounds[r] = row_bounds[r];
    for (int r = 0; r < P2P_MAX_WORLD; ++r) P.rows[r] = r < world ? d_rows[r] : nullptr;
    k_scatter_rows<<<(unsigned)P.n_tiles * 8, 256, 0, h->stream>>>(P);
    KRD_CUDA(h, cudaGetLastError());
    return KRD_OK;
}

// The rank's packed tiles -> the caller's n x n HOST matrix with the COPY ENGINE (one 2-D copy per tile and one per mirror),
// on a stream of the caller's choice — so that the copies of one tile share run under the pair kernel of the next one (the
// pair kernel fills every SM, a copy kernel could not run beside it; the copy engines can).
krd_status krd_tile_range_to_host_async(krd_handle* h, int64_t t0, int64_t t1, const double* d_packed, double* d_scratch,
                                        double* host_matrix, void* copy_stream)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!h->have_masses) return fail(h, KRD_ERR_STATE, "no masses set");
    if (!d_packed || !d_scratch || !host_matrix || t0 < 0 || t1 < t0 || t1 > krd_tile_count(h->N))
        return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_tile_range_to_host_async: bad arguments");
    KRD_CUDA(h, cudaSetDevice(h->device));
    if (h->piped) { krd_status fs = pipe_flatten(h); if (fs != KRD_OK) return fs; }
    const krd::Tile* d_tiles = nullptr; int64_t u0 = 0, u1 = 0;
    krd_status st = rank_tile_range(h, 0, 1, &d_tiles, &u0, &u1);
    if (st != KRD_OK) return st;
    if (t1 == t0) return KRD_OK;
    cudaStream_t cs = static_cast<cudaStream_t>(copy_stream);
    // transposes on the compute stream (they follow the epilogue that wrote d_packed), then the copies on `cs`
    k_tile_transpose<<<dim3((unsigned)(t1 - t0), 16), 256, 0, h->stream>>>(d_packed, d_tiles + t0, d_scratch);
    KRD_CUDA(h, cudaGetLastError());
    if (cs != h->stream) {
        KRD_CUDA(h, cudaEventRecord(h->ev_copy, h->stream));
        KRD_CUDA(h, cudaStreamWaitEvent(cs, h->ev_copy, 0));
    }
    const int64_t N = h->N;
    const size_t pitch = (size_t)N * 8, tp = (size_t)TILE * 8;
    for (int64_t t = t0; t < t1; ++t) {
        const krd::Tile tile = h->h_tiles[(size_t)t];
        const int64_t i0 = (int64_t)tile.i * TILE, j0 = (int64_t)tile.j * TILE;
        const size_t hh = (size_t)std::min<int64_t>(TILE, N - i0), ww = (size_t)std::min<int64_t>(TILE, N - j0);
        const double* P = d_packed + (size_t)(t - t0) * (TILE * TILE);
        const double* T = d_scratch + (size_t)(t - t0) * (TILE * TILE);
        if (tile.i == tile.j) {
            KRD_CUDA(h, cudaMemcpy2DAsync(host_matrix + i0 * N + j0, pitch, T, tp, ww * 8, hh, cudaMemcpyDeviceToHost, cs));
        } else {
            KRD_CUDA(h, cudaMemcpy2DAsync(host_matrix + i0 * N + j0, pitch, P, tp, ww * 8, hh, cudaMemcpyDeviceToHost, cs));
            KRD_CUDA(h, cudaMemcpy2DAsync(host_matrix + j0 * N + i0, pitch, T, tp, hh * 8, ww, cudaMemcpyDeviceToHost, cs));
        }
    }
    return KRD_OK;
}

krd_status krd_shard_to_host_async(krd_handle* h, int32_t rank, int32_t world, const double* d_packed, double* d_scratch,
                                   double* host_matrix, void* copy_stream)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!h->have_masses) return fail(h, KRD_ERR_STATE, "no masses set");
    if (world < 1 || rank < 0 || rank >= world) return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_shard_to_host_async: bad rank/world");
    int64_t t0 = 0, t1 = 0;
    krd_rank_tiles(krd_tile_count(h->N), rank, world, &t0, &t1);
    return krd_tile_range_to_host_async(h, t0, t1, d_packed, d_scratch, host_matrix, copy_stream);
}

krd_status krd_set_late_blocks(krd_handle* h, int32_t n_late_blocks, int64_t* t_split)
{
    if (!h) return KRD_ERR_INVALID_ARGUMENT;
    if (!h->have_masses) return fail(h, KRD_ERR_STATE, "no masses set");
    if (n_late_blocks < 0) return fail(h, KRD_ERR_INVALID_ARGUMENT, "krd_set_late_blocks: n_late_blocks must be >= 0");
    KRD_CUDA(h, cudaSetDevice(h->device));
    if (h->piped) { krd_status fs = pipe_flatten(h); if (fs != KRD_OK) return fs; }
    return set_late_blocks(h, n_late_blocks, t_split);
}

krd_status krd_copy_rect_to_host_async(double* host_dst, int64_t host_pitch_doubles, const double* d_src, int64_t src_pitch_doubles,
                                       int64_t cols, int64_t rows, void* cuda_stream)
{
    if (rows <= 0 || cols <= 0) return KRD_OK;
    if (!host_dst || !d_src || host_pitch_doubles < cols || src_pitch_doubles < cols) return fail(nullptr, KRD_ERR_INVALID_ARGUMENT, "krd_copy_rect_to_host_async: bad arguments");
    const cudaError_t e = cudaMemcpy2DAsync(host_dst, (size_t)host_pitch_doubles * 8, d_src, (size_t)src_pitch_doubles * 8, (size_t)cols * 8, (size_t)rows,
                                            cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(cuda_stream));
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaMemcpy2DAsync");
    return KRD_OK;
}

krd_status krd_p2p_alloc(int32_t device, int64_t bytes, void** d_ptr, unsigned char* ipc_handle_64)
{
    if (!d_ptr || bytes <= 0) return KRD_ERR_INVALID_ARGUMENT;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaSetDevice");
    krd_status st = alloc_sized(nullptr, d_ptr, (size_t)bytes, "a row block of the distance matrix (peer-shared)");
    if (st != KRD_OK) return st;
    if (ipc_handle_64) {
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle is 64 bytes");
        cudaIpcMemHandle_t hd;
        if ((e = cudaIpcGetMemHandle(&hd, *d_ptr)) != cudaSuccess) { cudaFree(*d_ptr); *d_ptr = nullptr; return cuda_fail(nullptr, e, "cudaIpcGetMemHandle"); }
        std::copy(reinterpret_cast<unsigned char*>(&hd), reinterpret_cast<unsigned char*>(&hd) + 64, ipc_handle_64);
    }
    return KRD_OK;
}

krd_status krd_p2p_open(int32_t device, const unsigned char* ipc_handle_64, void** d_ptr)
{
    if (!d_ptr || !ipc_handle_64) return KRD_ERR_INVALID_ARGUMENT;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaSetDevice");
    cudaIpcMemHandle_t hd;
    std::copy(ipc_handle_64, ipc_handle_64 + 64, reinterpret_cast<unsigned char*>(&hd));
    if ((e = cudaIpcOpenMemHandle(d_ptr, hd, cudaIpcMemLazyEnablePeerAccess)) != cudaSuccess) return cuda_fail(nullptr, e, "cudaIpcOpenMemHandle");
    return KRD_OK;
}

void krd_p2p_close(void* d_ptr) { if (d_ptr) { cudaIpcCloseMemHandle(d_ptr); cudaGetLastError(); } }
void krd_p2p_free(void* d_ptr) { if (d_ptr) { cudaFree(d_ptr); cudaGetLastError(); } }

krd_status krd_p2p_enable_peer(int32_t device, int32_t peer)
{
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaSetDevice");
    int can = 0;
    if ((e = cudaDeviceCanAccessPeer(&can, device, peer)) != cudaSuccess) return cuda_fail(nullptr, e, "cudaDeviceCanAccessPeer");
    if (!can) return fail(nullptr, KRD_ERR_STATE, "krd_p2p_enable_peer: the two devices have no peer access");
    e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(nullptr, e, "cudaDeviceEnablePeerAccess");
    cudaGetLastError();
    return KRD_OK;
}

}  // extern "C"
