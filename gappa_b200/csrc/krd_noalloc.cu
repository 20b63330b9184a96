// krd_matrix_from_F_noalloc — the pair reduction for callers that cannot call cudaMalloc (code under CUDA graph capture,
// tight serving loops, harnesses that forbid allocation).  No counterpart in the reference.
// It computes what krd_compute_device computes — KR_p(i,j) = (sum_s len[s] |F[s][i] - F[s][j]|^p)^(min(1,1/p))
// (genesis tree/mass_tree/emd.cpp:154,166,176-178) — from a dense cumulative-mass matrix the CALLER already holds on the
// device, for small cohorts, and allocates nothing: the only scratch, the per-chunk partial sums, lives in a file-scope
// static __device__ array of fixed size (below).
//
// Consequences of the static array, stated in the header too:
//  * every GPU that loads libkrd_b200.so holds the 8 MiB until the process exits;
//  * all calls share it and it keeps the last call's contents.  Nothing in it has to start at zero: every (chunk, pair)
//    cell that k_na_final reads is written by k_na_part of the same call first (empty chunks write 0), so no clearing pass
//    is needed — but two calls that overlap on the GPU (different streams, or different host threads) RACE on it.  Calls
//    on one stream are ordered and safe.
//  * in host code the array's name is a host-side stand-in, not a device address; the device address is taken once per
//    device with cudaGetSymbolAddress (with that device current) and handed to the kernels.
#include <algorithm>
#include <cstdio>
#include <mutex>

#include "krd_common.cuh"

namespace krd {

constexpr int NA_MAX_SAMPLES = KRD_NOALLOC_MAX_SAMPLES;   // 128
constexpr int NA_CHUNKS = 64;                             // segment chunks, summed in fixed order (deterministic)
constexpr int NA_ROWS = 32;                               // segment rows staged in shared memory at a time
constexpr size_t NA_SCRATCH_DOUBLES = (size_t)NA_CHUNKS * NA_MAX_SAMPLES * NA_MAX_SAMPLES;   // 1,048,576 doubles = 8 MiB

static __device__ __align__(256) double g_na_scratch[NA_SCRATCH_DOUBLES];

// part[chunk][i * n + j] (i < j) = sum over the chunk's segments of len[s] * |F[s][i] - F[s][j]|^p
__global__ void __launch_bounds__(256)
k_na_part(const double* __restrict__ F, const double* __restrict__ len, int64_t S, int64_t ld, int n, double p,
          int64_t rows_per_chunk, double* __restrict__ part)
{
    __shared__ double sF[NA_ROWS][NA_MAX_SAMPLES];
    __shared__ double sL[NA_ROWS];
    const int64_t s0 = (int64_t)blockIdx.x * rows_per_chunk, s1 = min(S, s0 + rows_per_chunk);
    double* out = part + (size_t)blockIdx.x * n * n;
    // the pairs of this thread: flat index q = i * n + j over the whole square, only i < j does work (n <= 128: 64 cells per thread)
    for (int q0 = 0; q0 < n * n; q0 += 256 * 8) {
        double acc[8];
        int pi[8], pj[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { const int q = q0 + u * 256 + threadIdx.x; acc[u] = 0.0; pi[u] = q / n; pj[u] = q % n; }
        for (int64_t r0 = s0; r0 < s1; r0 += NA_ROWS) {
            const int nr = (int)min((int64_t)NA_ROWS, s1 - r0);
            __syncthreads();
            for (int k = threadIdx.x; k < nr * n; k += 256) sF[k / n][k % n] = F[(r0 + k / n) * ld + k % n];
            if (threadIdx.x < nr) sL[threadIdx.x] = len[r0 + threadIdx.x];
            __syncthreads();
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                if (q0 + u * 256 + threadIdx.x >= n * n || pi[u] >= pj[u]) continue;
                double a = acc[u];
                for (int r = 0; r < nr; ++r) {
                    const double d = fabs(sF[r][pi[u]] - sF[r][pj[u]]);
                    const double v = (p == 1.0) ? d : (p == 2.0) ? d * d : (p == 0.5) ? sqrt(d) : pow(d, p);
                    a = fma(sL[r], v, a);
                }
                acc[u] = a;
            }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int q = q0 + u * 256 + threadIdx.x;
            if (q < n * n) out[q] = acc[u];          // (cells with i >= j: 0 — every cell of the call's range is rewritten)
        }
    }
}

__global__ void k_na_final(const double* __restrict__ part, int n, int n_chunks, double p, double divisor, double* __restrict__ out)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n * n) return;
    const int i = q / n, j = q % n;
    if (i > j) return;
    if (i == j) { out[q] = 0.0; return; }                                   // emd.cpp:193: zero diagonal
    double v = 0.0;
    for (int c = 0; c < n_chunks; ++c) v += part[(size_t)c * n * n + q];
    if (p > 1.0) v = (p == 2.0) ? sqrt(v) : pow(v, 1.0 / p);               // emd.cpp:176-178
    v = v / divisor;
    out[(size_t)i * n + j] = v; out[(size_t)j * n + i] = v;
}

// device address of the scratch array, resolved once per device with that device current
static krd_status na_scratch_ptr(int device, double** out)
{
    static std::mutex mu;
    static double* ptr[64] = {nullptr};
    if (device < 0 || device >= 64) return fail(nullptr, KRD_ERR_INVALID_ARGUMENT, "krd_matrix_from_F_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[device]) {
        double* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_na_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(nullptr, KRD_ERR_CUDA, std::string("krd_matrix_from_F_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
        }
        ptr[device] = p;
    }
    *out = ptr[device];
    return KRD_OK;
}

}  // namespace krd

using namespace krd;

extern "C" krd_status krd_matrix_from_F_noalloc(int32_t device, const double* d_F, const double* d_len, int64_t n_segments,
                                                 int32_t n_samples, int64_t ld, double p, double divisor, double* d_out,
                                                 void* cuda_stream)
{
    if (!(p > 0.0))   // emd.cpp:186-190
        return fail(nullptr, KRD_ERR_INVALID_EXPONENT, "Invalid exponent value p for earth mover's distance calculation. Has to be > 0.0.");
    if (n_samples < 0 || n_segments < 0 || ld < n_samples) return fail(nullptr, KRD_ERR_INVALID_ARGUMENT, "krd_matrix_from_F_noalloc: bad sizes");
    if (n_samples > NA_MAX_SAMPLES) {
        char msg[200];
        std::snprintf(msg, sizeof msg, "krd_matrix_from_F_noalloc: %d samples exceed the fixed scratch (at most %d samples; use krd_compute* for larger cohorts)",
                      n_samples, NA_MAX_SAMPLES);
        return fail(nullptr, KRD_ERR_INVALID_ARGUMENT, msg);
    }
    if (divisor == 0.0 || divisor != divisor) return fail(nullptr, KRD_ERR_INVALID_ARGUMENT, "divisor must be a non-zero number");
    if (n_samples == 0) return KRD_OK;
    if (!d_out || (n_segments > 0 && (!d_F || !d_len))) return fail(nullptr, KRD_ERR_INVALID_ARGUMENT, "krd_matrix_from_F_noalloc: null pointer");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaSetDevice");
    double* scratch = nullptr;
    krd_status st = na_scratch_ptr(device, &scratch);
    if (st != KRD_OK) return st;
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    const int64_t rpc = std::max<int64_t>(1, (n_segments + NA_CHUNKS - 1) / NA_CHUNKS);
    k_na_part<<<NA_CHUNKS, 256, 0, stream>>>(d_F, d_len, n_segments, ld, n_samples, p, rpc, scratch);
    k_na_final<<<(n_samples * n_samples + 255) / 256, 256, 0, stream>>>(scratch, n_samples, NA_CHUNKS, p, divisor, d_out);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "krd_matrix_from_F_noalloc launch");
    return KRD_OK;
}
