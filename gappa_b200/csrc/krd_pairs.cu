// K2/K3: all-pairs weighted-Lp reduction over the dense matrix F[S_pad][ld] (segment-major):
//     W_p(i,j) = sum_s len[s] * |F[s][i] - F[s][j]|^p          (genesis emd.cpp:154,166)
//     KR_p     = W_p            if p <= 1
//              = W_p^(1/p)      if p  > 1                       (genesis emd.cpp:176-178)
// for every i < j, written symmetrically with a zero diagonal (emd.cpp:193,214-215).
//
// Shape: a "GEMM with a different inner operator".  CTA tile 128 x 128 pairs, 256 threads, each
// thread an 8 x 8 register tile (64 FP64 accumulators).  The segment axis is streamed through a
// shared-memory ring filled by the TMA engine (cp.async.bulk, 1 KB rows, completion counted on a
// "full" mbarrier per stage; consumer warps release a stage on its "empty" mbarrier, so warps drift
// instead of meeting at a CTA barrier).
//
// Scheduling is persistent "stream-K": the (tile, chunk) iteration space of the upper triangle is
// cut into one contiguous, equal-cost span per SM (diagonal tiles are cheaper: their all-lower
// 16x32 blocks are skipped warp-uniformly).  A span is a short list of work items (tile, chunk
// range); each item leaves a partial 128x128 tile in a workspace and a small deterministic epilogue
// kernel adds the partials of a tile in fixed order, applies the outer root / divisor and mirrors.
// FP64 CUDA cores only: per (pair, segment) p = 1 costs DADD + DFMA, the floor for this operator.
#include <algorithm>
#include <cstdlib>
#include <cmath>
#include <type_traits>

#include "krd_common.cuh"
#include "krd_ops.cuh"

namespace krd {

constexpr int ST = 6;          // ring stages
constexpr int LAG = 2;         // a stage is refilled LAG chunks after it was consumed
constexpr int NT = 256;        // threads per CTA (8 warps; all compute, the TMA-issue role rotates)
constexpr int STAGE_DOUBLES = 2 * KC * TILE + 2 * KC;  // A rows, B rows, len, centre (Gram mode)
constexpr size_t SMEM_BYTES = (size_t)ST * STAGE_DOUBLES * sizeof(double) + 2 * ST * sizeof(uint64_t);
// Long inner operators (p = 1/2, integer p >= 5, general p) keep only HALF of the thread's 8 x 8 accumulators in
// registers: every chunk is walked twice (rows 0-3, rows 4-7) and the idle half waits in shared memory.  With all 64
// accumulators resident the compiler has ~45 doubles left for the operators' temporaries, interleaves 4-5 dependency
// chains and the FP64 latency shows (computing only four rows ran 1.41x faster per operator, profiles/r1o_long_ops.txt).
// These modes need a fraction of the operand bandwidth, so the ring shrinks to 4 stages and leaves room for the spill.
#ifndef KRD_PHALF_HALVED
#define KRD_PHALF_HALVED 1
#endif
constexpr bool mode_is_halved(int mode) { return (mode == MODE_PHALF && KRD_PHALF_HALVED) || mode == MODE_GENERIC || mode == MODE_LIBPOW; }
constexpr int stages_for(int mode) { return mode_is_halved(mode) ? 4 : ST; }
constexpr size_t SPILL_BYTES = (size_t)32 * NT * sizeof(double);
constexpr size_t smem_bytes_for(int mode)
{
    return (size_t)stages_for(mode) * STAGE_DOUBLES * sizeof(double) + 2 * stages_for(mode) * sizeof(uint64_t) + (mode_is_halved(mode) ? SPILL_BYTES : 0);
}
constexpr int FIX_CAP = 1 << 20;    // capacity of the Gram path's fix-up list (pairs)
constexpr double DIAG_COST_SKIP = 0.705;  // measured (profiles/cta_balance.py) cost of a diagonal-tile chunk relative to a full one (5/8 of the math, same pipeline overheads)

struct WorkItem { int32_t tile, c0, c1, pad; };   // chunks [c0, c1) of tile `tile` (index into the rank's tile list)

struct PairParams {
    const double* F;
    const double* len;     // len or sqrt(len) (MODE_P2)
    const Tile* tiles;
    const WorkItem* items;
    const int32_t* cta_item_off;   // gridDim.x + 1
    double* ws;            // [item][TILE*TILE]
    int64_t ld;
    double p;
    int N;                 // samples (columns >= N are zero padding: ragged last block)
    PowCoef pc;            // general p: p times the log2 polynomial's coefficients (constant-bank operands of the DFMAs)
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
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
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

// Tuning aid, compiled in only with -DKRD_TIMING (make EXTRA=-DKRD_TIMING ..., profiles/cta_balance.py): per-CTA
// (start, end, first item, last item) and per-item (start, end) of the last pair launch in %globaltimer ns.
#ifdef KRD_TIMING
__device__ unsigned long long g_cta_time[4 * 1024];
__device__ unsigned long long g_item_time[2 * 4096];
#define KRD_STAMP(dst) do { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); (dst) = t_; } while (0)
#endif

// One chunk (KC segments) of one tile for one thread; column blocks q < QMIN are skipped
// (diagonal tiles: blocks that lie entirely below the diagonal).
#ifndef KRD_P12_UNROLL
#define KRD_P12_UNROLL 2
#endif
#ifndef KRD_P34_UNROLL
#define KRD_P34_UNROLL 2
#endif
// unroll factor of the segment loop: the short operators (p = 1, 2) are unrolled 4x; the long ones keep ONE segment
// (64 independent operators, fixed accumulator registers) as the loop body so that it stays inside the instruction cache
#ifndef KRD_PHALF_UNROLL
#define KRD_PHALF_UNROLL 1
#endif
#ifndef KRD_GENERIC_UNROLL
#define KRD_GENERIC_UNROLL 1
#endif
// segment-loop unrolling of the halved modes' four-row body (chunk_math_rows4)
constexpr int rows4_unroll(int mode) { return mode == MODE_PHALF ? KRD_PHALF_UNROLL : mode == MODE_GENERIC ? KRD_GENERIC_UNROLL : 1; }
constexpr int mode_unroll(int mode) { return (mode == MODE_PHALF || mode == MODE_INT || mode == MODE_GENERIC || mode == MODE_LIBPOW) ? 1 : (mode == MODE_P3 || mode == MODE_P4) ? KRD_P34_UNROLL : KRD_P12_UNROLL; }

// Integer exponents 5..16 with the exponent as a template parameter: one segment = 64 independent, fully static operator
// chains that the compiler interleaves freely (the run-time ladder kr_row_powi had to stop at every warp-uniform branch:
// p = 5 ran at 64 % of its FP64 issue bound, p = 3 with a static body at 90 %).  Only the body of the exponent in use
// (and its diagonal-tile variants) is hot, so the instruction cache sees one ~7 KB loop, not twelve.
template <int QMIN, int N>
__device__ __forceinline__ void chunk_math_int(const double* __restrict__ pa, const double* __restrict__ pb,
                                            const double* __restrict__ sL, double (&acc)[8][8])
{
#pragma unroll 1
    for (int kk = 0; kk < KC; ++kk) {
        double a[8], b[8];
        const double l = sL[kk];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double2 va = *reinterpret_cast<const double2*>(pa + kk * TILE + 2 * q);
            a[2 * q] = va.x; a[2 * q + 1] = va.y;
        }
#pragma unroll
        for (int q = QMIN; q < 4; ++q) {
            const double2 vb = *reinterpret_cast<const double2*>(pb + kk * TILE + 32 * q);
            b[2 * q] = vb.x; b[2 * q + 1] = vb.y;
        }
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int cc = 2 * QMIN; cc < 8; ++cc) acc[r][cc] = fma(l, kr_pow_n<N>(fabs(a[r] - b[cc])), acc[r][cc]);
    }
}

// Short operators (p = 1, 2, 3, 4), OFF-DIAGONAL tiles: the thread's eight columns are 16 q + tx, q = 0..7 — one column per
// lane and step — so that whole steps can be skipped CTA-uniformly at a granularity of 16 columns: steps q >= QMAX on the
// tiles of a ragged last column block (1,000 samples: 104 of 128 columns -> 7 of 8 steps instead of 8).  (QMIN is kept
// for experiments; diagonal tiles use the 32-column-block body chunk_math<MODE, QMIN>.)  B operand fetch is one LDS.64 per step (16 lanes
// x 8 B contiguous, the other half-warp reads the same addresses), A stays LDS.128 broadcast.
#ifndef KRD_FINE
#define KRD_FINE 1      // 0: every tile on the 32-column-block body (A/B measurements: make EXTRA=-DKRD_FINE=0 ...)
#endif
constexpr bool mode_is_fine(int mode) { return KRD_FINE && (mode == MODE_P1 || mode == MODE_P2 || mode == MODE_P2S || mode == MODE_P3 || mode == MODE_P4); }

#ifndef KRD_FINE_UNROLL
#define KRD_FINE_UNROLL 2
#endif
constexpr int fine_unroll(int mode) { return mode_unroll(mode) < KRD_FINE_UNROLL ? mode_unroll(mode) : KRD_FINE_UNROLL; }
template <int MODE, int QMIN, int QMAX>
__device__ __forceinline__ void chunk_math16(const double* __restrict__ pa, const double* __restrict__ pb,
                                             const double* __restrict__ sL, double (&acc)[8][8])
{
#pragma unroll (fine_unroll(MODE))
    for (int kk = 0; kk < KC; ++kk) {
        double a[8], b[8];
        const double l = sL[kk];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double2 va = *reinterpret_cast<const double2*>(pa + kk * TILE + 2 * q);
            a[2 * q] = va.x; a[2 * q + 1] = va.y;
        }
#pragma unroll
        for (int q = QMIN; q < QMAX; ++q) b[q] = pb[kk * TILE + 16 * q];
        if (MODE == MODE_P2) {
#pragma unroll
            for (int q = 0; q < 8; ++q) a[q] *= l;
#pragma unroll
            for (int q = QMIN; q < QMAX; ++q) b[q] *= l;
        }
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int q = QMIN; q < QMAX; ++q) acc[r][q] = pair_op<MODE, 1>(a[r], b[q], l, 0.0, nullptr, acc[r][q]);
    }
}

template <int MODE>
__device__ __forceinline__ void chunk_math16_dispatch(int qmax, const double* __restrict__ pa, const double* __restrict__ pb,
                                                      const double* __restrict__ sL, double (&acc)[8][8])
{
    switch (qmax) {   // CTA-uniform; < 8 only on the off-diagonal tiles of a ragged last column block
        case 8: chunk_math16<MODE, 0, 8>(pa, pb, sL, acc); break;
        case 1: chunk_math16<MODE, 0, 1>(pa, pb, sL, acc); break;
        case 2: chunk_math16<MODE, 0, 2>(pa, pb, sL, acc); break;
        case 3: chunk_math16<MODE, 0, 3>(pa, pb, sL, acc); break;
        case 4: chunk_math16<MODE, 0, 4>(pa, pb, sL, acc); break;
        case 5: chunk_math16<MODE, 0, 5>(pa, pb, sL, acc); break;
        case 6: chunk_math16<MODE, 0, 6>(pa, pb, sL, acc); break;
        default: chunk_math16<MODE, 0, 7>(pa, pb, sL, acc); break;
    }
}

template <int MODE, int QMIN>
__device__ __forceinline__ void chunk_math(const double* __restrict__ pa, const double* __restrict__ pb,
                                           const double* __restrict__ sL, double p, const double* __restrict__ tab,
                                           double (&acc)[8][8])
{
    if (MODE == MODE_INT) {
        switch ((int)p) {   // warp-uniform
            case 5: chunk_math_int<QMIN, 5>(pa, pb, sL, acc); return;
            case 6: chunk_math_int<QMIN, 6>(pa, pb, sL, acc); return;
            case 7: chunk_math_int<QMIN, 7>(pa, pb, sL, acc); return;
            case 8: chunk_math_int<QMIN, 8>(pa, pb, sL, acc); return;
            case 9: chunk_math_int<QMIN, 9>(pa, pb, sL, acc); return;
            case 10: chunk_math_int<QMIN, 10>(pa, pb, sL, acc); return;
            case 11: chunk_math_int<QMIN, 11>(pa, pb, sL, acc); return;
            case 12: chunk_math_int<QMIN, 12>(pa, pb, sL, acc); return;
            case 13: chunk_math_int<QMIN, 13>(pa, pb, sL, acc); return;
            case 14: chunk_math_int<QMIN, 14>(pa, pb, sL, acc); return;
            case 15: chunk_math_int<QMIN, 15>(pa, pb, sL, acc); return;
            case 16: chunk_math_int<QMIN, 16>(pa, pb, sL, acc); return;
            default: return;   // (mode_for_exponent sends nothing else here)
        }
    }
#pragma unroll (mode_unroll(MODE))
    for (int kk = 0; kk < KC; ++kk) {
        double a[8], b[8];
        const double l = sL[kk];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double2 va = *reinterpret_cast<const double2*>(pa + kk * TILE + 2 * q);
            a[2 * q] = va.x; a[2 * q + 1] = va.y;
        }
#pragma unroll
        for (int q = QMIN; q < 4; ++q) {
            const double2 vb = *reinterpret_cast<const double2*>(pb + kk * TILE + 32 * q);
            b[2 * q] = vb.x; b[2 * q + 1] = vb.y;
        }
        if (MODE == MODE_P2) {
#pragma unroll
            for (int q = 0; q < 8; ++q) a[q] *= l;
#pragma unroll
            for (int q = 2 * QMIN; q < 8; ++q) b[q] *= l;
        }
        if (MODE == MODE_PHALF) {
#pragma unroll
            for (int r = 0; r < 8; ++r) kr_row_sqrt<8 - 2 * QMIN>(a[r], b + 2 * QMIN, l, &acc[r][2 * QMIN]);
        } else {
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int cc = 2 * QMIN; cc < 8; ++cc) acc[r][cc] = pair_op<MODE, POW_TS_PAIR>(a[r], b[cc], l, p, tab, acc[r][cc]);
        }
    }
}

// Four rows (4 * half .. 4 * half + 3 of the thread's eight) of one chunk, for the halved modes.
template <int MODE, int QMIN>
__device__ __forceinline__ void chunk_math_rows4(const double* __restrict__ pa, const double* __restrict__ pb,
                                                 const double* __restrict__ sL, double p, const double* __restrict__ tab,
                                                 int half, double (&acc)[4][8], const PowCoef& pc)
{
    constexpr int NC = 8 - 2 * QMIN;
#pragma unroll (rows4_unroll(MODE))
    for (int kk = 0; kk < KC; ++kk) {
        double a[4], b[8];
        const double l = sL[kk];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const double2 va = *reinterpret_cast<const double2*>(pa + kk * TILE + 4 * half + 2 * q);
            a[2 * q] = va.x; a[2 * q + 1] = va.y;
        }
#pragma unroll
        for (int q = QMIN; q < 4; ++q) {
            const double2 vb = *reinterpret_cast<const double2*>(pb + kk * TILE + 32 * q);
            b[2 * q] = vb.x; b[2 * q + 1] = vb.y;
        }
        if (MODE == MODE_PHALF) {
#pragma unroll
            for (int r = 0; r < 4; ++r) kr_row_sqrt<NC>(a[r], b + 2 * QMIN, l, &acc[r][2 * QMIN]);
        } else if (MODE == MODE_GENERIC && KRD_POW_FAST16) {
            const uint32_t tb = (uint32_t)__cvta_generic_to_shared(tab);
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int cc = 2 * QMIN; cc < 8; ++cc) kr_pow_acc16(a[r] - b[cc], p, l, tb, acc[r][cc], pc);
        } else {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int cc = 2 * QMIN; cc < 8; ++cc) acc[r][cc] = pair_op<MODE, POW_TS_PAIR>(a[r], b[cc], l, p, tab, acc[r][cc]);
        }
    }
}

// ---- Gram form on the FP64 tensor cores (p = 2) ---------------------------------------------------
// Warp tile 32 x 64 of the 128 x 128 CTA tile = 4 x 8 DMMA m8n8k4 accumulators (64 doubles per
// thread, as in the CUDA-core modes).  Fragments (PTX ISA, mma.m8n8k4 .f64): A row-major 8x4 -> the
// thread holds A[lane/4][lane%4]; B column-major 4x8 -> B[lane%4][lane/4]; C/D 8x8 -> row lane/4,
// columns 2*(lane%4) + {0,1}.  Both operands are F[k][sample] rows of the same shared-memory ring.
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// The operands arrive pre-centred and pre-scaled (Ft = sqrt(len) (F - c), written once per F by k_gram_prescale): the loop
// is LDS + DMMA only.  FP64 vector instructions and DMMA share one datapath (profiles/r1q_pipes_overlap.txt), so the 4 DADD +
// 8 DADD + 8 DMUL per 32 DMMA that round 1 spent on centring and scaling inside this loop cost tensor cycles one for one.
__device__ __forceinline__ void chunk_gram(const double* __restrict__ sA, const double* __restrict__ sB,
                                           int wr, int wc, int lane, double (&acc)[8][8])
{
    const int kq = lane & 3, rq = lane >> 2;
#pragma unroll
    for (int kk = 0; kk < KC; kk += 4) {
        const double* pa = sA + (kk + kq) * TILE + wr + rq;
        const double* pb = sB + (kk + kq) * TILE + wc + rq;
        double a[4], b[8];
#pragma unroll
        for (int mi = 0; mi < 4; ++mi) a[mi] = pa[8 * mi];
#pragma unroll
        for (int ni = 0; ni < 8; ++ni) b[ni] = pb[8 * ni];
#pragma unroll
        for (int mi = 0; mi < 4; ++mi)
#pragma unroll
            for (int ni = 0; ni < 8; ++ni) dmma884(acc[2 * mi][ni], acc[2 * mi + 1][ni], a[mi], b[ni]);
    }
}

// ---- the persistent tile kernel -------------------------------------------------------------------

template <int MODE>
__global__ void __launch_bounds__(NT, 1) k_pairs(const PairParams P)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int STM = stages_for(MODE);   // ring depth of this mode
    double* stage_base = reinterpret_cast<double*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)STM * STAGE_DOUBLES * sizeof(double));
    uint64_t* empty = full + STM;
    // halved modes: the thread's idle half of the accumulators, [32][NT] (conflict-free: consecutive threads, consecutive words)
    double* spill = reinterpret_cast<double*>(smem_raw + (size_t)STM * STAGE_DOUBLES * sizeof(double) + 2 * STM * sizeof(uint64_t)) + threadIdx.x;

    // general p: the pow tables, every entry POW_TS_PAIR times (lane L reads copy L & 15: conflict-free for any index)
    // kr_pow_acc16 forms entry addresses with OR: the tables must start at a multiple of 8 KB of the shared window.  An
    // alignment attribute does not give that (the first 1 KB of the window is reserved and the section starts behind it),
    // so the array is 8 KB longer and the start is rounded up at run time.  The exp entries carry hi - (i << 15).
    constexpr int TAB_PAD = (MODE == MODE_GENERIC && KRD_POW_FAST16) ? 8192 / 8 : 0;
    __shared__ __align__(16) double s_tab_raw[MODE == MODE_GENERIC ? POW_TAB * POW_TS_PAIR + TAB_PAD : 1];
    double* const s_tab_all = s_tab_raw + (TAB_PAD ? ((8192u - ((uint32_t)__cvta_generic_to_shared(s_tab_raw) & 8191u)) & 8191u) / 8 : 0);
    if (MODE == MODE_GENERIC) for (int k = threadIdx.x; k < POW_TAB * POW_TS_PAIR; k += NT) {   // visible after the barrier below
        const int ent = k / POW_TS_PAIR;
        double v = g_pow_tab[ent];
        if (KRD_POW_FAST16 && KRD_POW_PSCALE && ent >= POW_LOGN && ent < 2 * POW_LOGN) v *= P.p;
        if (KRD_POW_FAST16 && ent >= 2 * POW_LOGN) v = __hiloint2double(__double2hiint(v) - ((ent - 2 * POW_LOGN) << 15), __double2loint(v));
        s_tab_all[k] = v;
    }
    const double* const s_tab = s_tab_all + (MODE == MODE_GENERIC ? (threadIdx.x & (POW_TS_PAIR - 1)) : 0);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tx = lane & 15;
    // row group of the warp: SM sub-partition s hosts warps s and s+4 -> row groups s and 7-s, which
    // together skip 3 of 8 column blocks on a diagonal tile (balanced FP64 pipes)
    const int rg = (warp < 4) ? warp : 11 - warp;
    const int row0 = 16 * rg + 8 * (lane >> 4);

    const int it0 = P.cta_item_off[blockIdx.x], it1 = P.cta_item_off[blockIdx.x + 1];
    if (it0 == it1) return;
#ifdef KRD_TIMING
    if (threadIdx.x == 0 && blockIdx.x < 1024) { KRD_STAMP(g_cta_time[4 * blockIdx.x]); g_cta_time[4 * blockIdx.x + 2] = it0; g_cta_time[4 * blockIdx.x + 3] = it1; }
#endif

    if (tid == 0) {
        for (int s = 0; s < STM; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NT / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    // ---- producer role, rotated over the 8 warps (chunk j of the flattened sequence is issued by
    // warp j % 8) so that no warp carries extra work and falls behind its sub-partition mate.  Every
    // warp walks the same (item, chunk) sequence; only the owner waits for the stage and issues.
    int p_item = it0, p_flat = 0;
    WorkItem p_w = P.items[it0];
    int p_chunk = p_w.c0;
    auto produce_one = [&]() {   // advance the flattened sequence by one chunk; issue it if it is ours
        if (p_item >= it1) return;
        if ((p_flat & 7) == warp) {
            const int s = p_flat % STM;
            if (p_flat >= STM) mbar_wait(&empty[s], (uint32_t)(((p_flat / STM) - 1) & 1));
            const Tile t = P.tiles[p_w.tile];
            const bool dg = (t.i == t.j);
            double* sA = stage_base + (size_t)s * STAGE_DOUBLES;
            double* sB = sA + KC * TILE;
            double* sL = sB + KC * TILE;
            const int64_t k0 = (int64_t)p_chunk * KC;
            if (lane == 0) mbar_expect_tx(&full[s], (uint32_t)((dg ? 1 : 2) * KC * TILE * 8 + KC * 8));
            __syncwarp();
            const int r = lane & 15;
            if (lane < 16) tma_load_1d(sA + r * TILE, P.F + (int64_t)t.i * TILE + (k0 + r) * P.ld, TILE * 8, &full[s]);
            else if (!dg) tma_load_1d(sB + r * TILE, P.F + (int64_t)t.j * TILE + (k0 + r) * P.ld, TILE * 8, &full[s]);
            if (lane == 0) tma_load_1d(sL, P.len + k0, KC * 8, &full[s]);
        }
        ++p_flat;
        if (++p_chunk >= p_w.c1) {
            if (++p_item < it1) { p_w = P.items[p_item]; p_chunk = p_w.c0; }
        }
    };
    for (int c = 0; c < STM - LAG; ++c) produce_one();

    int flat = 0;
    for (int item = it0; item < it1; ++item) {
        const WorkItem w = P.items[item];
        const Tile t = P.tiles[w.tile];
        const bool diag = (t.i == t.j);
        // Diagonal tiles keep the 32-column-block layout for every mode (four skip variants at once): in the 16-column layout
        // their chunks cost 0.85 of a full one instead of 0.705 (measured at cfg2, eight variants thrashed the instruction
        // cache outright: 3.5 -> 4.7 ms), which ate the whole gain of the ragged-block skip.  `fine` is per work item.
        const bool fine_item = mode_is_fine(MODE) && !diag;
        const int qmin = (diag && MODE != MODE_P2G && MODE != MODE_GENERIC && MODE != MODE_LIBPOW) ? (rg >> 1) : 0;
        // fine modes: column steps of 16 at or beyond the last sample are skipped (off-diagonal tiles of the ragged last block)
        const int qmax = fine_item ? min(8, (P.N - t.j * TILE + 15) >> 4) : 8;
#ifdef KRD_TIMING
        if (threadIdx.x == 0 && item < 4096) KRD_STAMP(g_item_time[2 * item]);
#endif
        // The item's chunk loop and write-out, instantiated once per accumulator LAYOUT (fine = 16-column steps, else
        // 32-column blocks) with its own accumulators: with both layouts in one body ptxas spilled (228 B, cfg2 3.5 -> 4.0 ms).
        auto run_item = [&](auto fine_c) {
        constexpr bool fine = decltype(fine_c)::value;
        double acc[8][8];
        double h4[4][8];   // halved modes: the four accumulator rows that are in registers
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < 8; ++c) acc[r][c] = 0.0;
        int act = 0;
        if (mode_is_halved(MODE)) {
#pragma unroll
            for (int q = 0; q < 32; ++q) { spill[q * NT] = 0.0; h4[q >> 3][q & 7] = 0.0; }
        }

        for (int c = w.c0; c < w.c1; ++c, ++flat) {
            produce_one();   // (owner warp) refills the stage consumed LAG chunks ago
            const int s = flat % STM;
            mbar_wait(&full[s], (uint32_t)((flat / STM) & 1));
            const double* sA = stage_base + (size_t)s * STAGE_DOUBLES;
            const double* sB = diag ? sA : sA + KC * TILE;
            const double* sL = sA + 2 * KC * TILE;
            const double* pa = sA + row0;
            const double* pb = sB + (fine ? tx : 2 * tx);
            if (mode_is_halved(MODE)) {
                // rows [4 act, 4 act + 4) are in registers (acc[0..3]), the other four in `spill`; one exchange per chunk
#pragma unroll 1
                for (int pass = 0; pass < 2; ++pass) {
                    if (pass == 1) {
#pragma unroll
                        for (int r = 0; r < 4; ++r)
#pragma unroll
                            for (int cc = 0; cc < 8; ++cc) { const double t = spill[(r * 8 + cc) * NT]; spill[(r * 8 + cc) * NT] = h4[r][cc]; h4[r][cc] = t; }
                        act ^= 1;
                    }
                    if (MODE == MODE_GENERIC || MODE == MODE_LIBPOW) { chunk_math_rows4<MODE, 0>(pa, pb, sL, P.p, s_tab, act, h4, P.pc); continue; }   // one code variant
                    switch (qmin) {
                        case 0: chunk_math_rows4<MODE, 0>(pa, pb, sL, P.p, s_tab, act, h4, P.pc); break;
                        case 1: chunk_math_rows4<MODE, 1>(pa, pb, sL, P.p, s_tab, act, h4, P.pc); break;
                        case 2: chunk_math_rows4<MODE, 2>(pa, pb, sL, P.p, s_tab, act, h4, P.pc); break;
                        default: chunk_math_rows4<MODE, 3>(pa, pb, sL, P.p, s_tab, act, h4, P.pc); break;
                    }
                }
            } else if (MODE == MODE_P2G) {
                chunk_gram(sA, sB, 32 * (warp & 3), 64 * (warp >> 2), lane, acc);
            } else if (fine) {
                chunk_math16_dispatch<MODE>(qmax, pa, pb, sL, acc);
            } else {
                switch (qmin) {
                    case 0: chunk_math<MODE, 0>(pa, pb, sL, P.p, s_tab, acc); break;
                    case 1: chunk_math<MODE, 1>(pa, pb, sL, P.p, s_tab, acc); break;
                    case 2: chunk_math<MODE, 2>(pa, pb, sL, P.p, s_tab, acc); break;
                    default: chunk_math<MODE, 3>(pa, pb, sL, P.p, s_tab, acc); break;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }

#ifdef KRD_TIMING
        if (threadIdx.x == 0 && item < 4096) KRD_STAMP(g_item_time[2 * item + 1]);
#endif
        // partial tile -> workspace; row = row0 + r, col = 32*q + 2*tx + {0,1}
        double* wout = P.ws + (size_t)item * (size_t)(TILE * TILE);
        if (MODE == MODE_P2G) {
            // acc[2*mi + h][ni] = C fragment half h of DMMA (mi, ni): row 8*mi + lane/4, columns 8*ni + 2*(lane%4) + {0,1}
            const int wr = 32 * (warp & 3), wc = 64 * (warp >> 2);
#pragma unroll
            for (int mi = 0; mi < 4; ++mi)
#pragma unroll
                for (int ni = 0; ni < 8; ++ni)
                    *reinterpret_cast<double2*>(wout + (wr + 8 * mi + (lane >> 2)) * TILE + wc + 8 * ni + 2 * (lane & 3)) =
                        make_double2(acc[2 * mi][ni], acc[2 * mi + 1][ni]);
            return;
        }
        if (mode_is_halved(MODE)) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    *reinterpret_cast<double2*>(wout + (row0 + 4 * act + r) * TILE + 32 * q + 2 * tx) = make_double2(h4[r][2 * q], h4[r][2 * q + 1]);
                    *reinterpret_cast<double2*>(wout + (row0 + 4 * (act ^ 1) + r) * TILE + 32 * q + 2 * tx) =
                        make_double2(spill[(r * 8 + 2 * q) * NT], spill[(r * 8 + 2 * q + 1) * NT]);
                }
            return;
        }
        if (fine) {    // columns 16 q + tx
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int q = 0; q < 8; ++q) wout[(row0 + r) * TILE + 16 * q + tx] = acc[r][q];
            return;
        }
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int q = 0; q < 4; ++q)
                *reinterpret_cast<double2*>(wout + (row0 + r) * TILE + 32 * q + 2 * tx) = make_double2(acc[r][2 * q], acc[r][2 * q + 1]);
        };   // run_item
        if (mode_is_fine(MODE) && fine_item) run_item(std::true_type{}); else run_item(std::false_type{});
    }
#ifdef KRD_TIMING
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x < 1024) KRD_STAMP(g_cta_time[4 * blockIdx.x + 1]);
#endif
}

// ---- epilogue: sum a tile's partials in fixed order, outer root, divisor, mirror ------------------

struct EpiParams {
    const double* ws;
    const Tile* tiles;
    const int32_t* tile_item_off;   // n_tiles + 1
    int n_tiles, N;
    double p, divisor;
    double* out;        // N x N row-major or nullptr
    double* packed;     // [n_tiles][TILE*TILE] or nullptr
    const double* norm; // Gram mode: |a_i - c|^2 per sample, else nullptr
    int4* fix;          // Gram mode: pairs whose Gram difference cancelled too much; fix[0].x = count, entries from fix[1]
    int fix_cap;
    int split;          // CTAs per tile (row bands of 128 / split rows ... 16384 / split cells)
    double* acc;        // streamed exact mode: per-tile sums carried across the slabs ([n_tiles][TILE*TILE]), else nullptr
    int acc_first, acc_last;   // first slab: acc is not read; last slab: the sum is finished (root, divisor, mirror) instead of stored
};

// |a-b|^2 = |a|^2 + |b|^2 - 2ab loses  eps * (|a|^2+|b|^2) / |a-b|^2  relative accuracy.  Cells below this
// ratio are recomputed directly by k_gram_fixup, which keeps the Gram path inside the 1e-10 bar.
constexpr double GRAM_CANCEL_RATIO = 1e-3;

// CTAs per tile: 8 (row bands of 16) when there are many tiles; a launch over a few tiles (a phase of the pipelined ingest:
// 1-8 tiles, each with up to 148 partials) is split further so that the partial sums are spread over all SMs.
constexpr int EPI_SPLIT_MIN = 8, EPI_SPLIT_MAX = 64;
__global__ void __launch_bounds__(256) k_epilogue(const EpiParams P)
{
    const int t = blockIdx.x / P.split, band = blockIdx.x % P.split;
    const Tile tile = P.tiles[t];
    const int i0 = P.tile_item_off[t], i1 = P.tile_item_off[t + 1];
    const double inv_p = 1.0 / P.p;
    const int BAND = TILE * TILE / P.split;
    for (int idx = band * BAND + threadIdx.x; idx < (band + 1) * BAND; idx += blockDim.x) {
        double v = (P.acc && !P.acc_first) ? P.acc[(size_t)t * (TILE * TILE) + idx] : 0.0;
        for (int s = i0; s < i1; ++s) v += P.ws[(size_t)s * (size_t)(TILE * TILE) + idx];
        if (P.acc && !P.acc_last) { P.acc[(size_t)t * (TILE * TILE) + idx] = v; continue; }
        const int r = idx / TILE, c = idx % TILE;
        const int64_t i = (int64_t)tile.i * TILE + r, j = (int64_t)tile.j * TILE + c;
        if (P.norm) {   // |a-b|^2 from the Gram entry
            const double nn = P.norm[i] + P.norm[j];
            v = fmax(nn - 2.0 * v, 0.0);
            if (i < j && j < P.N && v < GRAM_CANCEL_RATIO * nn) {
                const int pos = atomicAdd(&P.fix[0].x, 1);
                if (pos < P.fix_cap) P.fix[1 + pos] = make_int4((int)i, (int)j, t, idx);
            }
        }
        if (P.p > 1.0) v = (P.p == 2.0) ? sqrt(v) : pow(v, inv_p);
        v = v / P.divisor;
        if (P.packed) P.packed[(size_t)t * (TILE * TILE) + idx] = (i < j) ? v : 0.0;
        if (P.out && i < P.N && j < P.N) {
            if (i < j) { P.out[i * P.N + j] = v; P.out[j * P.N + i] = v; }
            else if (i == j) P.out[i * P.N + i] = 0.0;
        }
    }
}

// Direct recomputation of the flagged pairs of the Gram path: one CTA per pair, threads stride over the
// segments, fixed-shape reduction (deterministic).  Normally a handful of pairs, often none.
__global__ void __launch_bounds__(256)
k_gram_fixup(const int4* __restrict__ fix, int fix_cap, const double* __restrict__ F, const double* __restrict__ len,
             int64_t ld, int64_t S_pad, int N, double divisor, double* __restrict__ out, double* __restrict__ packed)
{
    __shared__ double red[8];
    const int count = min(fix[0].x, fix_cap);
    for (int q = blockIdx.x; q < count; q += gridDim.x) {
        const int4 e = fix[1 + q];
        double acc = 0.0;
        for (int64_t s = threadIdx.x; s < S_pad; s += 256) {
            const double d = F[s * ld + e.x] - F[s * ld + e.y];
            acc = fma(len[s] * d, d, acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double v = 0.0;
            for (int w = 0; w < 8; ++w) v += red[w];
            v = sqrt(v) / divisor;
            if (packed) packed[(size_t)e.z * (TILE * TILE) + e.w] = v;
            if (out) { out[(int64_t)e.x * N + e.y] = v; out[(int64_t)e.y * N + e.x] = v; }
        }
        __syncthreads();
    }
}

// gathered packed tiles of all ranks -> full matrix
__global__ void __launch_bounds__(256) k_unpack(const double* __restrict__ gathered, const Tile* __restrict__ tiles,
                                                int64_t n_tiles, int N, double* __restrict__ out)
{
    // blockIdx.x = global tile number t; rank r holds tiles [r*per_rank, (r+1)*per_rank) contiguously,
    // and the all-gather concatenates ranks in order, so tile t sits at offset t.
    const int64_t t = blockIdx.x;
    if (t >= n_tiles) return;
    const Tile tile = tiles[t];
    const double* src = gathered + (size_t)t * (TILE * TILE);
    for (int idx = threadIdx.x; idx < TILE * TILE; idx += blockDim.x) {
        const int r = idx / TILE, c = idx % TILE;
        const int64_t i = (int64_t)tile.i * TILE + r, j = (int64_t)tile.j * TILE + c;
        if (i < N && j < N) {
            if (i < j) { const double v = src[idx]; out[i * N + j] = v; out[j * N + i] = v; }
            else if (i == j) out[i * N + i] = 0.0;
        }
    }
}

// Same for a ROW RANGE of the matrix: out is (row1 - row0) x N, row i of the matrix at out + (i - row0) * N.  One process per
// GPU: after the all-gather every rank unpacks only the rows it then copies to the host over its own PCIe link.
// Both passes write coalesced: the mirrored half is read transposed (from L2) instead of written transposed.
__global__ void __launch_bounds__(256) k_unpack_rows(const double* __restrict__ gathered, const Tile* __restrict__ tiles,
                                                     int64_t n_tiles, int N, int row0, int row1, double* __restrict__ out)
{
    const int64_t t = blockIdx.x;
    if (t >= n_tiles) return;
    const Tile tile = tiles[t];
    const int i0 = tile.i * TILE, j0 = tile.j * TILE;
    const double* src = gathered + (size_t)t * (TILE * TILE);
    if (i0 < row1 && i0 + TILE > row0) {                       // rows of block i: (i, j) for j in block j
        for (int idx = threadIdx.x; idx < TILE * TILE; idx += blockDim.x) {
            const int r = idx / TILE, c = idx % TILE;
            const int i = i0 + r, j = j0 + c;
            if (i < row0 || i >= row1 || i >= N || j >= N) continue;
            if (i < j) out[(int64_t)(i - row0) * N + j] = src[idx];
            else if (i == j) out[(int64_t)(i - row0) * N + i] = 0.0;
        }
    }
    if (j0 < row1 && j0 + TILE > row0) {                       // rows of block j: the mirror (j, i) for i in block i
        for (int idx = threadIdx.x; idx < TILE * TILE; idx += blockDim.x) {
            const int c = idx / TILE, r = idx % TILE;          // r fastest: consecutive threads write consecutive columns i
            const int i = i0 + r, j = j0 + c;
            if (j < row0 || j >= row1 || i >= N || j >= N) continue;
            if (i < j) out[(int64_t)(j - row0) * N + i] = src[r * TILE + c];
        }
    }
}

// ---- Gram mode: centred + scaled copy of F and per-sample squared norms ------------------------------
// Ft[s][i] = sqrt(len[s]) (F[s][i] - F[s][0]) and the partial sums of Ft^2 over a row range, one pass over F.  Column 0 is
// the reference column c: it cancels exactly in a - b and only keeps the norms (hence the cancellation) small.
__global__ void __launch_bounds__(128)
k_gram_prescale(const double* __restrict__ F, const double* __restrict__ sqrt_len, int64_t ld, int64_t S_pad,
                int64_t rows_per_chunk, double* __restrict__ Ft, double* __restrict__ part)
{
    const int64_t i = (int64_t)blockIdx.x * 128 + threadIdx.x;
    const int64_t s0 = (int64_t)blockIdx.y * rows_per_chunk, s1 = min(S_pad, s0 + rows_per_chunk);
    double acc = 0.0;
    for (int64_t s = s0; s < s1; ++s) {
        const double v = sqrt_len[s] * (F[s * ld + i] - F[s * ld]);
        Ft[s * ld + i] = v;
        acc = fma(v, v, acc);
    }
    part[(int64_t)blockIdx.y * ld + i] = acc;
}

__global__ void k_norm_reduce(const double* __restrict__ part, int64_t ld, int n_chunks, double* __restrict__ norm)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ld) return;
    double acc = 0.0;
    for (int ch = 0; ch < n_chunks; ++ch) acc += part[(int64_t)ch * ld + i];
    norm[i] = acc;
}

static krd_status ensure_norms(krd_handle* h)
{
    if (h->norms_valid) return KRD_OK;
    const int NCH = 128;
    krd_status st;
    if ((st = reserve(h, (void**)&h->d_norm, (size_t)h->ld * 8)) != KRD_OK) return st;
    if ((st = reserve(h, (void**)&h->d_norm_part, (size_t)NCH * h->ld * 8)) != KRD_OK) return st;
    if ((st = reserve(h, (void**)&h->d_Fg, (size_t)h->S_pad * h->ld * 8)) != KRD_OK) return st;   // (the caller falls back to the direct path)
    const int64_t rpc = (h->S_pad + NCH - 1) / NCH;
    k_gram_prescale<<<dim3((unsigned)(h->ld / 128), NCH), 128, 0, h->stream>>>(h->d_F, h->d_sqrt_len, h->ld, h->S_pad, rpc, h->d_Fg, h->d_norm_part);
    k_norm_reduce<<<(unsigned)((h->ld + 255) / 256), 256, 0, h->stream>>>(h->d_norm_part, h->ld, NCH, h->d_norm);
    KRD_CUDA(h, cudaGetLastError());
    h->norms_valid = true;
    h->stats.launches += 2;
    return KRD_OK;
}

// ---- host side ------------------------------------------------------------------------------------

int64_t tiles_per_rank(int64_t n_tiles, int world) { return (n_tiles + world - 1) / world; }

// Tile order of the handle.  Default: krd_tile_list (diagonal-major).  With late_blocks = L > 0 (krd_set_late_blocks) the list
// has two parts: first every tile whose blocks both lie before the last L sample blocks — in the default order of that smaller
// matrix — then the tiles that touch the last L blocks, column by column.  After the first part all cells (i, j) with i and j
// outside the late blocks are final, so their copy to the host can run under the pair launch of the (short) second part.
static void build_tile_order(int N, int late, std::vector<Tile>& out, int64_t* split)
{
    const int T = (N + TILE - 1) / TILE;
    late = std::max(0, std::min(late, T - 1));
    const int Ta = T - late;
    const int64_t na = (int64_t)Ta * (Ta + 1) / 2, n = (int64_t)T * (T + 1) / 2;
    out.resize((size_t)n);
    std::vector<int32_t> ti((size_t)std::max<int64_t>(na, 1)), tj((size_t)std::max<int64_t>(na, 1));
    krd_tile_list(Ta * TILE, ti.data(), tj.data());
    for (int64_t t = 0; t < na; ++t) out[(size_t)t] = Tile{ti[(size_t)t], tj[(size_t)t]};
    int64_t k = na;
    for (int j = Ta; j < T; ++j) {
        out[(size_t)k++] = Tile{j, j};
        for (int i = 0; i < j; ++i) out[(size_t)k++] = Tile{i, j};
    }
    *split = na;
}

static krd_status ensure_tiles(krd_handle* h)
{
    const int T = (h->N + TILE - 1) / TILE;   // (ld may be wider than N: squash clustering appends cluster columns)
    const int64_t n_tiles = (int64_t)T * (T + 1) / 2;
    if (h->n_tiles_alloc == n_tiles && h->d_tiles) return KRD_OK;
    if (h->custom_tiles) return fail(h, KRD_ERR_STATE, "pipelined tile list is out of date");
    build_tile_order(h->N, h->late_blocks, h->h_tiles, &h->late_split);
    cudaFree(h->d_tiles); h->d_tiles = nullptr;
    KRD_CUDA(h, cudaMalloc(&h->d_tiles, (size_t)std::max<int64_t>(n_tiles, 1) * sizeof(Tile)));
    KRD_CUDA(h, cudaMemcpyAsync(h->d_tiles, h->h_tiles.data(), (size_t)n_tiles * sizeof(Tile), cudaMemcpyHostToDevice, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    h->n_tiles_alloc = n_tiles;
    h->stats.n_tiles = n_tiles;
    h->custom_tiles = false;
    for (auto& kv : h->scheds) cudaFree(kv.second.d);
    h->scheds.clear();
    h->d_sched = nullptr; h->sched_key = -1;
    return KRD_OK;
}

krd_status set_late_blocks(krd_handle* h, int n_late, int64_t* split)
{
    const int T = (h->N + TILE - 1) / TILE;
    h->late_blocks = std::max(0, std::min(n_late, T - 1));
    h->n_tiles_alloc = -1;                    // rebuild the list (and drop the cached schedules) in the new order
    krd_status st = ensure_tiles(h);
    if (st == KRD_OK && split) *split = h->late_split;
    return st;
}

static void drop_schedules(krd_handle* h)
{
    for (auto& kv : h->scheds) cudaFree(kv.second.d);
    h->scheds.clear();
    h->d_sched = nullptr; h->sched_key = -1;
}

// Tile list in an order of the caller's choice (the pipelined ingest orders it by phase); run_pairs then works on the
// range [force_t0, force_t1) of it.
krd_status set_custom_tiles(krd_handle* h, const std::vector<Tile>& tiles)
{
    const int64_t n_tiles = (int64_t)tiles.size();
    h->h_tiles = tiles;
    cudaFree(h->d_tiles); h->d_tiles = nullptr;
    KRD_CUDA(h, cudaMalloc(&h->d_tiles, (size_t)std::max<int64_t>(n_tiles, 1) * sizeof(Tile)));
    KRD_CUDA(h, cudaMemcpyAsync(h->d_tiles, h->h_tiles.data(), (size_t)n_tiles * sizeof(Tile), cudaMemcpyHostToDevice, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    h->n_tiles_alloc = n_tiles;
    h->stats.n_tiles = n_tiles;
    h->custom_tiles = true;
    drop_schedules(h);
    return KRD_OK;
}

// Stream-K schedule: cut the cost axis (tiles x chunks, diagonal tiles cheaper) into `G` equal spans.
static void make_schedule(const Tile* tiles, int n_tiles, int n_chunks, int G, double DIAG_COST, double ragged_cost, int ragged_block,
                          std::vector<WorkItem>& items, std::vector<int32_t>& cta_off, std::vector<int32_t>& tile_off)
{
    items.clear(); cta_off.assign(G + 1, 0); tile_off.assign(n_tiles + 1, 0);
    std::vector<double> pre(n_tiles + 1, 0.0), wt(n_tiles, 1.0);
    for (int t = 0; t < n_tiles; ++t) {
        wt[t] = (tiles[t].i == tiles[t].j) ? DIAG_COST : (tiles[t].j == ragged_block) ? ragged_cost : 1.0;
        pre[t + 1] = pre[t] + wt[t] * n_chunks;
    }
    const double total = pre[n_tiles];
    const int min_chunks = 8;   // do not create slivers: snap cuts that close to a tile border
    int g = 0;                  // CTA that owns the current position
    std::vector<int32_t> item_cta;
    for (int t = 0; t < n_tiles; ++t) {
        const double w = wt[t];
        int c = 0;
        while (c < n_chunks) {
            // end of CTA g's span, in chunks of this tile
            const double bound = total * (g + 1) / G;
            int ce = (g + 1 >= G) ? n_chunks : (int)std::llround((bound - pre[t]) / w);
            if (ce >= n_chunks - min_chunks) ce = n_chunks;
            if (ce <= c + min_chunks && ce < n_chunks) ce = std::min(n_chunks, c + min_chunks);
            if (ce <= c) { ++g; continue; }
            items.push_back(WorkItem{t, c, ce, 0});
            item_cta.push_back(g);
            c = ce;
            if (c < n_chunks || pre[t + 1] >= bound - 1e-9) { if (g + 1 < G) ++g; }
        }
        tile_off[t + 1] = (int32_t)items.size();
    }
    // items are in (tile, chunk) order and CTA ids are non-decreasing
    for (size_t k = 0; k < items.size(); ++k) cta_off[item_cta[k] + 1] = (int32_t)(k + 1);
    for (int q = 1; q <= G; ++q) cta_off[q] = std::max(cta_off[q], cta_off[q - 1]);
}

template <int MODE>
static cudaError_t launch_pairs(const PairParams& P, int grid, cudaStream_t stream)
{
    static std::atomic<unsigned long long> configured{0};
    cudaError_t e = allow_dynamic_smem(k_pairs<MODE>, (int)smem_bytes_for(MODE), configured);
    if (e != cudaSuccess) return e;
    k_pairs<MODE><<<grid, NT, smem_bytes_for(MODE), stream>>>(P);
    return cudaGetLastError();
}

static krd_status run_pairs_once(krd_handle* h, double p, double divisor, int rank, int world,
                                 double* d_out_matrix, double* d_out_packed, bool allow_gram, const SlabArgs* slab = nullptr)
{
    if (slab) allow_gram = false;    // (the Gram path needs whole-column norms)
    krd_status st = ensure_tiles(h);
    if (st != KRD_OK) return st;
    const int64_t n_tiles_all = h->n_tiles_alloc;
    const int64_t per_rank = tiles_per_rank(n_tiles_all, world);
    int64_t t0 = 0, t1 = 0;
    krd_rank_tiles(n_tiles_all, rank, world, &t0, &t1);
    if (h->force_t0 >= 0) { t0 = h->force_t0; t1 = h->force_t1; }   // pipelined ingest: one phase of the phase-major list
    const int n_tiles = (int)(t1 - t0);
    if (d_out_packed && n_tiles < per_rank && h->force_t0 < 0 && (!slab || slab->last))   // ragged tail: keep the all-gather payload defined
        KRD_CUDA(h, cudaMemsetAsync(d_out_packed + (size_t)n_tiles * TILE * TILE, 0, (size_t)(per_rank - n_tiles) * TILE * TILE * 8, h->stream));
    h->stats.pair_ctas = 0;
    if (n_tiles == 0) { cudaEventRecord(h->ev[2], h->stream); return KRD_OK; }

    int mode = mode_for_exponent(p, allow_gram && h->cfg.p2_mode != 1 && !h->len_negative);   // p = 2: auto = Gram on the tensor cores + direct fix-ups
    if (mode == MODE_P2 && h->len_negative) mode = MODE_P2S;          // sqrt(len) pre-scaling needs len >= 0
    const bool full_diag = (mode == MODE_P2G) || mode == MODE_GENERIC || mode == MODE_LIBPOW;   // these modes do not skip the lower blocks of diagonal tiles
    // cost of a diagonal-tile chunk relative to a full one: 5/8 of the math, but the warps of a diagonal tile carry unequal
    // shares (the top row groups skip nothing).  p = 1, 2 are FP64-pipe-bound, so the lighter warps free the pipe for the
    // others (measured 0.705); the long operators are bound by each warp's own dependency chains, so the tile takes almost
    // as long as a full one (measured: p = 3, 4 0.8; halved modes p = 1/2 0.72, integer p 0.85; profiles/cta_balance.py).  Diagonal tiles also run FOUR
    // code variants at once (one per skip count), so the unrolled bodies must stay small: with the p = 3 body unrolled 4x
    // the diagonal tiles ran up to 50 % slower than full ones (instruction-cache misses); unrolled 2x they do not.  KRD_DIAG_COST overrides (tuning aid).
    double diag_cost = full_diag ? 1.0 : (mode == MODE_PHALF) ? 0.72 : (mode == MODE_P3 || mode == MODE_P4) ? 0.8 : DIAG_COST_SKIP;
    // short operators skip at 16-column granularity: the off-diagonal tiles of a ragged last column block run qmax of 8 steps
    int ragged_block = -1; double ragged_cost = 1.0;
    if (mode_is_fine(mode) && h->N % TILE != 0) {
        const int qmax = (h->N % TILE + 15) / 16;
        if (qmax < 8) { ragged_block = h->N / TILE; ragged_cost = 0.06 + 0.94 * qmax / 8.0; }
    }
    if (const char* ov = std::getenv("KRD_RAGGED_COST")) { const double v = std::atof(ov); if (v > 0.05 && v <= 1.5) ragged_cost = v; }
    if (mode == MODE_INT) {   // static bodies (chunk_math_int): 0.80 measured for 3 multiplications (p = 5), 0.85-0.875 for 6 (p = 15)
        int mults = 0;
        for (int n = (int)p; n > 1; n = (n & 1) ? n - 1 : n / 2) ++mults;
        diag_cost = 0.725 + 0.025 * mults;
    }
    if (const char* ov = std::getenv("KRD_DIAG_COST")) { const double v = std::atof(ov); if (v > 0.1 && v <= 1.5) diag_cost = v; }
    if (mode == MODE_P2G && ensure_norms(h) != KRD_OK) {   // no room for the scaled copy of F: the direct path needs none
        cudaGetLastError();
        return run_pairs_once(h, p, divisor, rank, world, d_out_matrix, d_out_packed, false, slab);
    }

    const int n_chunks = (int)((slab ? slab->rows : h->S_pad) / KC);
    const int G = h->cfg.split > 0 ? h->cfg.split : h->num_sms;
    const int64_t key = ((((t0 * 1000003 + t1) * 1000003 + n_chunks) * 1024 + G) * 4096 + (int64_t)(diag_cost * 2048.0)) * 4099 +
                        (ragged_block < 0 ? 0 : (int64_t)(ragged_cost * 2048.0) + 1);
    if (key != h->sched_key) {
        auto it = h->scheds.find(key);
        if (it == h->scheds.end()) {
            if (h->capturing) return fail(h, KRD_ERR_STATE, "schedule not cached inside a captured phase");   // (re-issued launch by launch)
            if (h->scheds.size() >= 64) drop_schedules(h);
            std::vector<WorkItem> items; std::vector<int32_t> cta_off, tile_off;
            make_schedule(h->h_tiles.data() + t0, n_tiles, n_chunks, G, diag_cost, ragged_cost, ragged_block, items, cta_off, tile_off);
            const size_t b_items = items.size() * sizeof(WorkItem), b_cta = cta_off.size() * 4, b_tile = tile_off.size() * 4;
            void* d = nullptr;
            KRD_CUDA(h, cudaMalloc(&d, b_items + b_cta + b_tile));
            char* base = static_cast<char*>(d);
            KRD_CUDA(h, cudaMemcpyAsync(base, items.data(), b_items, cudaMemcpyHostToDevice, h->stream));
            KRD_CUDA(h, cudaMemcpyAsync(base + b_items, cta_off.data(), b_cta, cudaMemcpyHostToDevice, h->stream));
            KRD_CUDA(h, cudaMemcpyAsync(base + b_items + b_cta, tile_off.data(), b_tile, cudaMemcpyHostToDevice, h->stream));
            KRD_CUDA(h, cudaStreamSynchronize(h->stream));   // the host vectors die here
            it = h->scheds.emplace(key, krd_handle::Sched{d, (int64_t)items.size(), (int64_t)b_cta}).first;
        }
        h->d_sched = it->second.d; h->sched_items = it->second.items; h->sched_cta_bytes = it->second.cta_bytes; h->sched_key = key;
    }
    const size_t b_items = (size_t)h->sched_items * sizeof(WorkItem);
    const WorkItem* d_items = reinterpret_cast<const WorkItem*>(h->d_sched);
    const int32_t* d_cta_off = reinterpret_cast<const int32_t*>(static_cast<char*>(h->d_sched) + b_items);
    const int32_t* d_tile_off = reinterpret_cast<const int32_t*>(static_cast<char*>(h->d_sched) + b_items + h->sched_cta_bytes);

    const size_t ws_need = (size_t)h->sched_items * TILE * TILE * sizeof(double);
    if (ws_need > h->ws_capacity) {
        if (h->capturing) return fail(h, KRD_ERR_STATE, "workspace growth inside a captured phase");
        cudaFree(h->d_ws); h->d_ws = nullptr; h->ws_capacity = 0;
        KRD_CUDA(h, cudaMalloc(&h->d_ws, ws_need));
        h->ws_capacity = ws_need;
    }

    PairParams P;
    P.F = slab ? slab->F : (mode == MODE_P2G) ? h->d_Fg : h->d_F;
    P.len = ((mode == MODE_P2) ? h->d_sqrt_len : h->d_len) + (slab ? slab->row0 : 0);
    P.tiles = h->d_tiles + t0; P.items = d_items; P.cta_item_off = d_cta_off; P.ws = h->d_ws; P.ld = h->ld; P.p = p; P.N = h->N; P.pc = pow_coef(p);
    cudaError_t e;
    switch (mode) {
        case MODE_P1: e = launch_pairs<MODE_P1>(P, G, h->stream); break;
        case MODE_P2: e = launch_pairs<MODE_P2>(P, G, h->stream); break;
        case MODE_P2S: e = launch_pairs<MODE_P2S>(P, G, h->stream); break;
        case MODE_PHALF: e = launch_pairs<MODE_PHALF>(P, G, h->stream); break;
        case MODE_P2G: e = launch_pairs<MODE_P2G>(P, G, h->stream); break;
        case MODE_INT: e = launch_pairs<MODE_INT>(P, G, h->stream); break;
        case MODE_P3: e = launch_pairs<MODE_P3>(P, G, h->stream); break;
        case MODE_P4: e = launch_pairs<MODE_P4>(P, G, h->stream); break;
        case MODE_LIBPOW: e = launch_pairs<MODE_LIBPOW>(P, G, h->stream); break;
        default:
            e = make_pow_tables();
            if (e == cudaSuccess) e = launch_pairs<MODE_GENERIC>(P, G, h->stream);
            break;
    }
    if (e != cudaSuccess) return cuda_fail(h, e, "k_pairs launch");
    if (!h->capturing) cudaEventRecord(h->ev[2], h->stream);

    EpiParams Q;
    Q.ws = h->d_ws; Q.tiles = h->d_tiles + t0; Q.tile_item_off = d_tile_off; Q.n_tiles = n_tiles; Q.N = h->N;
    Q.p = p; Q.divisor = divisor; Q.out = d_out_matrix; Q.packed = d_out_packed;
    Q.norm = nullptr; Q.fix = nullptr; Q.fix_cap = 0;
    Q.acc = nullptr; Q.acc_first = Q.acc_last = 1;
    if (slab && !(slab->first && slab->last)) {
        krd_status sa = reserve(h, (void**)&h->d_acc, (size_t)n_tiles * TILE * TILE * sizeof(double));
        if (sa != KRD_OK) return sa;
        Q.acc = h->d_acc; Q.acc_first = slab->first; Q.acc_last = slab->last;
    }
    if (mode == MODE_P2G) {
        krd_status sf = reserve(h, (void**)&h->d_fix, (size_t)(FIX_CAP + 1) * sizeof(int4));
        if (sf != KRD_OK) return sf;
        KRD_CUDA(h, cudaMemsetAsync(h->d_fix, 0, sizeof(int4), h->stream));
        Q.norm = h->d_norm; Q.fix = static_cast<int4*>(h->d_fix); Q.fix_cap = FIX_CAP;
    }
    Q.split = EPI_SPLIT_MIN;
    while (Q.split < EPI_SPLIT_MAX && (int64_t)n_tiles * Q.split < 4 * (int64_t)h->num_sms) Q.split *= 2;
    k_epilogue<<<(unsigned)n_tiles * Q.split, 256, 0, h->stream>>>(Q);
    if (mode == MODE_P2G) {
        k_gram_fixup<<<2 * h->num_sms, 256, 0, h->stream>>>(static_cast<const int4*>(h->d_fix), FIX_CAP, h->d_F, h->d_len, h->ld, h->S_pad,
                                                            h->N, divisor, d_out_matrix, d_out_packed);
        h->stats.launches += 1;
        h->fix_pending = true;
    }
    KRD_CUDA(h, cudaGetLastError());
    h->stats.split = (int32_t)((h->sched_items + n_tiles - 1) / n_tiles);
    h->stats.pair_ctas = G;
    h->stats.launches += 2;
    return KRD_OK;
}

// The Gram path's epilogue lists the cells whose subtraction cancelled; when that list overflows (more than FIX_CAP
// near-duplicate pairs) the launch is repeated on the direct FP64-core path — here, so that EVERY entry point (host,
// device, sharded, squash) returns cells inside the tolerance.  Costs one 4-byte read-back per p = 2 Gram launch.
krd_status run_pairs(krd_handle* h, double p, double divisor, int rank, int world, double* d_out_matrix, double* d_out_packed)
{
    krd_status st = run_pairs_once(h, p, divisor, rank, world, d_out_matrix, d_out_packed, true);
    if (st != KRD_OK || !h->fix_pending) return st;
    h->fix_pending = false;
    int cnt = 0;
    KRD_CUDA(h, cudaMemcpyAsync(&cnt, h->d_fix, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    KRD_CUDA(h, cudaStreamSynchronize(h->stream));
    h->stats.gram_fixups = (cnt > FIX_CAP) ? -1 : cnt;
    if (cnt <= FIX_CAP) return KRD_OK;
    return run_pairs_once(h, p, divisor, rank, world, d_out_matrix, d_out_packed, false);
}

krd_status run_pairs_slab(krd_handle* h, double p, double divisor, int rank, int world, double* d_out_matrix, double* d_out_packed,
                          const SlabArgs& slab)
{
    return run_pairs_once(h, p, divisor, rank, world, d_out_matrix, d_out_packed, false, &slab);
}

krd_status unpack_shards(krd_handle* h, int world, const double* d_gathered, double* d_out)
{
    krd_status st = ensure_tiles(h);
    if (st != KRD_OK) return st;
    const int64_t n_tiles = h->n_tiles_alloc;
    if (n_tiles == 0) return KRD_OK;
    (void)world;
    k_unpack<<<(unsigned)n_tiles, 256, 0, h->stream>>>(d_gathered, h->d_tiles, n_tiles, h->N, d_out);
    KRD_CUDA(h, cudaGetLastError());
    return KRD_OK;
}

krd_status rank_tile_range(krd_handle* h, int rank, int world, const Tile** d_tiles, int64_t* t0, int64_t* t1)
{
    krd_status st = ensure_tiles(h);
    if (st != KRD_OK) return st;
    krd_rank_tiles(h->n_tiles_alloc, rank, world, t0, t1);
    *d_tiles = h->d_tiles;
    return KRD_OK;
}

krd_status unpack_shard_rows(krd_handle* h, const double* d_gathered, int row0, int row1, double* d_out_rows)
{
    krd_status st = ensure_tiles(h);
    if (st != KRD_OK) return st;
    const int64_t n_tiles = h->n_tiles_alloc;
    if (n_tiles == 0 || row1 <= row0) return KRD_OK;
    k_unpack_rows<<<(unsigned)n_tiles, 256, 0, h->stream>>>(d_gathered, h->d_tiles, n_tiles, h->N, row0, row1, d_out_rows);
    KRD_CUDA(h, cudaGetLastError());
    return KRD_OK;
}

}  // namespace krd

extern "C" {

int64_t krd_tile_count(int32_t n_samples)
{
    const int64_t T = ((int64_t)n_samples + krd::TILE - 1) / krd::TILE;
    return T * (T + 1) / 2;
}

// The two-part order of krd_set_late_blocks, host-only (what a handle with n_late_blocks late blocks uses).
krd_status krd_tile_list_late(int32_t n_samples, int32_t n_late_blocks, int32_t* tile_i, int32_t* tile_j, int64_t* t_split)
{
    if (n_samples < 0 || n_late_blocks < 0 || !tile_i || !tile_j) return KRD_ERR_INVALID_ARGUMENT;
    std::vector<krd::Tile> tiles;
    int64_t split = 0;
    krd::build_tile_order(n_samples, n_late_blocks, tiles, &split);
    for (size_t t = 0; t < tiles.size(); ++t) { tile_i[t] = tiles[t].i; tile_j[t] = tiles[t].j; }
    if (t_split) *t_split = split;
    return KRD_OK;
}

// Off-diagonal tiles in diagonal-major order ((i,i+1) for all i, then (i,i+2), ...: every contiguous
// range mixes tile rows and columns evenly), with the T cheaper diagonal tiles spread evenly through
// the list — so that equal-COUNT contiguous shares (one per rank) are also equal-COST shares.
krd_status krd_tile_list(int32_t n_samples, int32_t* tile_i, int32_t* tile_j)
{
    if (n_samples < 0 || !tile_i || !tile_j) return KRD_ERR_INVALID_ARGUMENT;
    const int64_t T = (n_samples + krd::TILE - 1) / krd::TILE;
    const int64_t n_off = T * (T - 1) / 2;
    int64_t k = 0, q = 0;          // q: off-diagonal tiles emitted so far
    int d = 1, i = 0;              // next off-diagonal tile (i, i+d)
    for (int64_t g = 0; g < T; ++g) {
        tile_i[k] = (int32_t)g; tile_j[k] = (int32_t)g; ++k;
        const int64_t end = n_off * (g + 1) / T;
        for (; q < end; ++q, ++k) {
            tile_i[k] = i; tile_j[k] = i + d;
            if (++i + d >= T) { i = 0; ++d; }
        }
    }
    return KRD_OK;
}

#ifdef KRD_TIMING
/* tuning aid (variant builds only, not part of the product surface): per-CTA start/end (ns) and item range of the last pair launch */
int krd_debug_item_times(unsigned long long* out, int n_items)
{
    return (int)cudaMemcpyFromSymbol(out, krd::g_item_time, sizeof(unsigned long long) * 2 * (size_t)std::min(n_items, 4096));
}

int krd_debug_cta_times(unsigned long long* out, int n_ctas)
{
    return (int)cudaMemcpyFromSymbol(out, krd::g_cta_time, sizeof(unsigned long long) * 4 * (size_t)std::min(n_ctas, 1024));
}
#endif

void krd_rank_tiles(int64_t n_tiles, int32_t rank, int32_t world, int64_t* t0, int64_t* t1)
{
    const int64_t per_rank = krd::tiles_per_rank(n_tiles, world);
    *t0 = std::min<int64_t>((int64_t)rank * per_rank, n_tiles);
    *t1 = std::min<int64_t>(*t0 + per_rank, n_tiles);
}

}  // extern "C"
