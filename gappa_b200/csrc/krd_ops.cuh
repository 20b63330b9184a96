// Inner operators of the KR reduction  acc += len * |a - b|^p  and the exponent -> mode mapping, shared by
// the all-pairs tile kernel (krd_pairs.cu) and the one-vs-all kernel of squash clustering (krd_squash.cu).
#pragma once
#include <cmath>

#include "krd_common.cuh"

namespace krd {

// FP64 sqrt and pow written for the FP64 pipe's issue rate (the library versions spend > 100
// instructions and a divergent slow path on the exact zeros that |F_i - F_j| is full of).
//
// sqrt(|d|): MUFU.RSQ64H seed r = (1 + q)/sqrt(x), |q| <= 2^-22, then ONE third-order step:
//   y = x r,  e = 1 - y r = -2q - q^2,  sqrt(x) = y / (1 + q) = y (1 + e/2 + 3/8 e^2 + O(e^3)) = y + (y e)(1/2 + 3/8 e)
// -> 5 FP64 instructions (DMUL, DFMA, DFMA, DMUL, DFMA), error 2^-64 before the final rounding.
// x = 0 (and subnormal x, which the seed flushes): the seed is +inf; its high word is replaced by 0 (integer compare +
// select on the integer pipe), which makes r vanishingly small and the result exactly 0 — no FP64 compare, no NaN.
// |.| is a bit mask for the seed and an operand modifier for the multiply.
__device__ __forceinline__ double kr_rsqrt_seed_abs(double d)   // ~ 1/sqrt(|d|), 0-ish for |d| < 2^-1022
{
    const int xh = __double2hiint(d) & 0x7fffffff;
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(__hiloint2double(xh, __double2loint(d))));   // MUFU.RSQ64H (reads the high word)
    return __hiloint2double(xh < 0x00100000 ? 0 : __double2hiint(r), __double2loint(r));
}

// Seed for the 3-instruction form below: instead of replacing an infinite seed by 0 (compare + select), its high word is
// clamped to 2^512 with ONE integer min.  Normal x >= 2^-1022 have r <= 2^511 and are untouched; x = 0 gives y = 0 * 2^512 = 0
// exactly; a subnormal x (seed +inf, flushed) gives y = x 2^512 < 2^-510, t in (-1, 3], result < 2^-508: zero for every purpose.
__device__ __forceinline__ double kr_rsqrt_seed_clamped(double d)
{
    const int xh = __double2hiint(d) & 0x7fffffff;
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(__hiloint2double(xh, __double2loint(d))));
    return __hiloint2double(min(__double2hiint(r), 0x5ff00000), __double2loint(r));
}

// Shorter forms: the reference's std::pow is correctly rounded, the bar is 1e-10, so ONE second-order step is enough.
// With r = (1 + q)/sqrt(x), |q| <= 2^-22, both forms below return sqrt(x) (1 - 3/2 q^2 - ...): relative error <= 8.6e-14.
//   KRD_SQRT_SHORT = 2            y = x r,  e = 1 - y r,  s = y + (y e)/2        4 FP64 (DMUL, DFMA, DMUL, DFMA)
//   KRD_SQRT_SHORT = 1            y = x r,  s = y + (x - y^2) h,  h = r/2        3 FP64 + the integer work for h
//   KRD_SQRT_SHORT = 0            the 5-instruction third-order step above (error 2^-64)
//   KRD_SQRT_SHORT = 3 (default)  y = x r,  t = 3 - y r,  2 s = y t                3 FP64 (DMUL, DFMA, DMUL); the 1/2 is folded
//                                 into the segment length (exact), only in kr_row_sqrt (the pair kernel's p = 1/2 body):
//                                 2 s = sqrt(x) (1 + q)(2 - 2q - q^2) = 2 sqrt(x) (1 - 3/2 q^2 - ...): the same error as form 2
// Measured at the cfg2 shape (pair kernel, p = 1/2): 12.9 ms / 14.1 ms / 14.8 ms.  Per operator the kernel issues
// DADD + sqrt + DFMA on the FP64 pipe (2 issue cycles each) plus MUFU, LOP3, ISETP, SEL and a MOV for the seed: form 2 is
// FP64-pipe-bound at 12 cycles, form 1 needs h as a second register pair (VIADD, SEL, MOV) and becomes issue-bound at ~15.
#ifndef KRD_SQRT_SHORT
#define KRD_SQRT_SHORT 3
#endif
__device__ __forceinline__ void kr_rsqrt_seed_abs2(double d, double& r, double& h)   // r ~ 1/sqrt(|d|), h = r/2; both 0-ish for |d| < 2^-1022
{
    const int xh = __double2hiint(d) & 0x7fffffff;
    double s;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(s) : "d"(__hiloint2double(xh, __double2loint(d))));   // MUFU.RSQ64H
    const int sh = __double2hiint(s), sl = __double2loint(s);
    const bool tiny = xh < 0x00100000;
    r = __hiloint2double(tiny ? 0 : sh, sl);
    h = __hiloint2double(tiny ? 0 : sh - 0x00100000, sl);
}

__device__ __forceinline__ double kr_sqrt_abs(double d)
{
#if KRD_SQRT_SHORT >= 2   // (form 3 exists only in kr_row_sqrt, which can fold the 1/2 into the segment length)
    const double r = kr_rsqrt_seed_abs(d);
    const double y = fabs(d) * r;
    return fma(y * fma(-y, r, 1.0), 0.5, y);
#elif KRD_SQRT_SHORT
    double r, h;
    kr_rsqrt_seed_abs2(d, r, h);
    const double y = fabs(d) * r;
    return fma(fma(-y, y, fabs(d)), h, y);
#else
    const double r = kr_rsqrt_seed_abs(d);
    const double y = fabs(d) * r;
    const double e = fma(-y, r, 1.0);
    const double t = fma(0.375, e, 0.5);
    return fma(y * e, t, y);
#endif
}

// One row of the register tile for p = 1/2, written stage by stage over the NC columns so that the NC independent
// dependency chains are interleaved in program order (ptxas otherwise pairs the chains two at a time and the FP64
// latency shows: 44 % of the stall samples were fixed-latency waits).
template <int NC>
__device__ __forceinline__ void kr_row_sqrt(double a, const double* __restrict__ b, double l, double* __restrict__ acc)
{
    double d[NC], r[NC], e[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) d[c] = a - b[c];
#if KRD_SQRT_SHORT == 3   // y = x r, t = 3 - y r, 2 sqrt = y t; the caller's l is halved here once per row
#pragma unroll
    for (int c = 0; c < NC; ++c) r[c] = kr_rsqrt_seed_clamped(d[c]);
#pragma unroll
    for (int c = 0; c < NC; ++c) d[c] = fabs(d[c]) * r[c];          // y
#pragma unroll
    for (int c = 0; c < NC; ++c) e[c] = fma(-d[c], r[c], 3.0);      // t
#pragma unroll
    for (int c = 0; c < NC; ++c) d[c] = d[c] * e[c];                // 2 sqrt
    l *= 0.5;
#elif KRD_SQRT_SHORT == 2   // second-order step in the e form: y = x r, e = 1 - y r, s = y + (y e)/2 -> 4 FP64, no h
#pragma unroll
    for (int c = 0; c < NC; ++c) r[c] = kr_rsqrt_seed_abs(d[c]);
#pragma unroll
    for (int c = 0; c < NC; ++c) d[c] = fabs(d[c]) * r[c];          // y
#pragma unroll
    for (int c = 0; c < NC; ++c) e[c] = fma(-d[c], r[c], 1.0);
#pragma unroll
    for (int c = 0; c < NC; ++c) e[c] = d[c] * e[c];
#pragma unroll
    for (int c = 0; c < NC; ++c) d[c] = fma(e[c], 0.5, d[c]);
#elif KRD_SQRT_SHORT
#pragma unroll
    for (int c = 0; c < NC; ++c) kr_rsqrt_seed_abs2(d[c], r[c], e[c]);   // e = h = r/2
#pragma unroll
    for (int c = 0; c < NC; ++c) r[c] = fabs(d[c]) * r[c];               // y
#pragma unroll
    for (int c = 0; c < NC; ++c) d[c] = fma(-r[c], r[c], fabs(d[c]));    // x - y^2
#pragma unroll
    for (int c = 0; c < NC; ++c) d[c] = fma(d[c], e[c], r[c]);           // sqrt
#else
#pragma unroll
    for (int c = 0; c < NC; ++c) r[c] = kr_rsqrt_seed_abs(d[c]);
#pragma unroll
    for (int c = 0; c < NC; ++c) d[c] = fabs(d[c]) * r[c];          // y
#pragma unroll
    for (int c = 0; c < NC; ++c) e[c] = fma(-d[c], r[c], 1.0);
#pragma unroll
    for (int c = 0; c < NC; ++c) r[c] = fma(0.375, e[c], 0.5);      // t
#pragma unroll
    for (int c = 0; c < NC; ++c) e[c] = d[c] * e[c];
#pragma unroll
    for (int c = 0; c < NC; ++c) d[c] = fma(e[c], r[c], d[c]);
#endif
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[c] = fma(l, d[c], acc[c]);
}

// pow(x, p) = exp2(p * log2 x), x >= 0, p > 0.  Table-driven, all in FP64, relative error < 1.3e-12 + 4e-14 p:
//   log2: x = 2^e * m, m in [1,2); j = top 6 mantissa bits, c_j = interval midpoint, r = m/c_j - 1 (|r| <= 2^-7),
//         log2 x = e + log2 c_j + r (A1 + r (A2 + ... r A5))            (truncation 6e-14)
//   exp2: t = k/32 + g, |g| <= 1/64; 2^t = 2^(k>>5) * 2^((k&31)/32) * (1 + g (C1 + ... g C4))   (truncation 1.25e-12)
// tab = [invc 64][logc 64][2^(i/32) 32] in shared memory (make_pow_tables() below), entry stride TS.
#ifndef KRD_POW_I2F
#define KRD_POW_I2F 1   // pair kernel (kr_pow_acc16): 1 = the exponent through I2F (41.5 ms at the cfg2 shape against 42.2 with the 2^52 trick), 2 = k as well (41.9)
#endif
constexpr int POW_LOGN = 64, POW_EXPN = 32;
constexpr int POW_TAB = 2 * POW_LOGN + POW_EXPN;
// Table stride of the pair kernel: every entry is stored 16 times, lane L reads copy L & 15, so that the bank pair of a
// lookup is fixed by the lane and not by the (data-dependent) index — with one copy the three lookups per operator ran
// into 4-6-way bank conflicts and the kernel was bound by the shared-memory pipe (48.9 ms; 40.1 ms with every lane
// reading entry 0).  64 log entries instead of 128 keep the replicated tables at 20 KB (degree 5 still truncates < 6e-14).
constexpr int POW_TS_PAIR = 16;
static __device__ double g_pow_tab[POW_TAB];   // one copy per translation unit

template <int TS>
__device__ __forceinline__ double kr_pow(double x, double p, const double* __restrict__ tab)   // x >= 0 (no sign bit); tab stride TS
{
    // exponent / table index / mantissa from the HIGH word with 32-bit integer operations (64-bit shifts cost three
    // instructions each); every clamp and the x = 0 case are integer compares — an FP64 compare occupies the FP64 pipe
    const int hi = __double2hiint(x), lo = __double2loint(x);
    const int e = (hi >> 20) - 1023;
    const int j = (hi >> 14) & (POW_LOGN - 1);
    const double m = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, lo);
    const double r = fma(m, tab[j * TS], -1.0);
    double q = fma(r, 0.28853900817779268, -0.36067376022224085);     // 1/(5 ln2), -1/(4 ln2)  (degree 5: the r^6 term is < 6e-14)
    q = fma(r, q, 0.48089834696298783);                                //  1/(3 ln2)
    q = fma(r, q, -0.72134752044448170);                               // -1/(2 ln2)
    q = fma(r, q, 1.4426950408889634);                                 //  1/ln2
    // e as a double without an I2F conversion: 2^52 + 2^31 + e is exact in the low mantissa bits
#if KRD_POW_I2F
    const double ed = (double)e;                                       // I2F.F64.S32 on the conversion unit (A/B switch)
#else
    const double ed = __hiloint2double(0x43300000, e ^ 0x80000000) - 4503601774854144.0;
#endif
    const double L = fma(r, q, tab[(POW_LOGN + j) * TS]) + ed;
    const double t = p * L;
    const double MAGIC = 6755399441055744.0;                           // 1.5 * 2^52: rounds t*32 to an integer in the low bits
    double kd = fma(t, 32.0, MAGIC);
    int k = __double2loint(kd);
    kd -= MAGIC;
    const double g = fma(kd, -0.03125, t);                             // |g| <= 1/64, from the UNclamped k
    double w = fma(g, 9.6181291076284772e-3, 5.5504108664821580e-2);   // ln2^4/24, ln2^3/6  (degree 4: truncation 1.25e-12 relative, the bar is 1e-10)
    w = fma(g, w, 2.4022650695910071e-1);                              // ln2^2/2
    w = fma(g, w, 6.9314718055994531e-1);                              // ln2
    // the scale 2^(k/32) is built in the exponent field: keep it a normal number.  |32 t| < 2^31 is assumed, i.e.
    // p <= POW_MAX_P at the smallest normal x (larger exponents run the library-pow body, MODE_LIBPOW).  A result below
    // the normal range is flushed to exactly 0 (std::pow would return a subnormal or 0; at most 2^-1022 per operator).
    const int under = (k + 1022 * 32) >> 31;                           // all ones when 2^(k/32) < 2^-1022
    k = min(max(k, -1021 * 32), 1023 * 32);
    const double s0 = tab[(2 * POW_LOGN + (k & (POW_EXPN - 1))) * TS];
    // x = 0 or subnormal -> exactly 0 (|F_i - F_j| is full of zeros): a MASK, not a select — a select on so common a
    // condition was compiled into a divergent branch
    const int keep = ~(((hi - 0x00100000) >> 31) | under);
    const double sc = __hiloint2double((__double2hiint(s0) + ((k >> 5) << 20)) & keep, __double2loint(s0) & keep);
    return fma(sc, g * w, sc);
}
// The pair kernel's form of  acc += l * |d|^p  (the arithmetic of kr_pow with fewer instructions around it — the general-p
// body is ISSUE-bound: 20 FP64 instructions per operator against 31 others before this form; 19 against 17 now):
//  * d comes in SIGNED: |d| is an operand modifier of the one FP64 instruction that reads it, and the integer side masks
//    the sign away (fabs() on a value that is then taken apart as integers costs a DADD on the FP64 pipe);
//  * the mantissa is not rebuilt as a double m in [1, 2): r = m / c - 1 = |d| * (2^-e / c) - 1, and 2^-e / c is the table
//    value with its exponent field lowered IN PLACE (mask + three-input add on the loaded high word, no register-pair move).
//    Exact (powers of two) as long as the field stays positive: e <= 1021, i.e. |d| < 2^1022 ~ 4.5e307;
//  * tables in shared memory at an 8 KB-aligned address `tb` (32-bit shared address + 8 * (lane & 15)): [invc 64][logc 64]
//    [exp 32], every entry 16 times (128 B per entry), so an entry address is (index bits) | tb — ONE lop3 after the shift,
//    no add — and logc / exp are reached with immediate offsets (+8192, +16384) of ld.shared;
//  * the exp entries carry their high word pre-compensated by -(i << 15): the scale's exponent field is then
//    hi' + k * 32768 (one IMAD) instead of ((k >> 5) << 20) + hi (shift, mask, add);
//  * results below 2^-1022 need no flag: k is clamped to -1023 * 32, where the scale's high word becomes
//    0x3ff00000 - (1023 << 20) = 0 and its low word (entry 0 = 1.0) is 0: the scale is exactly +0.  (For k in between, the
//    exponent field is 0 and the scale some subnormal: an absolute error below 2^-1022 per operator, as before.)
//    d = 0 / subnormal selects the same clamp value (one compare, one select) -> contributes exactly 0 for every p.
#ifndef KRD_POW_FAST16
#define KRD_POW_FAST16 1
#endif
#ifndef KRD_POW_PSCALE
#define KRD_POW_PSCALE 1   // p folded into the log2 polynomial (coefficients = kernel parameters) and the logc table: one FP64 instruction less
#endif
struct PowCoef { double c5, c4, c3, c2, c1; };   // p / (5 ln2), -p / (4 ln2), p / (3 ln2), -p / (2 ln2), p / ln2
static inline PowCoef pow_coef(double p)
{
    return PowCoef{0.28853900817779268 * p, -0.36067376022224085 * p, 0.48089834696298783 * p, -0.72134752044448170 * p, 1.4426950408889634 * p};
}
__device__ __forceinline__ void kr_pow_acc16(double d, double p, double l, uint32_t tb, double& acc, const PowCoef& pc)
{
    const int hi = __double2hiint(d);
    const int eb = hi & 0x7ff00000;                                     // biased exponent, in place
    uint32_t a1;
    asm("lop3.b32 %0, %1, 0x1f80, %2, 0xEA;" : "=r"(a1) : "r"((uint32_t)hi >> 7), "r"(tb));   // (j * 128) | tb, j = top 6 mantissa bits
    double invc, logc;
    asm("ld.shared.f64 %0, [%1];" : "=d"(invc) : "r"(a1));
    asm("ld.shared.f64 %0, [%1+8192];" : "=d"(logc) : "r"(a1));
    const double sinvc = __hiloint2double(__double2hiint(invc) - eb + 0x3ff00000, __double2loint(invc));   // 2^-e / c
    const double r = fma(fabs(d), sinvc, -1.0);
#if KRD_POW_PSCALE   // p folded into the polynomial and the logc table (filled as p * logc): t = p e + (r q' + logc') — one FP64 less
    double q = fma(r, pc.c5, pc.c4);      // coefficients from the kernel parameters: constant-bank operands (as registers they
    q = fma(r, q, pc.c3);                 // made every DFMA of the chain a three-register-source instruction: slower)
    q = fma(r, q, pc.c2);
    q = fma(r, q, pc.c1);
#else
    double q = fma(r, 0.28853900817779268, -0.36067376022224085);
    q = fma(r, q, 0.48089834696298783);
    q = fma(r, q, -0.72134752044448170);
    q = fma(r, q, 1.4426950408889634);
#endif
#if KRD_POW_I2F
    const double ed = (double)((eb >> 20) - 1023);
#else
    const double ed = __hiloint2double(0x43300000, (eb >> 20) + 0x7ffffc01) - 4503601774854144.0;   // (e ^ 0x80000000 = e + 2^31)
#endif
#if KRD_POW_PSCALE
    const double t = fma(p, ed, fma(r, q, logc));
#else
    const double L = fma(r, q, logc) + ed;
    const double t = p * L;
#endif
    const double MAGIC = 6755399441055744.0;
    double kd = fma(t, 32.0, MAGIC);
    const int k = __double2loint(kd);
#if KRD_POW_I2F >= 2
    kd = (double)k;                                                    // = kd - MAGIC (|32 t| < 2^31), on the conversion unit
#else
    kd -= MAGIC;
#endif
    const double g = fma(kd, -0.03125, t);
    double w = fma(g, 9.6181291076284772e-3, 5.5504108664821580e-2);
    w = fma(g, w, 2.4022650695910071e-1);
    w = fma(g, w, 6.9314718055994531e-1);
    int kc = min(max(k, -1023 * 32), 1023 * 32);
    asm("{\n\t.reg .pred z;\n\tsetp.eq.s32 z, %1, 0;\n\tselp.s32 %0, -32736, %0, z;\n\t}" : "+r"(kc) : "r"(eb));   // d = 0 or subnormal
    uint32_t a2;
    asm("lop3.b32 %0, %1, 0xf80, %2, 0xEA;" : "=r"(a2) : "r"((uint32_t)kc << 7), "r"(tb));    // ((k & 31) * 128) | tb
    double s0;
    asm("ld.shared.f64 %0, [%1+16384];" : "=d"(s0) : "r"(a2));
    const double sc = __hiloint2double((int)((uint32_t)__double2hiint(s0) + (uint32_t)kc * 32768u), __double2loint(s0));
    acc = fma(l, fma(sc, g * w, sc), acc);
}

__device__ __forceinline__ double kr_powi(double x, int n)   // n >= 1, warp-uniform
{
    double r = (n & 1) ? x : 1.0, base = x;
    while (n >>= 1) { base *= base; if (n & 1) r *= base; }
    return r;
}

// x^N with the multiplication chain fixed at compile time (N = 7: x2, x3, x6, x7)
template <int N>
__device__ __forceinline__ double kr_pow_n(double x)
{
    if constexpr (N == 1) return x;
    else if constexpr (N % 2 == 0) { const double h = kr_pow_n<N / 2>(x); return h * h; }
    else return kr_pow_n<N - 1>(x) * x;
}

template <int MODE, int TS = 1>
__device__ __forceinline__ double pair_op(double a, double b, double l, double p, const double* __restrict__ tab, double acc)
{
    if (MODE == MODE_P1) return fma(l, fabs(a - b), acc);
    if (MODE == MODE_P2) { double d = a - b; return fma(d, d, acc); }   // operands pre-scaled by sqrt(len)
    if (MODE == MODE_P2S) { const double d = a - b; return fma(l * d, d, acc); }   // signed len (exact mode, pos outside the branch)
    if (MODE == MODE_PHALF) return fma(l, kr_sqrt_abs(a - b), acc);
    if (MODE == MODE_INT) return fma(l, kr_powi(fabs(a - b), (int)p), acc);
    if (MODE == MODE_P3) { const double d = fabs(a - b); return fma(l * d, d * d, acc); }
    if (MODE == MODE_P4) { const double d = a - b, d2 = d * d; return fma(l * d2, d2, acc); }
    if (MODE == MODE_LIBPOW) return fma(l, pow(fabs(a - b), p), acc);   // exponents beyond kr_pow's range: slow, but any p > 0
    return fma(l, kr_pow<TS>(fabs(a - b), p, tab), acc);
}

// Host: fill this translation unit's copy of the tables (cheap; done before every GENERIC launch).
static inline cudaError_t make_pow_tables()
{
    double tab[POW_TAB];
    for (int j = 0; j < POW_LOGN; ++j) {
        const long double c = 1.0L + ((long double)j + 0.5L) / (long double)POW_LOGN;
        const double invc = (double)(1.0L / c);
        tab[j] = invc;
        tab[POW_LOGN + j] = (double)(-log2l((long double)invc));   // consistent with the ROUNDED 1/c
    }
    for (int i = 0; i < POW_EXPN; ++i) tab[2 * POW_LOGN + i] = (double)exp2l((long double)i / (long double)POW_EXPN);
    return cudaMemcpyToSymbol(g_pow_tab, tab, sizeof tab);
}


constexpr double POW_MAX_P = 30000.0;   // kr_pow builds 2^(p log2 x) from the integer part of 32 p log2 x: |.| < 2^31 needs p < 6e4

static inline int mode_for_exponent(double p, bool gram_for_p2)
{
    if (p > POW_MAX_P) return MODE_LIBPOW;
    if (p == 1.0) return MODE_P1;
    if (p == 2.0) return gram_for_p2 ? MODE_P2G : MODE_P2;   // (MODE_P2S instead of MODE_P2 when a segment length is negative)
    if (p == 0.5) return MODE_PHALF;
    if (p == 3.0) return MODE_P3;
    if (p == 4.0) return MODE_P4;
    if (p == std::floor(p) && p >= 5.0 && p <= 16.0) return MODE_INT;
    return MODE_GENERIC;
}

}  // namespace krd
