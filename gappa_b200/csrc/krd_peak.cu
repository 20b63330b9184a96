// FP64 roofline denominators measured on the device itself (MEASURED_PEAKS.json has no FP64 entry).
// Register-resident loops, no memory traffic: what the FP64 pipe can issue, nothing else.
#include <cuda_runtime.h>

#include "../../include/krd_b200.h"

namespace {

constexpr int ILP = 16;
constexpr int INNER = 256;

// mode 0: DFMA only.  mode 1: DADD + DFMA pairs (the p = 1 inner operator).
template <int MODE>
__global__ void __launch_bounds__(256) k_fp64_peak(double seed, int iters, double* sink)
{
    double x[ILP], y[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) { x[i] = seed + i + threadIdx.x; y[i] = seed * 0.5 + i; }
    const double a = seed * 1e-9, b = 1.0 - seed * 1e-12;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < INNER; ++k) {
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                if (MODE == 0) {
                    x[i] = fma(x[i], b, a);
                } else if (MODE == 3) {
                    x[i] = x[i] + fabs(y[(i + k) % ILP]);
                } else if (MODE == 4) {
                    double d = x[i] - y[(i + k) % ILP];
                    x[i] = x[i] + fabs(d);
                } else if (MODE == 5) {
                    x[i] = fma(a, b, x[i]);
                } else if (MODE == 6) {       // MUFU.RSQ64H: the FP64 reciprocal-square-root seed
                    asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(x[i]) : "d"(x[i]));
                } else if (MODE == 7) {       // MUFU.RSQ: the FP32 seed
                    float f = __int_as_float(__double2hiint(x[i]));
                    asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(f) : "f"(f));
                    x[i] = __hiloint2double(__float_as_int(f), __double2loint(x[i]));
                } else if (MODE == 8) {       // F2F.F32.F64 + F2F.F64.F32 round trip
                    float f;
                    asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(f) : "d"(x[i]));
                    asm volatile("cvt.f64.f32 %0, %1;" : "=d"(x[i]) : "f"(f));
                } else {
                    double d = x[i] - y[(i + k) % ILP];
                    x[i] = fma(a, fabs(d), x[i]);
                }
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i];
    if (s == 12345.678) *sink = s;
}

// mode 2: FP64 tensor-core path, mma.sync m8n8k4 (DMMA). 8*8*4 FMA per warp instruction.
__global__ void __launch_bounds__(256) k_dmma_peak(double seed, int iters, double* sink)
{
    double c[ILP][2];
#pragma unroll
    for (int i = 0; i < ILP; ++i) { c[i][0] = seed + i; c[i][1] = seed - i; }
    const double a = seed * 1e-9 + threadIdx.x * 1e-12, b = 1.0 - seed * 1e-12;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < INNER / 4; ++k) {
#pragma unroll
            for (int i = 0; i < ILP; ++i) {
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1];
    if (s == 12345.678) *sink = s;
}

}  // namespace

extern "C" double krd_measure_fp64_peak(int device, int mode, double ms_target)
{
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return -1.0; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1.0;
    double* sink = nullptr;
    if (cudaMalloc(&sink, 8) != cudaSuccess) return -1.0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = prop.multiProcessorCount * 4, block = 256;
    auto launch = [&](int iters) {
        if (mode == 0) k_fp64_peak<0><<<grid, block>>>(1.000001, iters, sink);
        else if (mode == 1) k_fp64_peak<1><<<grid, block>>>(1.000001, iters, sink);
        else if (mode == 3) k_fp64_peak<3><<<grid, block>>>(1.000001, iters, sink);
        else if (mode == 4) k_fp64_peak<4><<<grid, block>>>(1.000001, iters, sink);
        else if (mode == 5) k_fp64_peak<5><<<grid, block>>>(1.000001, iters, sink);
        else if (mode == 6) k_fp64_peak<6><<<grid, block>>>(1.000001, iters, sink);
        else if (mode == 7) k_fp64_peak<7><<<grid, block>>>(1.000001, iters, sink);
        else if (mode == 8) k_fp64_peak<8><<<grid, block>>>(1.000001, iters, sink);
        else k_dmma_peak<<<grid, block>>>(1.000001, iters, sink);
    };
    auto flops = [&](int iters) -> double {
        const double threads = (double)grid * block;
        if (mode == 0 || mode == 5) return threads * iters * (double)INNER * ILP * 2.0;
        if (mode == 3 || mode == 6 || mode == 7) return threads * iters * (double)INNER * ILP * 1.0;   // 1 op each
        if (mode == 8) return threads * iters * (double)INNER * ILP * 2.0;            // two conversions
        if (mode == 4) return threads * iters * (double)INNER * ILP * 2.0;            // DADD + DADD
        if (mode == 1) return threads * iters * (double)INNER * ILP * 3.0;        // DADD (1) + DFMA (2)
        return (threads / 32.0) * iters * (double)(INNER / 4) * ILP * (8 * 8 * 4 * 2.0);
    };
    launch(4);  // warm-up
    cudaDeviceSynchronize();
    int iters = 16;
    float ms = 0.f;
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0);
        launch(iters);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
        cudaEventElapsedTime(&ms, e0, e1);
        double tf = flops(iters) / (ms * 1e-3) / 1e12;
        if (ms > 0.2 * ms_target && tf > best) best = tf;
        if (ms < ms_target) iters = (int)(iters * (ms_target / (ms > 0.01f ? ms : 0.01f)) * 1.05) + 1;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(sink);
    cudaGetLastError();
    return best;
}
