// Do the FP64 tensor sub-pipe (DMMA m8n8k4) and the FP64 vector pipe (DADD/DFMA) of a B200 SM run concurrently?
// Register-resident loops: A = DMMA only, B = DADD+DFMA only, C = both interleaved in every warp, D = warps
// specialised (even warps DMMA, odd warps vector).  Concurrent pipes: t(C) ~ max(t(A), t(B)); one shared datapath:
// t(C) ~ t(A) + t(B).   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pipes_overlap pipes_overlap.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int NA = 16;   // accumulator pairs (DMMA) / accumulators (vector)

template <int NM, int NV, bool SPLIT>
__global__ void __launch_bounds__(256) k(double seed, int iters, double* sink)
{
    double c[NA][2], x[NA], y[NA];
#pragma unroll
    for (int i = 0; i < NA; ++i) { c[i][0] = seed + i; c[i][1] = seed - i; x[i] = seed + i + threadIdx.x; y[i] = seed * 0.5 + i; }
    const double a = seed * 1e-9 + threadIdx.x * 1e-12, b = 1.0 - seed * 1e-12;
    const bool do_m = !SPLIT || ((threadIdx.x >> 5) & 1) == 0;
    const bool do_v = !SPLIT || ((threadIdx.x >> 5) & 1) == 1;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) {
            if (SPLIT) {
                if (do_m) {
#pragma unroll
                    for (int i = 0; i < NM; ++i)
                        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                                     : "+d"(c[i % NA][0]), "+d"(c[i % NA][1]) : "d"(a), "d"(b));
                }
                if (do_v) {
#pragma unroll
                    for (int i = 0; i < NV; ++i) { const double d = x[i % NA] - y[(i + k2) % NA]; x[i % NA] = fma(a, fabs(d), x[i % NA]); }
                }
            } else {
                // interleave: after every DMMA, NV/NM vector pairs
#pragma unroll
                for (int i = 0; i < (NM > 0 ? NM : 1); ++i) {
                    if (NM > 0)
                        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                                     : "+d"(c[i % NA][0]), "+d"(c[i % NA][1]) : "d"(a), "d"(b));
                    constexpr int PER = NM > 0 ? NV / NM : NV;
#pragma unroll
                    for (int j = 0; j < PER; ++j) {
                        const int q = (i * PER + j) % NA;
                        const double d = x[q] - y[(q + k2) % NA];
                        x[q] = fma(a, fabs(d), x[q]);
                    }
                }
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NA; ++i) s += c[i][0] + c[i][1] + x[i];
    if (s == 12345.678) *sink = s;
}

template <int NM, int NV, bool SPLIT>
static void run(const char* name, int sms)
{
    double* sink; cudaMalloc(&sink, 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = sms * 4, iters = 20000;
    k<NM, NV, SPLIT><<<grid, 256>>>(1.000001, 100, sink);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        cudaEventRecord(e0);
        k<NM, NV, SPLIT><<<grid, 256>>>(1.000001, iters, sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double warps = (double)grid * 8 * (SPLIT ? 0.5 : 1.0);
    const double dmma = warps * iters * 8.0 * NM, vec = warps * iters * 8.0 * NV * 32.0;   // warp DMMAs; thread-level (DADD+DFMA) pairs
    printf("%-34s %8.3f ms   DMMA %6.2f TFLOP/s   vector %6.2f TFLOP/s (3-flop pairs) = %5.2f e12 instr/s\n", name, best,
           dmma * 512.0 / (best * 1e-3) / 1e12, vec * 3.0 / (best * 1e-3) / 1e12, vec * 2.0 / (best * 1e-3) / 1e12);
    cudaFree(sink);
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    printf("%s, %d SMs\n", p.name, sms);
    run<16, 0, false>("A  DMMA only (16/iter)", sms);
    run<0, 32, false>("B  DADD+DFMA only (32 pairs/iter)", sms);
    run<16, 32, false>("C  interleaved 16 DMMA + 32 pairs", sms);
    run<16, 16, false>("C2 interleaved 16 DMMA + 16 pairs", sms);
    run<8, 32, false>("C3 interleaved 8 DMMA + 32 pairs", sms);
    run<16, 32, true>("D  split warps: 16 DMMA | 32 pairs", sms);
    return 0;
}
