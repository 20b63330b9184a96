// Are sequences of small kernel launches slower while an H2D copy saturates PCIe, and does a CUDA graph avoid it?
#include <cstdio>
#include <chrono>
#include <cuda_runtime.h>
__global__ void tiny(double* p, int n) { int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) p[i] = p[i] * 1.0000001 + 1.0; }
int main()
{
    const size_t BYTES = 400u << 20; void* hsrc; cudaMallocHost(&hsrc, BYTES); void* ddst; cudaMalloc(&ddst, BYTES);
    double* d; cudaMalloc(&d, 1 << 20); cudaMemset(d, 0, 1 << 20);
    cudaStream_t s, c; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&c, cudaStreamNonBlocking);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int NK = 30, REP = 8;
    cudaGraph_t g; cudaGraphExec_t ge;
    cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed);
    for (int k = 0; k < NK; ++k) tiny<<<64, 256, 0, s>>>(d, 16384);
    cudaStreamEndCapture(s, &g); cudaGraphInstantiate(&ge, g, 0); cudaGraphUpload(ge, s); cudaStreamSynchronize(s);
    for (int load = 0; load < 2; ++load)
        for (int mode = 0; mode < 2; ++mode) {
            float best = 1e9, sum = 0;
            for (int it = 0; it < 3; ++it) {
                cudaDeviceSynchronize();
                if (load) cudaMemcpyAsync(ddst, hsrc, BYTES, cudaMemcpyHostToDevice, c);
                cudaEventRecord(e0, s);
                for (int r = 0; r < REP; ++r) {
                    if (mode == 0) for (int k = 0; k < NK; ++k) tiny<<<64, 256, 0, s>>>(d, 16384);
                    else cudaGraphLaunch(ge, s);
                    cudaStreamSynchronize(s);   // a host sync per "phase", as in the pipelined ingest
                }
                cudaEventRecord(e1, s); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); sum += ms; if (ms < best) best = ms;
            }
            printf("%s  %s : %d x %d kernels  best %.3f ms  (%.1f us per kernel)\n", load ? "H2D running" : "idle PCIe  ", mode ? "graph" : "eager", REP, NK, best, best * 1e3 / (REP * NK));
        }
    return 0;
}
