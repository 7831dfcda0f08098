// Probe: FP64 throughput of a B200 SM through (a) DFMA chains, (b) mma.sync m8n8k4 f64 (DMMA) chains,
// (c) both interleaved.  Prints FMA lane-ops/s (one fp64 multiply-add = 1).  Decides whether the fp64
// pair kernel's dot product may move to DMMA (separate pipe?) - see DESIGN.md K3-S64.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_rate.bin dmma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int MODE>
__global__ void __launch_bounds__(256) probe(double* out, int iters, double a, double b) {
    double c[8][2];
    double f[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { c[k][0] = threadIdx.x; c[k][1] = k; f[k] = k + threadIdx.x; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (MODE & 1) dmma(c[k][0], c[k][1], a, b);
            if (MODE & 2) {
                f[k] = fma(f[k], a, b);
                if (MODE & 4) f[k] = fma(f[k], a, b);
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += c[k][0] + c[k][1] + f[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, double* out, int blocks) {
    const int iters = 20000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<MODE><<<blocks, 256>>>(out, 100, 1.0000001, 1e-9);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    probe<MODE><<<blocks, 256>>>(out, iters, 1.0000001, 1e-9);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warps = (double)blocks * 8;
    const double dmma_ops = (MODE & 1) ? warps * iters * 8.0 * 256.0 : 0.0;
    const double dfma_ops = (MODE & 2) ? warps * iters * 8.0 * 32.0 * ((MODE & 4) ? 2 : 1) : 0.0;
    printf("%-28s %8.3f ms  DMMA %.3e fma/s  DFMA %.3e fma/s  total %.3e\n", name, ms, dmma_ops / (ms * 1e-3),
           dfma_ops / (ms * 1e-3), (dmma_ops + dfma_ops) / (ms * 1e-3));
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int blocks = p.multiProcessorCount * 4;
    double* out; cudaMalloc(&out, sizeof(double) * blocks * 256);
    run<2>("DFMA only", out, blocks);
    run<6>("DFMA only (2 per slot)", out, blocks);
    run<1>("DMMA m8n8k4 only", out, blocks);
    run<3>("DMMA + 1 DFMA per DMMA", out, blocks);
    run<7>("DMMA + 2 DFMA per DMMA", out, blocks);
    printf("cuda error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
