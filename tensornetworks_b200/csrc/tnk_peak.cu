// ALU roofline probe: the denominator of the pair kernel's roofline fraction.
// MEASURED_PEAKS.json holds HBM and tensor peaks only; the pair kernel is bound by the
// FP32 (fp32 mode) or FP64 (fp64 mode) FMA issue rate, so that rate is measured here with
// independent dependent-chain FMAs on register operands (16 chains per thread, enough to
// cover the 4-cycle FMA latency at any occupancy).
#include "tnk_common.cuh"

namespace {

template <typename T>
__global__ void __launch_bounds__(256)
fma_chain_kernel(int iters, T a, T b, double* sink) {
    T x[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) x[c] = (T)(threadIdx.x + c) * (T)1e-3;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep)
#pragma unroll
            for (int c = 0; c < 16; ++c) x[c] = fma(x[c], a, b);
    }
    T s = (T)0;
#pragma unroll
    for (int c = 0; c < 16; ++c) s += x[c];
    if (s == (T)123.456) sink[0] = (double)s;  // keeps the chains alive, never true in practice
}

// packed variant: one FFMA2 (fma.rn.f32x2) per chain step = 2 lane-ops per instruction
__global__ void __launch_bounds__(256)
fma2_chain_kernel(int iters, float a, float b, double* sink) {
    float2 x[16];
    const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
#pragma unroll
    for (int c = 0; c < 16; ++c) x[c] = make_float2((threadIdx.x + c) * 1e-3f, (threadIdx.x + c) * 2e-3f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep)
#pragma unroll
            for (int c = 0; c < 16; ++c) x[c] = __ffma2_rn(x[c], a2, b2);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) s += x[c].x + x[c].y;
    if (s == 123.456f) sink[0] = (double)s;
}

}  // namespace

extern "C" int tnk_alu_peak(int precision, int iters, double* scratch, double* lane_ops_per_s, float* ms_out, void* stream) {
    TNK_REQUIRE(precision == TNK_PREC_FP64 || precision == TNK_PREC_FP32 || precision == TNK_PEAK_FP32X2,
                "tnk_alu_peak: precision must be 0 (DFMA), 1 (FFMA) or 3 (packed FFMA2)");
    TNK_REQUIRE(iters > 0 && scratch && lane_ops_per_s, "tnk_alu_peak: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0, sms = 0;
    TNK_CUDA_OK(cudaGetDevice(&dev));
    TNK_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int blocks = sms * 8;  // 8 CTAs of 256 threads per SM = full occupancy
    cudaEvent_t e0, e1;
    TNK_CUDA_OK(cudaEventCreate(&e0));
    TNK_CUDA_OK(cudaEventCreate(&e1));
    for (int pass = 0; pass < 2; ++pass) {  // pass 0 warms up
        TNK_CUDA_OK(cudaEventRecord(e0, st));
        if (precision == TNK_PREC_FP32) fma_chain_kernel<float><<<blocks, 256, 0, st>>>(iters, 0.999f, 1e-3f, scratch);
        else if (precision == TNK_PEAK_FP32X2) fma2_chain_kernel<<<blocks, 256, 0, st>>>(iters, 0.999f, 1e-3f, scratch);
        else fma_chain_kernel<double><<<blocks, 256, 0, st>>>(iters, 0.999, 1e-3, scratch);
        TNK_CUDA_OK(cudaGetLastError());
        TNK_CUDA_OK(cudaEventRecord(e1, st));
        TNK_CUDA_OK(cudaEventSynchronize(e1));
    }
    float ms = 0.f;
    TNK_CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    const double ops = (double)blocks * 256.0 * (double)iters * 64.0 * (precision == TNK_PEAK_FP32X2 ? 2.0 : 1.0);
    *lane_ops_per_s = ops / ((double)ms * 1e-3);
    if (ms_out) *ms_out = ms;
    return 0;
}
