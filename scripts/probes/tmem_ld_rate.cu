// Probe: TMEM -> register read-out rate of a B200 SM for the tcgen05.ld shapes the tile kernels could use, alone and with
// packed fp32 arithmetic running beside it in the same warps or in other warps of the same scheduler.  Answers two
// questions of DESIGN.md K3-TCS: what the read-out roof of a 128 x 128 fp32 accumulator tile is (B/clk/SM), and
// whether a tcgen05.ld in flight overlaps with FFMA2 work on the same scheduler.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_ld_rate.bin tmem_ld_rate.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define LD_16x256b_x8(v, addr)                                                                                  \
    asm volatile(                                                                                               \
        "tcgen05.ld.sync.aligned.16x256b.x8.b32 "                                                               \
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                               \
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"               \
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),       \
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), \
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), \
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]) \
        : "r"(addr))
#define LD_32x32b_x32(v, addr)                                                                                  \
    asm volatile(                                                                                               \
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                               \
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                               \
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"               \
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),       \
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), \
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), \
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]) \
        : "r"(addr))

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;"
        : "=l"(reinterpret_cast<unsigned long long&>(d))
        : "l"(reinterpret_cast<unsigned long long&>(a)), "l"(reinterpret_cast<unsigned long long&>(b)),
          "l"(reinterpret_cast<unsigned long long&>(c)));
    return d;
}

// MODE 0: 8 warps read a 128 x 128 fp32 tile per iteration as in the tile kernel (two 16x256b.x8 per warp)
// MODE 1: the same bytes as two 32x32b.x32 per warp
// MODE 2: no loads, 96 FFMA2 per warp and iteration (the fold's packed instruction count)
// MODE 3: loads (16x256b) then 96 FFMA2 on the loaded registers, per warp (the tile kernel's order: serial per warp)
// MODE 4: warps 0-3 only load, warps 4-7 only compute (two warps per scheduler: one loads, one computes)
// MODE 5: loads for the NEXT iteration issued before the 96 FFMA2 on the current registers (register double buffer)
// MODE 6: 96 FFMA2 per warp on warps 0-3 only (one warp per scheduler): the per-warp issue cadence of FFMA2
// MODE 8: as MODE 0 with the tcgen05 fences of the tile kernel around the loads (after_thread_sync before, before_thread_sync after)
// MODE 9: as MODE 8 plus __syncwarp and an mbarrier arrive by lane 0 (the stage release of the tile kernel)
// MODE 7: warps 0-3 load and compute on even iterations, warps 4-7 on odd ones (two warps per scheduler, alternate tiles)
template <int MODE>
__global__ void __launch_bounds__(256, 1) probe(float* out, int iters, long long* cycles) {
    __shared__ uint32_t slot;
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)), "r"(8));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot;
    const uint32_t lq = (uint32_t)(warp & 3) * 32u;
    const uint32_t col = (uint32_t)(warp >> 2) * 64u;
    const uint32_t addr = base + (lq << 16) + col;
    uint32_t v0[32], v1[32], w0[32], w1[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v0[i] = v1[i] = w0[i] = w1[i] = 0x3f800000u + i + lane;
    float2 acc[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) acc[i] = make_float2(0.f, 1.f * i);
    const float2 mul = make_float2(1.0001f, 0.9999f);
    auto compute = [&](uint32_t (&a)[32], uint32_t (&b)[32]) {
#pragma unroll
        for (int rep = 0; rep < 3; ++rep)
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float2 x = make_float2(__uint_as_float(a[2 * i]), __uint_as_float(a[2 * i + 1]));
                const float2 y = make_float2(__uint_as_float(b[2 * i]), __uint_as_float(b[2 * i + 1]));
                acc[(i + rep) % 12] = ffma2(x, mul, acc[(i + rep) % 12]);
                acc[(i + rep + 5) % 12] = ffma2(y, mul, acc[(i + rep + 5) % 12]);
            }
    };
    auto load = [&](uint32_t (&a)[32], uint32_t (&b)[32]) {
        if (MODE == 1) {
            LD_32x32b_x32(a, addr);
            LD_32x32b_x32(b, addr + 32u);
        } else {
            LD_16x256b_x8(a, addr);
            LD_16x256b_x8(b, addr + (16u << 16));
        }
    };
    __syncthreads();
    const long long t0 = clock64();
    if (MODE == 5) load(v0, v1);
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0 || MODE == 1) {
            load(v0, v1);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        } else if (MODE == 2) {
            compute(v0, v1);
        } else if (MODE == 3) {
            load(v0, v1);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            compute(v0, v1);
        } else if (MODE == 4) {
            if (warp < 4) {
                load(v0, v1);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            } else {
                compute(v0, v1);
            }
        } else if (MODE == 8 || MODE == 9) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            load(v0, v1);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            if (MODE == 9) {
                __syncwarp();
                if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)) : "memory");
            }
        } else if (MODE == 6) {
            if (warp < 4) compute(v0, v1);
        } else if (MODE == 7) {
            if (((warp >> 2) ^ it) & 1) {
                load(v0, v1);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                compute(v0, v1);
            }
        } else if (MODE == 5) {
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (it & 1) { load(v0, v1); compute(w0, w1); } else { load(w0, w1); compute(v0, v1); }
        }
    }
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 12; ++i) s += acc[i].x + acc[i].y;
#pragma unroll
    for (int i = 0; i < 32; ++i) s += __uint_as_float(v0[i] ^ v1[i] ^ w0[i] ^ w1[i]);
    if (s == 123.456f) out[0] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(512) : "memory");
}

template <int MODE>
void run(const char* name, double bytes_per_iter, double ffma2_per_iter) {
    float* out;
    long long* cyc;
    cudaMalloc(&out, 4);
    cudaMalloc(&cyc, 148 * sizeof(long long));
    const int iters = 20000;
    probe<MODE><<<148, 256>>>(out, 100, cyc);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a);
    probe<MODE><<<148, 256>>>(out, iters, cyc);
    cudaEventRecord(b);
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double c = 0;
    for (int i = 0; i < 148; ++i) c += (double)h[i];
    c /= 148.0 * iters;
    printf("{\"mode\": \"%s\", \"status\": \"%s\", \"ms\": %.3f, \"cycles_per_iter\": %.1f, \"tmem_bytes_per_clk_per_SM\": %.1f, "
           "\"ffma2_warp_instr_per_clk_per_scheduler\": %.3f}\n",
           name, cudaGetErrorString(e), ms, c, bytes_per_iter / c, ffma2_per_iter / 4.0 / c);
    cudaFree(out);
    cudaFree(cyc);
}

int main() {
    // per iteration and SM: 8 warps x 2 loads x 4 KB = 64 KB; compute: 96 FFMA2 per warp
    run<0>("read 128x128 fp32 tile, 16x256b.x8 x2 per warp, 8 warps", 65536.0, 0.0);
    run<1>("read 128x128 fp32 tile, 32x32b.x32 x2 per warp, 8 warps", 65536.0, 0.0);
    run<2>("96 FFMA2 per warp, 8 warps, no loads", 0.0, 8 * 96.0);
    run<3>("load then 96 FFMA2 per warp (serial per warp), 8 warps", 65536.0, 8 * 96.0);
    run<4>("warps 0-3 load (32 KB), warps 4-7 compute", 32768.0, 4 * 96.0);
    run<5>("register double buffer: next loads in flight during 96 FFMA2", 65536.0, 8 * 96.0);
    run<8>("read tile with tcgen05.fence after/before around the loads", 65536.0, 0.0);
    run<9>("read tile with fences, __syncwarp and mbarrier arrive", 65536.0, 0.0);
    run<6>("96 FFMA2 per warp, warps 0-3 only (one warp per scheduler)", 0.0, 4 * 96.0);
    run<7>("alternate iterations: warps 0-3 even, warps 4-7 odd, load (32 KB) then 96 FFMA2", 32768.0, 4 * 96.0);
    return 0;
}
