// Probe: register <-> (lane, column) mapping of tcgen05.ld.16x256b.x8 against data written with
// tcgen05.st.32x32b (thread = lane, register = column).  Build: nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(128) probe(uint32_t* out) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(64) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot;
    const int lq = warp * 32;
    const int L = lq + lane;
    for (int c0 = 0; c0 < 64; c0 += 8) {
        uint32_t v[8];
        for (int e = 0; e < 8; ++e) v[e] = (uint32_t)(L * 1000 + c0 + e);
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                     ::"r"(base + ((uint32_t)lq << 16) + (uint32_t)c0),
                       "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int hrow = 0; hrow < 2; ++hrow) {
        uint32_t r[32];
        const uint32_t taddr = base + ((uint32_t)(lq + 16 * hrow) << 16);
        asm volatile(
            "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
              "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
              "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
            : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int k = 0; k < 32; ++k) out[((warp * 2 + hrow) * 32 + lane) * 32 + k] = r[k];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(64) : "memory");
}

int main() {
    uint32_t* d;
    const int n = 4 * 2 * 32 * 32;
    cudaMalloc(&d, n * 4);
    cudaMemset(d, 0xff, n * 4);
    probe<<<1, 128>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    static uint32_t h[4 * 2 * 32 * 32];
    cudaMemcpy(h, d, n * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int w = 0; w < 4; ++w)
        for (int hr = 0; hr < 2; ++hr)
            for (int t = 0; t < 32; ++t)
                for (int k = 0; k < 32; ++k) {
                    const uint32_t v = h[((w * 2 + hr) * 32 + t) * 32 + k];
                    const int m = k >> 2, r = (k >> 1) & 1, e = k & 1;
                    const int lane = 32 * w + 16 * hr + t / 4 + 8 * r, col = 8 * m + 2 * (t % 4) + e;
                    if (v != (uint32_t)(lane * 1000 + col)) {
                        if (bad < 40) printf("w%d hr%d t%d k%d: got lane %u col %u, expected lane %d col %d\n", w, hr, t, k, v / 1000, v % 1000, lane, col);
                        ++bad;
                    }
                }
    printf("mismatches: %d of %d\n", bad, n);
    // dump thread 0, 1, 5 of warp 1 hrow 1 for inspection
    const int ts[3] = {0, 1, 5};
    for (int x = 0; x < 3; ++x) {
        printf("w1 hr1 t%d:", ts[x]);
        for (int k = 0; k < 12; ++k) { const uint32_t v = h[((1 * 2 + 1) * 32 + ts[x]) * 32 + k]; printf(" (%u,%u)", v / 1000, v % 1000); }
        printf("\n");
    }
    return 0;
}
