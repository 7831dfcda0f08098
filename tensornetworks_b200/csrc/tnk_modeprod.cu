// Mode-product (Kronecker) evaluation of the same quadratic form: u = K_p q in O(n^2 2^n).
//
// The Hamming base kernel factorises over bits, K = M^{(x)n} with M = [[1,a],[a,1]]
// (reference stein_utils.py:30-55: k = exp(-h/(n l)) = prod_p a^{z_p xor z'_p}), and the
// single-bit-flip differences of Eq. 13 (stein_utils.py:166-194) are (I - X_p) K, X_p = flip
// of bit p, which commutes with K.  With W_p = K (q .* s_p), w = K q, r_p = 1 - s_p:
//
//   u[z] = sum_p [ -r_p(z) W_p[z] + W_p[z^e_p] + (r_p(z) + 1) (w[z] - w[z^e_p]) ]
//
// i.e. n+1 transforms by K (each n butterfly stages over a 2^n vector) and one gather pass.
// This is the same number the O(4^n) pair kernel produces (tests compare both with the oracle);
// it is HBM/L2-bound instead of ALU-bound.  All arithmetic is fp64.
//
// Kernels:  mp_pack (build the n_loc+1 input vectors), mp_transform (one pass = up to 12 bits
// of butterflies on a shared-memory tile, radix-8 in registers), mp_combine (+ block partials
// of q.u), mp_scalar (fixed-order final sum).
#include "tnk_common.cuh"

namespace {

constexpr int kMpThreads = 256;
constexpr int kTileLog = 12;               // elements per shared-memory tile (2^12 doubles + padding)
constexpr int kMinLogW = 3;                // strided passes keep >= 8 contiguous doubles (64 B) per row

__device__ __forceinline__ int pad_idx(int e) { return e + (e >> 3); }  // 1 pad per 8: conflict-free strides

// ---- MP1: V[lv][z] = q[z] * s_p(z) for p = p0 + lv (lv < n_loc); V[n_loc][z] = q[z]
template <typename Q>
__global__ void __launch_bounds__(kMpThreads)
mp_pack_kernel(int np, int64_t N, int p0, int n_loc, const Q* __restrict__ q,
               const double* __restrict__ ST, double* __restrict__ V) {
    const int64_t z = (int64_t)blockIdx.x * kMpThreads + threadIdx.x;
    if (z >= N) return;
    const double qz = (double)q[z];
    const double* s = ST + z * np + p0;
    for (int lv = 0; lv < n_loc; ++lv) V[(int64_t)lv * N + z] = qz * s[lv];
    V[(int64_t)n_loc * N + z] = qz;
}

// ---- MP2: butterflies for bits [b0, b0+g) on tiles of 2^g rows x 2^logW contiguous elements.
// Element (m, w) of tile (hi, lob) sits at  hi << (b0+g) | m << b0 | lob << logW | w.
// b0 == 0 requires logW == 0 (the tile is one contiguous run).
__global__ void __launch_bounds__(kMpThreads)
mp_transform_kernel(double* __restrict__ V, int64_t N, int b0, int g, int logW, double a) {
    extern __shared__ double tile[];
    const int tile_elems = 1 << (g + logW);
    const int W = 1 << logW;
    const int64_t nlob = (b0 > logW) ? ((int64_t)1 << (b0 - logW)) : 1;
    const int64_t t = blockIdx.x;
    const int64_t lob = t % nlob, hi = t / nlob;
    double* vec = V + (int64_t)blockIdx.y * N + ((hi << (b0 + g)) | (lob << logW));

    for (int e = threadIdx.x; e < tile_elems; e += kMpThreads) {
        const int m = e >> logW, w = e & (W - 1);
        tile[pad_idx(e)] = vec[((int64_t)m << b0) + w];
    }
    __syncthreads();

    for (int k0 = 0; k0 < g; k0 += 3) {
        const int rb = (g - k0 < 3) ? (g - k0) : 3;      // bits handled in registers this round
        const int cnt = 1 << rb;
        const int units = tile_elems >> rb;
        for (int uidx = threadIdx.x; uidx < units; uidx += kMpThreads) {
            const int w = uidx & (W - 1), mm = uidx >> logW;
            const int mlow = mm & ((1 << k0) - 1), mhigh = mm >> k0;
            const int mbase = (mhigh << (k0 + rb)) | mlow;
            double x[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (j < cnt) x[j] = tile[pad_idx(((mbase | (j << k0)) << logW) + w)];
#pragma unroll
            for (int b = 0; b < 3; ++b) {
                if (b < rb) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (!(j & (1 << b)) && (j | (1 << b)) < cnt) {
                            const double lo = x[j], hi2 = x[j | (1 << b)];
                            x[j] = fma(a, hi2, lo);
                            x[j | (1 << b)] = fma(a, lo, hi2);
                        }
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (j < cnt) tile[pad_idx(((mbase | (j << k0)) << logW) + w)] = x[j];
        }
        __syncthreads();
    }

    for (int e = threadIdx.x; e < tile_elems; e += kMpThreads) {
        const int m = e >> logW, w = e & (W - 1);
        vec[((int64_t)m << b0) + w] = tile[pad_idx(e)];
    }
}

// ---- MP3: u[z] = sum_p [ -r_p W_p[z] + W_p[z^e_p] + (r_p + 1)(w[z] - w[z^e_p]) ],  block partials of q.u
template <typename Q>
__global__ void __launch_bounds__(kMpThreads)
mp_combine_kernel(int np, int64_t N, int p0, int n_loc, const Q* __restrict__ q,
                  const double* __restrict__ ST, const double* __restrict__ V,
                  double* __restrict__ u, double* __restrict__ blockpart) {
    __shared__ double red[kMpThreads];
    const int64_t z = (int64_t)blockIdx.x * kMpThreads + threadIdx.x;
    double contrib = 0.0;
    if (z < N) {
        const double* s = ST + z * np + p0;
        const double* wq = V + (int64_t)n_loc * N;
        const double wz = wq[z];
        double acc = 0.0;
        for (int lv = 0; lv < n_loc; ++lv) {
            const int64_t zf = z ^ ((int64_t)1 << (p0 + lv));
            const double* Wp = V + (int64_t)lv * N;
            const double r = 1.0 - s[lv];
            acc += fma(-r, Wp[z], Wp[zf]) + (r + 1.0) * (wz - wq[zf]);
        }
        u[z] = acc;
        contrib = (double)q[z] * acc;
    }
    red[threadIdx.x] = contrib;
    __syncthreads();
    for (int off = kMpThreads / 2; off > 0; off >>= 1) {
        if (threadIdx.x < off) red[threadIdx.x] += red[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) blockpart[blockIdx.x] = red[0];
}

__global__ void __launch_bounds__(256)
mp_scalar_kernel(const double* __restrict__ blockpart, int64_t nblocks, double* __restrict__ out) {
    __shared__ double red[256];
    double acc = 0.0;
    for (int64_t b = threadIdx.x; b < nblocks; b += 256) acc += blockpart[b];
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if (threadIdx.x < off) red[threadIdx.x] += red[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = red[0];
}

struct MpPlan {
    int64_t N, blocks;
    int n_loc;
    size_t off_v, off_block, total;
};

int mp_plan(int n, int p_begin, int p_end, MpPlan* p) {
    if (n < 0 || n > TNK_MAX_LATENT || p_begin < 0 || p_end > n || p_begin > p_end) return 1;
    p->N = (int64_t)1 << n;
    p->n_loc = p_end - p_begin;
    p->blocks = (p->N + kMpThreads - 1) / kMpThreads;
    size_t off = 0;
    p->off_v = off;     off += ((size_t)(p->n_loc + 1) * p->N * sizeof(double) + 255) / 256 * 256;
    p->off_block = off; off += ((size_t)(p->blocks + 1) * sizeof(double) + 255) / 256 * 256;
    p->total = off;
    return 0;
}

}  // namespace

extern "C" size_t tnk_ksd_modeproduct_workspace_bytes(int n_latent, int bit_begin, int bit_end) {
    MpPlan p;
    if (mp_plan(n_latent, bit_begin, bit_end, &p)) return 0;
    return p.total;
}

extern "C" int tnk_ksd_modeproduct(int n_latent, double length_scale, const void* q, int q_dtype,
                                   const double* ST64, int bit_begin, int bit_end,
                                   double* u, double* ksd_part,
                                   void* workspace, size_t workspace_bytes, int* n_launch, void* stream) {
    MpPlan p;
    TNK_REQUIRE(mp_plan(n_latent, bit_begin, bit_end, &p) == 0, "tnk_ksd_modeproduct: bad arguments n=%d bits=[%d,%d)",
                n_latent, bit_begin, bit_end);
    TNK_REQUIRE(q && ST64 && u && ksd_part && workspace, "tnk_ksd_modeproduct: NULL pointer argument");
    TNK_REQUIRE(q_dtype == 0 || q_dtype == 1, "tnk_ksd_modeproduct: q_dtype must be 0 (double) or 1 (float)");
    TNK_REQUIRE(workspace_bytes >= p.total, "tnk_ksd_modeproduct: workspace %zu < required %zu", workspace_bytes, p.total);
    TNK_REQUIRE(((uintptr_t)workspace & 255) == 0, "tnk_ksd_modeproduct: workspace must be 256-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const int n = n_latent;
    int launches = 0;
    if (n == 0 || p.n_loc == 0) {  // n == 0: k_p == 0 (stein_utils.py:151-152); empty bit range: zero partial
        TNK_CUDA_OK(cudaMemsetAsync(ksd_part, 0, sizeof(double), st));
        TNK_CUDA_OK(cudaMemsetAsync(u, 0, (size_t)p.N * sizeof(double), st));
        if (n_launch) *n_launch = 0;
        return 0;
    }
    TNK_REQUIRE((double)n * length_scale != 0.0, "tnk_ksd_modeproduct: num_vars*length_scale must be non-zero");
    TnkGeom geo;
    TNK_REQUIRE(tnk_geom(n, TNK_PREC_FP64, &geo) == 0, "tnk_ksd_modeproduct: unsupported n=%d", n);
    const double a = exp(-1.0 / ((double)n * length_scale));
    unsigned char* ws = (unsigned char*)workspace;
    double* V = (double*)(ws + p.off_v);
    double* blockpart = (double*)(ws + p.off_block);
    const unsigned blocks = (unsigned)p.blocks;
    const int nvec = p.n_loc + 1;

    if (q_dtype == 0) mp_pack_kernel<double><<<blocks, kMpThreads, 0, st>>>(geo.np, p.N, bit_begin, p.n_loc, (const double*)q, ST64, V);
    else mp_pack_kernel<float><<<blocks, kMpThreads, 0, st>>>(geo.np, p.N, bit_begin, p.n_loc, (const float*)q, ST64, V);
    TNK_CUDA_OK(cudaGetLastError());
    ++launches;

    // pass plan: one contiguous pass of min(n,12) bits, then strided passes of <= 9 bits each
    const size_t smem = ((size_t)(1 << kTileLog) * 9 / 8 + 8) * sizeof(double);
    TNK_CUDA_OK(cudaFuncSetAttribute(mp_transform_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int b0 = 0;
    {
        const int g = n < kTileLog ? n : kTileLog;
        dim3 grid((unsigned)(p.N >> g), (unsigned)nvec, 1);
        mp_transform_kernel<<<grid, kMpThreads, smem, st>>>(V, p.N, 0, g, 0, a);
        TNK_CUDA_OK(cudaGetLastError());
        ++launches;
        b0 = g;
    }
    const int rem = n - b0;
    if (rem > 0) {
        const int gmax = kTileLog - kMinLogW;
        const int passes = (rem + gmax - 1) / gmax;
        int left = rem;
        for (int ps = 0; ps < passes; ++ps) {
            const int g = (left + (passes - ps) - 1) / (passes - ps);
            int logW = kTileLog - g;
            if (logW > b0) logW = b0;
            dim3 grid((unsigned)(p.N >> (g + logW)), (unsigned)nvec, 1);
            mp_transform_kernel<<<grid, kMpThreads, smem, st>>>(V, p.N, b0, g, logW, a);
            TNK_CUDA_OK(cudaGetLastError());
            ++launches;
            b0 += g;
            left -= g;
        }
    }

    if (q_dtype == 0) mp_combine_kernel<double><<<blocks, kMpThreads, 0, st>>>(geo.np, p.N, bit_begin, p.n_loc, (const double*)q, ST64, V, u, blockpart);
    else mp_combine_kernel<float><<<blocks, kMpThreads, 0, st>>>(geo.np, p.N, bit_begin, p.n_loc, (const float*)q, ST64, V, u, blockpart);
    TNK_CUDA_OK(cudaGetLastError());
    mp_scalar_kernel<<<1, 256, 0, st>>>(blockpart, p.blocks, ksd_part);
    TNK_CUDA_OK(cudaGetLastError());
    launches += 2;
    if (n_launch) *n_launch = launches;
    return 0;
}
