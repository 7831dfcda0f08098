// K3-TC: tensor-core tile variant of the Stein pair kernel.
//
// The bracket of k_p(z,z') is bilinear in per-state feature vectors,
//   q_j * [ s(i).s(j) + C(i) + C(j) + g*T1(i,j) + g*T2(i,j) ] = U(i) . V(j),
// so a 128 x 128 tile of brackets is one GEMM tile: tcgen05.mma (kind::tf32) with the
// accumulator in TMEM.  fp32-grade accuracy (the 1e-4 parity mode) comes from the usual
// 3xTF32 split x = hi + lo (hi = top 10 mantissa bits, lo = x - hi exactly): every real
// product x*y is issued as hi*hi + hi*lo + lo*hi; factors that are exact in tf32 (bits, 1)
// need no split.  The a^popc(i^j) weight does NOT factorise into a low-rank form, so it is
// applied per pair in the epilogue: each epilogue thread owns one row (TMEM lane), pulls 32
// accumulator columns at a time with tcgen05.ld and folds them with a 16-entry register table
// a^popc((i^c)&15) (one FFMA per pair) times one shared-memory LUT value per 16 columns.
// Every ordered pair of the rows x columns handed to the kernel is evaluated (4^n in total).
//
// Operand tables (built per call by tc_pack_kernel, K-major "interleaved" no-swizzle UMMA
// layout, one 128-state tile contiguous so that a whole tile is one cp.async.bulk):
//   tile t, chunk c, row r:  float4 at ((t*K + c)*128 + r)       (chunk = 4 consecutive k)
// A (rows i):     s_hi | s_lo | bits | g_hi | g_lo | [1,1,C_hi,C_hi] | [C_lo,0,0,0]
// B (columns j):  t_hi | t_lo | r_hi | r_lo | bq_hi | bq_lo | [(qC)_hi,(qC)_lo,q_hi,q_lo] | [q_hi,0,0,0]
//   t = q_j s(j),  g = gamma*rho(i),  r = gamma*q_j*rho(j),  bq = b(j) q_j,  rho_p = (1-s_p)(1-2b_p)
// One MMA K-step (K = 8 tf32) consumes two (A-chunk, B-chunk) products; the host planner
// pairs the products so that both chunk offsets increase inside a step (they become the
// descriptor's leading-dimension byte offsets) - shared chunks (s_hi, t_hi, bits, ...) are
// stored once and referenced by several steps.
//
// Roles in a CTA of 192 threads (one CTA per SM, ~190 KB shared memory):
//   warp 0   producer: A tile once, then B column tiles through a 2-stage ring (bulk copy + mbarrier)
//   warp 1   allocates TMEM (2 x 128 columns), one lane issues the MMAs and commits
//   warps 2-5 epilogue: TMEM -> registers, Hamming weights, row sums in fp32 per tile / fp64 across tiles
#include "tnk_common.cuh"
#include "tnk_tc.cuh"

namespace {

constexpr int kTcThreads = 192;
constexpr int kTile = TNK_TC_TILE;            // rows per CTA = columns per stage = 128
constexpr int kChunkBytes = kTile * 16;       // one chunk of one tile
constexpr int kTmemCols = 2 * kTile;          // two accumulator stages

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// Bounded wait: a protocol error traps (launch failure) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    const long long t0 = clock64();
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) return;
        if (clock64() - t0 > 8000000000LL) __trap();
    }
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] . B[smem]^T, one K-step of 8 tf32, issued by one thread for the CTA
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// shared-memory matrix descriptor, SWIZZLE_NONE, K-major: 8 rows x 16 B core matrices (128 B,
// contiguous), SBO = 128 B between 8-row groups, LBO = distance between the two 16-byte K chunks
__device__ __forceinline__ uint64_t smem_desc(uint32_t byte_addr, uint32_t lbo_bytes) {
    const uint32_t lo = ((byte_addr >> 4) & 0x3FFFu) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
    const uint32_t hi = (128u >> 4) | (1u << 14);   // SBO, descriptor version 1 (sm_100)
    return ((uint64_t)hi << 32) | lo;
}

// instruction descriptor: D f32, A/B tf32, both K-major, M = 128, N = 128
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kTile >> 3) << 17) | ((uint32_t)(kTile >> 4) << 24);

struct TcParams {
    int n;
    int64_t row0;        // first row of blockIdx.x == 0 (multiple of 128)
    int64_t n_tiles;     // column tiles in total (2^n / 128)
    int col_splits;      // gridDim.y
    int64_t part_stride; // rows covered by the launch
    float wl[32];        // w[h] = exp(-h/(n l))
};

__device__ __forceinline__ float tf_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }
__device__ __forceinline__ float tf_lo(float x) { return x - tf_hi(x); }

// ---- operand tables.  One thread = one float4 (4 consecutive k of one state).
template <bool IS_B>
__global__ void __launch_bounds__(256)
tc_pack_kernel(int n, int np, int64_t first_state, int64_t n_tiles, float gamma,
               const float* __restrict__ ST, const float* __restrict__ Cst, const float* __restrict__ qT,
               float4* __restrict__ out) {
    const int nc = np >> 2;
    const int K = IS_B ? 6 * nc + 2 : 5 * nc + 2;
    const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int row = (int)(idx & (kTile - 1));
    const int64_t t = idx >> 7;
    const int chunk = (int)(t % K);
    const int64_t tile = t / K;
    if (tile >= n_tiles) return;
    const int64_t z = first_state + tile * kTile + row;
    const int nfield = IS_B ? 6 : 5;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (chunk < nfield * nc) {
        const int f = chunk / nc, c = chunk - f * nc;
        const float4 s4 = *reinterpret_cast<const float4*>(ST + z * np + 4 * c);
        const float s[4] = {s4.x, s4.y, s4.z, s4.w};
        const float q = IS_B ? qT[z] : 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int p = 4 * c + e;
            const bool valid = p < n;
            const bool bit = valid && ((z >> p) & 1);
            const float rho = valid ? (bit ? -(1.f - s[e]) : (1.f - s[e])) : 0.f;
            float x;
            if (!IS_B) {
                const float g = gamma * rho;
                x = f == 0 ? tf_hi(s[e]) : f == 1 ? tf_lo(s[e]) : f == 2 ? (bit ? 1.f : 0.f) : f == 3 ? tf_hi(g) : tf_lo(g);
            } else {
                const float ts = q * s[e];
                const float r = (gamma * q) * rho;
                x = f == 0 ? tf_hi(ts) : f == 1 ? tf_lo(ts) : f == 2 ? tf_hi(r) : f == 3 ? tf_lo(r)
                  : f == 4 ? (bit ? tf_hi(q) : 0.f) : (bit ? tf_lo(q) : 0.f);
            }
            v[e] = x;
        }
    } else {
        const float C = Cst[z];
        const bool first = chunk == nfield * nc;
        if (!IS_B) {
            if (first) { v[0] = 1.f; v[1] = 1.f; v[2] = tf_hi(C); v[3] = tf_hi(C); }
            else { v[0] = tf_lo(C); }
        } else {
            const float q = qT[z];
            const float qc = q * C;
            if (first) { v[0] = tf_hi(qc); v[1] = tf_lo(qc); v[2] = tf_hi(q); v[3] = tf_lo(q); }
            else { v[0] = tf_hi(q); }
        }
    }
    out[idx] = make_float4(v[0], v[1], v[2], v[3]);
}

// ---- the tile kernel, NC = 4-float chunks per field (np / 4)
template <int NC>
__global__ void __launch_bounds__(kTcThreads, 1)
tc_rows_kernel(const __grid_constant__ TcParams prm, const float* __restrict__ A_tab, const float* __restrict__ B_tab,
               double* __restrict__ part) {
    extern __shared__ __align__(1024) unsigned char smem[];
    constexpr TnkTcSteps P = tnk_tc_steps(NC);
    static_assert(P.ok, "no K-step plan for this chunk count");
    constexpr uint32_t a_bytes = (uint32_t)(5 * NC + 2) * kChunkBytes;
    constexpr uint32_t b_bytes = (uint32_t)(6 * NC + 2) * kChunkBytes;
    unsigned char* sA = smem;
    unsigned char* sB = smem + a_bytes;
    float* sWl = reinterpret_cast<float*>(smem + a_bytes + 2 * b_bytes);
    uint64_t* bars = reinterpret_cast<uint64_t*>(sWl + 32);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);
    uint64_t* bar_afull = bars;
    uint64_t* bar_full = bars + 1;    // [2] B stage landed
    uint64_t* bar_empty = bars + 3;   // [2] B stage consumed by the MMAs
    uint64_t* bar_tfull = bars + 5;   // [2] accumulator stage complete
    uint64_t* bar_tempty = bars + 7;  // [2] accumulator stage drained by the 4 epilogue warps

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x < 32) sWl[threadIdx.x] = prm.wl[threadIdx.x];
    if (threadIdx.x == 0) {
        mbar_init(bar_afull, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(bar_full + s, 1);
            mbar_init(bar_empty + s, 1);
            mbar_init(bar_tfull + s, 1);
            mbar_init(bar_tempty + s, 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int64_t n_local = (prm.n_tiles - blockIdx.y + prm.col_splits - 1) / prm.col_splits;

    if (warp == 0) {
        if (lane == 0) {
            // ===== producer
            const unsigned char* srcA = reinterpret_cast<const unsigned char*>(A_tab) + (size_t)blockIdx.x * a_bytes;
            mbar_expect_tx(bar_afull, a_bytes);
            for (uint32_t off = 0; off < a_bytes; off += 8 * kChunkBytes) {
                const uint32_t len = min(a_bytes - off, (uint32_t)(8 * kChunkBytes));
                bulk_g2s(sA + off, srcA + off, len, bar_afull);
            }
            for (int64_t it = 0; it < n_local; ++it) {
                const int s = (int)(it & 1);
                const uint32_t ph = (uint32_t)((it >> 1) & 1);
                const int64_t J = (int64_t)blockIdx.y + it * prm.col_splits;
                mbar_wait(bar_empty + s, ph ^ 1u);
                const unsigned char* srcB = reinterpret_cast<const unsigned char*>(B_tab) + (size_t)J * b_bytes;
                unsigned char* dstB = sB + (size_t)s * b_bytes;
                mbar_expect_tx(bar_full + s, b_bytes);
                for (uint32_t off = 0; off < b_bytes; off += 8 * kChunkBytes) {
                    const uint32_t len = min(b_bytes - off, (uint32_t)(8 * kChunkBytes));
                    bulk_g2s(dstB + off, srcB + off, len, bar_full + s);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer.  Every descriptor is the tile's base descriptor plus a compile-time
            // offset (address field and LBO field in 16-byte units), so a K-step costs a few uniform
            // integer instructions around the UTCHMMA.
            const uint64_t a_desc = smem_desc(smem_u32(sA), 0);
            mbar_wait(bar_afull, 0);
            for (int64_t it = 0; it < n_local; ++it) {
                const int s = (int)(it & 1);
                const uint32_t ph = (uint32_t)((it >> 1) & 1);
                mbar_wait(bar_tempty + s, ph ^ 1u);
                mbar_wait(bar_full + s, ph);
                tc_fence_after();
                const uint64_t b_desc = smem_desc(smem_u32(sB) + (uint32_t)s * b_bytes, 0);
                const uint32_t d_addr = tmem_base + (uint32_t)(s * kTile);
#pragma unroll
                for (int ks = 0; ks < P.nsteps; ++ks) {
                    constexpr uint32_t cu = kChunkBytes >> 4;   // one chunk in descriptor address units
                    const uint64_t da = a_desc + (uint64_t)((uint32_t)P.a0[ks] * cu + (((uint32_t)P.alb[ks] * cu) << 16));
                    const uint64_t db = b_desc + (uint64_t)((uint32_t)P.b0[ks] * cu + (((uint32_t)P.blb[ks] * cu) << 16));
                    umma_tf32(d_addr, da, db, kIdesc, ks > 0 ? 1u : 0u);
                }
                umma_commit(bar_empty + s);   // B stage free once these MMAs have read it
                umma_commit(bar_tfull + s);   // accumulator stage ready
            }
        }
    } else {
        // ===== epilogue: thread <-> TMEM lane <-> row
        const int lq = (warp & 3) * 32;
        const int L = lq + lane;
        const int64_t i = prm.row0 + (int64_t)blockIdx.x * kTile + L;
        float W[16];
#pragma unroll
        for (int cc = 0; cc < 16; ++cc) W[cc] = sWl[__popc(((uint32_t)i ^ (uint32_t)cc) & 15u)];
        const uint32_t imid = ((uint32_t)i >> 4) & 7u;
        const uint32_t ihi = (uint32_t)(i >> 7);
        double udbl = 0.0;
        for (int64_t it = 0; it < n_local; ++it) {
            const int s = (int)(it & 1);
            const uint32_t ph = (uint32_t)((it >> 1) & 1);
            const int64_t J = (int64_t)blockIdx.y + it * prm.col_splits;
            const int xhi = __popc(ihi ^ (uint32_t)J);
            mbar_wait(bar_tfull + s, ph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)lq << 16) + (uint32_t)(s * kTile);
            float utile = 0.f;
#pragma unroll
            for (int g2 = 0; g2 < kTile / 32; ++g2) {
                uint32_t v[32];
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                      "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                      "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr + (uint32_t)(g2 * 32)));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
                    for (int cc = 0; cc < 16; cc += 2) {
                        acc0 = fmaf(W[cc], __uint_as_float(v[half * 16 + cc]), acc0);
                        acc1 = fmaf(W[cc + 1], __uint_as_float(v[half * 16 + cc + 1]), acc1);
                    }
                    const int g = g2 * 2 + half;
                    utile = fmaf(sWl[xhi + __popc(imid ^ (uint32_t)g)], acc0 + acc1, utile);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty + s);
            udbl += (double)utile;
        }
        part[(size_t)blockIdx.y * prm.part_stride + (size_t)(i - prm.row0)] = udbl;
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

}  // namespace

// Host planner: chunk products (A-chunk x B-chunk) of the bracket, two per MMA K-step.
int tnk_tc_make_plan(int n, TnkTcPlan* p) {
    if (n < TNK_TC_MIN_LATENT || n > TNK_TC_MAX_LATENT) return 1;
    const int np = ((n + 3) / 4) * 4;
    const int nc = np / 4;
    p->np = np; p->nc = nc; p->ka = 5 * nc + 2; p->kb = 6 * nc + 2;
    p->s = tnk_tc_steps(nc);
    return p->s.ok ? 0 : 1;
}

extern "C" int tnk_tc_plan(int n_latent, int* np, int* ka, int* kb, int* nsteps, int32_t* steps) {
    TnkTcPlan p;
    TNK_REQUIRE(tnk_tc_make_plan(n_latent, &p) == 0, "tnk_tc_plan: no tensor-core plan for n=%d (supported %d..%d)",
                n_latent, TNK_TC_MIN_LATENT, TNK_TC_MAX_LATENT);
    if (np) *np = p.np;
    if (ka) *ka = p.ka;
    if (kb) *kb = p.kb;
    if (nsteps) *nsteps = p.s.nsteps;
    if (steps)
        for (int ks = 0; ks < p.s.nsteps; ++ks) {
            steps[4 * ks + 0] = p.s.a0[ks]; steps[4 * ks + 1] = p.s.alb[ks];
            steps[4 * ks + 2] = p.s.b0[ks]; steps[4 * ks + 3] = p.s.blb[ks];
        }
    return 0;
}

size_t tnk_tc_table_bytes(const TnkTcPlan& p, int64_t states, bool is_b) {
    return (size_t)states * (size_t)(is_b ? p.kb : p.ka) * 16;
}

namespace {
template <int NC>
int launch_tiles(const TcParams& prm, dim3 grid, const float* A_tab, const float* B_tab, double* part, cudaStream_t st) {
    const size_t smem = (size_t)((5 * NC + 2) + 2 * (6 * NC + 2)) * kChunkBytes + 32 * sizeof(float) + 9 * sizeof(uint64_t) + 16;
    auto kern = tc_rows_kernel<NC>;
    TNK_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, kTcThreads, smem, st>>>(prm, A_tab, B_tab, part);
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}
}  // namespace

// Builds both operand tables and launches the tile kernel.  qT: packed fp32 q [>= 2^n].
int tnk_tc_launch(int n, double a, double gamma, const double* wl, const float* ST, const float* Cst, const float* qT,
                  int64_t row0, int64_t launch_rows, int splits, float* A_tab, float* B_tab, double* part,
                  cudaEvent_t ev_start, cudaEvent_t ev_stop, int* launches, cudaStream_t st) {
    (void)a;
    TnkTcPlan p;
    TNK_REQUIRE(tnk_tc_make_plan(n, &p) == 0, "tnk_ksd_rows: no tensor-core plan for n=%d", n);
    const int64_t N = (int64_t)1 << n;
    const int64_t row_tiles = launch_rows / kTile, col_tiles = N / kTile;
    {
        const int64_t items = row_tiles * p.ka * kTile;
        tc_pack_kernel<false><<<(unsigned)((items + 255) / 256), 256, 0, st>>>(n, p.np, row0, row_tiles, (float)gamma, ST, Cst, qT,
                                                                              reinterpret_cast<float4*>(A_tab));
        TNK_CUDA_OK(cudaGetLastError());
    }
    {
        const int64_t items = col_tiles * p.kb * kTile;
        tc_pack_kernel<true><<<(unsigned)((items + 255) / 256), 256, 0, st>>>(n, p.np, 0, col_tiles, (float)gamma, ST, Cst, qT,
                                                                             reinterpret_cast<float4*>(B_tab));
        TNK_CUDA_OK(cudaGetLastError());
    }
    TcParams prm;
    prm.n = n; prm.row0 = row0; prm.n_tiles = col_tiles; prm.col_splits = splits; prm.part_stride = launch_rows;
    for (int h = 0; h < 32; ++h) prm.wl[h] = (float)wl[h];
    if (ev_start) TNK_CUDA_OK(cudaEventRecord(ev_start, st));
    dim3 grid((unsigned)row_tiles, (unsigned)splits, 1);
    int rc = 2;
    switch (p.nc) {
        case 2: rc = launch_tiles<2>(prm, grid, A_tab, B_tab, part, st); break;
        case 3: rc = launch_tiles<3>(prm, grid, A_tab, B_tab, part, st); break;
        case 4: rc = launch_tiles<4>(prm, grid, A_tab, B_tab, part, st); break;
        case 5: rc = launch_tiles<5>(prm, grid, A_tab, B_tab, part, st); break;
        case 6: rc = launch_tiles<6>(prm, grid, A_tab, B_tab, part, st); break;
        default: tnk_set_error("tnk_ksd_rows: no tensor-core kernel instance for %d chunks", p.nc); break;
    }
    if (rc) return rc;
    if (ev_stop) TNK_CUDA_OK(cudaEventRecord(ev_stop, st));
    if (launches) *launches += 3;
    return 0;
}
