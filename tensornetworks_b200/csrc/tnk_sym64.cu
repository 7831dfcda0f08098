// K3-S64: symmetric fp64 pair kernel (tnk_ksd_sym64) - the 1e-6 mode at half the pair evaluations.
//
// k_p(z,z') = k_p(z',z) (reference stein_utils.py:161-196 is symmetric in its arguments), so the pair space is cut
// into 128 x 128-state tiles and every unordered pair of tiles (I <= J) is evaluated once: its values feed the rows of
// tile I and, mirrored, the rows of tile J.  Per pair (closed form of tnk_ksd.cu / DESIGN.md section 2)
//     Y(i,j) = a^popc(i^j) * [ s(i).s(j) + C(i) + C(j) + T1(i,j) + T2(i,j) ],   T1 = sum_p g_p(i) b_p(j) = T2(j,i),
// with g_p(z) = gamma * (1 - s_p(z)) * (1 - 2 b_p(z)).  Only the rank-n term s(i).s(j) needs the pair space: summed over
// all j, the other terms are Kronecker-structured (C(i) K(q)_i + K(C.q)_i + sum_p [g_p(i) K(b_p.q)_i + b_p(i) K(g_p.q)_i],
// K = M^{(x)n}) and are added for all rows by n + 2 fp64 transforms before the pair kernel (tnk_hibits_pass,
// tnk_modeprod.cu: O(n^2 2^n)).
//
// Arithmetic of the pair kernel.  The score dot product runs as fp64 tensor-core MMAs (mma.sync.m8n8k4.f64 = DMMA:
// NP/4 K-steps per 8 x 8 pairs; the probe scripts/probes/dmma_rate.cu and ncu's pipe_shared counter show DMMA and DFMA
// share one pipe at about the same FMA rate, so the gain is issue slots, register and shared-memory traffic, not
// flops).  A warp owns 32 rows (A fragments resident in registers for a whole row tile) and walks its 64 columns in
// 8 batches of 8; the accumulator fragment gives thread t the rows t/4 + 8a (a = 0..3) and the columns 2(t%4) + e
// + 8b of a 32-column block.  The Hamming weight is one multiply by a^popc(a^b) (a compile-time constant after
// unrolling, 1 for a quarter of the pairs) - the rest of the weight is folded into q_j and into the column sums.
// Per executed pair: 4 (NP/4) DMMA lane-FMAs + 0.75 DMUL + 2 DFMA.
//
// Schedule and determinism.  The T = N/128 tiles are grouped into B <= 128 panels of s tiles; a work unit is a
// panel pair (pa <= pb) handled by one CTA: for every row tile I of pa it sweeps the column tiles J of pb (J >= I on
// the diagonal).  Row sums stay in registers over a sweep and are stored once into RowPart[pb][rows of I]; the
// mirrored column sums of a tile are added over the CTA's 32 row groups in shared memory in a fixed order and then
// accumulated into ColPart[pa][cols of J] by the same thread every time (first a plain store, afterwards
// red.global.add.f64 from that one thread: same-address program order).  The finish kernel adds the slots in a
// fixed order: the result is bit-reproducible run to run (the tcgen05 tile kernel's is not).
#include "tnk_common.cuh"
#include "tnk_tc.cuh"
#include <stdlib.h>

namespace {

constexpr int kTile = 128;
constexpr int kStageStride = 136;      // doubles per row group of the column-sum stage (128 + 8: conflict-free STS.128)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// D(8x8) += A(8x4) * B(4x8), fp64.  A: row t/4, k t%4;  B: k t%4, col t/4;  D: row t/4, cols 2(t%4)+{0,1}.
__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
        : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void red_add_f64(double* addr, double v) {
    asm volatile("red.global.add.f64 [%0], %1;" ::"l"(addr), "d"(v) : "memory");
}

struct Sym64Params {
    int n;
    int T;               // tiles of 128 states
    int B;               // panels
    int s;               // tiles per panel
    int unit_offset, unit_stride, n_units;   // this launch owns units unit_offset + k*unit_stride, k < n_units
    double a, gamma;
    double wl[32];       // exp(-h/(n l))
};

// unit k (row-major over the upper triangle incl. the diagonal) -> (pa, pb)
__host__ __device__ inline void unit_to_panels(int k, int B, int* pa, int* pb) {
    int a = 0;
    while (k >= B - a) { k -= B - a; ++a; }
    *pa = a;
    *pb = a + k;
}
__host__ __device__ inline int64_t panels_to_unit(int pa, int pb, int B) {
    return (int64_t)pa * B - (int64_t)pa * (pa - 1) / 2 + (pb - pa);
}

// Shared-memory layout.  RB = 8-row blocks per warp (4: 8 warps, 32 rows each; 2: 16 warps, 16 rows each).
template <int NP, int RB>
struct Smem64 {
    static constexpr int DOT = NP / 4;                  // K-steps of the score dot product
    static constexpr int STAGES = 3;                    // column-tile buffers: computing t, copies of t+1 and t+2 in flight
    static constexpr int NWR = 16 / RB;                 // row groups (warps along the rows) of the 128-row tile
    static constexpr int GROUPS = 8 * NWR;              // partial column sums per column: one per (row group, r)
    static constexpr size_t raw_bytes = (size_t)kTile * NP * sizeof(double);
    static constexpr size_t buf_bytes = raw_bytes + kTile * sizeof(double);          // raw | q
    static constexpr size_t stage_bytes = (size_t)GROUPS * kStageStride * sizeof(double);
    static constexpr size_t off_buf = 0;
    static constexpr size_t off_stage = off_buf + STAGES * buf_bytes;                 // one stage (arrive/wait protocol)
    static constexpr size_t off_rowred = off_stage + stage_bytes;                     // [2][128]
    static constexpr size_t off_wl = off_rowred + (size_t)2 * kTile * sizeof(double);      // [32]
    static constexpr size_t off_bar = off_wl + 32 * sizeof(double);
    static constexpr size_t total = off_bar + 8 * (STAGES + 1);
    static_assert(total <= 227 * 1024, "shared memory of the symmetric fp64 kernel");
};

template <int NP, int RB>
__global__ void __launch_bounds__(64 * (16 / RB), 1)
sym64_kernel(const Sym64Params prm, const double* __restrict__ ST, const double* __restrict__ qT,
             double* __restrict__ RowPart, double* __restrict__ ColPart) {
    using SM = Smem64<NP, RB>;
    constexpr int DOT = SM::DOT, STAGES = SM::STAGES, NWR = SM::NWR, GROUPS = SM::GROUPS;
    constexpr int THREADS = 64 * NWR;
    static_assert(RB == 2 || RB == 4, "row blocks per warp");
    extern __shared__ __align__(128) unsigned char smem[];
    double* stage = reinterpret_cast<double*>(smem + SM::off_stage);
    double* rowred = reinterpret_cast<double*>(smem + SM::off_rowred);
    double* sWl = reinterpret_cast<double*>(smem + SM::off_wl);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SM::off_bar);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wr = warp >> 1, wc = warp & 1;       // row group (8 RB rows) and column half (64 columns) of the warp
    const int r = lane >> 2, k4 = lane & 3;        // row inside an 8-row block; K index / column pair of the fragments
    const int rowbase = wr * 8 * RB + r;           // + 8a: this thread's rows inside the tile
    const int n = prm.n;
    const int64_t N = (int64_t)1 << n;
    const double av = prm.a;

    int pa, pb;
    unit_to_panels(prm.unit_offset + (int)blockIdx.x * prm.unit_stride, prm.B, &pa, &pb);
    const int s = prm.s;
    const int I_begin = pa * s, J_end = (pb + 1) * s;
    const bool diag = pa == pb;
    double* rowpart = RowPart + (size_t)pb * N;
    double* colpart = ColPart + (size_t)pa * N;
    const uint32_t n_tiles = diag ? (uint32_t)(s * (s + 1) / 2) : (uint32_t)(s * s);

    if (tid < 32) sWl[tid] = prm.wl[tid];
    if (tid == 0) {
#pragma unroll
        for (int b = 0; b < STAGES; ++b) mbar_init(&bars[b], 1);
        mbar_init(&bars[STAGES], THREADS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // the CTA's tile sequence: row tiles of panel pa in order, each sweeping the column tiles of panel pb (J >= I on the diagonal)
    auto advance = [&](int& I, int& J) {
        if (++J == J_end) { ++I; J = diag ? I : pb * s; }
    };
    auto issue_tile = [&](int J, uint32_t t) {      // one thread: two bulk copies on the buffer's barrier
        const int buf = (int)(t % STAGES);
        unsigned char* dst = smem + SM::off_buf + (size_t)buf * SM::buf_bytes;
        mbar_expect_tx(&bars[buf], (uint32_t)SM::buf_bytes);
        bulk_g2s(dst, ST + (size_t)J * kTile * NP, (uint32_t)SM::raw_bytes, &bars[buf]);
        bulk_g2s(dst + SM::raw_bytes, qT + (size_t)J * kTile, kTile * sizeof(double), &bars[buf]);
    };
    // Column sums of a finished tile: THREADS/128 threads per column, each adding 16 of the GROUPS row groups as four
    // independent chains; fixed order -> deterministic.
    auto flush_stage = [&](int J, bool first) {
        constexpr int TPC = THREADS / kTile;                          // 2 or 4 threads per column
        const int col = tid / TPC, part = tid % TPC;
        const double* src = stage + (16 * part) * kStageStride + col;
        double p4[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int g = 0; g < 16; ++g) p4[g & 3] += src[g * kStageStride];
        double acc = (p4[0] + p4[1]) + (p4[2] + p4[3]);
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        if (TPC == 4) acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        if (part == 0) {
            double* dst = colpart + (size_t)J * kTile + col;
            if (first) *dst = acc; else red_add_f64(dst, acc);
        }
    };
    static_assert(GROUPS == 16 * (THREADS / kTile), "flush covers every row group");

    // thread constants of the Hamming weight: bits 1,2 of i^j, and the e-dependent bit 0
    const int pt = __popc((uint32_t)(((r >> 1) & 3) ^ k4));
    const double wt = pt == 0 ? 1.0 : (pt == 1 ? av : av * av);
    const double we0 = (r & 1) ? av : 1.0, we1 = (r & 1) ? 1.0 : av;     // a^((r&1)^e)
    const double w1 = av, w2 = av * av;

    // ---- prologue: copies of tiles 0 and 1 in flight
    int I = I_begin, J = diag ? I_begin : pb * s;     // tile t
    int I1 = I, J1 = J;                               // tile t+1
    advance(I1, J1);
    if (tid == 0) {
        issue_tile(J, 0);
        if (n_tiles > 1) issue_tile(J1, 1);
    }

    double Af[DOT][RB];
    double qi[RB];
    double rsum[RB];
#pragma unroll
    for (int a = 0; a < RB; ++a) rsum[a] = 0.0;
    int I_loaded = -1;
    int pend_J = -1;                                 // column tile whose stage still has to be flushed
    bool pend_first = false;

    for (uint32_t t = 0; t < n_tiles; ++t) {
        __syncthreads();      // S_t: tile t-1 fully consumed; its stage complete; buffer (t+2)%3 free
        if (pend_J >= 0) { flush_stage(pend_J, pend_first); pend_J = -1; }
        mbar_arrive(&bars[STAGES]);                  // this thread no longer reads the stage (phase t of the stage barrier)
        int I2 = I1, J2 = J1;                         // tile t+2
        advance(I2, J2);
        if (tid == 0 && t + 2 < n_tiles) issue_tile(J2, t + 2);

        if (I != I_loaded) {
            // ---- row tile switch: store the finished row sums, then load the row operands of tile I
            if (I_loaded >= 0) {
#pragma unroll
                for (int a = 0; a < RB; ++a) {
                    double v = rsum[a];
                    v += __shfl_xor_sync(0xffffffffu, v, 1);
                    v += __shfl_xor_sync(0xffffffffu, v, 2);
                    if (k4 == 0) rowred[wc * kTile + rowbase + 8 * a] = v;
                    rsum[a] = 0.0;
                }
            }
#pragma unroll
            for (int a = 0; a < RB; ++a) {
                const int64_t i = (int64_t)I * kTile + rowbase + 8 * a;
                const double* si = ST + i * NP;
#pragma unroll
                for (int tt = 0; tt < DOT; ++tt) Af[tt][a] = si[4 * tt + k4];
                qi[a] = qT[i];
            }
            if (I_loaded >= 0) {
                __syncthreads();
                if (tid < kTile) rowpart[(size_t)I_loaded * kTile + tid] = rowred[tid] + rowred[kTile + tid];
            }
            I_loaded = I;
        }

        const int buf = (int)(t % STAGES);
        mbar_wait(&bars[buf], (t / STAGES) & 1);
        const unsigned char* bufp = smem + SM::off_buf + (size_t)buf * SM::buf_bytes;
        const double* raw = reinterpret_cast<const double*>(bufp);
        const double* cq = reinterpret_cast<const double*>(bufp + SM::raw_bytes);
        const bool mirrored = J != I;
        const int hh = __popc((uint32_t)(I ^ J));

        // ---- 8 batches of 8 columns, software-pipelined: the MMAs of batch b8+1 are issued before the epilogue of batch b8.
        //      batch b8: 32-column block cbi = b8 >> 2, 8-column group b = b8 & 3
        double acc[2][RB][2];
        double qe[2][2];
        double wst[2][2];                                                 // weights of the column sums of the batch in flight
        auto start_batch = [&](int b8, double (&ac)[RB][2], double (&qv2)[2], double (&ws2)[2]) {
            const int cb = 2 * wc + (b8 >> 2), b = b8 & 3;
            const int cfrag = cb * 32 + 8 * b + r;                     // column this lane supplies to the B fragment
            const double* sc = raw + cfrag * NP;
            double Bf[DOT];
#pragma unroll
            for (int tt = 0; tt < DOT; ++tt) Bf[tt] = sc[4 * tt + k4];
            const int cacc = cb * 32 + 8 * b + 2 * k4;                 // first of the two columns of the accumulator
            const double2 qv = *reinterpret_cast<const double2*>(cq + cacc);
            // weight of the warp-uniform bits of i^j: tile bits, bits 5,6 and (16-row warps) bit 4
            const int hb = hh + (RB == 4 ? __popc((uint32_t)(wr ^ cb)) : __popc((uint32_t)((wr >> 1) ^ cb)) + ((wr & 1) ^ (b >> 1)));
            const double wb = sWl[hb] * wt;
            ws2[0] = wb * we0;                                          // weight of everything but the bits held by a
            ws2[1] = wb * we1;
            qv2[0] = qv.x * ws2[0];                                     // ... folded into q_j for the row sums
            qv2[1] = qv.y * ws2[1];
#pragma unroll
            for (int a = 0; a < RB; ++a) { ac[a][0] = 0.0; ac[a][1] = 0.0; }
#pragma unroll
            for (int tt = 0; tt < DOT; ++tt)
#pragma unroll
                for (int a = 0; a < RB; ++a) dmma(ac[a][0], ac[a][1], Af[tt][a], Bf[tt]);
        };
        start_batch(0, acc[0], qe[0], wst[0]);
#pragma unroll
        for (int b8 = 0; b8 < 8; ++b8) {
            const int cur = b8 & 1, b = b8 & 3;
            if (b8 + 1 < 8) start_batch(b8 + 1, acc[cur ^ 1], qe[cur ^ 1], wst[cur ^ 1]);
            double cs0 = 0.0, cs1 = 0.0;
#pragma unroll
            for (int a = 0; a < RB; ++a) {
                const int x = RB == 4 ? (a ^ b) : (a ^ (b & 1));      // the row bits held by a against column bits 3(,4)
                const int pc = (x & 1) + (x >> 1);
                const double v0 = pc == 0 ? acc[cur][a][0] : acc[cur][a][0] * (pc == 1 ? w1 : w2);
                const double v1 = pc == 0 ? acc[cur][a][1] : acc[cur][a][1] * (pc == 1 ? w1 : w2);
                rsum[a] = fma(v0, qe[cur][0], rsum[a]);
                rsum[a] = fma(v1, qe[cur][1], rsum[a]);
                cs0 = fma(v0, qi[a], cs0);
                cs1 = fma(v1, qi[a], cs1);
            }
            if (mirrored) {
                if (b8 == 0) mbar_wait(&bars[STAGES], t & 1);     // every thread has flushed tile t-1 (long since: no stall)
                const int cacc = (2 * wc + (b8 >> 2)) * 32 + 8 * b + 2 * k4;
                *reinterpret_cast<double2*>(stage + (wr * 8 + r) * kStageStride + cacc) =
                    make_double2(cs0 * wst[cur][0], cs1 * wst[cur][1]);
            }
        }
        if (mirrored) { pend_J = J; pend_first = (I == I_begin); }   // a column tile is first met by the panel's first row tile
        I = I1; J = J1; I1 = I2; J1 = J2;
    }

    // ---- epilogue: the last row tile's sums and the last tile's column sums
#pragma unroll
    for (int a = 0; a < RB; ++a) {
        double v = rsum[a];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        if (k4 == 0) rowred[wc * kTile + rowbase + 8 * a] = v;
    }
    __syncthreads();
    if (tid < kTile) rowpart[(size_t)I_loaded * kTile + tid] = rowred[tid] + rowred[kTile + tid];
    if (pend_J >= 0) flush_stage(pend_J, pend_first);
}

// u[z] = sum of the row-part and column-part slots of the units this launch owned (fixed order); block partials of q.u
template <typename Q>
__global__ void __launch_bounds__(256)
sym64_finish_kernel(const Sym64Params prm, const double* __restrict__ RowPart, const double* __restrict__ ColPart,
                    const double* __restrict__ extra, const Q* __restrict__ q, double* __restrict__ u,
                    double* __restrict__ blockpart) {
    __shared__ double red[256];
    const int64_t N = (int64_t)1 << prm.n;
    const int64_t z = (int64_t)blockIdx.x * 256 + threadIdx.x;
    double contrib = 0.0;
    if (z < N) {
        const int Iz = (int)(z >> 7);
        const int pz = Iz / prm.s;
        auto owned = [&](int pa, int pb) -> bool {
            const int64_t k = panels_to_unit(pa, pb, prm.B) - prm.unit_offset;
            return k >= 0 && k % prm.unit_stride == 0 && k / prm.unit_stride < prm.n_units;
        };
        double acc = extra ? extra[z] : 0.0;     // C(z) + C(z') and the T1/T2 terms of all pairs (Kronecker pass), unit 0's owner
        for (int pb = pz; pb < prm.B; ++pb)
            if (owned(pz, pb)) acc += RowPart[(size_t)pb * N + z];
        for (int pa = 0; pa <= pz; ++pa)
            if (owned(pa, pz) && (pa < pz || Iz > pz * prm.s)) acc += ColPart[(size_t)pa * N + z];
        u[z] = acc;
        contrib = (double)q[z] * acc;
    }
    red[threadIdx.x] = contrib;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if (threadIdx.x < off) red[threadIdx.x] += red[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) blockpart[blockIdx.x] = red[0];
}

__global__ void __launch_bounds__(256)
sym64_scalar_kernel(const double* __restrict__ blockpart, int64_t nblocks, double* __restrict__ out) {
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

template <typename Q>
__global__ void sym64_qpack_kernel(const Q* __restrict__ q, int64_t N, double* __restrict__ qT) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) qT[i] = (double)q[i];
}

struct Sym64Plan {
    int64_t N, finish_blocks;
    int T, B, s, units_total, n_units, np;
    size_t off_q, off_row, off_col, off_block, off_extra, off_hib, total;
};

int make_plan64(int n, int unit_offset, int unit_stride, Sym64Plan* p) {
    if (n < TNK_SYM64_MIN_LATENT || n > TNK_SYM64_MAX_LATENT || unit_stride < 1 || unit_offset < 0) return 1;
    p->N = (int64_t)1 << n;
    p->T = (int)(p->N / kTile);
    int B = p->T < 128 ? p->T : 128;
    // both slot arrays are B x N doubles: keep them within 16 GB (n = 23: 64 panels, n = 24: 32)
    while (B > 1 && 2.0 * (double)B * (double)p->N * 8.0 > 16.0 * 1073741824.0) B /= 2;
    p->B = B;
    p->s = p->T / B;
    p->units_total = B * (B + 1) / 2;
    p->n_units = unit_offset < p->units_total ? (p->units_total - unit_offset + unit_stride - 1) / unit_stride : 0;
    p->np = ((n + 3) / 4) * 4;
    p->finish_blocks = (p->N + 255) / 256;
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t at = off; off += (bytes + 255) / 256 * 256; return at; };
    p->off_q = take((size_t)p->N * sizeof(double));
    p->off_row = take((size_t)B * p->N * sizeof(double));
    p->off_col = take((size_t)B * p->N * sizeof(double));
    p->off_block = take((size_t)(p->finish_blocks + 1) * sizeof(double));
    // every call adds its share of the terms that left the pair kernel (Kronecker pass: bit positions offset, offset + stride,
    // ...; the C terms with offset 0)
    p->off_extra = take((size_t)p->N * sizeof(double));
    p->off_hib = take(tnk_hibits_ws_bytes(n, 0, unit_offset == 0, unit_offset, unit_stride));
    p->total = off;
    return 0;
}

template <int NP, int RB>
int launch64(const Sym64Params& prm, int n_units, const double* ST, const double* qT,
             double* RowPart, double* ColPart, cudaStream_t st) {
    using SM = Smem64<NP, RB>;
    auto kern = sym64_kernel<NP, RB>;
    TNK_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM::total));
    kern<<<(unsigned)n_units, 64 * (16 / RB), SM::total, st>>>(prm, ST, qT, RowPart, ColPart);
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}

// warps per CTA: 8 (32 rows per warp, 255 registers) by default; TNK_SYM64_WARPS=16 selects the 16-rows-per-warp instance
// (128 registers, twice the warps: measured slower - every B fragment then feeds two MMAs instead of four and the
// shared-memory pipe saturates; kept for A/B runs)
int sym64_row_blocks() {
    static int rb = 0;
    if (rb == 0) {
        const char* e = getenv("TNK_SYM64_WARPS");
        rb = (e && atoi(e) == 16) ? 2 : 4;
    }
    return rb;
}

template <int NP>
int dispatch64(const Sym64Params& prm, int n_units, const double* ST, const double* qT,
               double* RowPart, double* ColPart, cudaStream_t st) {
    return sym64_row_blocks() == 4 ? launch64<NP, 4>(prm, n_units, ST, qT, RowPart, ColPart, st)
                                   : launch64<NP, 2>(prm, n_units, ST, qT, RowPart, ColPart, st);
}

}  // namespace

extern "C" size_t tnk_ksd_sym64_workspace_bytes(int n_latent, int unit_offset, int unit_stride) {
    Sym64Plan p;
    if (make_plan64(n_latent, unit_offset, unit_stride, &p)) return 0;
    return p.total;
}

extern "C" int tnk_ksd_sym64_units(int n_latent) {
    Sym64Plan p;
    if (make_plan64(n_latent, 0, 1, &p)) return 0;
    return p.units_total;
}

extern "C" int tnk_ksd_sym64(int n_latent, double length_scale, const void* q, int q_dtype,
                             const double* ST, const double* Cst, int unit_offset, int unit_stride,
                             double* u_partial, double* ksd_part,
                             void* workspace, size_t workspace_bytes, int* n_launch, void* stream) {
    Sym64Plan p;
    TNK_REQUIRE(make_plan64(n_latent, unit_offset, unit_stride, &p) == 0,
                "tnk_ksd_sym64: bad arguments n=%d (supported %d..%d) unit_offset=%d unit_stride=%d", n_latent,
                TNK_SYM64_MIN_LATENT, TNK_SYM64_MAX_LATENT, unit_offset, unit_stride);
    TNK_REQUIRE(q && ST && Cst && u_partial && ksd_part && workspace, "tnk_ksd_sym64: NULL pointer argument");
    TNK_REQUIRE(q_dtype == 0 || q_dtype == 1, "tnk_ksd_sym64: q_dtype must be 0 (double) or 1 (float)");
    TNK_REQUIRE(workspace_bytes >= p.total, "tnk_ksd_sym64: workspace %zu < required %zu", workspace_bytes, p.total);
    TNK_REQUIRE(((uintptr_t)workspace & 255) == 0, "tnk_ksd_sym64: workspace must be 256-byte aligned");
    TNK_REQUIRE((double)n_latent * length_scale != 0.0, "tnk_ksd_sym64: num_vars*length_scale must be non-zero");
    cudaStream_t st = (cudaStream_t)stream;
    const int n = n_latent;
    unsigned char* ws = (unsigned char*)workspace;
    double* qT = (double*)(ws + p.off_q);
    double* RowPart = (double*)(ws + p.off_row);
    double* ColPart = (double*)(ws + p.off_col);
    double* blockpart = (double*)(ws + p.off_block);
    int launches = 0;

    Sym64Params prm;
    prm.n = n; prm.T = p.T; prm.B = p.B; prm.s = p.s;
    prm.unit_offset = unit_offset; prm.unit_stride = unit_stride; prm.n_units = p.n_units;
    prm.a = exp(-1.0 / ((double)n * length_scale));
    prm.gamma = prm.a - 1.0 / prm.a;
    for (int h = 0; h < 32; ++h) prm.wl[h] = (h <= n) ? exp(-(double)h / ((double)n * length_scale)) : 0.0;

    const unsigned qblocks = (unsigned)((p.N + 255) / 256);
    if (q_dtype == 0) sym64_qpack_kernel<double><<<qblocks, 256, 0, st>>>((const double*)q, p.N, qT);
    else sym64_qpack_kernel<float><<<qblocks, 256, 0, st>>>((const float*)q, p.N, qT);
    TNK_CUDA_OK(cudaGetLastError());
    ++launches;

    double* extra = nullptr;
    if (unit_offset == 0 || tnk_hibits_count(n, 0, unit_offset, unit_stride) > 0) {
        extra = (double*)(ws + p.off_extra);
        const int rc_hi = tnk_hibits_pass(n, 0, unit_offset == 0, unit_offset, unit_stride, prm.a, prm.gamma, ST, Cst, qT, 8, extra,
                                          (double*)(ws + p.off_hib), &launches, st);
        if (rc_hi) return rc_hi;
    }
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    tnk_timing_events_get(&ev0, &ev1);
    if (ev0) TNK_CUDA_OK(cudaEventRecord(ev0, st));
    if (p.n_units > 0) {
        int rc = 0;
        switch (p.np) {
            case 8:  rc = dispatch64<8>(prm, p.n_units, ST, qT, RowPart, ColPart, st); break;
            case 12: rc = dispatch64<12>(prm, p.n_units, ST, qT, RowPart, ColPart, st); break;
            case 16: rc = dispatch64<16>(prm, p.n_units, ST, qT, RowPart, ColPart, st); break;
            case 20: rc = dispatch64<20>(prm, p.n_units, ST, qT, RowPart, ColPart, st); break;
            case 24: rc = dispatch64<24>(prm, p.n_units, ST, qT, RowPart, ColPart, st); break;
            default: tnk_set_error("tnk_ksd_sym64: no kernel instance for padded length %d", p.np); return 2;
        }
        if (rc) return rc;
        ++launches;
    }
    if (ev1) TNK_CUDA_OK(cudaEventRecord(ev1, st));

    if (q_dtype == 0)
        sym64_finish_kernel<double><<<(unsigned)p.finish_blocks, 256, 0, st>>>(prm, RowPart, ColPart, extra, (const double*)q, u_partial, blockpart);
    else
        sym64_finish_kernel<float><<<(unsigned)p.finish_blocks, 256, 0, st>>>(prm, RowPart, ColPart, extra, (const float*)q, u_partial, blockpart);
    TNK_CUDA_OK(cudaGetLastError());
    sym64_scalar_kernel<<<1, 256, 0, st>>>(blockpart, p.finish_blocks, ksd_part);
    TNK_CUDA_OK(cudaGetLastError());
    launches += 2;
    if (n_launch) *n_launch = launches;
    return 0;
}
