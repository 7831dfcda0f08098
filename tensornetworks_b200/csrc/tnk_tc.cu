// K3-TC: tensor-core tile variant of the Stein pair kernel.
//
// With j the column state, the bracket of k_p splits into
//   q_j * [ s(i).s(j) + g*T2(i,j) + C(j) ]           bilinear in per-state feature vectors -> tensor cores
//   q_j * [ C(i) + g*T1(i,j) ],  T1 = sum_p rho_p(i) b_p(j)   depends on j only through its BITS -> ALU
// A 128 x 128 tile of the first part is one GEMM tile: tcgen05.mma with the accumulator in TMEM.
// fp32-grade accuracy (the 1e-4 parity mode) comes from a split x = hi + lo (hi = top 10 mantissa
// bits, lo = x - hi exactly): a real product x*y is issued as hi*hi (kind::tf32) plus the
// second-order cross terms hi*lo + lo*hi, which tolerate 8-bit mantissas and are issued as bf16
// (kind::f16, K = 16 per step, half the tensor time per element) into the SAME accumulator;
// factors that are exact (bits, 1) need no split.  Measured error against the fp64 kernel <= 2e-6.
// The a^popc(i^j) weight does not factorise into a low-rank form, so it is applied per pair in
// the epilogue: a thread owns one row (TMEM lane) and 64 of the tile's columns, pulls 32
// accumulator columns at a time with tcgen05.ld and folds each group of 16 columns with a register
// table W[c] = a^popc((i^c)&15) (packed FFMA2) times one LUT value per group.  The second part
// needs per group only Q0 = sum_c W[c] q_c and Q1 = sum_{p<4} g_p(i) sum_c W[c] b_p(c) q_c, which
// depend on the row only through its low nibble: tc_qtab_kernel tabulates them once per column
// tile and the epilogue looks them up; C(i) + g*T1 is constant over a group up to the 4 low
// column bits: contribution = w_group * (main + E(i,J,group)*Q0 + Q1).
// Every ordered pair of the rows x columns handed to the kernel is evaluated (4^n in total).
//
// Column operand table B (built per call by tc_pack_kernel, K-major "interleaved" no-swizzle UMMA
// layout, one 128-state tile contiguous so that a whole tile is one run of cp.async.bulk):
//   tile t, chunk c, row r:  float4 at ((t*K + c)*128 + r)       (chunk = 4 consecutive k)
//   tf32 chunks (4 floats): t_hi | r_hi | [(qC)_hi,(qC)_lo,0,0] | 0
//   bf16 chunks (8 values): bf(t_lo) | bf(t) | bf(r_lo) | 0...
//   t = q_j s(j),  r = gamma*q_j*rho(j),  rho_p = (1-s_p)(1-2b_p)
// Row operand A: tf32 s_hi | bits | [1,1,0,0] | 0 and bf16 bf(s) | bf(s_lo) | bits, written by the
// epilogue threads straight into TMEM (tcgen05.st), 8 columns per K-step; the MMAs are the TS form
// (A from TMEM, B from shared memory).  A tf32 K-step (K = 8) consumes two (A-chunk, B-chunk)
// products of 4 floats, paired by the constexpr planner (tnk_tc.cuh) so that the B chunk offsets
// increase inside a step (they become the descriptor's leading-dimension byte offset); a bf16
// K-step (K = 16) consumes two consecutive 8-value chunks.  n = 20: 6 tf32 + 4 bf16 = 10 K-steps.
//
// Roles in a CTA of 320 threads (one CTA per SM):
//   warp 0    producer: B column tiles and their q tables through a ring of shared-memory stages
//             (bulk copy + mbarrier tx-count)
//   warp 1    allocates TMEM (2 x 128 accumulator columns + the A operand), one elect.sync lane
//             issues the MMAs and commits
//   warps 2-9 epilogue: A -> TMEM once, then per tile TMEM -> registers, Hamming weights, row sums
//             in fp32 per tile / compensated fp32 across tiles / one fp64 add at the end
#include <cuda_bf16.h>
#include "tnk_common.cuh"
#include <stdlib.h>
#include "tnk_tc.cuh"

namespace {

constexpr int kTcThreads = 320;
constexpr int kMaxStages = 4;                 // B ring depth (NC = 6: 4 x 48 KB + tables = 213 KB of shared memory)
constexpr int kQSlots = 8;                    // q-table ring depth >= kStages + 2 + 1 (see the producer comment)
constexpr int kQTab = 5 * 8 * 16;             // floats per column tile: [Q0,Qb0..Qb3][8 groups][16 row variants]
constexpr int kTile = TNK_TC_TILE;            // rows per CTA = columns per stage = 128
constexpr int kChunkBytes = kTile * 16;       // one chunk of one tile
constexpr int kTmemCols = 512;                // 2 accumulator stages (2 x 128 columns) + the A operand (8 columns per K-step)
constexpr int kTmemA = 2 * kTile;             // first TMEM column of the A operand

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
__device__ __forceinline__ uint32_t mbar_try(uint32_t addr, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    return done;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    if (mbar_try(addr, parity)) return;          // the common case costs no clock read
    const long long t0 = clock64();
    for (;;) {
        if (mbar_try(addr, parity)) return;
        if (clock64() - t0 > 8000000000LL) __trap();
    }
}
// one lane of a converged warp (elect.sync): lets ptxas issue UTCHMMA/UTCBAR without a per-lane loop
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[tmem] . B[smem]^T, one K-step of 8 tf32, issued by one thread for the CTA;
// A operand in TMEM: lane = row, one 32-bit column per tf32 element
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// same accumulator, kind::f16 with bf16 operands: one K-step of 16, A packed two per TMEM column
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
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
constexpr uint32_t kIdescBf16 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kTile >> 3) << 17) | ((uint32_t)(kTile >> 4) << 24);
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kTile >> 3) << 17) | ((uint32_t)(kTile >> 4) << 24);

struct TcParams {
    int n;
    int64_t row0;        // first row of blockIdx.x == 0 (multiple of 128)
    int64_t n_tiles;     // column tiles in total (2^n / 128)
    int col_splits;      // gridDim.y
    int64_t part_stride; // rows covered by the launch
    float gamma;         // a - 1/a
    float wl[32];        // w[h] = exp(-h/(n l))
};

__device__ __forceinline__ float tf_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }
__device__ __forceinline__ float tf_lo(float x) { return x - tf_hi(x); }

__device__ __forceinline__ uint32_t bf16_bits(float x) {   // round to nearest even
    return (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(x));
}
__device__ __forceinline__ uint32_t bf16_pack(float lo, float hi) { return bf16_bits(lo) | (bf16_bits(hi) << 16); }

// ---- column operand table.  One thread = one 16-byte chunk of one state: 4 floats (tf32 part) or
// 8 bf16 (cross-term part).  Layout: tile t, chunk c, row r at 16-byte slot ((t*K + c)*128 + r).
__global__ void __launch_bounds__(256)
tc_pack_kernel(int n, int np, int64_t n_tiles, float gamma,
               const float* __restrict__ ST, const float* __restrict__ Cst, const float* __restrict__ qT,
               uint4* __restrict__ out) {
    const int nc = np >> 2;
    const int K32 = tnk_tc_k32(nc), K = tnk_tc_kb(nc);
    const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int row = (int)(idx & (kTile - 1));
    const int64_t t = idx >> 7;
    const int chunk = (int)(t % K);
    const int64_t tile = t / K;
    if (tile >= n_tiles) return;
    const int64_t z = tile * kTile + row;
    const float q = qT[z];
    const float* srow = ST + z * np;
    auto col_terms = [&](int p, float& ts, float& r) {   // t_p = q s_p, r_p = gamma q rho_p
        const float sp = srow[p];
        const bool bit = (z >> p) & 1;
        ts = q * sp;
        r = (p < n) ? (gamma * q) * (bit ? -(1.f - sp) : (1.f - sp)) : 0.f;
    };
    uint32_t w[4] = {0u, 0u, 0u, 0u};
    if (chunk < 2 * nc) {                              // t_hi | r_hi
        const int f = chunk / nc, c = chunk - f * nc;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float ts, r;
            col_terms(4 * c + e, ts, r);
            w[e] = __float_as_uint(tf_hi(f == 0 ? ts : r));
        }
    } else if (chunk == 2 * nc) {                      // [(qC)_hi, (qC)_lo, 0, 0]
        const float qc = q * Cst[z];
        w[0] = __float_as_uint(tf_hi(qc));
        w[1] = __float_as_uint(tf_lo(qc));
    } else if (chunk >= K32) {                         // bf16 region: bf(t_lo) | bf(t) | bf(r_lo) | 0
        const int m = chunk - K32;
        float v[8];
#pragma unroll
        for (int x = 0; x < 8; ++x) {
            const int e = 8 * m + x;
            const int f = e / np, p = e - f * np;
            float val = 0.f;
            if (f < 3) {
                float ts, r;
                col_terms(p, ts, r);
                val = f == 0 ? tf_lo(ts) : f == 1 ? ts : tf_lo(r);
            }
            v[x] = val;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) w[e] = bf16_pack(v[2 * e], v[2 * e + 1]);
    }   // chunk == 2*nc + 1: the zero chunk that pads the odd tf32 product count
    out[idx] = make_uint4(w[0], w[1], w[2], w[3]);
}

// ---- per-column-tile q tables of the ALU part.  For group g (16 columns) and row variant v = i & 15:
//   Q0[g][v]   = sum_c a^popc(v^c) q_c          Qb_p[g][v] = sum_c a^popc(v^c) b_p(c) q_c   (p = 0..3)
// so that a row needs per group only Q0 and Q1 = sum_{p<4} g_p(i) Qb_p.  They depend on the column
// tile alone, hence are built once per evaluation: grid = column tiles, 128 threads = (g, v).
__global__ void __launch_bounds__(128)
tc_qtab_kernel(const float* __restrict__ qT, float w1, float* __restrict__ out) {
    const int g = threadIdx.x >> 4, v = threadIdx.x & 15;
    const float* q = qT + (size_t)blockIdx.x * kTile + g * 16;
    float wp[5];
    wp[0] = 1.f;
#pragma unroll
    for (int h = 1; h < 5; ++h) wp[h] = wp[h - 1] * w1;     // a^h, h <= 4
    float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 16; ++c) {
        const float wq = wp[__popc((unsigned)(v ^ c))] * q[c];
        acc[0] += wq;
#pragma unroll
        for (int p = 0; p < 4; ++p) acc[1 + p] += ((c >> p) & 1) ? wq : 0.f;
    }
    float* o = out + (size_t)blockIdx.x * kQTab;
#pragma unroll
    for (int k = 0; k < 5; ++k) o[(k * 8 + g) * 16 + v] = acc[k];
}

// ---- the tile kernel, NC = 4-float chunks per field (np / 4)
template <int NC>
struct TcSmem {
    static constexpr int stages = kMaxStages;
    static constexpr uint32_t b_bytes = (uint32_t)tnk_tc_kb(NC) * kChunkBytes;
    static constexpr uint32_t off_b = 0;
    static constexpr uint32_t off_q = off_b + stages * b_bytes;           // [kQSlots][kQTab] float
    static constexpr uint32_t off_wl = off_q + kQSlots * kQTab * 4;       // [32] float
    static constexpr uint32_t off_red = off_wl + 32 * 4;                  // [128] double
    static constexpr uint32_t off_bar = off_red + kTile * 8;              // mbarriers
    static constexpr int n_bars = 1 + 2 * kMaxStages + 4 + kQSlots;
    static constexpr uint32_t off_tmem = off_bar + n_bars * 8;
    static constexpr uint32_t total = off_tmem + 16;
};

template <int NC>
__global__ void __launch_bounds__(kTcThreads, 1)
tc_rows_kernel(const __grid_constant__ TcParams prm, const float* __restrict__ B_tab,
               const float* __restrict__ ST, const float* __restrict__ Cst, const float* __restrict__ Qtab,
               double* __restrict__ part) {
    extern __shared__ __align__(1024) unsigned char smem[];
    using SM = TcSmem<NC>;
    constexpr TnkTcSteps P = tnk_tc_steps(NC);
    static_assert(P.ok, "no K-step plan for this chunk count");
    constexpr int NP = 4 * NC;
    constexpr int K32 = tnk_tc_k32(NC), S16 = tnk_tc_steps16(NC);
    constexpr int kStages = SM::stages;
    constexpr uint32_t b_bytes = SM::b_bytes;
    unsigned char* sB = smem + SM::off_b;
    float* sQ = reinterpret_cast<float*>(smem + SM::off_q);
    float* sWl = reinterpret_cast<float*>(smem + SM::off_wl);
    double* sRed = reinterpret_cast<double*>(smem + SM::off_red);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SM::off_bar);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + SM::off_tmem);
    uint64_t* bar_aready = bars;                        // A operand stored to TMEM by the 8 epilogue warps
    uint64_t* bar_full = bars + 1;                      // [kStages] B stage landed
    uint64_t* bar_empty = bar_full + kStages;           // [kStages] B stage consumed by the MMAs
    uint64_t* bar_tfull = bar_empty + kStages;          // [2] accumulator stage complete
    uint64_t* bar_tempty = bar_tfull + 2;               // [2] accumulator stage drained by the 8 epilogue warps
    uint64_t* bar_qfull = bar_tempty + 2;               // [kQSlots] q tile landed

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x < 32) sWl[threadIdx.x] = prm.wl[threadIdx.x];
    if (threadIdx.x == 0) {
        mbar_init(bar_aready, 8);
        for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(bar_tfull + s, 1); mbar_init(bar_tempty + s, 8); }
        for (int s = 0; s < kQSlots; ++s) mbar_init(bar_qfull + s, 1);
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

    // Column split y owns a CONTIGUOUS range of column tiles and the grid is row-tile-major, so all
    // CTAs resident at one time stream the same slice of the B table (sized by the host to fit L2).
    const int64_t n_local = prm.n_tiles / prm.col_splits;
    const int64_t j_first = (int64_t)blockIdx.y * n_local;

    if (warp == 0) {
        if (lane == 0) {
            // ===== producer (B column tiles and their q tables).  It runs at most kStages tiles ahead of the MMAs, which run at most 2 tiles
            // (accumulator stages) ahead of the epilogue: while epilogue(it) still reads its q tile the
            // producer can be loading tile it + kStages + 1 at most, so kQSlots = 8 slots never alias.
            int s = 0;
            uint32_t ph = 0;
            for (int64_t it = 0; it < n_local; ++it) {
                const int64_t J = j_first + it;
                mbar_wait(bar_empty + s, ph ^ 1u);
                const int qs = (int)(it & (kQSlots - 1));
                mbar_expect_tx(bar_qfull + qs, kQTab * 4);
                bulk_g2s(sQ + qs * kQTab, Qtab + J * kQTab, kQTab * 4, bar_qfull + qs);
                const unsigned char* srcB = reinterpret_cast<const unsigned char*>(B_tab) + (size_t)J * b_bytes;
                unsigned char* dstB = sB + (size_t)s * b_bytes;
                mbar_expect_tx(bar_full + s, b_bytes);
                for (uint32_t off = 0; off < b_bytes; off += 8 * kChunkBytes) {
                    const uint32_t len = min(b_bytes - off, (uint32_t)(8 * kChunkBytes));
                    bulk_g2s(dstB + off, srcB + off, len, bar_full + s);
                }
                if (++s == kStages) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        {
            // ===== MMA issuer.  A comes from TMEM (8 columns per K-step, written once by the epilogue
            // warps), B from shared memory: its descriptor is the stage's base descriptor plus a
            // compile-time offset (address and LBO fields in 16-byte units), so a K-step costs a few
            // uniform integer instructions around the UTCHMMA and reads only 4 KB of shared memory.
            mbar_wait(bar_aready, 0);
            tc_fence_after();
            int s = 0;
            uint32_t ph = 0;
            for (int64_t it = 0; it < n_local; ++it) {
                const int ts = (int)(it & 1);
                const uint32_t tph = (uint32_t)((it >> 1) & 1);
                mbar_wait(bar_tempty + ts, tph ^ 1u);
                mbar_wait(bar_full + s, ph);
                tc_fence_after();
                const uint64_t b_desc = smem_desc(smem_u32(sB) + (uint32_t)s * b_bytes, 0);
                const uint32_t d_addr = tmem_base + (uint32_t)(ts * kTile);
                if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < P.nsteps; ++ks) {
                    constexpr uint32_t cu = kChunkBytes >> 4;   // one chunk in descriptor address units
                    const uint64_t db = b_desc + (uint64_t)((uint32_t)P.b0[ks] * cu + (((uint32_t)P.blb[ks] * cu) << 16));
                    umma_tf32_ts(d_addr, tmem_base + (uint32_t)(kTmemA + 8 * ks), db, kIdesc, ks > 0 ? 1u : 0u);
                }
#pragma unroll
                for (int k = 0; k < S16; ++k) {               // cross terms in bf16: chunks 2k, 2k+1 of the bf16 region
                    constexpr uint32_t cu = kChunkBytes >> 4;
                    const uint64_t db = b_desc + (uint64_t)((uint32_t)(K32 + 2 * k) * cu + (cu << 16));
                    umma_bf16_ts(d_addr, tmem_base + (uint32_t)(kTmemA + 8 * (P.nsteps + k)), db, kIdescBf16, 1u);
                }
                umma_commit(bar_empty + s);    // B stage free once these MMAs have read it
                umma_commit(bar_tfull + ts);   // accumulator stage ready
                }
                __syncwarp();
                if (++s == kStages) { s = 0; ph ^= 1u; }
            }
        }
    } else {
        // ===== epilogue: thread <-> TMEM lane <-> row; warps w and w+4 share a lane quarter and split the columns
        const int lq = (warp & 3) * 32;
        const int half = (warp - 2) >> 2;                  // columns [64*half, 64*half+64) of every tile
        const int L = lq + lane;
        const int64_t i = prm.row0 + (int64_t)blockIdx.x * kTile + L;
        const int n = prm.n;
        // row-side constants of the ALU part: g_p = gamma*rho_p(i)
        float gp[NP];
        {
            float sv[NP];
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const float4 s4 = *reinterpret_cast<const float4*>(ST + i * NP + 4 * c);
                sv[4 * c] = s4.x; sv[4 * c + 1] = s4.y; sv[4 * c + 2] = s4.z; sv[4 * c + 3] = s4.w;
            }
            // A operand of this row -> TMEM: every K-step occupies 8 columns from kTmemA (tf32: its two
            // 4-float chunks side by side; bf16: 16 values packed two per column, low half = even k);
            // the two warps of a lane quarter write alternate steps.
            // A32 chunks: s_hi | bits | [1,1,0,0] | 0.   A16 vector: bf(s) | bf(s_lo) | bits | 0.
            auto a32 = [&](int chunk, int e) -> float {
                if (chunk >= 2 * NC) return (chunk == 2 * NC && e < 2) ? 1.f : 0.f;
                const int f = chunk / NC, p = 4 * (chunk - f * NC) + e;
                return f == 0 ? tf_hi(sv[p]) : ((p < n && ((i >> p) & 1)) ? 1.f : 0.f);
            };
            auto a16 = [&](int e) -> float {
                const int f = e / NP, p = e - f * NP;
                if (f >= 3) return 0.f;
                const float x = sv[p];
                return f == 0 ? x : f == 1 ? tf_lo(x) : ((p < n && ((i >> p) & 1)) ? 1.f : 0.f);
            };
#pragma unroll
            for (int ks = 0; ks < P.nsteps + S16; ++ks) {
                if ((ks & 1) == half) {
                    uint32_t av[8];
                    if (ks < P.nsteps) {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            av[e] = __float_as_uint(a32(P.a0[ks < P.nsteps ? ks : 0], e));
                            av[4 + e] = __float_as_uint(a32(P.a0[ks < P.nsteps ? ks : 0] + P.alb[ks < P.nsteps ? ks : 0], e));
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < 8; ++c)
                            av[c] = bf16_pack(a16(16 * (ks - P.nsteps) + 2 * c), a16(16 * (ks - P.nsteps) + 2 * c + 1));
                    }
                    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                                 ::"r"(tmem_base + ((uint32_t)lq << 16) + (uint32_t)(kTmemA + 8 * ks)),
                                   "r"(av[0]), "r"(av[1]), "r"(av[2]), "r"(av[3]), "r"(av[4]), "r"(av[5]), "r"(av[6]), "r"(av[7])
                                 : "memory");
                }
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_aready);
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                const float v = prm.gamma * (1.f - sv[p]);
                gp[p] = (p < n) ? (((i >> p) & 1) ? -v : v) : 0.f;
            }
        }
        const float Ci = Cst[i];
        float2 W2[8];                                      // W[c] = a^popc((i^c)&15), packed for FFMA2
#pragma unroll
        for (int cc = 0; cc < 16; ++cc) {
            const float w = sWl[__popc(((uint32_t)i ^ (uint32_t)cc) & 15u)];
            if (cc & 1) W2[cc >> 1].y = w; else W2[cc >> 1].x = w;
        }
        float Gmid[4];                                     // column bits 4..6 of this warp's 4 groups
#pragma unroll
        for (int gl = 0; gl < 4; ++gl) {
            const int g = half * 4 + gl;
            float v = 0.f;
#pragma unroll
            for (int b = 0; b < 3; ++b) v += ((g >> b) & 1) ? gp[4 + b] : 0.f;
            Gmid[gl] = v;
        }
        int pmid[4];                                       // Hamming distance on column bits 4..6
#pragma unroll
        for (int gl = 0; gl < 4; ++gl) pmid[gl] = __popc((((uint32_t)i >> 4) & 7u) ^ (uint32_t)(half * 4 + gl));
        const uint32_t ihi = (uint32_t)(i >> 7);
        // Stepping to the next column tile flips tile bits 0..k-1 to 0 and sets bit k (k = ctz(J)):
        // dE[k] is the change of sum_{p>=7} g_p b_p(J) for k < 4; every 16th tile recomputes exactly.
        float dE[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float v = (7 + k < NP) ? gp[(7 + k < NP) ? 7 + k : 0] : 0.f;
#pragma unroll
            for (int b = 0; b < k; ++b) v -= (7 + b < NP) ? gp[(7 + b < NP) ? 7 + b : 0] : 0.f;
            dE[k] = v;
        }
        float Ehi = 0.f;                                   // C(i) + sum_{p>=7} g_p b_p(J)
        const int tab_off = (half * 4) * 16 + (int)(i & 15);   // [k][g][v] -> k*128 + g*16 + v
        // Row sum across tiles in compensated fp32 (Fast2Sum): the FP64 pipe stalls for hundreds of
        // cycles behind the tensor pipe (ncu: math-pipe throttle on a per-tile DADD), so fp64 is
        // touched only once, after the sweep.
        float uhi = 0.f, ulo = 0.f;
        const uint32_t n_loc = (uint32_t)n_local, j0 = (uint32_t)j_first;   // < 2^17 column tiles
        for (uint32_t it = 0; it < n_loc; ++it) {
            const int ts = (int)(it & 1);
            const uint32_t tph = (it >> 1) & 1;
            const int qs = (int)(it & (kQSlots - 1));
            const uint32_t qph = (it / kQSlots) & 1;
            const uint32_t J = j0 + it;
            const int xhi = __popc(ihi ^ J);
            if (it == 0 || (J & 15) == 0) {
                Ehi = Ci;
#pragma unroll
                for (int p = 7; p < NP; ++p) Ehi += ((J >> (p - 7)) & 1) ? gp[p] : 0.f;
            } else {
                const int k = __ffs((int)J) - 1;
                Ehi += (k == 0) ? dE[0] : (k == 1) ? dE[1] : (k == 2) ? dE[2] : dE[3];
            }
            mbar_wait(bar_qfull + qs, qph);
            mbar_wait(bar_tfull + ts, tph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)lq << 16) + (uint32_t)(ts * kTile + half * 64);
            uint32_t v[64];
#pragma unroll
            for (int g2 = 0; g2 < 2; ++g2) {
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[g2 * 32 + 0]), "=r"(v[g2 * 32 + 1]), "=r"(v[g2 * 32 + 2]), "=r"(v[g2 * 32 + 3]),
                      "=r"(v[g2 * 32 + 4]), "=r"(v[g2 * 32 + 5]), "=r"(v[g2 * 32 + 6]), "=r"(v[g2 * 32 + 7]),
                      "=r"(v[g2 * 32 + 8]), "=r"(v[g2 * 32 + 9]), "=r"(v[g2 * 32 + 10]), "=r"(v[g2 * 32 + 11]),
                      "=r"(v[g2 * 32 + 12]), "=r"(v[g2 * 32 + 13]), "=r"(v[g2 * 32 + 14]), "=r"(v[g2 * 32 + 15]),
                      "=r"(v[g2 * 32 + 16]), "=r"(v[g2 * 32 + 17]), "=r"(v[g2 * 32 + 18]), "=r"(v[g2 * 32 + 19]),
                      "=r"(v[g2 * 32 + 20]), "=r"(v[g2 * 32 + 21]), "=r"(v[g2 * 32 + 22]), "=r"(v[g2 * 32 + 23]),
                      "=r"(v[g2 * 32 + 24]), "=r"(v[g2 * 32 + 25]), "=r"(v[g2 * 32 + 26]), "=r"(v[g2 * 32 + 27]),
                      "=r"(v[g2 * 32 + 28]), "=r"(v[g2 * 32 + 29]), "=r"(v[g2 * 32 + 30]), "=r"(v[g2 * 32 + 31])
                    : "r"(taddr + (uint32_t)(g2 * 32)));
            }
            // ALU part of the four groups (table lookups) overlaps the TMEM load latency
            const float* tab = sQ + qs * kQTab + tab_off;
            float alu[4], wg[4];
#pragma unroll
            for (int gl = 0; gl < 4; ++gl) {
                const float Q0 = tab[gl * 16];
                float Q1 = gp[0] * tab[128 + gl * 16];
                Q1 = fmaf(gp[1], tab[256 + gl * 16], Q1);
                Q1 = fmaf(gp[2], tab[384 + gl * 16], Q1);
                Q1 = fmaf(gp[3], tab[512 + gl * 16], Q1);
                alu[gl] = fmaf(Ehi + Gmid[gl], Q0, Q1);
                wg[gl] = sWl[xhi + pmid[gl]];
            }
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty + ts);   // all 64 accumulator values are in registers: stage is free
            float utile = 0.f;
#pragma unroll
            for (int gl = 0; gl < 4; ++gl) {
                float2 m0 = make_float2(alu[gl], 0.f), m1 = make_float2(0.f, 0.f);
#pragma unroll
                for (int k = 0; k < 8; k += 2) {
                    m0 = __ffma2_rn(W2[k], make_float2(__uint_as_float(v[gl * 16 + 2 * k]), __uint_as_float(v[gl * 16 + 2 * k + 1])), m0);
                    m1 = __ffma2_rn(W2[k + 1], make_float2(__uint_as_float(v[gl * 16 + 2 * k + 2]), __uint_as_float(v[gl * 16 + 2 * k + 3])), m1);
                }
                utile = fmaf(wg[gl], (m0.x + m0.y) + (m1.x + m1.y), utile);
            }
            {
                const float sum = uhi + utile;              // TwoSum (no magnitude assumption)
                const float bb = sum - uhi;
                ulo += (uhi - (sum - bb)) + (utile - bb);
                uhi = sum;
            }
        }
        const double udbl = (double)uhi + (double)ulo;
        if (half == 1) sRed[L] = udbl;
        asm volatile("bar.sync 1, 256;" ::: "memory");     // the 8 epilogue warps only
        if (half == 0)
            part[(size_t)blockIdx.y * prm.part_stride + (size_t)(i - prm.row0)] = udbl + sRed[L];
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}


// ======================================================================================================
// K3-TCS: symmetric tile variant ("N(N+1)/2 if the symmetry k_p(z,z') = k_p(z',z)
// is exploited").  A tile (I, J) of 128 x 128 states with I < J is evaluated ONCE and used twice:
//   u_i += sum_j Y(i,j) q_j  (row part)       u_j += sum_i Y(i,j) q_i  (mirrored column part)
// with Y = a^popc(i^j) * bracket(i,j); diagonal tiles (I = J) contribute their row part only.  The whole
// bracket is bilinear in q-free feature vectors (App. A.4) and goes through the tensor cores:
//   A32 (rows):    s_hi | (g rho)_hi | bits | [C_hi, C_lo, 1, 1] | 0      g rho_p = gamma (1-s_p)(1-2 b_p)
//   B32 (columns): s_hi | bits | (g rho)_hi | [1, 1, C_hi, C_lo] | 0      -> chunk c of A meets chunk c of B
//   A16 (rows):    bf(s) | bf(s_lo) | bf((g rho)_lo) | bits               bf16 cross terms, K = 16 per step
//   B16 (columns): bf(s_lo) | bf(s) | bits | bf((g rho)_lo)
// n = 20: 8 tf32 + 5 bf16 = 13 K-steps per tile, i.e. 6.5 per ordered tile pair.  The column table is
// q-independent.  The epilogue reads the accumulator with tcgen05.ld.16x256b (the mma C-fragment
// distribution: thread t of a warp holds rows t/4 + 8k, k = 0..3, and columns 8m + 2(t%4) + {0,1}, m = 0..7
// of its warp's 32 x 64 block), so a thread owns a 4 x 16 patch: the mirrored column sums are reduced over
// its 4 rows in registers and over the 8 row groups of the warp with a transposing exchange (14 SHFL),
// then added to ucol[] with one fp64 RED per column, warp and tile (the 4 lane quarters meet in L2).  The Hamming weight of the 7 in-tile bits factorises over the patch: per-thread constant x
// (compile-time power of a from the unrolled row/column slots) -> 6 packed weights in registers.
// Row sums: fp32 per chunk of 8 tiles, compensated fp32 across chunks, fp64 at the end, added to the
// same accumulator u[] with one fp64 RED per row and CTA (no per-split partial buffer: at n = 24 it would be 68 GB).
// Both kinds of sums arrive by atomics, so the last bits of u depend on the arrival order (observed: <= 2e-16).
// ======================================================================================================
// Operand format 0 (above): tf32 high parts + bf16 cross terms, ~2^-19 per product.  Format 1: every real factor
// as TWO bf16 pieces x = x1 + x2 (x1 = bf(x), x2 = bf(x - x1), residual <= 2^-18 |x|), all K-steps kind::f16:
//   A16 (rows):    s1 | s1 | s2 | g1 | g2 | bits | bits | c1 c2 c3 1 1 1        (g = gamma rho, c = C in three pieces)
//   B16 (columns): s1 | s2 | s1 | bits | bits | g1 | g2 | 1 1 1 c1 c2 c3
// 7 np + 6 values: 10 K-steps at n = 20 instead of 13, 320 instead of 416 table bytes per state, ~2^-17 per product.
// Format 2 ("tile"): as format 0, but the T1/T2 terms of the bit positions p >= 7 - constant over a 128-state tile on
// one side - leave the pair kernel altogether.  Summed over ALL columns they are Kronecker-structured,
//     sum_j a^h(i,j) [ g_p(i) b_p(j) + b_p(i) g_p(j) ] q_j  =  g_p(i) K(b_p .* q)_i + b_p(i) K(g_p .* q)_i,     K = M^{(x)n},
// and K(b_p .* q) follows elementwise from w = K q, so one pass of n - 6 fp64 Kronecker transforms (tnk_hibits_pass in
// tnk_modeprod.cu, O(n^2 2^n): 0.1 ms at n = 20) seeds the accumulator u before the tile kernel adds the rest.
//   A32: s_hi (nc) | g_hi[0..7] | bits[0..7] | [C_hi, C_lo, 1, 1]      B32: s_hi | bits[0..7] | g_hi[0..7] | [1, 1, C_hi, C_lo]
//   A16: bf(s) | bf(s_lo) | bf(g_lo)[0..7] | bits[0..7]                B16: bf(s_lo) | bf(s) | bits[0..7] | bf(g_lo)[0..7]
// (positions 7 of the low fields are zero).  n = 20: 5 tf32 + 4 bf16 = 9 K-steps instead of 13; n = 24: 10 instead of 16.
// Format 3 ("dot"): only the score dot product stays: A32 = B32 = s_hi (nc), A16 = bf(s) | bf(s_lo), B16 = bf(s_lo) | bf(s);
// C(i) + C(j) and the T1/T2 terms of every bit position go through the Kronecker pass (p_lo = 0, with_c).
__host__ __device__ constexpr int tcs_k32(int nc, int fmt) {
    return fmt == 1 ? 0 : fmt == 2 ? (nc + 5 + 1) & ~1 : fmt == 3 ? (nc + 1) & ~1 : (3 * nc + 2) & ~1;   // chunks, padded to even
}
__host__ __device__ constexpr int tcs_steps32(int nc, int fmt) { return tcs_k32(nc, fmt) / 2; }
__host__ __device__ constexpr int tcs_steps16(int nc, int fmt) {
    return fmt == 1 ? (7 * 4 * nc + 6 + 15) / 16 : fmt == 2 ? (2 * 4 * nc + 16 + 15) / 16 : fmt == 3 ? (2 * 4 * nc + 15) / 16 : nc;
}
__host__ __device__ constexpr int tcs_kb(int nc, int fmt) { return tcs_k32(nc, fmt) + 2 * tcs_steps16(nc, fmt); }

__device__ __forceinline__ float bf_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

struct TcsParams {
    int n;
    int tile_offset, tile_stride;  // row tile of blockIdx.x = tile_offset + tile_stride * blockIdx.x
    int64_t n_tiles;               // 2^n / 128
    int col_splits;                // gridDim.y
    int row_tiles;                 // row tiles of this launch (tcs_tilesN_kernel: RT per CTA)
    int ablate;                    // diagnostics only (TNK_TCS_ABLATE, wrong results): 1 = no mirrored part, 2 = mirrored sums not stored;
                                   // + 4 / 12 / 16 select the DIAG instances of tcs_tilesN_kernel<5, 3, RT> (scripts/sym_limits.py)
    float gamma;
    float wl[32];
};

// column operand table of the symmetric variant: q-independent, same slot layout as tc_pack_kernel
__global__ void __launch_bounds__(256)
tcs_pack_kernel(int n, int np, int fmt, int64_t n_tiles, float gamma,
                const float* __restrict__ ST, const float* __restrict__ Cst, uint4* __restrict__ out) {
    const int nc = np >> 2;
    const int K32 = tcs_k32(nc, fmt), K = tcs_kb(nc, fmt);
    const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int row = (int)(idx & (kTile - 1));
    const int64_t t = idx >> 7;
    const int chunk = (int)(t % K);
    const int64_t tile = t / K;
    if (tile >= n_tiles) return;
    const int64_t z = tile * kTile + row;
    const float* srow = ST + z * np;
    auto grho = [&](int p) -> float {
        if (p >= n) return 0.f;
        const float v = gamma * (1.f - srow[p]);
        return ((z >> p) & 1) ? -v : v;
    };
    auto bit = [&](int p) -> float { return (p < n && ((z >> p) & 1)) ? 1.f : 0.f; };
    uint32_t w[4] = {0u, 0u, 0u, 0u};
    if (fmt == 3) {
        if (chunk < nc) {
#pragma unroll
            for (int e = 0; e < 4; ++e) w[e] = __float_as_uint(tf_hi(srow[4 * chunk + e]));
        } else if (chunk >= K32) {
            const int m = chunk - K32;
            float v[8];
#pragma unroll
            for (int x = 0; x < 8; ++x) {
                const int e = 8 * m + x;
                v[x] = e < np ? tf_lo(srow[e]) : e < 2 * np ? srow[e - np] : 0.f;
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) w[e] = bf16_pack(v[2 * e], v[2 * e + 1]);
        }
    } else if (fmt == 2) {
        auto low = [&](int p, bool want_bit) -> float {          // low fields: positions 0..6 only
            if (p >= 7) return 0.f;
            return want_bit ? bit(p) : grho(p);
        };
        if (chunk < nc) {
#pragma unroll
            for (int e = 0; e < 4; ++e) w[e] = __float_as_uint(tf_hi(srow[4 * chunk + e]));
        } else if (chunk < nc + 4) {
            const bool want_bit = chunk < nc + 2;
            const int c2 = (chunk - nc) & 1;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float val = low(4 * c2 + e, want_bit);
                w[e] = __float_as_uint(want_bit ? val : tf_hi(val));
            }
        } else if (chunk == nc + 4) {
            const float C = Cst[z];
            w[0] = w[1] = __float_as_uint(1.f);
            w[2] = __float_as_uint(tf_hi(C));
            w[3] = __float_as_uint(tf_lo(C));
        } else if (chunk >= K32) {
            const int m = chunk - K32;
            float v[8];
#pragma unroll
            for (int x = 0; x < 8; ++x) {
                const int e = 8 * m + x;
                float val = 0.f;
                if (e < np) val = tf_lo(srow[e]);
                else if (e < 2 * np) val = srow[e - np];
                else if (e < 2 * np + 8) val = low(e - 2 * np, true);
                else if (e < 2 * np + 16) val = tf_lo(low(e - 2 * np - 8, false));
                v[x] = val;
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) w[e] = bf16_pack(v[2 * e], v[2 * e + 1]);
        }   // nc + 4 < chunk < K32: zero padding chunk
    } else if (fmt) {
        float v[8];
#pragma unroll
        for (int x = 0; x < 8; ++x) {
            const int e = 8 * chunk + x;
            const int f = e / np, p = e - f * np;
            float val = 0.f;
            if (f < 7) {
                if (f == 3 || f == 4) val = bit(p);
                else {
                    const float full = f < 3 ? srow[p] : grho(p);
                    const float x1 = bf_round(full);
                    val = (f == 1 || f == 6) ? bf_round(full - x1) : x1;
                }
            } else {
                const int k = e - 7 * np;
                if (k < 3) val = 1.f;
                else if (k < 6) {
                    const float C = Cst[z];
                    const float c1 = bf_round(C), c2 = bf_round(C - c1);
                    val = k == 3 ? c1 : k == 4 ? c2 : bf_round(C - c1 - c2);
                }
            }
            v[x] = val;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) w[e] = bf16_pack(v[2 * e], v[2 * e + 1]);
    } else if (chunk < 3 * nc) {
        const int f = chunk / nc, c = chunk - f * nc;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int p = 4 * c + e;
            w[e] = __float_as_uint(f == 0 ? tf_hi(srow[p]) : f == 1 ? bit(p) : tf_hi(grho(p)));
        }
    } else if (chunk == 3 * nc) {
        const float C = Cst[z];
        w[0] = w[1] = __float_as_uint(1.f);
        w[2] = __float_as_uint(tf_hi(C));
        w[3] = __float_as_uint(tf_lo(C));
    } else if (chunk >= K32) {
        const int m = chunk - K32;
        float v[8];
#pragma unroll
        for (int x = 0; x < 8; ++x) {
            const int e = 8 * m + x;
            const int f = e / np, p = e - f * np;
            v[x] = f == 0 ? tf_lo(srow[p]) : f == 1 ? srow[p] : f == 2 ? bit(p) : tf_lo(grho(p));
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) w[e] = bf16_pack(v[2 * e], v[2 * e + 1]);
    }   // 3nc < chunk < K32: zero padding chunk
    out[idx] = make_uint4(w[0], w[1], w[2], w[3]);
}

template <int NC, int FMT>
struct TcsSmem {
    static constexpr uint32_t b_bytes = (uint32_t)tcs_kb(NC, FMT) * kChunkBytes;
    static constexpr int stages = (4u * b_bytes <= 216u * 1024u) ? 4 : 3;
    static constexpr uint32_t off_b = 0;
    static constexpr uint32_t off_q = off_b + stages * b_bytes;           // [kQSlots][128] float: q of the column tile
    static constexpr uint32_t off_wl = off_q + kQSlots * kTile * 4;       // [32] float
    static constexpr uint32_t off_red = off_wl + 32 * 4;                  // [128] double
    static constexpr uint32_t off_bar = off_red + 4 * kTile * 8;          // mbarriers (red: up to 3 column slices / 4 row tiles)
    static constexpr int n_bars = 1 + 2 * kMaxStages + 4 + kQSlots;
    static constexpr uint32_t off_tmem = off_bar + n_bars * 8;
    static constexpr uint32_t total = off_tmem + 16;
    static_assert(total <= 227u * 1024u, "shared memory of the symmetric tile kernel");
    static_assert(kQSlots >= stages + 3, "q ring depth");
};

#define TNK_LD_16x256b_x8(v, addr)                                                                              \
    asm volatile(                                                                                               \
        "tcgen05.ld.sync.aligned.16x256b.x8.b32 "                                                               \
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                               \
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"               \
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),       \
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), \
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), \
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]) \
        : "r"(addr))

#define TNK_LD_16x256b_x4(v, addr)                                                                              \
    asm volatile(                                                                                               \
        "tcgen05.ld.sync.aligned.16x256b.x4.b32 "                                                               \
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"                        \
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),       \
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])  \
        : "r"(addr))

template <int NC, int FMT, int CS>
__global__ void __launch_bounds__(64 + 128 * CS, 1)
tcs_tiles_kernel(const __grid_constant__ TcsParams prm, const float* __restrict__ B_tab,
                 const float* __restrict__ ST, const float* __restrict__ Cst, const float* __restrict__ qT,
                 double* __restrict__ ucol) {
    extern __shared__ __align__(1024) unsigned char smem[];
    using SM = TcsSmem<NC, FMT>;
    constexpr int NP = 4 * NC;
    constexpr int MC = 16 / CS;          // 8-column atoms per thread and tile: a 4 x 2MC patch
    static_assert(CS == 2 || CS == 4, "column slices per tile");
    constexpr int K32 = tcs_k32(NC, FMT), S32 = tcs_steps32(NC, FMT), S16 = tcs_steps16(NC, FMT);
    constexpr int kStages = SM::stages;
    constexpr uint32_t b_bytes = SM::b_bytes;
    static_assert(2 * kTile + 8 * (S32 + S16) <= kTmemCols, "TMEM columns");

    // Row tile I meets the column tiles J >= I of this column split; nothing to do when the split lies below I.
    const int64_t I = (int64_t)prm.tile_offset + (int64_t)prm.tile_stride * blockIdx.x;
    const int64_t n_per = prm.n_tiles / prm.col_splits;
    const int64_t j_split = (int64_t)blockIdx.y * n_per;
    const int64_t j_begin = j_split > I ? j_split : I;
    const int64_t n_local = j_split + n_per - j_begin;
    if (n_local <= 0) return;

    unsigned char* sB = smem + SM::off_b;
    float* sQ = reinterpret_cast<float*>(smem + SM::off_q);
    float* sWl = reinterpret_cast<float*>(smem + SM::off_wl);
    double* sRed = reinterpret_cast<double*>(smem + SM::off_red);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SM::off_bar);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + SM::off_tmem);
    uint64_t* bar_aready = bars;
    uint64_t* bar_full = bars + 1;
    uint64_t* bar_empty = bar_full + kMaxStages;
    uint64_t* bar_tfull = bar_empty + kMaxStages;
    uint64_t* bar_tempty = bar_tfull + 2;
    uint64_t* bar_qfull = bar_tempty + 2;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x < 32) sWl[threadIdx.x] = prm.wl[threadIdx.x];
    if (threadIdx.x == 0) {
        mbar_init(bar_aready, 4 * CS);
        for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(bar_tfull + s, 1); mbar_init(bar_tempty + s, 4 * CS); }
        for (int s = 0; s < kQSlots; ++s) mbar_init(bar_qfull + s, 1);
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

    if (warp == 0) {
        if (lane == 0) {
            // ===== producer: column tile of the operand table + its 128 q values (same ring protocol as tc_rows_kernel)
            int s = 0;
            uint32_t ph = 0;
            for (int64_t it = 0; it < n_local; ++it) {
                const int64_t J = j_begin + it;
                mbar_wait(bar_empty + s, ph ^ 1u);
                const int qs = (int)(it & (kQSlots - 1));
                mbar_expect_tx(bar_qfull + qs, kTile * 4);
                bulk_g2s(sQ + qs * kTile, qT + J * kTile, kTile * 4, bar_qfull + qs);
                const unsigned char* srcB = reinterpret_cast<const unsigned char*>(B_tab) + (size_t)J * b_bytes;
                unsigned char* dstB = sB + (size_t)s * b_bytes;
                mbar_expect_tx(bar_full + s, b_bytes);
                for (uint32_t off = 0; off < b_bytes; off += 8 * kChunkBytes) {
                    const uint32_t len = min(b_bytes - off, (uint32_t)(8 * kChunkBytes));
                    bulk_g2s(dstB + off, srcB + off, len, bar_full + s);
                }
                if (++s == kStages) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: K-step ks multiplies chunks (2ks, 2ks+1) of A (TMEM) and B (shared memory)
        mbar_wait(bar_aready, 0);
        tc_fence_after();
        int s = 0;
        uint32_t ph = 0;
        for (int64_t it = 0; it < n_local; ++it) {
            const int ts = (int)(it & 1);
            const uint32_t tph = (uint32_t)((it >> 1) & 1);
            mbar_wait(bar_tempty + ts, tph ^ 1u);
            mbar_wait(bar_full + s, ph);
            tc_fence_after();
            const uint64_t b_desc = smem_desc(smem_u32(sB) + (uint32_t)s * b_bytes, 0);
            const uint32_t d_addr = tmem_base + (uint32_t)(ts * kTile);
            if (elect_one()) {
                constexpr uint32_t cu = kChunkBytes >> 4;
#pragma unroll
                for (int ks = 0; ks < S32; ++ks) {
                    const uint64_t db = b_desc + (uint64_t)((uint32_t)(2 * ks) * cu + (cu << 16));
                    umma_tf32_ts(d_addr, tmem_base + (uint32_t)(kTmemA + 8 * ks), db, kIdesc, ks > 0 ? 1u : 0u);
                }
#pragma unroll
                for (int k = 0; k < S16; ++k) {
                    const uint64_t db = b_desc + (uint64_t)((uint32_t)(K32 + 2 * k) * cu + (cu << 16));
                    umma_bf16_ts(d_addr, tmem_base + (uint32_t)(kTmemA + 8 * (S32 + k)), db, kIdescBf16, (S32 + k) > 0 ? 1u : 0u);
                }
                umma_commit(bar_empty + s);
                umma_commit(bar_tfull + ts);
            }
            __syncwarp();
            if (++s == kStages) { s = 0; ph ^= 1u; }
        }
    } else {
        // ===== epilogue warps 2..: lane quarter lqi = warp & 3, column slice (128/CS columns) of every tile = (warp - 2) >> 2
        const int lqi = warp & 3;
        const int lq = lqi * 32;
        const int cs = (warp - 2) >> 2;
        const int colbase = cs * 8 * MC;
        const int n = prm.n;
        {
            // A operand: thread <-> TMEM lane <-> row L of the tile (32x32b stores), the two warps of a lane
            // quarter write alternate K-steps
            const int L = lq + lane;
            const int64_t i = I * kTile + L;
            float sv[NP], gr[NP];
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const float4 s4 = *reinterpret_cast<const float4*>(ST + i * NP + 4 * c);
                sv[4 * c] = s4.x; sv[4 * c + 1] = s4.y; sv[4 * c + 2] = s4.z; sv[4 * c + 3] = s4.w;
            }
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                const float v = prm.gamma * (1.f - sv[p]);
                gr[p] = (p < n) ? (((i >> p) & 1) ? -v : v) : 0.f;
            }
            const float Ci = Cst[i];
            auto bitf = [&](int p) -> float { return (p < n && ((i >> p) & 1)) ? 1.f : 0.f; };
            auto a32 = [&](int chunk, int e) -> float {
                if (FMT == 3) return chunk < NC ? tf_hi(sv[4 * chunk + e]) : 0.f;
                if (FMT == 2) {
                    if (chunk < NC) return tf_hi(sv[4 * chunk + e]);
                    if (chunk < NC + 4) {
                        const int p = 4 * ((chunk - NC) & 1) + e;
                        if (p >= 7) return 0.f;
                        return chunk < NC + 2 ? tf_hi(gr[p]) : bitf(p);
                    }
                    if (chunk == NC + 4) return e == 0 ? tf_hi(Ci) : e == 1 ? tf_lo(Ci) : 1.f;
                    return 0.f;
                }
                if (chunk > 3 * NC) return 0.f;
                if (chunk == 3 * NC) return e == 0 ? tf_hi(Ci) : e == 1 ? tf_lo(Ci) : 1.f;
                const int f = chunk / NC, p = 4 * (chunk - f * NC) + e;
                return f == 0 ? tf_hi(sv[p]) : f == 1 ? tf_hi(gr[p]) : bitf(p);
            };
            auto a16 = [&](int e) -> float {
                if (FMT == 3) return e < NP ? sv[e] : e < 2 * NP ? tf_lo(sv[e - NP]) : 0.f;
                if (FMT == 2) {
                    if (e < NP) return sv[e];
                    if (e < 2 * NP) return tf_lo(sv[e - NP]);
                    const int p = (e - 2 * NP) & 7;
                    if (e >= 2 * NP + 16 || p >= 7) return 0.f;
                    return e < 2 * NP + 8 ? tf_lo(gr[p]) : bitf(p);
                }
                const int f = e / NP, p = e - f * NP;
                if (FMT == 0) return f == 0 ? sv[p] : f == 1 ? tf_lo(sv[p]) : f == 2 ? tf_lo(gr[p]) : bitf(p);
                if (f < 7) {
                    if (f >= 5) return bitf(p);
                    const float full = f < 3 ? sv[p] : gr[p];
                    const float x1 = bf_round(full);
                    return (f == 2 || f == 4) ? bf_round(full - x1) : x1;
                }
                const int k = e - 7 * NP;
                if (k >= 6) return 0.f;
                if (k >= 3) return 1.f;
                const float c1 = bf_round(Ci), c2 = bf_round(Ci - c1);
                return k == 0 ? c1 : k == 1 ? c2 : bf_round(Ci - c1 - c2);
            };
#pragma unroll
            for (int ks = 0; ks < S32 + S16; ++ks) {
                if ((ks % CS) == cs) {
                    uint32_t av[8];
                    if (ks < S32) {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            av[e] = __float_as_uint(a32(2 * ks, e));
                            av[4 + e] = __float_as_uint(a32(2 * ks + 1, e));
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < 8; ++c)
                            av[c] = bf16_pack(a16(16 * (ks - S32) + 2 * c), a16(16 * (ks - S32) + 2 * c + 1));
                    }
                    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                                 ::"r"(tmem_base + ((uint32_t)lq << 16) + (uint32_t)(kTmemA + 8 * ks)),
                                   "r"(av[0]), "r"(av[1]), "r"(av[2]), "r"(av[3]), "r"(av[4]), "r"(av[5]), "r"(av[6]), "r"(av[7])
                                 : "memory");
                }
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_aready);
        }
        // Main loop: thread t owns rows rq + 8k (k = 0..3) of its lane quarter and columns 8m + 2(t&3) + {0,1}
        // (m = 0..7) of its column half.  In-tile state bits: row bits 0-2 = t>>2, 3-4 = k, 5-6 = lqi;
        // column bit 0 = e, bits 1-2 = t&3, bits 3-5 = m, bit 6 = half.
        const int t = lane;
        const int rq = lq + (t >> 2);
        float2 T2[3][2];                                   // a^(in-tile Hamming distance), packed over e
        {
            const int i0 = (t >> 2) & 1;
            const int c0 = __popc((unsigned)(((t >> 3) & 3) ^ (t & 3)));
#pragma unroll
            for (int h = 0; h < 3; ++h)
#pragma unroll
                for (int mh = 0; mh < 2; ++mh) {
                    const int e0 = c0 + h + __popc((unsigned)((((colbase >> 5) + mh) ^ lqi) & 3));   // column bits 5-6
                    T2[h][mh] = make_float2(sWl[e0 + i0], sWl[e0 + (i0 ^ 1)]);
                }
        }
        float qi[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) qi[k] = qT[I * kTile + rq + 8 * k];
        float uhi[4] = {0.f, 0.f, 0.f, 0.f}, ulo[4] = {0.f, 0.f, 0.f, 0.f}, uch[4] = {0.f, 0.f, 0.f, 0.f};
        const bool b4 = (t & 16) != 0, b3 = (t & 8) != 0, b2 = (t & 4) != 0;
        const uint32_t n_loc = (uint32_t)n_local, I32 = (uint32_t)I, j0 = (uint32_t)j_begin;   // < 2^17 tiles
        for (uint32_t it = 0; it < n_loc; ++it) {
            const int ts = (int)(it & 1);
            const uint32_t tph = (it >> 1) & 1;
            const int qs = (int)(it & (kQSlots - 1));
            const uint32_t qph = (it / kQSlots) & 1;
            const uint32_t J = j0 + it;
            const float wg = sWl[__popc(I32 ^ J)];
            mbar_wait(bar_qfull + qs, qph);
            mbar_wait(bar_tfull + ts, tph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)lq << 16) + (uint32_t)(ts * kTile + colbase);
            uint32_t v0[4 * MC], v1[4 * MC];
            if constexpr (MC == 8) {
                TNK_LD_16x256b_x8(v0, taddr);
                TNK_LD_16x256b_x8(v1, taddr + (16u << 16));
            } else {
                TNK_LD_16x256b_x4(v0, taddr);
                TNK_LD_16x256b_x4(v1, taddr + (16u << 16));
            }
            float2 qj[MC];
            {
                const float2* q2 = reinterpret_cast<const float2*>(sQ + qs * kTile + colbase + 2 * (t & 3));
#pragma unroll
                for (int m = 0; m < MC; ++m) qj[m] = q2[4 * m];
            }
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty + ts);
            float2 racc[4], cacc[MC];
#pragma unroll
            for (int k = 0; k < 4; ++k) racc[k] = make_float2(0.f, 0.f);
#pragma unroll
            for (int m = 0; m < MC; ++m) cacc[m] = make_float2(0.f, 0.f);
            auto fold = [&](const uint32_t (&v)[4 * MC], const int hr) {
#pragma unroll
                for (int m = 0; m < MC; ++m)
#pragma unroll
                    for (int r = 0; r < 2; ++r) {
                        const int k = 2 * hr + r;
                        const int d = (k ^ m) & 3;            // row bits 3-4 against column bits 3-4
                        const float2 acc = make_float2(__uint_as_float(v[4 * m + 2 * r]), __uint_as_float(v[4 * m + 2 * r + 1]));
                        const float2 p = __fmul2_rn(T2[(d & 1) + (d >> 1)][m >> 2], acc);
                        racc[k] = __ffma2_rn(p, qj[m], racc[k]);
                        cacc[m] = __ffma2_rn(p, make_float2(qi[k], qi[k]), cacc[m]);
                    }
            };
            fold(v0, 0);
            fold(v1, 1);
            // row sums: plain fp32 over a chunk of 8 tiles, TwoSum across chunks
#pragma unroll
            for (int k = 0; k < 4; ++k) uch[k] = fmaf(wg, racc[k].x + racc[k].y, uch[k]);
            if ((it & 7) == 7 || it + 1 == n_loc) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float sum = uhi[k] + uch[k];      // TwoSum
                    const float bb = sum - uhi[k];
                    ulo[k] += (uhi[k] - (sum - bb)) + (uch[k] - bb);
                    uhi[k] = sum;
                    uch[k] = 0.f;
                }
            }
            if (J != I32 && !(prm.ablate & 1)) {
                // mirrored part: sum over the warp's 32 rows (8 row groups = lane bits 2-4), exchanging halves of the
                // column set per level; with 4 atoms per thread the last level is a plain exchange-add
                auto xadd = [&](const float2 keep, const float2 send, const int mask) {
                    return __fadd2_rn(keep, make_float2(__shfl_xor_sync(0xffffffffu, send.x, mask),
                                                        __shfl_xor_sync(0xffffffffu, send.y, mask)));
                };
                float2 c3;
                int col;
                if constexpr (MC == 8) {
                    float2 c1[4], c2[2];
#pragma unroll
                    for (int k = 0; k < 4; ++k) c1[k] = xadd(b4 ? cacc[k + 4] : cacc[k], b4 ? cacc[k] : cacc[k + 4], 16);
#pragma unroll
                    for (int k = 0; k < 2; ++k) c2[k] = xadd(b3 ? c1[k + 2] : c1[k], b3 ? c1[k] : c1[k + 2], 8);
                    c3 = xadd(b2 ? c2[1] : c2[0], b2 ? c2[0] : c2[1], 4);
                    col = 2 * t;                                  // atom t>>2: columns 8(t>>2) + 2(t&3) + e
                } else {
                    float2 c1[2];
#pragma unroll
                    for (int k = 0; k < 2; ++k) c1[k] = xadd(b4 ? cacc[k + 2] : cacc[k], b4 ? cacc[k] : cacc[k + 2], 16);
                    const float2 c2 = xadd(b3 ? c1[1] : c1[0], b3 ? c1[0] : c1[1], 8);
                    c3 = xadd(c2, c2, 4);                         // both lanes of a pair hold the sum; b2 == 0 adds it
                    col = 8 * (t >> 3) + 2 * (t & 3);             // atom t>>3
                }
                // one fp64 RED per column, warp and tile (fire and forget; the four lane quarters meet in L2 - measured
                // faster than combining them in shared memory behind a named barrier)
                if ((MC == 8 || !b2) && !(prm.ablate & 2)) {
                    double* dst = ucol + (size_t)J * kTile + colbase + col;
                    atomicAdd(dst, (double)(wg * c3.x));
                    atomicAdd(dst + 1, (double)(wg * c3.y));
                }
            }
        }
        // row sums: over the 4 threads sharing a row, then over the CS column slices
        double ud[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            double x = (double)uhi[k] + (double)ulo[k];
            x += __shfl_xor_sync(0xffffffffu, x, 1);
            x += __shfl_xor_sync(0xffffffffu, x, 2);
            ud[k] = x;
        }
        if (cs > 0 && (t & 3) == 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) sRed[(cs - 1) * kTile + rq + 8 * k] = ud[k];
        }
        asm volatile("bar.sync 1, %0;" ::"n"(128 * CS) : "memory");   // the epilogue warps only
        if (cs == 0 && (t & 3) == 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                double x = ud[k];
#pragma unroll
                for (int c = 0; c < CS - 1; ++c) x += sRed[c * kTile + rq + 8 * k];
                atomicAdd(ucol + I * kTile + rq + 8 * k, x);    // row sums of this column split
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}


// ---- RT row tiles per CTA.  A column tile's operand stage (B, 16 * K bytes per state) and q values are fetched once
// and multiplied against the A operands of RT row tiles, all resident in TMEM (RT x 8 columns per K-step next to the
// two accumulator stages): 1/RT of the shared-memory fills per executed tile, and the mirrored column sums of the RT
// tiles are added in registers BEFORE the transposing exchange, so there is one exchange and one RED pair per column
// tile and warp instead of one per executed tile.  Row tile r of CTA x is tile_offset + tile_stride * (RT x + r); the
// tiles of a CTA are increasing, so the ones meeting column tile J (J >= I_r) are a prefix r < nact(J).
// DIAG (timing-only instances of scripts/sym_limits.py, results wrong): 4 = accumulators read from TMEM but not folded,
// 12 = not read either (MMAs alone), 16 = no MMAs (the epilogue alone), 20 = no MMAs and no fold (the read-out loop alone).
template <int NC, int FMT, int RT, int DIAG = 0>
__global__ void __launch_bounds__(64 + 128 * 2, 1)
tcs_tilesN_kernel(const __grid_constant__ TcsParams prm, const float* __restrict__ B_tab,
                  const float* __restrict__ ST, const float* __restrict__ Cst, const float* __restrict__ qT,
                  double* __restrict__ ucol) {
    extern __shared__ __align__(1024) unsigned char smem[];
    using SM = TcsSmem<NC, FMT>;
    constexpr int NP = 4 * NC;
    constexpr int CS = 2, MC = 8;        // 8 epilogue warps, 4 x 16 patches
    constexpr int K32 = tcs_k32(NC, FMT), S32 = tcs_steps32(NC, FMT), S16 = tcs_steps16(NC, FMT);
    constexpr int KS = S32 + S16;
    constexpr int kStages = SM::stages;
    constexpr uint32_t b_bytes = SM::b_bytes;
    static_assert(RT >= 2 && RT <= 4, "row tiles per CTA");
    static_assert(2 * kTile + RT * 8 * KS <= kTmemCols, "TMEM columns");

    const int r_first = RT * (int)blockIdx.x;
    const int nvalid = min(RT, prm.row_tiles - r_first);
    const int64_t I0 = (int64_t)prm.tile_offset + (int64_t)prm.tile_stride * r_first;
    const int64_t n_per = prm.n_tiles / prm.col_splits;
    const int64_t j_split = (int64_t)blockIdx.y * n_per;
    const int64_t j_begin = j_split > I0 ? j_split : I0;
    const int64_t n_local = j_split + n_per - j_begin;
    if (n_local <= 0) return;
    // row tiles meeting column tile J: r < nvalid with I0 + stride r <= J; J only grows, so the count is kept
    // incrementally (tile indices < 2^17: 32-bit arithmetic)
    const uint32_t I0u = (uint32_t)I0, strideu = (uint32_t)prm.tile_stride;
    auto grow_active = [&](int nact, uint32_t J) -> int {
        while (nact < nvalid && I0u + strideu * (uint32_t)nact <= J) ++nact;
        return nact;
    };

    unsigned char* sB = smem + SM::off_b;
    float* sQ = reinterpret_cast<float*>(smem + SM::off_q);
    float* sWl = reinterpret_cast<float*>(smem + SM::off_wl);
    double* sRed = reinterpret_cast<double*>(smem + SM::off_red);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SM::off_bar);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + SM::off_tmem);
    uint64_t* bar_aready = bars;
    uint64_t* bar_full = bars + 1;
    uint64_t* bar_empty = bar_full + kMaxStages;
    uint64_t* bar_tfull = bar_empty + kMaxStages;
    uint64_t* bar_tempty = bar_tfull + 2;
    uint64_t* bar_qfull = bar_tempty + 2;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x < 32) sWl[threadIdx.x] = prm.wl[threadIdx.x];
    if (threadIdx.x == 0) {
        mbar_init(bar_aready, 4 * CS);
        for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + s, 1); mbar_init(bar_empty + s, 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(bar_tfull + s, 1); mbar_init(bar_tempty + s, 4 * CS); }
        for (int s = 0; s < kQSlots; ++s) mbar_init(bar_qfull + s, 1);
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

    if (warp == 0) {
        if (lane == 0) {
            // ===== producer: one operand stage and one q tile per COLUMN tile
            int s = 0;
            uint32_t ph = 0;
            for (int64_t it = 0; it < n_local; ++it) {
                const int64_t J = j_begin + it;
                mbar_wait(bar_empty + s, ph ^ 1u);
                const int qs = (int)(it & (kQSlots - 1));
                mbar_expect_tx(bar_qfull + qs, kTile * 4);
                bulk_g2s(sQ + qs * kTile, qT + J * kTile, kTile * 4, bar_qfull + qs);
                const unsigned char* srcB = reinterpret_cast<const unsigned char*>(B_tab) + (size_t)J * b_bytes;
                unsigned char* dstB = sB + (size_t)s * b_bytes;
                mbar_expect_tx(bar_full + s, b_bytes);
                for (uint32_t off = 0; off < b_bytes; off += 8 * kChunkBytes) {
                    const uint32_t len = min(b_bytes - off, (uint32_t)(8 * kChunkBytes));
                    bulk_g2s(dstB + off, srcB + off, len, bar_full + s);
                }
                if (++s == kStages) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the accumulator stages alternate over EXECUTED tiles (vt), the B stage over column tiles
        mbar_wait(bar_aready, 0);
        tc_fence_after();
        int s = 0;
        uint32_t ph = 0, vt = 0;
        int nact = 0;
        for (uint32_t it = 0; it < (uint32_t)n_local; ++it) {
            nact = grow_active(nact, (uint32_t)j_begin + it);
            mbar_wait(bar_full + s, ph);
            const uint64_t b_desc = smem_desc(smem_u32(sB) + (uint32_t)s * b_bytes, 0);
            for (int r = 0; r < nact; ++r, ++vt) {
                const int ts = (int)(vt & 1);
                mbar_wait(bar_tempty + ts, ((vt >> 1) & 1) ^ 1u);
                tc_fence_after();
                const uint32_t d_addr = tmem_base + (uint32_t)(ts * kTile);
                const uint32_t a_addr = tmem_base + (uint32_t)(kTmemA + r * 8 * KS);
                if (elect_one()) {
                    constexpr uint32_t cu = kChunkBytes >> 4;
                    if constexpr (!(DIAG & 16)) {
#pragma unroll
                        for (int ks = 0; ks < S32; ++ks) {
                            const uint64_t db = b_desc + (uint64_t)((uint32_t)(2 * ks) * cu + (cu << 16));
                            umma_tf32_ts(d_addr, a_addr + (uint32_t)(8 * ks), db, kIdesc, ks > 0 ? 1u : 0u);
                        }
#pragma unroll
                        for (int k = 0; k < S16; ++k) {
                            const uint64_t db = b_desc + (uint64_t)((uint32_t)(K32 + 2 * k) * cu + (cu << 16));
                            umma_bf16_ts(d_addr, a_addr + (uint32_t)(8 * (S32 + k)), db, kIdescBf16, (S32 + k) > 0 ? 1u : 0u);
                        }
                    }
                    if (r + 1 == nact) umma_commit(bar_empty + s);      // the stage is free once the LAST row tile has read it
                    umma_commit(bar_tfull + ts);
                }
                __syncwarp();
            }
            if (++s == kStages) { s = 0; ph ^= 1u; }
        }
    } else {
        // ===== epilogue warps 2..9: lane quarter lqi = warp & 3, column half cs = (warp - 2) >> 2
        const int lqi = warp & 3;
        const int lq = lqi * 32;
        const int cs = (warp - 2) >> 2;
        const int colbase = cs * 8 * MC;
        const int n = prm.n;
        for (int r = 0; r < nvalid; ++r) {
            // A operand of row tile r (same construction as tcs_tiles_kernel), TMEM columns kTmemA + r 8 KS ...
            const int L = lq + lane;
            const int64_t i = (I0 + (int64_t)prm.tile_stride * r) * kTile + L;
            float sv[NP], gr[NP];
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const float4 s4 = *reinterpret_cast<const float4*>(ST + i * NP + 4 * c);
                sv[4 * c] = s4.x; sv[4 * c + 1] = s4.y; sv[4 * c + 2] = s4.z; sv[4 * c + 3] = s4.w;
            }
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                const float v = prm.gamma * (1.f - sv[p]);
                gr[p] = (p < n) ? (((i >> p) & 1) ? -v : v) : 0.f;
            }
            const float Ci = (FMT == 3) ? 0.f : Cst[i];
            auto bitf = [&](int p) -> float { return (p < n && ((i >> p) & 1)) ? 1.f : 0.f; };
            auto a32 = [&](int chunk, int e) -> float {
                if (FMT == 3) return chunk < NC ? tf_hi(sv[4 * chunk + e]) : 0.f;
                if (FMT == 2) {
                    if (chunk < NC) return tf_hi(sv[4 * chunk + e]);
                    if (chunk < NC + 4) {
                        const int p = 4 * ((chunk - NC) & 1) + e;
                        if (p >= 7) return 0.f;
                        return chunk < NC + 2 ? tf_hi(gr[p]) : bitf(p);
                    }
                    if (chunk == NC + 4) return e == 0 ? tf_hi(Ci) : e == 1 ? tf_lo(Ci) : 1.f;
                    return 0.f;
                }
                if (chunk > 3 * NC) return 0.f;
                if (chunk == 3 * NC) return e == 0 ? tf_hi(Ci) : e == 1 ? tf_lo(Ci) : 1.f;
                const int f = chunk / NC, p = 4 * (chunk - f * NC) + e;
                return f == 0 ? tf_hi(sv[p]) : f == 1 ? tf_hi(gr[p]) : bitf(p);
            };
            auto a16 = [&](int e) -> float {
                if (FMT == 3) return e < NP ? sv[e] : e < 2 * NP ? tf_lo(sv[e - NP]) : 0.f;
                if (FMT == 2) {
                    if (e < NP) return sv[e];
                    if (e < 2 * NP) return tf_lo(sv[e - NP]);
                    const int p = (e - 2 * NP) & 7;
                    if (e >= 2 * NP + 16 || p >= 7) return 0.f;
                    return e < 2 * NP + 8 ? tf_lo(gr[p]) : bitf(p);
                }
                const int f = e / NP, p = e - f * NP;
                if (FMT == 0) return f == 0 ? sv[p] : f == 1 ? tf_lo(sv[p]) : f == 2 ? tf_lo(gr[p]) : bitf(p);
                if (f < 7) {
                    if (f >= 5) return bitf(p);
                    const float full = f < 3 ? sv[p] : gr[p];
                    const float x1 = bf_round(full);
                    return (f == 2 || f == 4) ? bf_round(full - x1) : x1;
                }
                const int k = e - 7 * NP;
                if (k >= 6) return 0.f;
                if (k >= 3) return 1.f;
                const float c1 = bf_round(Ci), c2 = bf_round(Ci - c1);
                return k == 0 ? c1 : k == 1 ? c2 : bf_round(Ci - c1 - c2);
            };
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) {
                if ((ks % CS) == cs) {
                    uint32_t av[8];
                    if (ks < S32) {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            av[e] = __float_as_uint(a32(2 * ks, e));
                            av[4 + e] = __float_as_uint(a32(2 * ks + 1, e));
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < 8; ++c)
                            av[c] = bf16_pack(a16(16 * (ks - S32) + 2 * c), a16(16 * (ks - S32) + 2 * c + 1));
                    }
                    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                                 ::"r"(tmem_base + ((uint32_t)lq << 16) + (uint32_t)(kTmemA + r * 8 * KS + 8 * ks)),
                                   "r"(av[0]), "r"(av[1]), "r"(av[2]), "r"(av[3]), "r"(av[4]), "r"(av[5]), "r"(av[6]), "r"(av[7])
                                 : "memory");
                }
            }
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_aready);

        const int t = lane;
        const int rq = lq + (t >> 2);
        float2 T2[3][2];                                   // a^(in-tile Hamming distance), packed over e
        {
            const int i0 = (t >> 2) & 1;
            const int c0 = __popc((unsigned)(((t >> 3) & 3) ^ (t & 3)));
#pragma unroll
            for (int h = 0; h < 3; ++h)
#pragma unroll
                for (int mh = 0; mh < 2; ++mh) {
                    const int e0 = c0 + h + __popc((unsigned)((((colbase >> 5) + mh) ^ lqi) & 3));
                    T2[h][mh] = make_float2(sWl[e0 + i0], sWl[e0 + (i0 ^ 1)]);
                }
        }
        float qi[RT][4];
        float uhi[RT][4], ulo[RT][4], uch[RT][4];
#pragma unroll
        for (int r = 0; r < RT; ++r)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                qi[r][k] = r < nvalid ? qT[(I0 + (int64_t)prm.tile_stride * r) * kTile + rq + 8 * k] : 0.f;
                uhi[r][k] = ulo[r][k] = uch[r][k] = 0.f;
            }
        const bool b4 = (t & 16) != 0, b3 = (t & 8) != 0, b2 = (t & 4) != 0;
        const uint32_t n_loc = (uint32_t)n_local, I32 = (uint32_t)I0, j0 = (uint32_t)j_begin, stride = (uint32_t)prm.tile_stride;
        uint32_t vt = 0;
        int nact = 0;
        for (uint32_t it = 0; it < n_loc; ++it) {
            const int qs = (int)(it & (kQSlots - 1));
            const uint32_t qph = (it / kQSlots) & 1;
            const uint32_t J = j0 + it;
            nact = grow_active(nact, J);
            mbar_wait(bar_qfull + qs, qph);
            float2 qj[MC];
            {
                const float2* q2 = reinterpret_cast<const float2*>(sQ + qs * kTile + colbase + 2 * (t & 3));
#pragma unroll
                for (int m = 0; m < MC; ++m) qj[m] = q2[4 * m];
            }
            float2 cacc[MC];
#pragma unroll
            for (int m = 0; m < MC; ++m) cacc[m] = make_float2(0.f, 0.f);
#pragma unroll
            for (int r = 0; r < RT; ++r) {
                if (r < nact) {
                    const int ts = (int)(vt & 1);
                    const uint32_t tph = (vt >> 1) & 1;
                    ++vt;
                    const uint32_t Ir = I32 + stride * (uint32_t)r;
                    const float wg = sWl[__popc(Ir ^ J)];
                    const float wm = (J != Ir) ? wg : 0.f;       // a diagonal tile has no mirrored part
                    float qw[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) qw[k] = wm * qi[r][k];
                    mbar_wait(bar_tfull + ts, tph);
                    tc_fence_after();
                    const uint32_t taddr = tmem_base + ((uint32_t)lq << 16) + (uint32_t)(ts * kTile + colbase);
                    uint32_t v0[4 * MC], v1[4 * MC];
                    if constexpr (!(DIAG & 8)) {
                        TNK_LD_16x256b_x8(v0, taddr);
                        TNK_LD_16x256b_x8(v1, taddr + (16u << 16));
                    } else {
#pragma unroll
                        for (int x = 0; x < 4 * MC; ++x) v0[x] = v1[x] = 0u;
                    }
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_tempty + ts);
                    float2 racc[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) racc[k] = make_float2(0.f, 0.f);
                    auto fold = [&](const uint32_t (&v)[4 * MC], const int hr) {
#pragma unroll
                        for (int m = 0; m < MC; ++m)
#pragma unroll
                            for (int rr = 0; rr < 2; ++rr) {
                                const int k = 2 * hr + rr;
                                const int d = (k ^ m) & 3;
                                const float2 acc = make_float2(__uint_as_float(v[4 * m + 2 * rr]), __uint_as_float(v[4 * m + 2 * rr + 1]));
                                const float2 p = __fmul2_rn(T2[(d & 1) + (d >> 1)][m >> 2], acc);
                                racc[k] = __ffma2_rn(p, qj[m], racc[k]);
                                cacc[m] = __ffma2_rn(p, make_float2(qw[k], qw[k]), cacc[m]);
                            }
                    };
                    if constexpr (!(DIAG & 4)) {
                        fold(v0, 0);
                        fold(v1, 1);
                    } else {
                        racc[0].x = __uint_as_float(v0[0] ^ v1[5]);      // keep the loads alive
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k) uch[r][k] = fmaf(wg, racc[k].x + racc[k].y, uch[r][k]);
                }
            }
            if ((it & 7) == 7 || it + 1 == n_loc) {
#pragma unroll
                for (int r = 0; r < RT; ++r)
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float sum = uhi[r][k] + uch[r][k];      // TwoSum
                        const float bb = sum - uhi[r][k];
                        ulo[r][k] += (uhi[r][k] - (sum - bb)) + (uch[r][k] - bb);
                        uhi[r][k] = sum;
                        uch[r][k] = 0.f;
                    }
            }
            if (!(nact == 1 && J == I32) && !(prm.ablate & 1)) {
                // mirrored sums of all row tiles of this column tile: one transposing exchange, one RED pair
                auto xadd = [&](const float2 keep, const float2 send, const int mask) {
                    return __fadd2_rn(keep, make_float2(__shfl_xor_sync(0xffffffffu, send.x, mask),
                                                        __shfl_xor_sync(0xffffffffu, send.y, mask)));
                };
                float2 c1[4], c2[2];
#pragma unroll
                for (int k = 0; k < 4; ++k) c1[k] = xadd(b4 ? cacc[k + 4] : cacc[k], b4 ? cacc[k] : cacc[k + 4], 16);
#pragma unroll
                for (int k = 0; k < 2; ++k) c2[k] = xadd(b3 ? c1[k + 2] : c1[k], b3 ? c1[k] : c1[k + 2], 8);
                const float2 c3 = xadd(b2 ? c2[1] : c2[0], b2 ? c2[0] : c2[1], 4);
                if (!(prm.ablate & 2)) {
                    double* dst = ucol + (size_t)J * kTile + colbase + 2 * t;
                    atomicAdd(dst, (double)c3.x);
                    atomicAdd(dst + 1, (double)c3.y);
                }
            }
        }
        // row sums: over the 4 threads sharing a row, then over the two column halves
#pragma unroll
        for (int r = 0; r < RT; ++r) {
            double ud[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                double x = (double)uhi[r][k] + (double)ulo[r][k];
                x += __shfl_xor_sync(0xffffffffu, x, 1);
                x += __shfl_xor_sync(0xffffffffu, x, 2);
                ud[k] = x;
            }
            if (cs > 0 && (t & 3) == 0) {
#pragma unroll
                for (int k = 0; k < 4; ++k) sRed[r * kTile + rq + 8 * k] = ud[k];
            }
            asm volatile("bar.sync 1, %0;" ::"n"(128 * CS) : "memory");   // the epilogue warps only
            if (cs == 0 && (t & 3) == 0 && r < nvalid) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    atomicAdd(ucol + (I0 + (int64_t)prm.tile_stride * r) * kTile + rq + 8 * k, ud[k] + sRed[r * kTile + rq + 8 * k]);
            }
        }
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
    p->np = np; p->nc = nc; p->k32 = tnk_tc_k32(nc); p->kb = tnk_tc_kb(nc); p->steps16 = tnk_tc_steps16(nc);
    p->s = tnk_tc_steps(nc);
    return p->s.ok ? 0 : 1;
}

extern "C" int tnk_tc_plan(int n_latent, int* np, int* k32, int* kb, int* nsteps, int* nsteps16, int32_t* steps) {
    TnkTcPlan p;
    TNK_REQUIRE(tnk_tc_make_plan(n_latent, &p) == 0, "tnk_tc_plan: no tensor-core plan for n=%d (supported %d..%d)",
                n_latent, TNK_TC_MIN_LATENT, TNK_TC_MAX_LATENT);
    if (np) *np = p.np;
    if (k32) *k32 = p.k32;
    if (kb) *kb = p.kb;
    if (nsteps) *nsteps = p.s.nsteps;
    if (nsteps16) *nsteps16 = p.steps16;
    if (steps)
        for (int ks = 0; ks < p.s.nsteps; ++ks) {
            steps[4 * ks + 0] = p.s.a0[ks]; steps[4 * ks + 1] = p.s.alb[ks];
            steps[4 * ks + 2] = p.s.b0[ks]; steps[4 * ks + 3] = p.s.blb[ks];
        }
    return 0;
}

// bytes of the column operand table (the row operand is built in TMEM by the kernel)
size_t tnk_tc_table_bytes(const TnkTcPlan& p, int64_t states) { return (size_t)states * (size_t)p.kb * 16; }

size_t tnk_tc_qtab_bytes(int64_t states) { return (size_t)(states / kTile) * kQTab * sizeof(float); }

namespace {
template <int NC>
int launch_tiles(const TcParams& prm, dim3 grid, const float* B_tab, const float* ST, const float* Cst,
                 const float* Qtab, double* part, cudaStream_t st) {
    const size_t smem = TcSmem<NC>::total;
    auto kern = tc_rows_kernel<NC>;
    TNK_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, kTcThreads, smem, st>>>(prm, B_tab, ST, Cst, Qtab, part);
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}
}  // namespace

// Builds both operand tables and launches the tile kernel.  qT: packed fp32 q [>= 2^n].
int tnk_tc_launch(int n, double a, double gamma, const double* wl, const float* ST, const float* Cst, const float* qT,
                  int64_t row0, int64_t launch_rows, int splits, float* B_tab, float* Qtab, double* part,
                  cudaEvent_t ev_start, cudaEvent_t ev_stop, int* launches, cudaStream_t st) {
    (void)a;
    TnkTcPlan p;
    TNK_REQUIRE(tnk_tc_make_plan(n, &p) == 0, "tnk_ksd_rows: no tensor-core plan for n=%d", n);
    const int64_t N = (int64_t)1 << n;
    const int64_t row_tiles = launch_rows / kTile, col_tiles = N / kTile;
    {
        const int64_t items = col_tiles * p.kb * kTile;
        tc_pack_kernel<<<(unsigned)((items + 255) / 256), 256, 0, st>>>(n, p.np, col_tiles, (float)gamma, ST, Cst, qT,
                                                                       reinterpret_cast<uint4*>(B_tab));
        TNK_CUDA_OK(cudaGetLastError());
    }
    tc_qtab_kernel<<<(unsigned)col_tiles, 128, 0, st>>>(qT, (float)wl[1], Qtab);
    TNK_CUDA_OK(cudaGetLastError());
    TcParams prm;
    prm.n = n; prm.row0 = row0; prm.n_tiles = col_tiles; prm.col_splits = splits; prm.part_stride = launch_rows;
    prm.gamma = (float)gamma;
    for (int h = 0; h < 32; ++h) prm.wl[h] = (float)wl[h];
    if (ev_start) TNK_CUDA_OK(cudaEventRecord(ev_start, st));
    dim3 grid((unsigned)row_tiles, (unsigned)splits, 1);
    int rc = 2;
    switch (p.nc) {
        case 2: rc = launch_tiles<2>(prm, grid, B_tab, ST, Cst, Qtab, part, st); break;
        case 3: rc = launch_tiles<3>(prm, grid, B_tab, ST, Cst, Qtab, part, st); break;
        case 4: rc = launch_tiles<4>(prm, grid, B_tab, ST, Cst, Qtab, part, st); break;
        case 5: rc = launch_tiles<5>(prm, grid, B_tab, ST, Cst, Qtab, part, st); break;
        case 6: rc = launch_tiles<6>(prm, grid, B_tab, ST, Cst, Qtab, part, st); break;
        default: tnk_set_error("tnk_ksd_rows: no tensor-core kernel instance for %d chunks", p.nc); break;
    }
    if (rc) return rc;
    if (ev_stop) TNK_CUDA_OK(cudaEventRecord(ev_stop, st));
    if (launches) *launches += 3;
    return 0;
}

// ---- symmetric variant: host side
size_t tnk_tcs_table_bytes(int n, int fmt, int64_t states) {
    const int nc = ((n + 3) / 4);
    return (size_t)states * (size_t)tcs_kb(nc, fmt) * 16;
}

namespace {
template <int NC, int FMT, int CS>
int launch_sym_cs(const TcsParams& prm, dim3 grid, const float* B_tab, const float* ST, const float* Cst,
                  const float* qT, double* ucol, cudaStream_t st) {
    const size_t smem = TcsSmem<NC, FMT>::total;
    auto kern = tcs_tiles_kernel<NC, FMT, CS>;
    TNK_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, 64 + 128 * CS, smem, st>>>(prm, B_tab, ST, Cst, qT, ucol);
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}
template <int NC, int FMT, int RT>
int launch_symN(const TcsParams& prm, dim3 grid, const float* B_tab, const float* ST, const float* Cst,
                const float* qT, double* ucol, cudaStream_t st) {
    const size_t smem = TcsSmem<NC, FMT>::total;
    auto kern = tcs_tilesN_kernel<NC, FMT, RT>;
    if constexpr (NC == 5 && FMT == 3) {
        // timing-only instances of scripts/sym_limits.py (n = 17..20, 6-step format): parts of the pipeline compiled out
        switch (prm.ablate & 28) {
            case 4: kern = tcs_tilesN_kernel<NC, FMT, RT, 4>; break;     // accumulators read, not folded
            case 12: kern = tcs_tilesN_kernel<NC, FMT, RT, 12>; break;   // accumulators not read: the MMAs alone
            case 16: kern = tcs_tilesN_kernel<NC, FMT, RT, 16>; break;   // no MMAs: the epilogue alone
            case 20: kern = tcs_tilesN_kernel<NC, FMT, RT, 20>; break;   // no MMAs, accumulators read but not folded: the read-out loop alone
            default: break;
        }
    }
    TNK_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    grid.x = (grid.x + RT - 1) / RT;
    kern<<<grid, 64 + 128 * 2, smem, st>>>(prm, B_tab, ST, Cst, qT, ucol);
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}
template <int NC, int FMT>
int launch_sym(const TcsParams& prm, dim3 grid, const float* B_tab, const float* ST, const float* Cst,
               const float* qT, double* ucol, cudaStream_t st) {
    // Row tiles per CTA: as many as fit in TMEM next to the two accumulator stages (4 for the dot format, else 2);
    // n = 20, one B200: tf32_dot 133.2 (1) / 114.0 (2) / 111.0 ms (4), 13-step format 166.8 (1) / 146.5 ms (2).
    // TNK_TCS_ROWTILES=1|2|4 overrides (1 = tcs_tiles_kernel; measurements in profiles/r02_sym_ablations_rt*.log).
    // Epilogue warps: CS = 2 (8 warps, 4 x 16 patches); CS = 4 measured slower on the one-tile kernel (183.1 vs 178.6 ms).
    constexpr int KS = tcs_steps32(NC, FMT) + tcs_steps16(NC, FMT);
    constexpr bool fits4 = 2 * kTile + 4 * 8 * KS <= kTmemCols;
    int rt = fits4 ? 4 : 2;
    if (const char* e = getenv("TNK_TCS_ROWTILES")) {
        const int v = atoi(e);
        if (v == 1 || v == 2 || (v == 4 && fits4)) rt = v;
    }
    if (rt == 1) return launch_sym_cs<NC, FMT, 2>(prm, grid, B_tab, ST, Cst, qT, ucol, st);
    if constexpr (fits4) {
        if (rt == 4) return launch_symN<NC, FMT, 4>(prm, grid, B_tab, ST, Cst, qT, ucol, st);
    }
    return launch_symN<NC, FMT, 2>(prm, grid, B_tab, ST, Cst, qT, ucol, st);
}
}  // namespace

// Builds the (q-independent) column table unless the caller vouches for the one in B_tab, zeroes ucol and
// launches the symmetric tile kernel for row tiles tile_offset + k*tile_stride, k < row_tiles.
int tnk_tcs_launch(int n, int fmt, double gamma, const double* wl, const float* ST, const float* Cst, const float* qT,
                   int tile_offset, int tile_stride, int row_tiles, int splits, float* B_tab, int reuse_table,
                   double* ucol, double* hibits_ws, cudaEvent_t ev_start, cudaEvent_t ev_stop, int* launches,
                   cudaStream_t st) {
    TNK_REQUIRE(n >= TNK_TC_MIN_LATENT && n <= TNK_TC_MAX_LATENT, "tnk_ksd_sym: n=%d outside %d..%d", n,
                TNK_TC_MIN_LATENT, TNK_TC_MAX_LATENT);
    const int np = ((n + 3) / 4) * 4, nc = np / 4;
    const int64_t N = (int64_t)1 << n;
    const int64_t col_tiles = N / kTile;
    if (!reuse_table) {
        const int64_t items = col_tiles * tcs_kb(nc, fmt) * kTile;
        tcs_pack_kernel<<<(unsigned)((items + 255) / 256), 256, 0, st>>>(n, np, fmt, col_tiles, (float)gamma, ST, Cst,
                                                                        reinterpret_cast<uint4*>(B_tab));
        TNK_CUDA_OK(cudaGetLastError());
        if (launches) *launches += 1;
    }
    if (fmt == 3 || (fmt == 2 && n > 7)) {
        // the terms that left the pair kernel, for ALL rows: this call's share of the bit positions (the calls of one stride
        // split the pass between them; offset 0 also takes the C terms), written into ucol before the tile kernel adds its part
        const int rc_hi = tnk_hibits_pass(n, fmt == 3 ? 0 : 7, fmt == 3 && tile_offset == 0, tile_offset, tile_stride, wl[1], gamma,
                                          ST, Cst, qT, 4, ucol, hibits_ws, launches, st);
        if (rc_hi) return rc_hi;
    } else {
        TNK_CUDA_OK(cudaMemsetAsync(ucol, 0, (size_t)N * sizeof(double), st));
    }
    if (row_tiles <= 0) return 0;
    TcsParams prm;
    prm.n = n; prm.tile_offset = tile_offset; prm.tile_stride = tile_stride; prm.n_tiles = col_tiles;
    prm.col_splits = splits; prm.row_tiles = row_tiles; prm.gamma = (float)gamma;
    {
        const char* e = getenv("TNK_TCS_ABLATE");      // ablation runs of scripts/sym_ablate.py: results are WRONG when set
        prm.ablate = e ? atoi(e) : 0;
        static bool warned = false;
        if (prm.ablate && !warned) {
            warned = true;
            fprintf(stderr, "libtnk_b200: TNK_TCS_ABLATE=%d - timing-only mode of tnk_ksd_sym, its results are WRONG\n", prm.ablate);
        }
    }
    for (int h = 0; h < 32; ++h) prm.wl[h] = (float)wl[h];
    if (ev_start) TNK_CUDA_OK(cudaEventRecord(ev_start, st));
    dim3 grid((unsigned)row_tiles, (unsigned)splits, 1);
    int rc = 2;
    switch (nc * 4 + fmt) {
        case 8: rc = launch_sym<2, 0>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 9: rc = launch_sym<2, 1>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 10: rc = launch_sym<2, 2>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 11: rc = launch_sym<2, 3>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 12: rc = launch_sym<3, 0>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 13: rc = launch_sym<3, 1>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 14: rc = launch_sym<3, 2>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 15: rc = launch_sym<3, 3>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 16: rc = launch_sym<4, 0>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 17: rc = launch_sym<4, 1>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 18: rc = launch_sym<4, 2>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 19: rc = launch_sym<4, 3>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 20: rc = launch_sym<5, 0>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 21: rc = launch_sym<5, 1>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 22: rc = launch_sym<5, 2>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 23: rc = launch_sym<5, 3>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 24: rc = launch_sym<6, 0>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 25: rc = launch_sym<6, 1>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 26: rc = launch_sym<6, 2>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        case 27: rc = launch_sym<6, 3>(prm, grid, B_tab, ST, Cst, qT, ucol, st); break;
        default: tnk_set_error("tnk_ksd_sym: no kernel instance for %d chunks", nc); break;
    }
    if (rc) return rc;
    if (ev_stop) TNK_CUDA_OK(cudaEventRecord(ev_stop, st));
    if (launches) *launches += 1;
    return 0;
}
