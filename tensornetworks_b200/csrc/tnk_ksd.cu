// K3: tiled on-the-fly Stein pair kernel  u[rows] = K_p[rows,:] q   (never materialises K_p)
// K4: deterministic finish  u = sum of column-split partials,  ksd_part = sum_i q_i u_i.
//
// Closed form evaluated per pair (derivation and reference mapping in DESIGN.md):
//   k_p(z,z') = a^h * [ s(z).s(z') + C(z) + C(z') + g*T1(z,z') + g*T2(z,z') ],  h = popc(z^z')
//   C(z)  = alpha*R(z) + g*rho0(z)            (per-state constant, packed by tnk_score)
//   T1    = sum_p rho_p(z)  b_p(z')           rho_p(z) = (1 - s_p(z)) * (1 - 2 b_p(z))
//   T2    = sum_p b_p(z)    rho_p(z')
// with a = exp(-1/(n l)), alpha = 1-a, g = a - 1/a.  This equals the four Eq.-13 terms of
// reference stein_utils.py:161-196 (tests compare against that restatement).
//
// Work decomposition.  A CTA (256 threads) owns RB = 256*R consecutive rows; thread t owns
// rows i0..i0+R-1 (i0 = block_row0 + t*R) and keeps their score vectors in registers for
// the whole sweep.  Columns are streamed in tiles of TC consecutive states: the raw score
// rows of the tile arrive in shared memory by one cp.async.bulk (TMA 1-D bulk copy) per
// tile, double buffered on mbarriers; every lane then reads the same column -> shared
// memory broadcasts.  Because row bits are fixed per thread/lane/warp/CTA and column bits
// are fixed per tile/group/unroll slot, T1 and T2 are O(1) per pair:
//   T2: per column a 32-entry lane table + 8-entry warp table (+ CTA-high bits folded in),
//       built once per tile, plus an R-entry table for the thread's own low bits;
//   T1: tile-high bits once per tile, group bits by a Gray-code walk over the groups of 8
//       columns, low 3 bits by a Gray-code walk inside the unrolled group.
// a^h: h = popc(upper bits) + popc(low bits); the low part is a compile-time power of a
// folded into q_j, the upper part is one LUT fetch per (thread, 8 columns).
#include "tnk_common.cuh"
#include "tnk_tc.cuh"
#include <stdlib.h>
#include <mutex>

namespace {

constexpr int kThreads = 256;
constexpr int kU = 8;  // columns per unrolled group

template <typename T> struct VecOf;
template <> struct VecOf<float> { using type = float4; static constexpr int n = 4; };
template <> struct VecOf<double> { using type = double2; static constexpr int n = 2; };

template <typename T, int NP>
__device__ __forceinline__ void load_vec(const T* __restrict__ src, T (&dst)[NP]) {
    using V = typename VecOf<T>::type;
    constexpr int VN = VecOf<T>::n;
    const V* s = reinterpret_cast<const V*>(src);
#pragma unroll
    for (int v = 0; v < NP / VN; ++v) {
        V x = s[v];
        const T* xs = reinterpret_cast<const T*>(&x);
#pragma unroll
        for (int e = 0; e < VN; ++e) dst[v * VN + e] = xs[e];
    }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
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

__host__ __device__ constexpr int ilog2(int v) { return v <= 1 ? 0 : 1 + ilog2(v / 2); }
__host__ __device__ constexpr int gray(int s) { return s ^ (s >> 1); }
// non-recursive so that the optimiser folds them for unrolled loop indices (s in 1..7, x in 0..7)
__host__ __device__ constexpr int ctz_c(int s) { return (s & 1) ? 0 : ((s & 2) ? 1 : 2); }
__host__ __device__ constexpr int popc_c(int x) { return (x & 1) + ((x >> 1) & 1) + ((x >> 2) & 1); }

template <typename T, int NP, int R, int TC>
struct Smem {
    static constexpr size_t raw_bytes = (size_t)TC * NP * sizeof(T);
    static constexpr size_t off_raw = 0;                                  // [2][TC][NP]
    static constexpr size_t off_tl = off_raw + 2 * raw_bytes;             // [TC][32]
    static constexpr size_t off_tw = off_tl + (size_t)TC * 32 * sizeof(T);// [TC][8]
    static constexpr size_t off_cm = off_tw + (size_t)TC * 8 * sizeof(T); // [TC][4]
    static constexpr size_t off_qa = off_cm + (size_t)TC * 4 * sizeof(T); // [TC][4]
    static constexpr size_t off_wl = off_qa + (size_t)TC * 4 * sizeof(T); // [40]
    static constexpr size_t off_bar = off_wl + 40 * sizeof(T);            // 2 mbarriers
    static constexpr size_t total = off_bar + 16;
};

struct KsdParams {
    int n;
    int64_t row0;        // first row of blockIdx.x == 0 (multiple of the row block)
    int64_t n_tiles;     // column tiles in total
    int col_splits;      // gridDim.y
    int64_t part_stride; // rows covered by the launch (stride between split partials)
    double a, gamma;
    double wl[32];       // w[h] = exp(-h/(n l)), h = 0..n
};

}  // namespace

namespace {

template <typename T, int NP, int R, int TC>
__global__ void __launch_bounds__(kThreads)
ksd_rows_kernel(const KsdParams prm, const T* __restrict__ ST, const T* __restrict__ Cst,
                const T* __restrict__ qT, double* __restrict__ part) {
    constexpr int RBITS = ilog2(R);
    constexpr int NV = kU / R;              // distinct upper-bit weights inside a group
    constexpr int GROUPS = TC / kU;
    constexpr int MID = ilog2(GROUPS);      // column bits 3 .. 3+MID-1 walk with the groups
    constexpr int TCB = ilog2(TC);          // column bits >= TCB are constant per tile
    constexpr int LANE0 = RBITS, WARP0 = RBITS + 5, CTA0 = RBITS + 8;
    using SM = Smem<T, NP, R, TC>;
    static_assert(R == 1 || R == 2 || R == 4, "rows per thread");
    static_assert(NP % 4 == 0 && NP >= 4, "padded score length");

    extern __shared__ __align__(128) unsigned char smem[];
    T* raw = reinterpret_cast<T*>(smem + SM::off_raw);
    T* sTl = reinterpret_cast<T*>(smem + SM::off_tl);
    T* sTw = reinterpret_cast<T*>(smem + SM::off_tw);
    T* sCm = reinterpret_cast<T*>(smem + SM::off_cm);
    T* sQa = reinterpret_cast<T*>(smem + SM::off_qa);
    T* sWl = reinterpret_cast<T*>(smem + SM::off_wl);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SM::off_bar);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = prm.n;
    const T gam = (T)prm.gamma;
    const int64_t blk_row0 = prm.row0 + (int64_t)blockIdx.x * (kThreads * R);
    const int64_t i0 = blk_row0 + (int64_t)tid * R;

    // ---- per-thread row operands, resident for the whole column sweep
    T srow[R][NP];
    T crow[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        load_vec<T, NP>(ST + (i0 + r) * NP, srow[r]);
        crow[r] = Cst[i0 + r];
    }
    auto row_grho = [&](int r, int p) -> T {  // gamma * rho_p(z_r); p must be a constant after unrolling
        const T v = gam * ((T)1 - srow[r][p < NP ? p : 0]);
        const bool neg = ((i0 + r) >> p) & 1;
        return (p < NP && p < n) ? (neg ? -v : v) : (T)0;
    };
    T glow[R][3], gmid[R][MID > 0 ? MID : 1];
#pragma unroll
    for (int r = 0; r < R; ++r) {
#pragma unroll
        for (int b = 0; b < 3; ++b) glow[r][b] = row_grho(r, b);
#pragma unroll
        for (int b = 0; b < MID; ++b) gmid[r][b] = row_grho(r, 3 + b);
    }

    if (tid < 32) sWl[tid] = (T)prm.wl[tid];
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int64_t n_local = (prm.n_tiles - blockIdx.y + prm.col_splits - 1) / prm.col_splits;
    auto tile_col0 = [&](int64_t it) -> int64_t { return ((int64_t)blockIdx.y + it * prm.col_splits) * TC; };
    if (tid == 0 && n_local > 0) {
        mbar_expect_tx(&bars[0], (uint32_t)SM::raw_bytes);
        bulk_g2s(raw, ST + tile_col0(0) * NP, (uint32_t)SM::raw_bytes, &bars[0]);
    }

    double udbl[R];
#pragma unroll
    for (int r = 0; r < R; ++r) udbl[r] = 0.0;

    for (int64_t it = 0; it < n_local; ++it) {
        const int buf = (int)(it & 1);
        const int64_t j0 = tile_col0(it);
        T* rawb = raw + (size_t)buf * TC * NP;
        __syncthreads();  // tile it-1 fully consumed: tables and raw[buf^1] are free
        if (tid == 0 && it + 1 < n_local) {
            mbar_expect_tx(&bars[buf ^ 1], (uint32_t)SM::raw_bytes);
            bulk_g2s(raw + (size_t)(buf ^ 1) * TC * NP, ST + tile_col0(it + 1) * NP, (uint32_t)SM::raw_bytes, &bars[buf ^ 1]);
        }
        mbar_wait(&bars[buf], (uint32_t)((it >> 1) & 1));

        // ---- per-tile column tables (T2 and the q-weights)
        for (int item = tid; item < TC * 2; item += kThreads) {
            const int c = item >> 1, half = item & 1;
            const int64_t j = j0 + c;
            const T* sc = rawb + c * NP;
            auto col_grho = [&](int p) -> T {  // gamma * rho_p(z'_j)
                if (p >= n) return (T)0;
                const T v = gam * ((T)1 - sc[p]);
                return ((j >> p) & 1) ? -v : v;
            };
            T gl[5];
#pragma unroll
            for (int b = 0; b < 5; ++b) gl[b] = col_grho(LANE0 + b);
#pragma unroll
            for (int l = 0; l < 16; ++l) {
                const int ll = half * 16 + l;
                T v = (T)0;
#pragma unroll
                for (int b = 0; b < 5; ++b) v += ((ll >> b) & 1) ? gl[b] : (T)0;
                sTl[c * 32 + ll] = v;
            }
            if (half == 0) {
                T e = Cst[j];
                for (int p = CTA0; p < n; ++p) e += ((blk_row0 >> p) & 1) ? col_grho(p) : (T)0;
                T gw[3];
#pragma unroll
                for (int b = 0; b < 3; ++b) gw[b] = col_grho(WARP0 + b);
#pragma unroll
                for (int w = 0; w < 8; ++w) {
                    T v = e;
#pragma unroll
                    for (int b = 0; b < 3; ++b) v += ((w >> b) & 1) ? gw[b] : (T)0;
                    sTw[c * 8 + w] = v;
                }
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    T v = (T)0;
#pragma unroll
                    for (int b = 0; b < RBITS; ++b) v += ((r >> b) & 1) ? col_grho(b) : (T)0;
                    sCm[c * 4 + r] = v;
                }
                const T a = (T)prm.a;
                T qv = qT[j];
#pragma unroll
                for (int e2 = 0; e2 < 4; ++e2) { sQa[c * 4 + e2] = qv; qv *= a; }
            }
        }
        __syncthreads();

        // ---- T1, tile-high part: column bits >= TCB are those of j0
        T G[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            T v = crow[r];
#pragma unroll
            for (int p = TCB; p < NP; ++p) v += ((j0 >> p) & 1) ? row_grho(r, p) : (T)0;
            G[r] = v;
        }
        T uacc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) uacc[r] = (T)0;

#pragma unroll 1
        for (int g = 0; g < GROUPS; ++g) {
            const int gg = g ^ (g >> 1);  // Gray order: one column bit changes per group
            if (g > 0) {
                const int kb = __ffs(g) - 1;
                const T sgn = ((gg >> kb) & 1) ? (T)1 : (T)-1;
#pragma unroll
                for (int b = 0; b < MID; ++b)
                    if (b == kb) {
#pragma unroll
                        for (int r = 0; r < R; ++r) G[r] = fma(sgn, gmid[r][b], G[r]);
                    }
            }
            const int cg = gg * kU;
            const uint32_t x = (uint32_t)((i0 ^ (j0 + cg)) >> RBITS);
            T wv[NV];
#pragma unroll
            for (int v = 0; v < NV; ++v) wv[v] = sWl[__popc(x ^ (uint32_t)v)];
            T tacc[NV][R];
#pragma unroll
            for (int v = 0; v < NV; ++v)
#pragma unroll
                for (int r = 0; r < R; ++r) tacc[v][r] = (T)0;
            T Y[R];
#pragma unroll
            for (int r = 0; r < R; ++r) Y[r] = G[r];

#pragma unroll
            for (int s = 0; s < kU; ++s) {
                const int jlo = gray(s);
                if (s > 0) {
                    const int kb = ctz_c(s);
                    const bool up = (jlo >> kb) & 1;
#pragma unroll
                    for (int r = 0; r < R; ++r) Y[r] = up ? Y[r] + glow[r][kb] : Y[r] - glow[r][kb];
                }
                const int c = cg + jlo;
                const T colbase = sTw[c * 8 + warp] + sTl[c * 32 + lane];
                T cm[4], qa[4], scol[NP];
                load_vec<T, 4>(sCm + c * 4, cm);
                load_vec<T, 4>(sQa + c * 4, qa);
                load_vec<T, NP>(rawb + c * NP, scol);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    T acc;
                    if constexpr (sizeof(T) == 4) {
                        // packed FFMA2 (fma.rn.f32x2): even/odd partial dot products, seeded with the
                        // two additive terms so that no separate add is needed for Y
                        float2 a2 = make_float2(r > 0 ? colbase + cm[r] : colbase, Y[r]);
#pragma unroll
                        for (int p = 0; p < NP; p += 2)
                            a2 = __ffma2_rn(make_float2(srow[r][p], srow[r][p + 1]), make_float2(scol[p], scol[p + 1]), a2);
                        acc = a2.x + a2.y;
                    } else {
                        acc = colbase + Y[r];
                        if (r > 0) acc += cm[r];
#pragma unroll
                        for (int p = 0; p < NP; ++p) acc = fma(srow[r][p], scol[p], acc);
                    }
                    const int cc = popc_c((jlo & (R - 1)) ^ r);
                    tacc[jlo >> RBITS][r] = fma(acc, qa[cc], tacc[jlo >> RBITS][r]);
                }
            }
#pragma unroll
            for (int v = 0; v < NV; ++v)
#pragma unroll
                for (int r = 0; r < R; ++r) uacc[r] = fma(wv[v], tacc[v][r], uacc[r]);
        }
#pragma unroll
        for (int r = 0; r < R; ++r) udbl[r] += (double)uacc[r];
    }

    double* dst = part + (size_t)blockIdx.y * prm.part_stride + (i0 - prm.row0);
#pragma unroll
    for (int r = 0; r < R; ++r) dst[r] = udbl[r];
}

}  // namespace

namespace {

// q (float or double, length N) -> T, zero padded to n_pad
template <typename T, typename Q>
__global__ void q_pack_kernel(const Q* __restrict__ q, int64_t N, int64_t n_pad, T* __restrict__ qT) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_pad) qT[i] = (i < N) ? (T)q[i] : (T)0;
}

// Peer sinks of the fused exchange (tnk_ksd_rows_peers): u buffers and partial-K slot arrays of
// every rank of the box, mapped into this process (NVLink peer / symmetric memory).
struct PeerSinks {
    int n, rank;
    double* u[TNK_MAX_PEERS];
    double* parts[TNK_MAX_PEERS];
};

// K4a: u[row] = sum over column splits (fixed order); per-block partial of sum q_i u_i.
// With peers.n > 0 the finished row is also stored straight into every peer's u buffer
// (st.global on peer-mapped addresses over NVLink) - the all-gather of SURVEY.md 8(e) fused
// into the reduction instead of a separate collective.
template <typename Q>
__global__ void __launch_bounds__(256)
finish_rows_kernel(const double* __restrict__ part, int splits, int64_t part_stride, int64_t row0,
                   int64_t row_begin, int64_t row_end, const Q* __restrict__ q,
                   double* __restrict__ u, double* __restrict__ blockpart, const PeerSinks peers) {
    __shared__ double red[256];
    const int64_t row = row_begin + (int64_t)blockIdx.x * 256 + threadIdx.x;
    double contrib = 0.0;
    if (row < row_end) {
        double acc = 0.0;
        for (int s = 0; s < splits; ++s) acc += part[(size_t)s * part_stride + (row - row0)];
        u[row] = acc;
        for (int p = 0; p < peers.n; ++p)
            if (p != peers.rank) peers.u[p][row] = acc;
        contrib = (double)q[row] * acc;
    }
    red[threadIdx.x] = contrib;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if (threadIdx.x < off) red[threadIdx.x] += red[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) blockpart[blockIdx.x] = red[0];
}

// K4b: one CTA adds the block partials in a fixed order -> deterministic scalar
__global__ void __launch_bounds__(256)
finish_scalar_kernel(const double* __restrict__ blockpart, int64_t nblocks, double* __restrict__ out,
                     const PeerSinks peers) {
    __shared__ double red[256];
    double acc = 0.0;
    for (int64_t b = threadIdx.x; b < nblocks; b += 256) acc += blockpart[b];
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if (threadIdx.x < off) red[threadIdx.x] += red[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        out[0] = red[0];
        for (int p = 0; p < peers.n; ++p) peers.parts[p][peers.rank] = red[0];  // slot `rank` of every rank
        __threadfence_system();
    }
}

// Cross-check kernel (TNK_PREC_FP64_SIMPLE): one thread per row, plain closed form, no tiling.
__global__ void __launch_bounds__(128)
ksd_rows_simple_kernel(KsdParams prm, int np, int64_t N, int64_t row_begin, int64_t row_end,
                       const double* __restrict__ ST, const double* __restrict__ Cst,
                       const double* __restrict__ qT, double* __restrict__ part) {
    const int64_t i = row_begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row_end) return;
    const int n = prm.n;
    const double* si = ST + i * np;
    double acc = 0.0;
    for (int64_t j = 0; j < N; ++j) {
        const double* sj = ST + j * np;
        const int64_t x = i ^ j;
        double br = Cst[i] + Cst[j];
        for (int p = 0; p < n; ++p) {
            br = fma(si[p], sj[p], br);
            const double ri = (1.0 - si[p]) * (((i >> p) & 1) ? -1.0 : 1.0);
            const double rj = (1.0 - sj[p]) * (((j >> p) & 1) ? -1.0 : 1.0);
            if ((j >> p) & 1) br += prm.gamma * ri;
            if ((i >> p) & 1) br += prm.gamma * rj;
        }
        acc += prm.wl[__popcll(x)] * br * qT[j];
    }
    part[i - prm.row0] = acc;
}

template <typename T, int NP, int R, int TC>
int launch_rows(const KsdParams& prm, dim3 grid, const void* ST, const void* Cst, const void* qT,
                double* part, cudaStream_t st) {
    using SM = Smem<T, NP, R, TC>;
    auto kern = ksd_rows_kernel<T, NP, R, TC>;
    TNK_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM::total));
    kern<<<grid, kThreads, SM::total, st>>>(prm, (const T*)ST, (const T*)Cst, (const T*)qT, part);
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}

struct Plan {
    TnkGeom g;
    int64_t N, row0, launch_rows, n_tiles, finish_blocks;
    int row_blocks, splits;
    bool tensor;          // TNK_PREC_TC32 with a tensor-core plan (7 <= n <= 24); else the ALU kernels
    TnkTcPlan tc;
    size_t off_q, off_part, off_block, off_b, off_qtab, total;
};

int make_plan(int n, int precision, int64_t row_begin, int64_t row_end, Plan* p) {
    if (tnk_geom(n, precision, &p->g)) return 1;
    p->N = (int64_t)1 << n;
    if (row_begin < 0 || row_end > p->N || row_begin > row_end) return 1;
    p->tensor = precision == TNK_PREC_TC32 && tnk_tc_make_plan(n, &p->tc) == 0;
    const int64_t RB = p->tensor ? (int64_t)TNK_TC_TILE : (int64_t)kThreads * p->g.R;
    p->row0 = (row_begin / RB) * RB;
    const int64_t row_hi = ((row_end + RB - 1) / RB) * RB;
    p->row_blocks = (int)((row_hi - p->row0) / RB);
    p->launch_rows = row_hi - p->row0;
    p->n_tiles = p->tensor ? p->N / TNK_TC_TILE : p->g.n_pad / p->g.TC;
    int sms = 148;
    int splits = 1;
    if (precision != TNK_PREC_FP64_SIMPLE) {
        while ((int64_t)p->row_blocks * splits < 4 * sms && splits * 2 <= p->n_tiles) splits *= 2;
    }
    if (p->tensor) {
        // one split's slice of the column table should stay L2-resident (126 MB L2; 32 MB target)
        const size_t b_bytes = tnk_tc_table_bytes(p->tc, p->N);
        while ((b_bytes / splits) > ((size_t)32 << 20) && splits * 2 <= p->n_tiles) splits *= 2;
    }
    p->splits = splits;
    p->finish_blocks = (row_end - row_begin + 255) / 256;
    size_t off = 0;
    p->off_q = off;      off += ((size_t)p->g.n_pad * p->g.elem + 255) / 256 * 256;
    p->off_part = off;   off += ((size_t)splits * p->launch_rows * sizeof(double) + 255) / 256 * 256;
    p->off_block = off;  off += ((size_t)(p->finish_blocks + 1) * sizeof(double) + 255) / 256 * 256;
    p->off_b = p->off_qtab = off;
    if (p->tensor) {
        p->off_qtab = off;  off += (tnk_tc_qtab_bytes(p->N) + 255) / 256 * 256;
        p->off_b = off;  off += (tnk_tc_table_bytes(p->tc, p->N) + 255) / 256 * 256;
    }
    p->total = off;
    return 0;
}

}  // namespace

extern "C" size_t tnk_ksd_workspace_bytes(int n_latent, int precision, int64_t row_begin, int64_t row_end) {
    Plan p;
    if (make_plan(n_latent, precision, row_begin, row_end, &p)) return 0;
    return p.total;
}

namespace {

template <typename T, int R, int TC>
int dispatch_np(int np, const KsdParams& prm, dim3 grid, const void* ST, const void* Cst, const void* qT,
                double* part, cudaStream_t st) {
    switch (np) {
        case 4:  return launch_rows<T, 4, R, TC>(prm, grid, ST, Cst, qT, part, st);
        case 8:  return launch_rows<T, 8, R, TC>(prm, grid, ST, Cst, qT, part, st);
        case 12: return launch_rows<T, 12, R, TC>(prm, grid, ST, Cst, qT, part, st);
        case 16: return launch_rows<T, 16, R, TC>(prm, grid, ST, Cst, qT, part, st);
        case 20: return launch_rows<T, 20, R, TC>(prm, grid, ST, Cst, qT, part, st);
        case 24: return launch_rows<T, 24, R, TC>(prm, grid, ST, Cst, qT, part, st);
        default: break;
    }
    tnk_set_error("tnk_ksd_rows: no kernel instance for padded length %d", np);
    return 2;
}

template <typename T, int R, int TC>
int dispatch_np_wide(int np, const KsdParams& prm, dim3 grid, const void* ST, const void* Cst, const void* qT,
                     double* part, cudaStream_t st) {
    switch (np) {
        case 28: return launch_rows<T, 28, R, TC>(prm, grid, ST, Cst, qT, part, st);
        case 32: return launch_rows<T, 32, R, TC>(prm, grid, ST, Cst, qT, part, st);
        default: break;
    }
    tnk_set_error("tnk_ksd_rows: no kernel instance for padded length %d", np);
    return 2;
}

}  // namespace

static thread_local cudaEvent_t g_ev_start = nullptr, g_ev_stop = nullptr;

void tnk_timing_events_get(cudaEvent_t* start, cudaEvent_t* stop) {
    *start = g_ev_start;
    *stop = g_ev_stop;
}

extern "C" int tnk_ksd_timing_events(void* start_event, void* stop_event) {
    g_ev_start = (cudaEvent_t)start_event;
    g_ev_stop = (cudaEvent_t)stop_event;
    return 0;
}

static int ksd_rows_impl(int n_latent, double length_scale, int precision,
                         const void* q, int q_dtype, const void* ST, const void* Cst,
                         int64_t row_begin, int64_t row_end,
                         double* u, double* ksd_part,
                         void* workspace, size_t workspace_bytes, int* n_launch, void* stream,
                         const PeerSinks& peers) {
    Plan p;
    TNK_REQUIRE(make_plan(n_latent, precision, row_begin, row_end, &p) == 0,
                "tnk_ksd_rows: bad arguments n=%d precision=%d rows=[%lld,%lld)", n_latent, precision,
                (long long)row_begin, (long long)row_end);
    TNK_REQUIRE(q && ST && Cst && u && ksd_part && workspace, "tnk_ksd_rows: NULL pointer argument");
    TNK_REQUIRE(q_dtype == 0 || q_dtype == 1, "tnk_ksd_rows: q_dtype must be 0 (double) or 1 (float)");
    TNK_REQUIRE(workspace_bytes >= p.total, "tnk_ksd_rows: workspace %zu < required %zu", workspace_bytes, p.total);
    TNK_REQUIRE(((uintptr_t)workspace & 255) == 0, "tnk_ksd_rows: workspace must be 256-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const int n = n_latent;
    int launches = 0;
    unsigned char* ws = (unsigned char*)workspace;
    void* qT = ws + p.off_q;
    double* part = (double*)(ws + p.off_part);
    double* blockpart = (double*)(ws + p.off_block);

    if (row_end == row_begin || n == 0) {
        // n == 0: k_p is identically 0 (reference stein_utils.py:151-152)
        TNK_CUDA_OK(cudaMemsetAsync(ksd_part, 0, sizeof(double), st));
        if (row_end > row_begin) TNK_CUDA_OK(cudaMemsetAsync(u + row_begin, 0, (row_end - row_begin) * sizeof(double), st));
        for (int pr = 0; pr < peers.n; ++pr) {
            TNK_CUDA_OK(cudaMemsetAsync(peers.parts[pr] + peers.rank, 0, sizeof(double), st));
            if (row_end > row_begin && pr != peers.rank)
                TNK_CUDA_OK(cudaMemsetAsync(peers.u[pr] + row_begin, 0, (row_end - row_begin) * sizeof(double), st));
        }
        if (n_launch) *n_launch = 0;
        return 0;
    }
    TNK_REQUIRE((double)n * length_scale != 0.0, "tnk_ksd_rows: num_vars*length_scale must be non-zero");

    KsdParams prm;
    prm.n = n;
    prm.row0 = p.row0;
    prm.n_tiles = p.n_tiles;
    prm.col_splits = p.splits;
    prm.part_stride = p.launch_rows;
    prm.a = exp(-1.0 / ((double)n * length_scale));
    prm.gamma = prm.a - 1.0 / prm.a;
    for (int h = 0; h < 32; ++h) prm.wl[h] = (h <= n) ? exp(-(double)h / ((double)n * length_scale)) : 0.0;

    const unsigned qblocks = (unsigned)((p.g.n_pad + 255) / 256);
    if (p.g.elem == 4) {
        if (q_dtype == 0) q_pack_kernel<float, double><<<qblocks, 256, 0, st>>>((const double*)q, p.N, p.g.n_pad, (float*)qT);
        else q_pack_kernel<float, float><<<qblocks, 256, 0, st>>>((const float*)q, p.N, p.g.n_pad, (float*)qT);
    } else {
        if (q_dtype == 0) q_pack_kernel<double, double><<<qblocks, 256, 0, st>>>((const double*)q, p.N, p.g.n_pad, (double*)qT);
        else q_pack_kernel<double, float><<<qblocks, 256, 0, st>>>((const float*)q, p.N, p.g.n_pad, (double*)qT);
    }
    TNK_CUDA_OK(cudaGetLastError());
    ++launches;

    int rc = 0;
    if (p.tensor) {
        rc = tnk_tc_launch(n, prm.a, prm.gamma, prm.wl, (const float*)ST, (const float*)Cst, (const float*)qT,
                           p.row0, p.launch_rows, p.splits, (float*)(ws + p.off_b), (float*)(ws + p.off_qtab), part,
                           g_ev_start, g_ev_stop, &launches, st);
        if (rc) return rc;
        --launches;   // the common "++launches" below counts the pair kernel
    } else {
    if (g_ev_start) TNK_CUDA_OK(cudaEventRecord(g_ev_start, st));
    if (precision == TNK_PREC_FP64_SIMPLE) {
        const unsigned blocks = (unsigned)((row_end - row_begin + 127) / 128);
        ksd_rows_simple_kernel<<<blocks, 128, 0, st>>>(prm, p.g.np, p.N, row_begin, row_end,
                                                       (const double*)ST, (const double*)Cst, (const double*)qT, part);
        TNK_CUDA_OK(cudaGetLastError());
    } else {
        dim3 grid((unsigned)p.row_blocks, (unsigned)p.splits, 1);
        if (precision == TNK_PREC_FP32 || precision == TNK_PREC_TC32) {
            if (p.g.R == 4) rc = dispatch_np<float, 4, 256>(p.g.np, prm, grid, ST, Cst, qT, part, st);
            else rc = dispatch_np_wide<float, 2, 256>(p.g.np, prm, grid, ST, Cst, qT, part, st);
        } else {
            if (p.g.R == 2) rc = dispatch_np<double, 2, 128>(p.g.np, prm, grid, ST, Cst, qT, part, st);
            else rc = dispatch_np_wide<double, 1, 128>(p.g.np, prm, grid, ST, Cst, qT, part, st);
        }
        if (rc) return rc;
    }
    if (g_ev_stop) TNK_CUDA_OK(cudaEventRecord(g_ev_stop, st));
    }
    ++launches;

    if (q_dtype == 0)
        finish_rows_kernel<double><<<(unsigned)p.finish_blocks, 256, 0, st>>>(part, p.splits, p.launch_rows, p.row0, row_begin, row_end,
                                                                             (const double*)q, u, blockpart, peers);
    else
        finish_rows_kernel<float><<<(unsigned)p.finish_blocks, 256, 0, st>>>(part, p.splits, p.launch_rows, p.row0, row_begin, row_end,
                                                                            (const float*)q, u, blockpart, peers);
    TNK_CUDA_OK(cudaGetLastError());
    finish_scalar_kernel<<<1, 256, 0, st>>>(blockpart, p.finish_blocks, ksd_part, peers);
    TNK_CUDA_OK(cudaGetLastError());
    launches += 2;
    if (n_launch) *n_launch = launches;
    return 0;
}

extern "C" int tnk_ksd_rows(int n_latent, double length_scale, int precision,
                            const void* q, int q_dtype, const void* ST, const void* Cst,
                            int64_t row_begin, int64_t row_end,
                            double* u, double* ksd_part,
                            void* workspace, size_t workspace_bytes, int* n_launch, void* stream) {
    PeerSinks none;
    none.n = 0;
    none.rank = 0;
    return ksd_rows_impl(n_latent, length_scale, precision, q, q_dtype, ST, Cst, row_begin, row_end, u, ksd_part,
                         workspace, workspace_bytes, n_launch, stream, none);
}

extern "C" int tnk_ksd_rows_peers(int n_latent, double length_scale, int precision,
                                  const void* q, int q_dtype, const void* ST, const void* Cst,
                                  int64_t row_begin, int64_t row_end,
                                  int n_peers, int my_rank, void* const* peer_u, void* const* peer_parts,
                                  double* ksd_part,
                                  void* workspace, size_t workspace_bytes, int* n_launch, void* stream) {
    TNK_REQUIRE(n_peers >= 1 && n_peers <= TNK_MAX_PEERS, "tnk_ksd_rows_peers: n_peers must be in [1,%d]", TNK_MAX_PEERS);
    TNK_REQUIRE(my_rank >= 0 && my_rank < n_peers, "tnk_ksd_rows_peers: my_rank %d outside [0,%d)", my_rank, n_peers);
    TNK_REQUIRE(peer_u && peer_parts, "tnk_ksd_rows_peers: NULL peer pointer table");
    PeerSinks peers;
    peers.n = n_peers;
    peers.rank = my_rank;
    for (int p = 0; p < n_peers; ++p) {
        TNK_REQUIRE(peer_u[p] && peer_parts[p], "tnk_ksd_rows_peers: NULL buffer for peer %d", p);
        peers.u[p] = (double*)peer_u[p];
        peers.parts[p] = (double*)peer_parts[p];
    }
    return ksd_rows_impl(n_latent, length_scale, precision, q, q_dtype, ST, Cst, row_begin, row_end,
                         peers.u[my_rank], ksd_part, workspace, workspace_bytes, n_launch, stream, peers);
}


// ------------------------------------------------------------------------------------
// Symmetric tensor-core evaluation (tnk_ksd_sym): every unordered pair of 128-state tiles is evaluated once.
// ------------------------------------------------------------------------------------
namespace {

struct SymPlan {
    int64_t N, T, finish_blocks;
    int row_tiles, splits;
    size_t off_q, off_col, off_block, off_b, off_hib, total;
};

int make_sym_plan(int n, int split, int tile_offset, int tile_stride, SymPlan* p) {
    if (n < TNK_TC_MIN_LATENT || n > TNK_TC_MAX_LATENT || tile_stride < 1 || tile_offset < 0) return 1;
    if (split < TNK_SYM_SPLIT_TF32 || split > TNK_SYM_SPLIT_TF32_DOT) return 1;
    p->N = (int64_t)1 << n;
    p->T = p->N / TNK_TC_TILE;
    p->row_tiles = tile_offset < p->T ? (int)((p->T - tile_offset + tile_stride - 1) / tile_stride) : 0;
    const size_t b_bytes = tnk_tcs_table_bytes(n, split, p->N);
    // column splits: enough CTAs (row-tile groups x splits) for the block scheduler to balance the triangle - a CTA of
    // the row-tile kernel owns up to 4 row tiles, and the longest one must stay a small part of an SM's share
    const int rt = split == TNK_SYM_SPLIT_TF32_DOT ? 4 : 2;
    const int64_t row_groups = (p->row_tiles + rt - 1) / rt;
    int splits = 1;
    while (row_groups * splits < 32 * 148 && splits * 2 <= p->T) splits *= 2;   // scripts/sym_splits.py: 1/8 stripe at n = 20 14.56 (2 splits) -> 13.87 ms (32)
    while ((b_bytes / splits) > ((size_t)32 << 20) && splits * 2 <= p->T) splits *= 2;
    if (const char* e = getenv("TNK_SYM_SPLITS")) {          // measurement override (scripts/sym_splits.py)
        const int v = atoi(e);
        if (v >= 1 && v <= p->T && (v & (v - 1)) == 0) splits = v;
    }
    p->splits = splits;
    p->finish_blocks = (p->N + 255) / 256;
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t at = off; off += (bytes + 255) / 256 * 256; return at; };
    p->off_q = take((size_t)p->N * sizeof(float));
    p->off_col = take((size_t)p->N * sizeof(double));
    p->off_block = take((size_t)(p->finish_blocks + 1) * sizeof(double));
    p->off_b = take(b_bytes);
    p->off_hib = off;
    // every call takes its share of the Kronecker side pass (bit positions offset, offset + stride, ...; C terms with offset 0)
    if (split == TNK_SYM_SPLIT_TF32_TILE) p->off_hib = take(tnk_hibits_ws_bytes(n, 7, 0, tile_offset, tile_stride));
    if (split == TNK_SYM_SPLIT_TF32_DOT) p->off_hib = take(tnk_hibits_ws_bytes(n, 0, tile_offset == 0, tile_offset, tile_stride));
    p->total = off;
    return 0;
}

// u_partial = the accumulator the tile kernel added its row sums and mirrored column sums to; block partials of q.u
template <typename Q>
__global__ void __launch_bounds__(256)
finish_sym_kernel(const double* __restrict__ uacc, const Q* __restrict__ q, int64_t N,
                  double* __restrict__ u, double* __restrict__ blockpart) {
    __shared__ double red[256];
    const int64_t z = (int64_t)blockIdx.x * 256 + threadIdx.x;
    double contrib = 0.0;
    if (z < N) {
        const double acc = uacc[z];
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

}  // namespace

extern "C" size_t tnk_ksd_sym_workspace_bytes(int n_latent, int split, int tile_offset, int tile_stride) {
    SymPlan p;
    if (make_sym_plan(n_latent, split, tile_offset, tile_stride, &p)) return 0;
    return p.total;
}

extern "C" int tnk_ksd_sym(int n_latent, double length_scale, int split, const void* q, int q_dtype,
                           const void* ST, const void* Cst, int tile_offset, int tile_stride, int reuse_table,
                           double* u_partial, double* ksd_part,
                           void* workspace, size_t workspace_bytes, int* n_launch, void* stream) {
    SymPlan p;
    TNK_REQUIRE(make_sym_plan(n_latent, split, tile_offset, tile_stride, &p) == 0,
                "tnk_ksd_sym: bad arguments n=%d (supported %d..%d) split=%d tile_offset=%d tile_stride=%d", n_latent,
                TNK_TC_MIN_LATENT, TNK_TC_MAX_LATENT, split, tile_offset, tile_stride);
    TNK_REQUIRE(q && ST && Cst && u_partial && ksd_part && workspace, "tnk_ksd_sym: NULL pointer argument");
    TNK_REQUIRE(q_dtype == 0 || q_dtype == 1, "tnk_ksd_sym: q_dtype must be 0 (double) or 1 (float)");
    TNK_REQUIRE(workspace_bytes >= p.total, "tnk_ksd_sym: workspace %zu < required %zu", workspace_bytes, p.total);
    TNK_REQUIRE(((uintptr_t)workspace & 255) == 0, "tnk_ksd_sym: workspace must be 256-byte aligned");
    TNK_REQUIRE((double)n_latent * length_scale != 0.0, "tnk_ksd_sym: num_vars*length_scale must be non-zero");
    cudaStream_t st = (cudaStream_t)stream;
    const int n = n_latent;
    unsigned char* ws = (unsigned char*)workspace;
    float* qT = (float*)(ws + p.off_q);
    double* ucol = (double*)(ws + p.off_col);
    double* blockpart = (double*)(ws + p.off_block);
    int launches = 0;
    const double a = exp(-1.0 / ((double)n * length_scale));
    const double gamma = a - 1.0 / a;
    double wl[32];
    for (int h = 0; h < 32; ++h) wl[h] = (h <= n) ? exp(-(double)h / ((double)n * length_scale)) : 0.0;

    const unsigned qblocks = (unsigned)((p.N + 255) / 256);
    if (q_dtype == 0) q_pack_kernel<float, double><<<qblocks, 256, 0, st>>>((const double*)q, p.N, p.N, qT);
    else q_pack_kernel<float, float><<<qblocks, 256, 0, st>>>((const float*)q, p.N, p.N, qT);
    TNK_CUDA_OK(cudaGetLastError());
    ++launches;
    const int rc = tnk_tcs_launch(n, split, gamma, wl, (const float*)ST, (const float*)Cst, qT, tile_offset, tile_stride,
                                  p.row_tiles, p.splits, (float*)(ws + p.off_b), reuse_table, ucol, (double*)(ws + p.off_hib),
                                  g_ev_start, g_ev_stop, &launches, st);
    if (rc) return rc;
    if (q_dtype == 0)
        finish_sym_kernel<double><<<(unsigned)p.finish_blocks, 256, 0, st>>>(ucol, (const double*)q, p.N, u_partial, blockpart);
    else
        finish_sym_kernel<float><<<(unsigned)p.finish_blocks, 256, 0, st>>>(ucol, (const float*)q, p.N, u_partial, blockpart);
    TNK_CUDA_OK(cudaGetLastError());
    PeerSinks none;
    none.n = 0;
    none.rank = 0;
    finish_scalar_kernel<<<1, 256, 0, st>>>(blockpart, p.finish_blocks, ksd_part, none);
    TNK_CUDA_OK(cudaGetLastError());
    launches += 2;
    if (n_launch) *n_launch = launches;
    return 0;
}

// tnk_ksd_sym + exchange over the communicator's peer-mapped buffers (tnk_comm.cu): the partial u and the partial
// KSD^2 are written straight into the send region, one all-reduce of N + 1 doubles follows on the same stream.
extern "C" int tnk_ksd_sym_allreduce(int n_latent, double length_scale, int split, const void* q, int q_dtype,
                                     const void* ST, const void* Cst, int reuse_table, tnk_comm* comm,
                                     double* u, double* ksd,
                                     void* workspace, size_t workspace_bytes, int* n_launch, void* stream) {
    TNK_REQUIRE(comm && u && ksd, "tnk_ksd_sym_allreduce: NULL argument");
    TNK_REQUIRE(n_latent >= TNK_TC_MIN_LATENT && n_latent <= TNK_TC_MAX_LATENT, "tnk_ksd_sym_allreduce: n=%d outside %d..%d",
                n_latent, TNK_TC_MIN_LATENT, TNK_TC_MAX_LATENT);
    const int64_t N = (int64_t)1 << n_latent;
    double* send = tnk_comm_sendbuf(comm);
    int l1 = 0, l2 = 0;
    int rc = tnk_ksd_sym(n_latent, length_scale, split, q, q_dtype, ST, Cst, tnk_comm_rank(comm), tnk_comm_world(comm),
                         reuse_table, send, send + N, workspace, workspace_bytes, &l1, stream);
    if (rc) return rc;
    rc = tnk_comm_allreduce_sum(comm, N + 1, nullptr, &l2, stream);
    if (rc) return rc;
    const double* recv = tnk_comm_recvbuf(comm);
    TNK_CUDA_OK(cudaMemcpyAsync(u, recv, (size_t)N * sizeof(double), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    TNK_CUDA_OK(cudaMemcpyAsync(ksd, recv + N, sizeof(double), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    if (n_launch) *n_launch = l1 + l2;
    return 0;
}

// ---- tnk_ksd_rows_static: tnk_ksd_rows for callers that cannot provide (or allocate) a workspace --------------------
// No counterpart in the reference; added for callers that must not call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The workspace of
// tnk_ksd_rows (packed q, per-split row partials, block partials, and for TNK_PREC_TC32 the operand tables) lives in a
// file-scope device array of TNK_STATIC_SCRATCH_BYTES instead; every GPU that loads the library holds that array until
// the process exits.  `g_static_scratch` is a host-side stand-in for the symbol in host code, not a device address:
// the address is looked up once per device with cudaGetSymbolAddress (that device current) and handed to the kernels.
// Nothing in the workspace accumulates across calls (every slot a kernel reads is written earlier in the same call),
// so nothing has to be cleared; all calls on one device share the array, so calls that overlap on the GPU (different
// streams, or graph replays running concurrently) race - serialise them.
static __device__ __align__(256) unsigned char g_static_scratch[TNK_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
void* g_static_addr[kMaxDevices];                 // device address of g_static_scratch, per device ordinal
std::mutex g_static_mutex;

int static_scratch_address(void** out) {
    int dev = -1;
    TNK_CUDA_OK(cudaGetDevice(&dev));
    TNK_REQUIRE(dev >= 0 && dev < kMaxDevices, "tnk_ksd_rows_static: device ordinal %d outside [0,%d)", dev, kMaxDevices);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_addr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        TNK_REQUIRE(e == cudaSuccess && p, "tnk_ksd_rows_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                    cudaGetErrorString(e));
        g_static_addr[dev] = p;
    }
    *out = g_static_addr[dev];
    return 0;
}
}  // namespace

extern "C" size_t tnk_ksd_rows_static_capacity(void) { return (size_t)TNK_STATIC_SCRATCH_BYTES; }

extern "C" int tnk_ksd_rows_static(int n_latent, double length_scale, int precision,
                                   const void* q, int q_dtype, const void* ST, const void* Cst,
                                   int64_t row_begin, int64_t row_end,
                                   double* u, double* ksd_part, int* n_launch, void* stream) {
    Plan p;
    TNK_REQUIRE(make_plan(n_latent, precision, row_begin, row_end, &p) == 0,
                "tnk_ksd_rows_static: bad arguments n=%d precision=%d rows=[%lld,%lld)", n_latent, precision,
                (long long)row_begin, (long long)row_end);
    TNK_REQUIRE(p.total <= (size_t)TNK_STATIC_SCRATCH_BYTES,
                "tnk_ksd_rows_static: n=%d precision=%d rows=[%lld,%lld) needs %zu bytes of scratch, the static array holds %zu "
                "(TNK_STATIC_SCRATCH_BYTES); use tnk_ksd_rows with a caller-provided workspace",
                n_latent, precision, (long long)row_begin, (long long)row_end, p.total, (size_t)TNK_STATIC_SCRATCH_BYTES);
    void* scratch = nullptr;
    if (const int rc = static_scratch_address(&scratch)) return rc;
    PeerSinks none;
    none.n = 0;
    none.rank = 0;
    return ksd_rows_impl(n_latent, length_scale, precision, q, q_dtype, ST, Cst, row_begin, row_end, u, ksd_part,
                         scratch, (size_t)TNK_STATIC_SCRATCH_BYTES, n_launch, stream, none);
}
