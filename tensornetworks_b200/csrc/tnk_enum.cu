// Enumeration kernels: K1 joint table p(x,z) and K2 score / packed-table builder.
// HBM-bound streaming kernels over the 2^n packed states (one thread per state).
#include "tnk_common.cuh"
#include <type_traits>

static thread_local char g_err[512] = "";

void tnk_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* tnk_last_error(void) { return g_err; }
extern "C" int tnk_version(void) { return TNK_ABI_VERSION; }

extern "C" int tnk_ksd_layout(int n_latent, int precision, int* np, int64_t* n_pad, int* elem) {
    TnkGeom g;
    TNK_REQUIRE(tnk_geom(n_latent, precision, &g) == 0, "tnk_ksd_layout: unsupported n=%d precision=%d", n_latent, precision);
    if (np) *np = g.np;
    if (n_pad) *n_pad = g.n_pad;
    if (elem) *elem = g.elem;
    return 0;
}

// ------------------------------------------------------------------------------------
// K1.  Replaces compute_prob_joint_xz (reference stein_utils.py:58-112) for every state.
// The network descriptor is tiny and identical for all threads: it is staged in shared
// memory once per CTA; each thread then walks the nodes in bn.nodes order.
// ------------------------------------------------------------------------------------
constexpr int kJointPaSlots = 4096;   // 16 KB of dynamic shared memory for the parent lists

struct BnDesc {
    int32_t kind[TNK_MAX_NODES];
    int32_t arg[TNK_MAX_NODES];
    int32_t pa_off[TNK_MAX_NODES + 1];
    int32_t cpt_off[TNK_MAX_NODES];
};

__global__ void __launch_bounds__(256)
bn_joint_kernel(int n_latent, int n_hidden, int n_nodes,
                const int32_t* __restrict__ node_kind, const int32_t* __restrict__ node_arg,
                const int32_t* __restrict__ pa_off, const int32_t* __restrict__ pa_idx,
                const int32_t* __restrict__ cpt_off, const double* __restrict__ cpt,
                double* __restrict__ P) {
    __shared__ BnDesc d;
    extern __shared__ int32_t s_pa[];  // parent lists
    for (int v = threadIdx.x; v < n_nodes; v += blockDim.x) {
        d.kind[v] = node_kind[v];
        d.arg[v] = node_arg[v];
        d.cpt_off[v] = cpt_off[v];
    }
    for (int v = threadIdx.x; v <= n_nodes; v += blockDim.x) d.pa_off[v] = pa_off[v];
    __syncthreads();
    // parent lists: staged in shared memory when they fit (kJointPaSlots entries), else read from global memory
    const int n_pa = d.pa_off[n_nodes];
    const int32_t* pa = pa_idx;
    if (n_pa <= kJointPaSlots) {
        for (int t = threadIdx.x; t < n_pa; t += blockDim.x) s_pa[t] = pa_idx[t];
        pa = s_pa;
    }
    __syncthreads();

    const int64_t N = (int64_t)1 << n_latent;
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N) return;

    // node values that do not depend on the hidden assignment
    uint64_t fixed = 0;
    for (int v = 0; v < n_nodes; ++v) {
        uint64_t bit = 0;
        if (d.kind[v] == TNK_NODE_LATENT) bit = (uint64_t)(idx >> d.arg[v]) & 1u;
        else if (d.kind[v] == TNK_NODE_CONST) bit = (uint64_t)(d.arg[v] & 1);
        fixed |= bit << v;
    }
    const uint32_t n_hid_cfg = 1u << n_hidden;
    double total = 0.0;
    for (uint32_t hcfg = 0; hcfg < n_hid_cfg; ++hcfg) {
        uint64_t assign = fixed;
        if (n_hidden) {
            for (int v = 0; v < n_nodes; ++v)
                if (d.kind[v] == TNK_NODE_HIDDEN) assign |= (uint64_t)((hcfg >> d.arg[v]) & 1u) << v;
        }
        double prob = 1.0;  // bayesian_network.py:129
        for (int v = 0; v < n_nodes; ++v) {
            int cfg = 0;
            for (int t = d.pa_off[v]; t < d.pa_off[v + 1]; ++t)
                cfg = (cfg << 1) | (int)((assign >> pa[t]) & 1u);
            const int val = (int)((assign >> v) & 1u);
            prob = __dmul_rn(prob, __ldg(&cpt[2 * ((int64_t)d.cpt_off[v] + cfg) + val]));  // :144
        }
        total = (n_hidden == 0) ? prob : __dadd_rn(total, prob);  // stein_utils.py:96-111
    }
    P[idx] = total;
}

extern "C" int tnk_bn_joint(int n_latent, int n_hidden, int n_nodes,
                            const int32_t* node_kind, const int32_t* node_arg,
                            const int32_t* pa_off, const int32_t* pa_idx,
                            const int32_t* cpt_off, const double* cpt,
                            double* P, void* stream) {
    TNK_REQUIRE(n_latent >= 0 && n_latent <= TNK_MAX_LATENT, "tnk_bn_joint: n_latent=%d out of range", n_latent);
    TNK_REQUIRE(n_hidden >= 0 && n_hidden <= 20, "tnk_bn_joint: n_hidden=%d out of range", n_hidden);
    TNK_REQUIRE(n_nodes >= 0 && n_nodes <= TNK_MAX_NODES, "tnk_bn_joint: n_nodes=%d exceeds %d", n_nodes, TNK_MAX_NODES);
    TNK_REQUIRE(P != nullptr, "tnk_bn_joint: P is NULL");
    const int64_t N = (int64_t)1 << n_latent;
    const unsigned blocks = (unsigned)((N + 255) / 256);
    bn_joint_kernel<<<blocks, 256, (size_t)kJointPaSlots * sizeof(int32_t), (cudaStream_t)stream>>>(
        n_latent, n_hidden, n_nodes, node_kind, node_arg, pa_off, pa_idx, cpt_off, cpt, P);
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------
// K2.  Replaces get_score_function_sp_for_z (reference stein_utils.py:115-136) for every
// state and emits the packed operands of the pair kernel.  flip_bit(z,k) is
// idx ^ (1 << (n-1-k)), so the n neighbour probabilities are gathers from P itself.
// Each CTA computes 256 rows into shared memory and writes them out coalesced.
// ------------------------------------------------------------------------------------
// a / p, correctly rounded, from the correctly rounded reciprocal rp = RN(1/p) shared by the n divisions of a
// state: q0 = a rp is faithful, and each residual step q += (a - q p) rp (exact residual by FMA) then returns
// RN(a/p) (Markstein's theorem; two steps for margin).  Probabilities keep every intermediate in the normal
// range (|p| >= 1e-12, 0 <= a <= 1).  The oracle divides with '/', the parity tests compare the tables with ==.
__device__ __forceinline__ double div_by(double a, double p, double rp) {
    double q = __dmul_rn(a, rp);
    q = __fma_rn(__fma_rn(-q, p, a), rp, q);
    q = __fma_rn(__fma_rn(-q, p, a), rp, q);
    return q;
}

// NP = padded score length (compile time: the loops over score components unroll and the shared-memory
// strides are constants), n <= NP at run time.  One fp64 staging tile serves both outputs (S64 in variable order,
// the packed table in bit order and compute type): 256 x (NP + 1) doubles = 43 KB at n = 20, five CTAs per SM, so the
// gather / divide / copy-out phases of different CTAs overlap.
constexpr int kScoreRows = 64;    // states per CTA of the score kernel
// NN = n when it is a compile-time constant of the instance (NP - 3 <= n <= NP, n >= 1: the k < n tests, the flip masks
// and the copy-out strides fold), -1 = run-time n (n = 0).
// KEEP: S64 is written too, so the staging tile holds doubles; without it the tile holds the packed table's own type
// (a third of the shared memory for fp32 tables: full occupancy for the gathers).
template <typename T, int NP, int NN, bool KEEP>
__global__ void __launch_bounds__(kScoreRows)
score_kernel(int n_rt, int64_t N, double alpha, double gamma,
             const double* __restrict__ P, double* __restrict__ S64,
             T* __restrict__ ST, T* __restrict__ Cst) {
    // Rows are staged with the ODD stride NP + 1 so that the per-thread row writes are bank-conflict-free; the
    // copy-outs then walk the tile linearly.
    constexpr int kStride = NP + 1;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using E = typename std::conditional<KEEP, double, T>::type;
    E* sS = reinterpret_cast<E*>(smem_raw);                                                    // [kScoreRows][NP+1]
    const int n = NN >= 0 ? NN : n_rt;
    const int tid = threadIdx.x;
    const int64_t base = (int64_t)blockIdx.x * kScoreRows;
    const int64_t i = base + tid;
    double Rsum = 0.0, rho0 = 0.0;
    if (i < N) {
        const double p = P[i];
        const bool dead = fabs(p) < 1e-12;  // stein_utils.py:126
        const double rp = dead ? 0.0 : __drcp_rn(p);
        const uint32_t iu = (uint32_t)i;     // n <= 30
        constexpr int G = (NP % 8 == 0) ? 8 : 4;     // flipped reads in flight; arithmetic in the original order
#pragma unroll
        for (int k0 = 0; k0 < NP; k0 += G) {
            double pf[G];
#pragma unroll
            for (int c = 0; c < G; ++c) {
                const int k = k0 + c;
                pf[c] = (k < n && !dead) ? P[iu ^ (1u << (n - 1 - k))] : 0.0;
            }
#pragma unroll
            for (int c = 0; c < G; ++c) {
                const int k = k0 + c;
                double s = 0.0;
                if (k < n) {
                    const int bit = n - 1 - k;
                    if (!dead) s = __dsub_rn(1.0, div_by(pf[c], p, rp));  // :134
                    const double r = 1.0 - s;
                    Rsum += r;
                    if ((iu >> bit) & 1u) rho0 += r;
                }
                sS[tid * kStride + k] = (E)s;                              // zero beyond n and in dead rows
            }
        }
        if (Cst) Cst[i] = (T)(alpha * Rsum + gamma * rho0);
    } else {
#pragma unroll
        for (int k = 0; k < NP; ++k) sS[tid * kStride + k] = (E)0;         // padding rows of the packed table
        if (Cst) Cst[i] = (T)0;
    }
    __syncthreads();
    if (KEEP && n > 0) {
        const int live = (int)(N - base < kScoreRows ? (N - base > 0 ? N - base : 0) : kScoreRows) * n;
        const int dq = kScoreRows / n, dr = kScoreRows % n;  // element t = tid + kScoreRows j sits in row t / n, column t % n
        int row = tid / n, col = tid % n;
        double* out = S64 + base * n;
#pragma unroll
        for (int t = tid; t < live; t += kScoreRows) {
            __stcs(out + t, (double)sS[row * kStride + col]);   // streaming stores: the tables must not evict P from L2
            row += dq; col += dr;
            if (col >= n) { col -= n; ++row; }
        }
    }
    if (ST) {
        // 16-byte stores: a row of NP elements is NP*sizeof(T)/16 chunks; column `bit` of the packed table is
        // variable n - 1 - bit, zero for bit >= n
        constexpr int kPer = 16 / (int)sizeof(T);           // elements per chunk
        constexpr int kChunksPerRow = NP / kPer;
        T* out = ST + base * NP;
#pragma unroll
        for (int c = 0; c < kChunksPerRow; ++c) {
            const int ch = tid + kScoreRows * c;                   // chunk index in the tile
            const int row = ch / kChunksPerRow, col = (ch % kChunksPerRow) * kPer;
            T v[kPer];
#pragma unroll
            for (int e = 0; e < kPer; ++e) {
                const int k = n - 1 - (col + e);
                v[e] = k >= 0 ? (T)sS[row * kStride + k] : (T)0;
            }
            __stcs(reinterpret_cast<uint4*>(out + (size_t)ch * kPer), *reinterpret_cast<const uint4*>(v));
        }
    }
}

template <typename T, int NP, int NN>
int launch_score_n(int n, int64_t N, int64_t rows, double alpha, double gamma, const double* P, double* S64,
                   T* ST, T* Cst, cudaStream_t st) {
    const unsigned blocks = (unsigned)((rows + kScoreRows - 1) / kScoreRows);
    if (S64) {
        const size_t smem = (size_t)kScoreRows * (NP + 1) * sizeof(double);
        TNK_CUDA_OK(cudaFuncSetAttribute(score_kernel<T, NP, NN, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        score_kernel<T, NP, NN, true><<<blocks, kScoreRows, smem, st>>>(n, N, alpha, gamma, P, S64, ST, Cst);
    } else {
        const size_t smem = (size_t)kScoreRows * (NP + 1) * sizeof(T);
        TNK_CUDA_OK(cudaFuncSetAttribute(score_kernel<T, NP, NN, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        score_kernel<T, NP, NN, false><<<blocks, kScoreRows, smem, st>>>(n, N, alpha, gamma, P, S64, ST, Cst);
    }
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}
template <typename T, int NP>
int launch_score(int n, int64_t N, int64_t rows, double alpha, double gamma, const double* P, double* S64,
                 T* ST, T* Cst, cudaStream_t st) {
    switch (NP - n) {
        case 0: return launch_score_n<T, NP, NP>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        case 1: return launch_score_n<T, NP, NP - 1>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        case 2: return launch_score_n<T, NP, NP - 2>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        case 3: return launch_score_n<T, NP, NP - 3>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        default: return launch_score_n<T, NP, -1>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
    }
}

template <typename T>
int dispatch_score(int np, int n, int64_t N, int64_t rows, double alpha, double gamma, const double* P, double* S64,
                   T* ST, T* Cst, cudaStream_t st) {
    switch (np) {
        case 4:  return launch_score<T, 4>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        case 8:  return launch_score<T, 8>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        case 12: return launch_score<T, 12>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        case 16: return launch_score<T, 16>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        case 20: return launch_score<T, 20>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        case 24: return launch_score<T, 24>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        case 28: return launch_score<T, 28>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        case 32: return launch_score<T, 32>(n, N, rows, alpha, gamma, P, S64, ST, Cst, st);
        default: tnk_set_error("tnk_score: no kernel instance for padded length %d", np); return 2;
    }
}

extern "C" int tnk_score(int n_latent, double length_scale, const double* P,
                         double* S64, int precision, void* ST, void* Cst, void* stream) {
    TnkGeom g;
    TNK_REQUIRE(tnk_geom(n_latent, precision, &g) == 0, "tnk_score: unsupported n=%d precision=%d", n_latent, precision);
    TNK_REQUIRE(P != nullptr, "tnk_score: P is NULL");
    TNK_REQUIRE((ST == nullptr) == (Cst == nullptr), "tnk_score: ST and Cst must both be given or both NULL");
    const int n = n_latent;
    const int64_t N = (int64_t)1 << n;
    double alpha = 0.0, gamma = 0.0;
    if (n > 0) {
        TNK_REQUIRE((double)n * length_scale != 0.0, "tnk_score: num_vars*length_scale must be non-zero");
        const double a = exp(-1.0 / ((double)n * length_scale));
        alpha = 1.0 - a;
        gamma = a - 1.0 / a;
    }
    const int64_t rows = ST ? g.n_pad : N;
    cudaStream_t st = (cudaStream_t)stream;
    const int rc = g.elem == 4
        ? dispatch_score<float>(g.np, n, N, rows, alpha, gamma, P, S64, (float*)ST, (float*)Cst, st)
        : dispatch_score<double>(g.np, n, N, rows, alpha, gamma, P, S64, (double*)ST, (double*)Cst, st);
    if (rc) return rc;
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}
