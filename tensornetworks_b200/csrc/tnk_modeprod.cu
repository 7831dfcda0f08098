// Mode-product (Kronecker) evaluation of the same quadratic form: u = K_p q in O(n^2 2^n).
//
// The Hamming base kernel factorises over bits, K = M^{(x)n} with M = [[1,a],[a,1]]
// (reference stein_utils.py:30-55: k = exp(-h/(n l)) = prod_p a^{z_p xor z'_p}), and the
// single-bit-flip differences of Eq. 13 (stein_utils.py:166-194) are (I - X_p) K, X_p = flip
// of bit p, which commutes with K.  With W_p = K (q .* s_p), w = K q, r_p = 1 - s_p:
//
//   u[z] = sum_p [ -r_p(z) W_p[z] + W_p[z^e_p] + (r_p(z) + 1) (w[z] - w[z^e_p]) ]
//
// Because X_p K = K X_p every flipped read can be moved in front of the transform or behind it.
// With y_p = q.*s_p + X_p q, W'_p = K y_p and Rs = n + sum_p r_p the sum regroups into
//
//   u = - sum_p r_p .* W'_p  +  w .* Rs  -  sum_p X_p K (q .* r_p)
//
// and the last term is taken per bit on whichever side is local:
//   "pre"  bits (below the last pass): folded into extra input vectors Y = -sum_pre X_p (q.*r_p),
//          u += K Y;
//   "post" bits (rows of the last pass tile): K(q.*r_p) = w - W'_p + X_p w, so
//          u[z] += W'_p[z^e_p] - w[z^e_p] - w[z], both flipped elements being in the tile that is
//          in shared memory at that moment.
//
// So the evaluation is n_loc + 1 + (#Y chunks) transforms by K and no gather pass: the first
// pass builds its input on the fly from q and the column-major score table (never materialised),
// the last pass folds every transformed vector into u while its tile is still in shared memory.
// HBM traffic per evaluation (fp64, 2 passes): read S twice, write + read the V vectors once.
// All arithmetic is fp64.
//
// Kernels:  mp_tables (one-off: column-major copy of the score table + row sums of r),
// mp_first (build input tile, butterflies on the low bits, write V), mp_transform (middle passes,
// only for n > 24), mp_last (butterflies on the high bits, fold into u, block partials of q.u),
// mp_scalar (fixed-order final sum).
#include "tnk_common.cuh"
#include "tnk_tc.cuh"

namespace {

constexpr int kMpThreads = 512;
constexpr int kTileLog = 13;               // elements per shared-memory tile (2^13 doubles)
constexpr int kMinLogW = 2;                // strided passes keep >= 4 contiguous doubles (one 32 B sector) per row
constexpr int kMpPerThread = (1 << kTileLog) / kMpThreads;
constexpr size_t kMpSmem = (size_t)(1 << kTileLog) * sizeof(double);

// Shared-memory swizzle of a tile index (8-byte elements, 16 per bank row): XOR-linear, so
// swz(a | b) = swz(a) ^ swz(b) for disjoint bit fields.  Bits 4..6 are folded into bits 0..3 so that
// the 16 lanes of a half-warp, which differ in the four lowest index bits outside the radix field
// of the round ({0,1,2,3}, {0,1,2,6}, {0,1,5,6}, {0,4,5,6} or {3,4,5,6}), always hit 16 different
// bank pairs - as do plain consecutive accesses.
__device__ __forceinline__ int swz(int idx) { return idx ^ (((idx >> 4) & 7) << 1) ^ ((idx >> 6) & 1); }

// In-register butterflies of 2^RB values: x[j] <-> x[j ^ (1<<b)] for every bit b of j.
template <int RB>
__device__ __forceinline__ void reg_butterflies(double (&x)[1 << RB], double a) {
#pragma unroll
    for (int b = 0; b < RB; ++b) {
#pragma unroll
        for (int j = 0; j < (1 << RB); ++j) {
            if (!(j & (1 << b))) {
                const double lo = x[j], hi = x[j | (1 << b)];
                x[j] = fma(a, hi, lo);
                x[j | (1 << b)] = fma(a, lo, hi);
            }
        }
    }
}

// One round: butterflies of tile index bits [sh, sh + RB) of a 2^total_log element tile held in
// swizzled shared memory, radix 2^RB in registers.  Ends with __syncthreads().
template <int RB>
__device__ __forceinline__ void smem_round(double* tile, int sh, int total_log, double a) {
    constexpr int CNT = 1 << RB;
    int cj[CNT];
#pragma unroll
    for (int j = 0; j < CNT; ++j) cj[j] = swz(j << sh);
    const int units = 1 << (total_log - RB);
    const int lowmask = (1 << sh) - 1;
    for (int uidx = threadIdx.x; uidx < units; uidx += kMpThreads) {
        const int pb = swz(((uidx & ~lowmask) << RB) | (uidx & lowmask));
        double x[CNT];
#pragma unroll
        for (int j = 0; j < CNT; ++j) x[j] = tile[pb ^ cj[j]];
        reg_butterflies<RB>(x, a);
#pragma unroll
        for (int j = 0; j < CNT; ++j) tile[pb ^ cj[j]] = x[j];
    }
    __syncthreads();
}

// Butterflies of tile index bits [sh0, sh0 + nbits), in rounds of at most 3 bits.
__device__ __forceinline__ void tile_butterflies(double* tile, int sh0, int nbits, int total_log, double a) {
    int k0 = 0;
    while (nbits - k0 >= 3 && nbits - k0 != 4) { smem_round<3>(tile, sh0 + k0, total_log, a); k0 += 3; }
    while (nbits - k0 >= 2) { smem_round<2>(tile, sh0 + k0, total_log, a); k0 += 2; }
    if (nbits - k0 == 1) smem_round<1>(tile, sh0 + k0, total_log, a);
}

// Address of state z in the layout the last pass reads.  Two-pass plans (13 < n <= 24) store V, the
// second score copy and r_sum "blocked": [z bits logW..12][z bits 13..n-1][z bits 0..logW-1], so that a
// last-pass tile (2^g rows x 2^logW states) is one contiguous 64 KB run instead of 2^g runs of
// 2^logW doubles 64 KB apart.  Other plans keep the natural order.
__device__ __forceinline__ int64_t last_layout(int64_t z, int blocked, int logW) {
    if (!blocked) return z;
    const int64_t zl = z & ((1 << kTileLog) - 1), zh = z >> kTileLog;
    return ((zl >> logW) << kTileLog) | (zh << logW) | (zl & ((1 << logW) - 1));
}

// ---- MP0 (one-off per table): S[lv][z] = s_p(z) for p = p0 + lv, the same in the last-pass layout
// (S2), and Rs[last_layout(z)] = n_loc + sum_lv (1 - s_p(z))
template <typename T>
__global__ void __launch_bounds__(256)
mp_tables_kernel(int np, int64_t N, int p0, int n_loc, int blocked, int logW, const double* __restrict__ ST,
                 T* __restrict__ S, T* __restrict__ S2, double* __restrict__ Rs) {
    const int64_t z = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (z >= N) return;
    const int64_t zb = last_layout(z, blocked, logW);
    const double* s = ST + z * np + p0;
    double rs = (double)n_loc;
    for (int lv = 0; lv < n_loc; ++lv) {
        const double v = s[lv];
        S[(int64_t)lv * N + z] = (T)v;
        S2[(int64_t)lv * N + zb] = (T)v;
        rs += 1.0 - (double)(T)v;            // the sum of the r_p the kernels will see
    }
    Rs[zb] = rs;
}

// ---- MP1: first pass.  blockIdx.x = tile of 2^T contiguous states, blockIdx.y = input vector:
//   y <  nchunk            Y_c[z] = -sum_{lv in chunk c, pre} q[z^e_p] r_p(z^e_p)      -> V slot n_loc+1+c
//   y <  nchunk + n_loc    y_p[z] = q[z] s_p(z) + q[z^e_p],  p = p0 + (y - nchunk)      -> V slot lv
//   y == nchunk + n_loc    q itself                                                    -> V slot n_loc
// then the butterflies of bits [0,T) and the store.  (The heavy Y tiles are scheduled first.)
template <typename Q, typename TS>
__global__ void __launch_bounds__(kMpThreads, 2)
mp_first_kernel(int T, int64_t N, int p0, int n_loc, int pre_cnt, int chunk, int nchunk, int blocked, int logW, int group_log,
                const Q* __restrict__ q, const TS* __restrict__ S, TS* __restrict__ V, double a) {
    extern __shared__ double tile[];
    const int tile_elems = 1 << T;
    // block order: [tile group][vector][tile in group] - neighbouring tiles of one vector run together
    // (their blocked stores fill whole lines in L2) and all vectors of a tile group reuse its q from L2
    const int nvec = n_loc + 1 + nchunk;
    const int64_t bid = blockIdx.x;
    const int64_t tg = bid / ((int64_t)nvec << group_log);
    const int rest = (int)(bid - tg * ((int64_t)nvec << group_log));
    const int y = rest >> group_log;
    const int64_t tile_id = (tg << group_log) | (rest & ((1 << group_log) - 1));
    const int64_t z0 = tile_id << T;
    const Q* qz = q + z0;
    int slot, lv0 = 0, lv1 = 0;
    if (y < nchunk) {
        slot = n_loc + 1 + y;
        lv0 = y * chunk;
        lv1 = (lv0 + chunk < pre_cnt) ? lv0 + chunk : pre_cnt;
    } else {
        slot = y - nchunk;           // < n_loc: component lv = slot;  == n_loc: q itself
    }
    // input element e of this tile: flipped partners inside the tile are addressed by e ^ fl,
    // partners in another tile by the same e from a shifted base
    auto partner = [&](int p, int64_t& zb, int& fl) {
        if (p < T) { zb = z0; fl = 1 << p; } else { zb = z0 ^ ((int64_t)1 << p); fl = 0; }
    };

    if (T == kTileLog) {
        // bits 10..12 in registers straight from global memory (8 loads 8 KB apart per thread)
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            const int w = threadIdx.x + h * kMpThreads;
            double x[8];
            if (y < nchunk) {
#pragma unroll
                for (int j = 0; j < 8; ++j) x[j] = 0.0;
                for (int lv = lv0; lv < lv1; ++lv) {
                    int64_t zb; int fl;
                    partner(p0 + lv, zb, fl);
                    const Q* qf = q + zb;
                    const TS* Sf = S + (int64_t)lv * N + zb;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int ef = (w + (j << 10)) ^ fl;
                        x[j] = fma(-(double)qf[ef], 1.0 - (double)Sf[ef], x[j]);
                    }
                }
            } else if (slot < n_loc) {
                int64_t zb; int fl;
                partner(p0 + slot, zb, fl);
                const Q* qf = q + zb;
                const TS* Sp = S + (int64_t)slot * N + z0;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int e = w + (j << 10);
                    x[j] = fma((double)qz[e], (double)Sp[e], (double)qf[e ^ fl]);
                }
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) x[j] = (double)qz[w + (j << 10)];
            }
            reg_butterflies<3>(x, a);
            const int pw = swz(w);
#pragma unroll
            for (int j = 0; j < 8; ++j) tile[pw + (j << 10)] = x[j];
        }
        __syncthreads();
        smem_round<3>(tile, 0, kTileLog, a);
        smem_round<3>(tile, 3, kTileLog, a);
        // bits 6..9 radix-16, straight to global memory (a warp stores 32 consecutive states per value)
        const int base = ((threadIdx.x >> 6) << 10) | (threadIdx.x & 63);
        const int pb = swz(base);
        double x[16];
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) x[jj] = tile[pb ^ ((jj << 6) ^ ((jj & 1) ? 9 : 0))];
        reg_butterflies<4>(x, a);
        TS* out = V + (int64_t)slot * N;
        if (blocked) {
            out += tile_id << logW;
            const int W = 1 << logW;
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) {
                const int e = base | (jj << 6);
                out[((int64_t)(e >> logW) << kTileLog) | (e & (W - 1))] = (TS)x[jj];
            }
        } else {
            out += z0;
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) out[base | (jj << 6)] = (TS)x[jj];
        }
        return;
    }

    // n < 13: a single small tile, everything through shared memory
    for (int e = threadIdx.x; e < tile_elems; e += kMpThreads) {
        double val;
        if (y < nchunk) {
            val = 0.0;
            for (int lv = lv0; lv < lv1; ++lv) {
                int64_t zb; int fl;
                partner(p0 + lv, zb, fl);
                val = fma(-(double)q[zb + (e ^ fl)], 1.0 - (double)S[(int64_t)lv * N + zb + (e ^ fl)], val);
            }
        } else if (slot < n_loc) {
            int64_t zb; int fl;
            partner(p0 + slot, zb, fl);
            val = fma((double)qz[e], (double)S[(int64_t)slot * N + z0 + e], (double)q[zb + (e ^ fl)]);
        } else {
            val = (double)qz[e];
        }
        tile[swz(e)] = val;
    }
    __syncthreads();
    tile_butterflies(tile, 0, T, T, a);
    TS* out = V + (int64_t)slot * N + z0;          // never blocked: one pass only
    for (int e = threadIdx.x; e < tile_elems; e += kMpThreads) out[e] = (TS)tile[swz(e)];
}

// ---- MP2: middle pass (n > kTileLog + 11 only): butterflies for bits [b0, b0+g) on tiles of
// 2^g rows x 2^logW contiguous elements, in place.
template <typename TS>
__global__ void __launch_bounds__(kMpThreads, 2)
mp_transform_kernel(TS* __restrict__ V, int64_t N, int b0, int g, int logW, double a) {
    extern __shared__ double tile[];
    const int tile_elems = 1 << (g + logW);
    const int W = 1 << logW;
    const int64_t nlob = (int64_t)1 << (b0 - logW);
    const int64_t t = blockIdx.x;
    const int64_t lob = t % nlob, hi = t / nlob;
    TS* vec = V + (int64_t)blockIdx.y * N + ((hi << (b0 + g)) | (lob << logW));
    for (int e = threadIdx.x; e < tile_elems; e += kMpThreads) {
        const int m = e >> logW, w = e & (W - 1);
        tile[swz(e)] = (double)vec[((int64_t)m << b0) + w];
    }
    __syncthreads();
    tile_butterflies(tile, logW, g, g + logW, a);
    for (int e = threadIdx.x; e < tile_elems; e += kMpThreads) {
        const int m = e >> logW, w = e & (W - 1);
        vec[((int64_t)m << b0) + w] = (TS)tile[swz(e)];
    }
}

// ---- MP3: last pass.  One CTA per tile of 2^g rows (bits [b0,b0+g)) x 2^logW contiguous states;
// it walks over all V vectors, finishes their butterflies in shared memory and folds each into
// the u values its threads own (element e = tid + 512 k).  g == 0: nothing left to transform
// (n <= kTileLog), the tile is a contiguous run and every bit is "pre".
// FAST: the two-pass case with a full tile (blocked layout, g >= 3), specialised at compile time.
template <typename Q, typename TS, bool FAST>
__global__ void __launch_bounds__(kMpThreads, 2)
mp_last_kernel(int64_t N, int b0, int g, int logW, int p0, int n_loc, int nchunk, int flip_split, int blocked_arg,
               const Q* __restrict__ q, const TS* __restrict__ S, const double* __restrict__ Rs,
               const TS* __restrict__ V, double* __restrict__ u, double* __restrict__ blockpart, double a,
               double* __restrict__ upart) {
    // gridDim.y > 1: the vectors are split over gridDim.y CTAs per tile (few tiles: small n); each writes its
    // partial of u to upart[blockIdx.y][.] and mp_usum_kernel adds them in a fixed order.
    extern __shared__ double tile[];
    __shared__ double red[kMpThreads / 32];
    const bool blocked = FAST || blocked_arg;
    const int tile_elems = FAST ? (1 << kTileLog) : (1 << (g + logW));
    const int W = 1 << logW;
    const int64_t nlob = (int64_t)1 << (b0 - logW);
    const int64_t t = blockIdx.x;
    const int64_t base = ((t / nlob) << (b0 + g)) | ((t % nlob) << logW);
    auto off = [&](int e) -> int64_t { return ((int64_t)(e >> logW) << b0) + (e & (W - 1)); };   // natural order
    // V, S and Rs in the last-pass layout: a contiguous tile when blocked
    const int64_t lbase = blocked ? (t << kTileLog) : base;
    auto loff = [&](int e) -> int64_t { return blocked ? (int64_t)e : off(e); };
    auto live = [&](int e) -> bool { return FAST || e < tile_elems; };
    const int post0 = (flip_split > p0) ? flip_split - p0 : 0;       // local components >= post0 are "post"
    const int n_post = (n_loc > post0) ? n_loc - post0 : 0;

    double acc[kMpPerThread];
#pragma unroll
    for (int k = 0; k < kMpPerThread; ++k) acc[k] = 0.0;

    const bool fused = FAST || ((g + logW == kTileLog) && g >= 3);   // full tile: top three row bits go through registers
    const int64_t jstride = blocked ? 1024 : ((int64_t)(1024 >> logW) << b0);   // +1024 tile elements in V
    const int ptid = swz(threadIdx.x);                     // swz(tid + 512 k) = ptid + 512 k
    const int nvec = n_loc + 1 + nchunk;
    const int v_begin = (int)((int64_t)nvec * blockIdx.y / gridDim.y), v_end = (int)((int64_t)nvec * (blockIdx.y + 1) / gridDim.y);
    for (int v = v_begin; v < v_end; ++v) {
        const TS* vec = V + (int64_t)v * N + lbase;
        if (fused) {
#pragma unroll 1
            for (int h = 0; h < 2; ++h) {
                const int w = threadIdx.x + h * kMpThreads;
                const TS* src = vec + loff(w);
                double x[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) x[j] = (double)src[j * jstride];
                reg_butterflies<3>(x, a);
                const int pw = swz(w);
#pragma unroll
                for (int j = 0; j < 8; ++j) tile[pw + (j << 10)] = x[j];
            }
            __syncthreads();
            tile_butterflies(tile, logW, g - 3, kTileLog, a);
        } else {
#pragma unroll
            for (int k = 0; k < kMpPerThread; ++k) {
                const int e = threadIdx.x + k * kMpThreads;
                if (live(e)) tile[ptid + k * kMpThreads] = (double)vec[loff(e)];
            }
            __syncthreads();
            tile_butterflies(tile, logW, g, g + logW, a);
        }
        if (v < n_loc) {
            const TS* Sp = S + (int64_t)v * N + lbase;
            const int fb = (v >= post0) ? swz(1 << (p0 + v - b0 + logW)) : 0;
#pragma unroll
            for (int k = 0; k < kMpPerThread; ++k) {
                const int e = threadIdx.x + k * kMpThreads;
                if (live(e)) {
                    const int pe = ptid + k * kMpThreads;
                    const double r = 1.0 - (double)Sp[loff(e)];
                    double x = fma(-r, tile[pe], acc[k]);
                    if (fb) x += tile[pe ^ fb];
                    acc[k] = x;
                }
            }
        } else if (v == n_loc) {
#pragma unroll
            for (int k = 0; k < kMpPerThread; ++k) {
                const int e = threadIdx.x + k * kMpThreads;
                if (live(e)) {
                    const int pe = ptid + k * kMpThreads;
                    double x = tile[pe] * (Rs[lbase + loff(e)] - (double)n_post);
                    for (int lv = post0; lv < n_loc; ++lv) x -= tile[pe ^ swz(1 << (p0 + lv - b0 + logW))];
                    acc[k] += x;
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < kMpPerThread; ++k) {
                const int e = threadIdx.x + k * kMpThreads;
                if (live(e)) acc[k] += tile[ptid + k * kMpThreads];
            }
        }
        __syncthreads();
    }

    if (gridDim.y > 1) {
        double* out = upart + (int64_t)blockIdx.y * N;
#pragma unroll
        for (int k = 0; k < kMpPerThread; ++k) {
            const int e = threadIdx.x + k * kMpThreads;
            if (live(e)) out[base + off(e)] = acc[k];
        }
        return;
    }
    double contrib = 0.0;
#pragma unroll
    for (int k = 0; k < kMpPerThread; ++k) {
        const int e = threadIdx.x + k * kMpThreads;
        if (live(e)) {
            const int64_t z = base + off(e);
            u[z] = acc[k];
            contrib = fma((double)q[z], acc[k], contrib);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = contrib;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
#pragma unroll
        for (int w2 = 0; w2 < kMpThreads / 32; ++w2) s += red[w2];
        blockpart[blockIdx.x] = s;
    }
}

// ---- MP3b (only when the vectors were split): u = sum_g upart[g] in group order, block partials of q.u
template <typename Q>
__global__ void __launch_bounds__(256)
mp_usum_kernel(int64_t N, int groups, const Q* __restrict__ q, const double* __restrict__ upart,
               double* __restrict__ u, double* __restrict__ blockpart) {
    __shared__ double red[8];
    const int64_t z = (int64_t)blockIdx.x * 256 + threadIdx.x;
    double contrib = 0.0;
    if (z < N) {
        double acc = 0.0;
        for (int gq = 0; gq < groups; ++gq) acc += upart[(int64_t)gq * N + z];
        u[z] = acc;
        contrib = (double)q[z] * acc;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) contrib += __shfl_xor_sync(0xffffffffu, contrib, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = contrib;
    __syncthreads();
    if (threadIdx.x == 0) {
        double sum = 0.0;
#pragma unroll
        for (int w2 = 0; w2 < 8; ++w2) sum += red[w2];
        blockpart[blockIdx.x] = sum;
    }
}

__global__ void __launch_bounds__(256)
mp_scalar_kernel(const double* __restrict__ blockpart, int64_t nblocks, double* __restrict__ out) {
    __shared__ double red[256];
    double acc = 0.0;
    for (int64_t b = threadIdx.x; b < nblocks; b += 256) acc += blockpart[b];
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = red[0];
}

// Row bits one strided pass may take (2^g rows x 2^(13-g) states per tile).  The default lets two passes
// reach n = 24; tnk_modeproduct_plan_limit lowers it so that tests reach the multi-pass plans at small n.
int g_max_row_bits = kTileLog - kMinLogW;

struct MpPass { int b0, g, logW; };

struct MpPlan {
    int64_t N, tiles_first, tiles_last;
    int n_loc, T, n_mid, pre_cnt, chunk, nchunk, nvec, flip_split, blocked, group_log, vgroups;
    int64_t part_blocks;
    MpPass mid[4], last;
    size_t off_v, off_block, off_upart, total;
};

int mp_plan(int n, int p_begin, int p_end, int storage, MpPlan* p) {
    if (n < 0 || n > TNK_MAX_LATENT || p_begin < 0 || p_end > n || p_begin > p_end) return 1;
    if (storage != TNK_MP_STORE_FP64 && storage != TNK_MP_STORE_FP32) return 1;
    const size_t elem = storage == TNK_MP_STORE_FP32 ? sizeof(float) : sizeof(double);
    p->N = (int64_t)1 << n;
    p->n_loc = p_end - p_begin;
    p->T = n < kTileLog ? n : kTileLog;
    p->tiles_first = p->N >> p->T;
    p->n_mid = 0;
    const int rem = n - p->T;
    if (rem == 0) {
        p->last = MpPass{n, 0, n};
        p->flip_split = n;
    } else {
        const int gmax = g_max_row_bits;
        const int passes = (rem + gmax - 1) / gmax;
        if (passes - 1 > 4) return 1;
        int b0 = p->T, left = rem;
        for (int ps = 0; ps < passes; ++ps) {
            const int g = (left + (passes - ps) - 1) / (passes - ps);
            int logW = kTileLog - g;
            if (logW > b0) logW = b0;
            const MpPass pass{b0, g, logW};
            if (ps + 1 < passes) p->mid[p->n_mid++] = pass; else p->last = pass;
            b0 += g;
            left -= g;
        }
        p->flip_split = p->last.b0;
    }
    p->tiles_last = p->N >> (p->last.g + p->last.logW);
    p->blocked = (rem > 0 && p->n_mid == 0) ? 1 : 0;        // two passes: V and the second score copy are blocked
    p->group_log = 0;
    while (p->group_log < 3 && ((int64_t)2 << p->group_log) <= p->tiles_first) ++p->group_log;
    int pre = p->flip_split - p_begin;
    if (pre < 0) pre = 0;
    if (pre > p->n_loc) pre = p->n_loc;
    p->pre_cnt = pre;
    p->chunk = (p->tiles_first >= 1024 || pre <= 4) ? (pre > 0 ? pre : 1) : 4;   // few tiles: spread the Y work over more CTAs
    p->nchunk = (pre + p->chunk - 1) / p->chunk;
    p->nvec = p->n_loc + 1 + p->nchunk;
    size_t off = 0;
    p->off_v = off;     off += ((size_t)p->nvec * p->N * elem + 255) / 256 * 256;
    // few last-pass tiles (small n): split the vectors of a tile over several CTAs to fill the SMs
    p->vgroups = 1;
    while (p->tiles_last * p->vgroups < 2 * 148 && p->vgroups * 2 <= p->nvec && p->vgroups < 8) p->vgroups *= 2;
    p->part_blocks = p->vgroups > 1 ? (p->N + 255) / 256 : p->tiles_last;
    p->off_block = off; off += ((size_t)(p->part_blocks + 1) * sizeof(double) + 255) / 256 * 256;
    p->off_upart = off; off += p->vgroups > 1 ? ((size_t)p->vgroups * p->N * sizeof(double) + 255) / 256 * 256 : 0;
    p->total = off;
    return 0;
}

// launches of one evaluation for q type Q and storage type TS
template <typename Q, typename TS>
int mp_run(const MpPlan& p, int bit_begin, double a, const Q* q, const TS* S_cols, const double* r_sum, TS* V,
           double* u, double* blockpart, double* upart, double* ksd_part, int* launches, cudaStream_t st) {
    TNK_CUDA_OK(cudaFuncSetAttribute(mp_first_kernel<Q, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMpSmem));
    TNK_CUDA_OK(cudaFuncSetAttribute(mp_last_kernel<Q, TS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMpSmem));
    TNK_CUDA_OK(cudaFuncSetAttribute(mp_last_kernel<Q, TS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMpSmem));
    TNK_CUDA_OK(cudaFuncSetAttribute(mp_transform_kernel<TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMpSmem));
    mp_first_kernel<Q, TS><<<(unsigned)(p.tiles_first * p.nvec), kMpThreads, kMpSmem, st>>>(
        p.T, p.N, bit_begin, p.n_loc, p.pre_cnt, p.chunk, p.nchunk, p.blocked, p.last.logW, p.group_log, q, S_cols, V, a);
    TNK_CUDA_OK(cudaGetLastError());
    ++*launches;
    for (int m = 0; m < p.n_mid; ++m) {
        const MpPass& ps = p.mid[m];
        dim3 grid((unsigned)(p.N >> (ps.g + ps.logW)), (unsigned)p.nvec, 1);
        mp_transform_kernel<TS><<<grid, kMpThreads, kMpSmem, st>>>(V, p.N, ps.b0, ps.g, ps.logW, a);
        TNK_CUDA_OK(cudaGetLastError());
        ++*launches;
    }
    const MpPass& ps = p.last;
    const TS* S2 = S_cols + (size_t)p.n_loc * p.N;       // the copy in the last-pass layout
    const dim3 grid((unsigned)p.tiles_last, (unsigned)p.vgroups, 1);
    if (p.blocked && ps.g >= 3)
        mp_last_kernel<Q, TS, true><<<grid, kMpThreads, kMpSmem, st>>>(p.N, ps.b0, ps.g, ps.logW, bit_begin, p.n_loc, p.nchunk,
                                                                     p.flip_split, p.blocked, q, S2, r_sum, V, u, blockpart, a, upart);
    else
        mp_last_kernel<Q, TS, false><<<grid, kMpThreads, kMpSmem, st>>>(p.N, ps.b0, ps.g, ps.logW, bit_begin, p.n_loc, p.nchunk,
                                                                      p.flip_split, p.blocked, q, S2, r_sum, V, u, blockpart, a, upart);
    TNK_CUDA_OK(cudaGetLastError());
    if (p.vgroups > 1) {
        mp_usum_kernel<Q><<<(unsigned)p.part_blocks, 256, 0, st>>>(p.N, p.vgroups, q, upart, u, blockpart);
        TNK_CUDA_OK(cudaGetLastError());
        ++*launches;
    }
    mp_scalar_kernel<<<1, 256, 0, st>>>(blockpart, p.part_blocks, ksd_part);
    TNK_CUDA_OK(cudaGetLastError());
    *launches += 2;
    return 0;
}

}  // namespace

// ------------------------------------------------------------------------------------
// High-bit pass of the symmetric tile kernel's format 2 (tnk_tc.cu): for the bit positions p >= 7
//   u[z] = sum_p [ g_p(z) K(b_p .* q)_z + b_p(z) K(g_p .* q)_z ],   g_p = gamma (1 - s_p)(1 - 2 b_p),  K = M^{(x)n}.
// K(b_p .* q) needs no transform of its own: with w = K q and q'_1 = K_{-p} q restricted to b_p = 1,
// w(b_p=0) = q'_0 + a q'_1, w(b_p=1) = a q'_0 + q'_1, so K(b_p .* q)_z = (b_p(z) ? 1 : a) (w[z|p] - a w[z&~p]) / (1 - a^2).
// ------------------------------------------------------------------------------------
namespace {

// V[0] = q, V[1 + p - p_lo] = g_p .* q for p = p_lo..n-1, and (with_c) one more vector C .* q  (fp64 [nvec][N]);
// g and C from the packed tables in their own type (fp32: exactly as the tile kernel forms them; fp64 for tnk_ksd_sym64)
template <typename TS>
__global__ void __launch_bounds__(256)
hib_build_kernel(int n, int np, int p0, int p_stride, int cnt, int with_c, int64_t N, TS gamma, const TS* __restrict__ ST,
                 const TS* __restrict__ Cst, const TS* __restrict__ qT, double* __restrict__ V) {
    const int64_t z = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (z >= N) return;
    const double q = (double)qT[z];
    V[z] = q;
    for (int k = 0; k < cnt; ++k) {                      // the owned bit positions p0, p0 + stride, ...
        const int p = p0 + k * p_stride;
        const TS v = gamma * ((TS)1 - ST[z * np + p]);
        V[(int64_t)(1 + k) * N + z] = (double)(((z >> p) & 1) ? -v : v) * q;
    }
    if (with_c) V[(int64_t)(1 + cnt) * N + z] = (double)Cst[z] * q;
}

template <typename TS>
__global__ void __launch_bounds__(256)
hib_fold_kernel(int n, int np, int p0, int p_stride, int cnt, int with_c, int64_t N, TS gamma, double a, const TS* __restrict__ ST,
                const TS* __restrict__ Cst, const double* __restrict__ V, double* __restrict__ u) {
    const int64_t z = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (z >= N) return;
    const double inv = 1.0 / (1.0 - a * a);
    double acc = 0.0;
    for (int k = 0; k < cnt; ++k) {
        const int p = p0 + k * p_stride;
        const int64_t bitp = (int64_t)1 << p;
        const bool on = (z & bitp) != 0;
        const TS v = gamma * ((TS)1 - ST[z * np + p]);
        const double g = (double)(on ? -v : v);
        const double A = (on ? 1.0 : a) * (V[z | bitp] - a * V[z & ~bitp]) * inv;      // K(b_p .* q)_z from w = V[0]
        acc += g * A;
        if (on) acc += V[(int64_t)(1 + k) * N + z];                                   // b_p(z) K(g_p .* q)_z
    }
    if (with_c) acc += (double)Cst[z] * V[z] + V[(int64_t)(1 + cnt) * N + z];          // sum_j a^h (C_z + C_j) q_j
    u[z] = acc;
}

}  // namespace

// bit positions p_lo + share_offset + k * share_stride < n: the share of one call (rank) of a strided evaluation
int tnk_hibits_count(int n, int p_lo, int share_offset, int share_stride) {
    const int first = p_lo + share_offset;
    return (share_offset < share_stride && first < n) ? (n - first + share_stride - 1) / share_stride : 0;
}

size_t tnk_hibits_ws_bytes(int n, int p_lo, int with_c, int share_offset, int share_stride) {
    const int cnt = tnk_hibits_count(n, p_lo, share_offset, share_stride);
    if (cnt == 0 && !with_c) return 0;
    return (size_t)(1 + cnt + (with_c ? 1 : 0)) * ((size_t)1 << n) * sizeof(double);
}

int tnk_hibits_pass(int n, int p_lo, int with_c, int share_offset, int share_stride, double a, double gamma, const void* ST,
                    const void* Cst, const void* qT, int elem_bytes, double* u, double* ws, int* launches, cudaStream_t st) {
    const int64_t N = (int64_t)1 << n;
    const int np = ((n + 3) / 4) * 4;
    TNK_REQUIRE(n >= 7 && n <= 24 && p_lo >= 0 && p_lo <= n && share_stride >= 1 && share_offset >= 0,
                "tnk_hibits_pass: bad arguments n=%d p_lo=%d share=(%d,%d)", n, p_lo, share_offset, share_stride);
    const int cnt = tnk_hibits_count(n, p_lo, share_offset, share_stride);
    const int p0 = p_lo + share_offset;
    if (cnt == 0 && !with_c) {           // nothing of the pass falls to this call: its partial starts at zero
        TNK_CUDA_OK(cudaMemsetAsync(u, 0, (size_t)N * sizeof(double), st));
        return 0;
    }
    TNK_REQUIRE(ws, "tnk_hibits_pass: NULL workspace");
    const int nvec = 1 + cnt + (with_c ? 1 : 0);
    const unsigned blocks = (unsigned)((N + 255) / 256);
    if (elem_bytes == 4)
        hib_build_kernel<float><<<blocks, 256, 0, st>>>(n, np, p0, share_stride, cnt, with_c, N, (float)gamma, (const float*)ST,
                                                        (const float*)Cst, (const float*)qT, ws);
    else
        hib_build_kernel<double><<<blocks, 256, 0, st>>>(n, np, p0, share_stride, cnt, with_c, N, gamma, (const double*)ST,
                                                         (const double*)Cst, (const double*)qT, ws);
    TNK_CUDA_OK(cudaGetLastError());
    TNK_CUDA_OK(cudaFuncSetAttribute(mp_transform_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMpSmem));
    // all n bits of every vector: bits [0, min(n,13)) on contiguous tiles, the rest on 2^g x 2^(13-g) tiles
    const int low = n < kTileLog ? n : kTileLog;
    {
        dim3 grid((unsigned)(N >> low), (unsigned)nvec, 1);
        mp_transform_kernel<double><<<grid, kMpThreads, kMpSmem, st>>>(ws, N, 0, low, 0, a);
        TNK_CUDA_OK(cudaGetLastError());
    }
    int nl = 2;
    if (n > kTileLog) {
        const int g = n - kTileLog;
        dim3 grid((unsigned)(N >> kTileLog), (unsigned)nvec, 1);
        mp_transform_kernel<double><<<grid, kMpThreads, kMpSmem, st>>>(ws, N, kTileLog, g, kTileLog - g, a);
        TNK_CUDA_OK(cudaGetLastError());
        ++nl;
    }
    if (elem_bytes == 4)
        hib_fold_kernel<float><<<blocks, 256, 0, st>>>(n, np, p0, share_stride, cnt, with_c, N, (float)gamma, a, (const float*)ST, (const float*)Cst, ws, u);
    else
        hib_fold_kernel<double><<<blocks, 256, 0, st>>>(n, np, p0, share_stride, cnt, with_c, N, gamma, a, (const double*)ST, (const double*)Cst, ws, u);
    TNK_CUDA_OK(cudaGetLastError());
    if (launches) *launches += nl + 1;
    return 0;
}

extern "C" int tnk_modeproduct_plan_limit(int max_row_bits) {
    const int prev = g_max_row_bits;
    if (max_row_bits >= 1 && max_row_bits <= kTileLog - kMinLogW) g_max_row_bits = max_row_bits;
    return prev;
}

extern "C" size_t tnk_ksd_modeproduct_workspace_bytes(int n_latent, int bit_begin, int bit_end, int storage) {
    MpPlan p;
    if (mp_plan(n_latent, bit_begin, bit_end, storage, &p)) return 0;
    return p.total;
}

extern "C" int tnk_modeproduct_tables(int n_latent, const double* ST64, int bit_begin, int bit_end, int storage,
                                      void* S_cols, double* r_sum, void* stream) {
    MpPlan p;
    TNK_REQUIRE(mp_plan(n_latent, bit_begin, bit_end, storage, &p) == 0,
                "tnk_modeproduct_tables: bad arguments n=%d bits=[%d,%d) storage=%d", n_latent, bit_begin, bit_end, storage);
    TNK_REQUIRE(ST64 && S_cols && r_sum, "tnk_modeproduct_tables: NULL pointer argument");
    TnkGeom geo;
    TNK_REQUIRE(tnk_geom(n_latent, TNK_PREC_FP64, &geo) == 0, "tnk_modeproduct_tables: unsupported n=%d", n_latent);
    const unsigned blocks = (unsigned)((p.N + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (storage == TNK_MP_STORE_FP32) {
        float* S = (float*)S_cols;
        mp_tables_kernel<float><<<blocks, 256, 0, st>>>(geo.np, p.N, bit_begin, p.n_loc, p.blocked, p.last.logW, ST64,
                                                        S, S + (size_t)p.n_loc * p.N, r_sum);
    } else {
        double* S = (double*)S_cols;
        mp_tables_kernel<double><<<blocks, 256, 0, st>>>(geo.np, p.N, bit_begin, p.n_loc, p.blocked, p.last.logW, ST64,
                                                         S, S + (size_t)p.n_loc * p.N, r_sum);
    }
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int tnk_ksd_modeproduct(int n_latent, double length_scale, const void* q, int q_dtype, int storage,
                                   const void* S_cols, const double* r_sum, int bit_begin, int bit_end,
                                   double* u, double* ksd_part,
                                   void* workspace, size_t workspace_bytes, int* n_launch, void* stream) {
    MpPlan p;
    TNK_REQUIRE(mp_plan(n_latent, bit_begin, bit_end, storage, &p) == 0,
                "tnk_ksd_modeproduct: bad arguments n=%d bits=[%d,%d) storage=%d", n_latent, bit_begin, bit_end, storage);
    TNK_REQUIRE(q && S_cols && r_sum && u && ksd_part && workspace, "tnk_ksd_modeproduct: NULL pointer argument");
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
    const double a = exp(-1.0 / ((double)n * length_scale));
    unsigned char* ws = (unsigned char*)workspace;
    double* blockpart = (double*)(ws + p.off_block);
    double* upart = (double*)(ws + p.off_upart);
    int rc;
    if (storage == TNK_MP_STORE_FP32) {
        float* V = (float*)(ws + p.off_v);
        rc = q_dtype == 0 ? mp_run<double, float>(p, bit_begin, a, (const double*)q, (const float*)S_cols, r_sum, V, u, blockpart, upart, ksd_part, &launches, st)
                          : mp_run<float, float>(p, bit_begin, a, (const float*)q, (const float*)S_cols, r_sum, V, u, blockpart, upart, ksd_part, &launches, st);
    } else {
        double* V = (double*)(ws + p.off_v);
        rc = q_dtype == 0 ? mp_run<double, double>(p, bit_begin, a, (const double*)q, (const double*)S_cols, r_sum, V, u, blockpart, upart, ksd_part, &launches, st)
                          : mp_run<float, double>(p, bit_begin, a, (const float*)q, (const double*)S_cols, r_sum, V, u, blockpart, upart, ksd_part, &launches, st);
    }
    if (rc) return rc;
    if (n_launch) *n_launch = launches;
    return 0;
}
