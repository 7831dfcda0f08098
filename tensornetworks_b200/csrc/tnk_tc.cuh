// Internal interface of the tensor-core tile variant of the pair kernel (tnk_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define TNK_TC_TILE 128        // rows per CTA (TMEM lanes) and columns per MMA / stage
#define TNK_TC_MAX_STEPS 32    // MMA K-steps per column tile
#define TNK_TC_MIN_LATENT 7    // one full 128-state tile
#define TNK_TC_MAX_LATENT 24   // shared memory: A tile + 2 B stages at np = 24 is 221 KB

// K-step list for `nc` 4-float chunks per field.  Step ks multiplies A chunks {a0, a0+alb} with
// B chunks {b0, b0+blb}.  constexpr so that the kernel (templated on nc) folds every descriptor
// offset into an immediate; the host uses the same function for table sizes and the C-ABI dump.
struct TnkTcSteps {
    int nsteps;
    bool ok;
    int a0[TNK_TC_MAX_STEPS], alb[TNK_TC_MAX_STEPS], b0[TNK_TC_MAX_STEPS], blb[TNK_TC_MAX_STEPS];
};

// perfect matching of the left-over chunk products into steps with strictly increasing offsets
__host__ __device__ constexpr bool tnk_tc_match(const int* ia, const int* ib, int count, bool* used, TnkTcSteps& p) {
    int x = -1;
    for (int k = 0; k < count; ++k)
        if (!used[k]) { x = k; break; }
    if (x < 0) return true;
    used[x] = true;
    for (int y = x + 1; y < count; ++y) {
        if (used[y]) continue;
        int lo = -1, hi = -1;
        if (ia[x] < ia[y] && ib[x] < ib[y]) { lo = x; hi = y; }
        else if (ia[y] < ia[x] && ib[y] < ib[x]) { lo = y; hi = x; }
        if (lo < 0) continue;
        used[y] = true;
        const int ks = p.nsteps++;
        p.a0[ks] = ia[lo]; p.alb[ks] = ia[hi] - ia[lo];
        p.b0[ks] = ib[lo]; p.blb[ks] = ib[hi] - ib[lo];
        if (tnk_tc_match(ia, ib, count, used, p)) return true;
        --p.nsteps;
        used[y] = false;
    }
    used[x] = false;
    return false;
}

// Operand structure (chunk = 16 bytes of one state).
//   tf32 part (4 floats per chunk), products with the high parts:
//     A32 (rows):    s_hi (nc) | bits (nc) | [1,1,0,0] | 0
//     B32 (columns): t_hi (nc) | r_hi (nc) | [(qC)_hi,(qC)_lo,0,0] | 0          -> 2*nc+2 chunks, nc+1 K-steps
//   bf16 part (8 bf16 per chunk), the second-order cross terms hi*lo + lo*hi, which tolerate 8-bit
//   mantissas (error ~2^-20 of the product):
//     A16 (rows):    bf(s) (np) | bf(s_lo) (np) | bits (np) | 0...
//     B16 (columns): bf(t_lo)   | bf(t)         | bf(r_lo)  | 0...             -> ceil(3*np/16) K-steps of K = 16
// The column table stores the B32 chunks followed by the B16 chunks of a tile.
__host__ __device__ constexpr int tnk_tc_k32(int nc) { return 2 * nc + 2; }
__host__ __device__ constexpr int tnk_tc_steps16(int nc) { return (3 * 4 * nc + 15) / 16; }
__host__ __device__ constexpr int tnk_tc_k16(int nc) { return 2 * tnk_tc_steps16(nc); }
__host__ __device__ constexpr int tnk_tc_kb(int nc) { return tnk_tc_k32(nc) + tnk_tc_k16(nc); }

// tf32 K-steps: chunk indices inside A32 / B32
__host__ __device__ constexpr TnkTcSteps tnk_tc_steps(int nc) {
    TnkTcSteps p{};
    int ra[16] = {}, rb[16] = {};
    int nrest = 0;
    for (int f = 0; f < 2; ++f) {            // s_hi x t_hi, bits x r_hi: same chunk index on both sides
        int c = 0;
        for (; c + 1 < nc; c += 2) {
            const int ks = p.nsteps++;
            p.a0[ks] = f * nc + c; p.alb[ks] = 1;
            p.b0[ks] = f * nc + c; p.blb[ks] = 1;
        }
        if (c < nc) { ra[nrest] = f * nc + c; rb[nrest] = f * nc + c; ++nrest; }
    }
    ra[nrest] = 2 * nc; rb[nrest] = 2 * nc; ++nrest;                      // [1,1,0,0] x [(qC)_hi,(qC)_lo,0,0]
    if (nrest & 1) { ra[nrest] = 2 * nc + 1; rb[nrest] = 2 * nc + 1; ++nrest; }   // zero x zero
    bool used[16] = {};
    p.ok = tnk_tc_match(ra, rb, nrest, used, p) && p.nsteps + tnk_tc_steps16(nc) <= TNK_TC_MAX_STEPS;
    return p;
}

struct TnkTcPlan {
    int np, nc;      // padded score length, 4-float chunks per field
    int k32, kb;     // tf32 chunks per state; total chunks per state of the column table (tf32 + bf16)
    int steps16;     // bf16 K-steps (K = 16), chunks 2k, 2k+1 of the bf16 region on both sides
    TnkTcSteps s;    // tf32 K-steps
};

int tnk_tc_make_plan(int n, TnkTcPlan* p);
size_t tnk_tc_table_bytes(const TnkTcPlan& p, int64_t states);
size_t tnk_tc_qtab_bytes(int64_t states);
int tnk_tc_launch(int n, double a, double gamma, const double* wl, const float* ST, const float* Cst, const float* qT,
                  int64_t row0, int64_t launch_rows, int splits, float* B_tab, float* Qtab, double* part,
                  cudaEvent_t ev_start, cudaEvent_t ev_stop, int* launches, cudaStream_t st);

// symmetric tile variant (tnk_ksd_sym): column-table bytes and launcher
size_t tnk_tcs_table_bytes(int n, int fmt, int64_t states);
int tnk_tcs_launch(int n, int fmt, double gamma, const double* wl, const float* ST, const float* Cst, const float* qT,
                   int tile_offset, int tile_stride, int row_tiles, int splits, float* B_tab, int reuse_table,
                   double* ucol, double* hibits_ws, cudaEvent_t ev_start, cudaEvent_t ev_stop, int* launches,
                   cudaStream_t st);

// Formats 2, 3 of the symmetric variant: u[z] = the terms of ALL pairs (z, j) that left the pair kernel - the T1/T2 terms of
// the bit positions p >= p_lo, g_p(z) K(b_p .* q)_z + b_p(z) K(g_p .* q)_z, and (with_c) C(z) K(q)_z + K(C .* q)_z -
// through 1 + n - p_lo (+1) fp64 Kronecker transforms (tnk_modeprod.cu).
// The pass is linear in its terms, so a strided evaluation splits it: the call with share (offset, stride) takes the bit
// positions p_lo + offset + k * stride (and, with_c, the C terms - the caller gives them to offset 0); every call needs
// w = K q itself.  offset >= stride or no bit left: the call only zeroes u.
// ws: tnk_hibits_ws_bytes bytes.  Overwrites u (the tile kernel then adds the remaining terms).
int tnk_hibits_count(int n, int p_lo, int share_offset, int share_stride);
size_t tnk_hibits_ws_bytes(int n, int p_lo, int with_c, int share_offset, int share_stride);
int tnk_hibits_pass(int n, int p_lo, int with_c, int share_offset, int share_stride, double a, double gamma, const void* ST,
                    const void* Cst, const void* qT, int elem_bytes, double* u, double* ws, int* launches, cudaStream_t st);
