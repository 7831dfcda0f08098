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

// Table widths in chunks.  A (rows): s_hi | s_lo | bits | [1,1,0,0] (| zero);
// B (columns): t_hi | t_lo | r_hi | r_lo | [(qC)_hi,(qC)_lo,0,0] (| zero).  The zero chunks exist only
// when the number of chunk products 5*nc+1 is odd (nc even) and pad the last K-step.
__host__ __device__ constexpr int tnk_tc_ka(int nc) { return 3 * nc + 1 + (((5 * nc + 1) & 1) ? 1 : 0); }
__host__ __device__ constexpr int tnk_tc_kb(int nc) { return 4 * nc + 1 + (((5 * nc + 1) & 1) ? 1 : 0); }

__host__ __device__ constexpr TnkTcSteps tnk_tc_steps(int nc) {
    TnkTcSteps p{};
    // field pairs: A {s_hi s_hi s_lo bits bits} x B {t_hi t_lo t_hi r_hi r_lo}
    const int FA[5] = {0, 0, 1, 2, 2};
    const int FB[5] = {0, 1, 0, 2, 3};
    int ra[16] = {}, rb[16] = {};
    int nrest = 0;
    for (int f = 0; f < 5; ++f) {
        int c = 0;
        for (; c + 1 < nc; c += 2) {
            const int ks = p.nsteps++;
            p.a0[ks] = FA[f] * nc + c; p.alb[ks] = 1;
            p.b0[ks] = FB[f] * nc + c; p.blb[ks] = 1;
        }
        if (c < nc) { ra[nrest] = FA[f] * nc + c; rb[nrest] = FB[f] * nc + c; ++nrest; }
    }
    ra[nrest] = 3 * nc; rb[nrest] = 4 * nc; ++nrest;                   // [1,1,0,0] x [(qC)_hi,(qC)_lo,0,0]
    if (nrest & 1) { ra[nrest] = 3 * nc + 1; rb[nrest] = 4 * nc + 1; ++nrest; }   // zero x zero
    bool used[16] = {};
    p.ok = tnk_tc_match(ra, rb, nrest, used, p) && p.nsteps <= TNK_TC_MAX_STEPS;
    return p;
}

struct TnkTcPlan {
    int np, nc;      // padded score length, 4-float chunks per field
    int ka, kb;      // chunks per state in the A (row) and B (column) tables
    TnkTcSteps s;
};

int tnk_tc_make_plan(int n, TnkTcPlan* p);
size_t tnk_tc_table_bytes(const TnkTcPlan& p, int64_t states, bool is_b);
size_t tnk_tc_qtab_bytes(int64_t states);
int tnk_tc_launch(int n, double a, double gamma, const double* wl, const float* ST, const float* Cst, const float* qT,
                  int64_t row0, int64_t launch_rows, int splits, float* A_tab, float* B_tab, float* Qtab, double* part,
                  cudaEvent_t ev_start, cudaEvent_t ev_stop, int* launches, cudaStream_t st);
