// Internal interface of the tensor-core tile variant of the pair kernel (tnk_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define TNK_TC_TILE 128        // rows per CTA (TMEM lanes) and columns per MMA / stage
#define TNK_TC_MAX_STEPS 32    // MMA K-steps per column tile
#define TNK_TC_MIN_LATENT 7    // one full 128-state tile
#define TNK_TC_MAX_LATENT 24   // shared memory: A tile + 2 B stages at np = 24 is 221 KB

struct TnkTcPlan {
    int np, nc;      // padded score length, 4-float chunks per field
    int ka, kb;      // chunks per state in the A (row) and B (column) tables
    int nsteps;      // MMA K-steps; step ks multiplies A chunks {a0, a0+alb} with B chunks {b0, b0+blb}
    uint16_t a0[TNK_TC_MAX_STEPS], alb[TNK_TC_MAX_STEPS], b0[TNK_TC_MAX_STEPS], blb[TNK_TC_MAX_STEPS];
};

int tnk_tc_make_plan(int n, TnkTcPlan* p);
size_t tnk_tc_table_bytes(const TnkTcPlan& p, int64_t states, bool is_b);
int tnk_tc_launch(int n, double a, double gamma, const double* wl, const float* ST, const float* Cst, const float* qT,
                  int64_t row0, int64_t launch_rows, int splits, float* A_tab, float* B_tab, double* part,
                  cudaEvent_t ev_start, cudaEvent_t ev_stop, int* launches, cudaStream_t st);
