// Shared helpers for the libtnk_b200 translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include "tnk.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libtnk_b200 is written for sm_100a (B200); build with -gencode arch=compute_100a,code=sm_100a"
#endif

void tnk_set_error(const char* fmt, ...);
// the event pair of tnk_ksd_timing_events for the calling thread (both NULL when switched off)
void tnk_timing_events_get(cudaEvent_t* start, cudaEvent_t* stop);

#define TNK_CUDA_OK(call)                                                              \
    do {                                                                               \
        cudaError_t e__ = (call);                                                      \
        if (e__ != cudaSuccess) {                                                      \
            tnk_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),     \
                          __FILE__, __LINE__);                                         \
            return 1;                                                                  \
        }                                                                              \
    } while (0)

#define TNK_REQUIRE(cond, ...)                                                         \
    do {                                                                               \
        if (!(cond)) {                                                                 \
            tnk_set_error(__VA_ARGS__);                                                \
            return 2;                                                                  \
        }                                                                              \
    } while (0)

// ---- geometry of the pair kernel, shared by tnk_score (packing) and tnk_ksd_rows ----
// One CTA = 256 threads, each owning R consecutive rows; a column tile has TC columns.
struct TnkGeom {
    int np;        // padded score length (multiple of 4)
    int R;         // rows per thread
    int TC;        // columns per staged tile
    int elem;      // bytes per element
    int64_t n_pad; // rows in the packed table
};

static inline int tnk_geom(int n, int precision, TnkGeom* g) {
    if (n < 0 || n > TNK_MAX_LATENT) return 1;
    int np = ((n + 3) / 4) * 4;
    if (np < 4) np = 4;
    if (np > 32) return 1;
    g->np = np;
    if (precision == TNK_PREC_FP32 || precision == TNK_PREC_TC32) {
        g->elem = 4; g->R = (np <= 24) ? 4 : 2; g->TC = 256;
    } else if (precision == TNK_PREC_FP64 || precision == TNK_PREC_FP64_SIMPLE) {
        g->elem = 8; g->R = (np <= 24) ? 2 : 1; g->TC = 128;
    } else {
        return 1;
    }
    int64_t N = (int64_t)1 << n;
    int64_t gran = 256 * 4;  // >= any row block (256*R) and any TC; keeps both loops unmasked
    g->n_pad = ((N + gran - 1) / gran) * gran;
    return 0;
}
