// Classical Born machine over 2^n logits: q = softmax(logits), H(q) = -sum q log clamp(q, 1e-10),
// and their backward pass, as HBM-bound streaming kernels.
//
// Replaces, for one unconditional distribution on the GPU, the tensor expressions of reference
// born_machine_classical_sim.py:74-99 (get_probabilities: softmax(raw - max)) and :177-181 (entropy):
// torch.softmax runs one thread block per row, i.e. one SM for the single row of 2^n logits, and the
// autograd graph of the two expressions is ~25 elementwise passes over 2^n floats per epoch.  Here
// forward = 2 passes after the max (sum of exp; probabilities + entropy partials), backward = 2 passes
// (dot of q with the incoming gradient; logits gradient).  fp32 storage like the reference's
// parameters, fp64 block partials summed in a fixed order (deterministic).
#include "tnk_common.cuh"

namespace {

constexpr int kBornThreads = 256;
constexpr int kBornUnroll = 8;
constexpr int kBornMaxBlocks = 148 * 8;
constexpr float kClampMin = 1e-10f;        // born_machine_classical_sim.py:180

int born_blocks(int64_t N) {
    int64_t b = (N + (int64_t)kBornThreads * kBornUnroll - 1) / ((int64_t)kBornThreads * kBornUnroll);
    if (b < 1) b = 1;
    if (b > kBornMaxBlocks) b = kBornMaxBlocks;
    return (int)b;
}

// fixed-order block sum: warp shuffles, then lane 0 of warp 0 adds the 8 warp totals
__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = 0.0;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int w = 0; w < kBornThreads / 32; ++w) s += red[w];
    }
    return s;      // valid in thread 0
}

// every block folds the previous kernel's block partials itself (<= 1184 doubles, L2-resident)
__device__ __forceinline__ double fold_partials(const double* __restrict__ partial, int count, double* red) {
    double v = 0.0;
    for (int i = threadIdx.x; i < count; i += kBornThreads) v += partial[i];
    const double s = block_sum(v, red);
    __shared__ double total;
    if (threadIdx.x == 0) total = s;
    __syncthreads();
    return total;
}

// ---- B1: partial[b] = sum over the block's elements of exp(l - max)
__global__ void __launch_bounds__(kBornThreads)
born_sumexp_kernel(const float* __restrict__ logits, int64_t N, const float* __restrict__ max_logit,
                   double* __restrict__ partial) {
    __shared__ double red[kBornThreads / 32];
    const float m = max_logit[0];
    const int64_t stride = (int64_t)gridDim.x * kBornThreads;
    double acc = 0.0;
    for (int64_t i0 = (int64_t)blockIdx.x * kBornThreads + threadIdx.x; i0 < N; i0 += stride * kBornUnroll) {
        float v[kBornUnroll];
#pragma unroll
        for (int k = 0; k < kBornUnroll; ++k) {
            const int64_t i = i0 + k * stride;
            v[k] = (i < N) ? expf(logits[i] - m) : 0.f;
        }
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < kBornUnroll; ++k) s += v[k];
        acc += (double)s;
    }
    const double tot = block_sum(acc, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

// ---- B2: q = exp(l - max) / Z,  h_partial[b] = -sum q log max(q, 1e-10)
__global__ void __launch_bounds__(kBornThreads)
born_probs_kernel(const float* __restrict__ logits, int64_t N, const float* __restrict__ max_logit,
                  const double* __restrict__ z_partial, int z_count, float* __restrict__ q,
                  double* __restrict__ h_partial, double* __restrict__ z_out) {
    __shared__ double red[kBornThreads / 32];
    const double Z = fold_partials(z_partial, z_count, red);
    if (blockIdx.x == 0 && threadIdx.x == 0) z_out[0] = Z;
    const float m = max_logit[0];
    const float zf = (float)Z;
    const int64_t stride = (int64_t)gridDim.x * kBornThreads;
    double acc = 0.0;
    for (int64_t i0 = (int64_t)blockIdx.x * kBornThreads + threadIdx.x; i0 < N; i0 += stride * kBornUnroll) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < kBornUnroll; ++k) {
            const int64_t i = i0 + k * stride;
            if (i < N) {
                const float p = expf(logits[i] - m) / zf;
                q[i] = p;
                s -= p * logf(fmaxf(p, kClampMin));
            }
        }
        acc += (double)s;
    }
    __syncthreads();
    const double tot = block_sum(acc, red);
    if (threadIdx.x == 0) h_partial[blockIdx.x] = tot;
}

// incoming gradient of element i: g_q[i] (may be absent) plus g_H * dH/dq_i,
// dH/dq = -(log max(q,1e-10) + [q >= 1e-10])  (clamp passes the gradient where q >= min)
template <typename G>
__device__ __forceinline__ float total_grad(const G* __restrict__ g_q, float gH, float p, int64_t i) {
    float g = g_q ? (float)g_q[i] : 0.f;
    if (gH != 0.f) g -= gH * (logf(fmaxf(p, kClampMin)) + (p >= kClampMin ? 1.f : 0.f));
    return g;
}

// ---- B3: partial[b] = sum q_i g_i
template <typename G>
__global__ void __launch_bounds__(kBornThreads)
born_dot_kernel(const float* __restrict__ q, const G* __restrict__ g_q, const double* __restrict__ g_H,
                int64_t N, double* __restrict__ partial) {
    __shared__ double red[kBornThreads / 32];
    const float gH = g_H ? (float)g_H[0] : 0.f;
    const int64_t stride = (int64_t)gridDim.x * kBornThreads;
    double acc = 0.0;
    for (int64_t i0 = (int64_t)blockIdx.x * kBornThreads + threadIdx.x; i0 < N; i0 += stride * kBornUnroll) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < kBornUnroll; ++k) {
            const int64_t i = i0 + k * stride;
            if (i < N) {
                const float p = q[i];
                s = fmaf(p, total_grad(g_q, gH, p, i), s);
            }
        }
        acc += (double)s;
    }
    const double tot = block_sum(acc, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

// ---- B4: g_logits = q (g - sum_j q_j g_j)      (softmax backward; the max shift has zero net gradient)
template <typename G>
__global__ void __launch_bounds__(kBornThreads)
born_backward_kernel(const float* __restrict__ q, const G* __restrict__ g_q, const double* __restrict__ g_H,
                     const double* __restrict__ dot_partial, int dot_count, int64_t N, float* __restrict__ g_logits) {
    __shared__ double red[kBornThreads / 32];
    const float dot = (float)fold_partials(dot_partial, dot_count, red);
    const float gH = g_H ? (float)g_H[0] : 0.f;
    const int64_t stride = (int64_t)gridDim.x * kBornThreads;
    for (int64_t i0 = (int64_t)blockIdx.x * kBornThreads + threadIdx.x; i0 < N; i0 += stride * kBornUnroll) {
#pragma unroll
        for (int k = 0; k < kBornUnroll; ++k) {
            const int64_t i = i0 + k * stride;
            if (i < N) {
                const float p = q[i];
                g_logits[i] = p * (total_grad(g_q, gH, p, i) - dot);
            }
        }
    }
}

}  // namespace

extern "C" int tnk_born_blocks(int64_t num_states) { return num_states > 0 ? born_blocks(num_states) : 0; }

extern "C" int tnk_born_forward(const float* logits, int64_t num_states, const float* max_logit,
                                float* q, double* entropy_partial, double* sum_exp, double* scratch, void* stream) {
    TNK_REQUIRE(num_states > 0, "tnk_born_forward: num_states must be positive");
    TNK_REQUIRE(logits && max_logit && q && entropy_partial && sum_exp && scratch, "tnk_born_forward: NULL pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int blocks = born_blocks(num_states);
    born_sumexp_kernel<<<blocks, kBornThreads, 0, st>>>(logits, num_states, max_logit, scratch);
    TNK_CUDA_OK(cudaGetLastError());
    born_probs_kernel<<<blocks, kBornThreads, 0, st>>>(logits, num_states, max_logit, scratch, blocks, q, entropy_partial, sum_exp);
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int tnk_born_backward(const float* q, const void* grad_q, int grad_dtype, const double* grad_entropy,
                                 int64_t num_states, float* grad_logits, double* scratch, void* stream) {
    TNK_REQUIRE(num_states > 0, "tnk_born_backward: num_states must be positive");
    TNK_REQUIRE(q && grad_logits && scratch, "tnk_born_backward: NULL pointer argument");
    TNK_REQUIRE(grad_dtype == 0 || grad_dtype == 1, "tnk_born_backward: grad_dtype must be 0 (double) or 1 (float)");
    cudaStream_t st = (cudaStream_t)stream;
    const int blocks = born_blocks(num_states);
    if (grad_dtype == 0) {
        born_dot_kernel<double><<<blocks, kBornThreads, 0, st>>>(q, (const double*)grad_q, grad_entropy, num_states, scratch);
        TNK_CUDA_OK(cudaGetLastError());
        born_backward_kernel<double><<<blocks, kBornThreads, 0, st>>>(q, (const double*)grad_q, grad_entropy, scratch, blocks,
                                                                     num_states, grad_logits);
    } else {
        born_dot_kernel<float><<<blocks, kBornThreads, 0, st>>>(q, (const float*)grad_q, grad_entropy, num_states, scratch);
        TNK_CUDA_OK(cudaGetLastError());
        born_backward_kernel<float><<<blocks, kBornThreads, 0, st>>>(q, (const float*)grad_q, grad_entropy, scratch, blocks,
                                                                    num_states, grad_logits);
    }
    TNK_CUDA_OK(cudaGetLastError());
    return 0;
}
