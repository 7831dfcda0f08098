/* tnk.h - C ABI of the B200 exact-KSD hot path (libtnk_b200.so).
 *
 * The reference (sozoluffy/TensorNetworks) is pure Python and has no FFI; the boundary it
 * offers is its module API.  Each entry point below replaces the body of one reference
 * function (or loop) on the path; the Python shim in tensornetworks_b200/ keeps the
 * reference signatures and calls these through ctypes.
 *
 * Conventions: every function returns 0 on success, non-zero on failure with the message
 * available from tnk_last_error() (thread-local).  Pointers marked "device" are CUDA
 * device pointers owned by the caller and valid on `stream`'s device; nothing is allocated
 * or freed behind the caller's back; all work is enqueued on `stream` (a cudaStream_t
 * passed as void*) and the call returns without synchronising unless stated.
 */
#ifndef TNK_H
#define TNK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* precision modes of the pair kernel */
#define TNK_PREC_FP64 0        /* DFMA pair arithmetic, fp64 accumulation (1e-6 parity mode) */
#define TNK_PREC_FP32 1        /* FFMA pair arithmetic, fp64 accumulation across tiles (1e-4 mode) */
#define TNK_PREC_FP64_SIMPLE 2 /* one-thread-per-row fp64 check kernel (debug / cross-check only) */

#define TNK_PREC_TC32 4        /* tensor-core tiles: tcgen05.mma kind::tf32 with a 3xTF32 split (fp32-grade, 1e-4 mode),
                                  Hamming weights and row sums on the ALU; 7 <= n <= 24, else the FP32 kernel runs */

#define TNK_PEAK_FP32X2 3      /* tnk_alu_peak only: packed fma.rn.f32x2 chain (2 lane-ops per instruction) */

/* node kinds of the flattened network (tnk_bn_joint) */
#define TNK_NODE_LATENT 0 /* value = bit `arg` of the packed state index            */
#define TNK_NODE_CONST 1  /* value = `arg` (clamped evidence)                       */
#define TNK_NODE_HIDDEN 2 /* value = bit `arg` of the hidden-assignment counter     */

#define TNK_MAX_NODES 64
#define TNK_MAX_LATENT 30
#define TNK_MAX_PEERS 8   /* GPUs of one NVSwitch box (tnk_ksd_rows_peers) */

#define TNK_ABI_VERSION 201   /* bumped whenever a signature below changes; tnk_version() returns the built value */

const char* tnk_last_error(void);
int tnk_version(void);

/* p(x,z) for all 2^n latent states, evidence clamped, hidden nodes summed out.
 * Replaces the 2^n calls of compute_prob_joint_xz (reference stein_utils.py:58-112), each of
 * which calls BayesianNetwork.get_joint_probability (bayesian_network.py:111-146) 2^m times.
 * Bit-exact: the product runs over nodes in bn.nodes order from 1.0 and the hidden sum over
 * generate_all_binary_outcomes(m) order from 0.0, in non-fused IEEE fp64.
 *   node_kind,node_arg [n_nodes] device int32;  pa_off [n_nodes+1], pa_idx [pa_off[n_nodes]]
 *   device int32 (parents in CPT-key order, most significant first);  cpt_off [n_nodes] device
 *   int32 (row offset of the node's table);  cpt device double [rows][2] = {P(v=0|cfg),P(v=1|cfg)}.
 *   P: device double [2^n_latent] (out). */
int tnk_bn_joint(int n_latent, int n_hidden, int n_nodes,
                 const int32_t* node_kind, const int32_t* node_arg,
                 const int32_t* pa_off, const int32_t* pa_idx,
                 const int32_t* cpt_off, const double* cpt,
                 double* P, void* stream);

/* Layout of the packed score table consumed by tnk_ksd_rows for (n, precision):
 *   *np      padded score-vector length (multiple of 4, >= n)
 *   *n_pad   number of rows of the packed table (2^n rounded up to the tile granularity)
 *   *elem    bytes per element (4 or 8) */
int tnk_ksd_layout(int n_latent, int precision, int* np, int64_t* n_pad, int* elem);

/* Discrete score function for all states from the joint table.
 * Replaces _precompute_all_s_p (reference ksd_vi.py:55-60) = 2^n calls of
 * get_score_function_sp_for_z (stein_utils.py:115-136): S[z][k] = 1 - P[z ^ (1<<(n-1-k))]/P[z],
 * the whole row zero where |P[z]| < 1e-12.  Non-fused fp64 division then subtraction (bit-exact).
 *   S64    device double [2^n][n], latent-variable order (may be NULL)
 *   ST     device [n_pad][np] of float/double per `precision`, indexed by BIT position
 *          (column p = latent variable n-1-p), zero padded (may be NULL)
 *   Cst    device [n_pad] same type: alpha*R(z) + gamma*rho0(z) (NULL iff ST NULL) */
int tnk_score(int n_latent, double length_scale, const double* P,
              double* S64, int precision, void* ST, void* Cst, void* stream);

/* Bytes of caller-provided workspace tnk_ksd_rows needs for this row range. */
size_t tnk_ksd_workspace_bytes(int n_latent, int precision, int64_t row_begin, int64_t row_end);

/* Rows [row_begin,row_end) of u = K_p q and the partial sum_{i in rows} q_i u_i.
 * Replaces the O(4^n) double loop of reference ksd_vi.py:114-131 (== ksd_vi_quantum.py:125-142),
 * i.e. 4^n calls of get_stein_kernel_kp_value (stein_utils.py:138-197), and the autograd pass
 * of ksd_vi.py:145: K = sum of the partials over all row ranges, dK/dq = 2u.
 *   q        device [2^n], q_dtype 0 = double, 1 = float
 *   ST,Cst   from tnk_score with the same (n, length_scale, precision)
 *   u        device double [2^n]; only rows [row_begin,row_end) are written
 *   ksd_part device double [1] (out)
 *   n_launch optional host int*: number of kernels enqueued (for gpu_launches accounting) */
int tnk_ksd_rows(int n_latent, double length_scale, int precision,
                 const void* q, int q_dtype, const void* ST, const void* Cst,
                 int64_t row_begin, int64_t row_end,
                 double* u, double* ksd_part,
                 void* workspace, size_t workspace_bytes, int* n_launch, void* stream);

/* tnk_ksd_rows without a caller-provided workspace and without any allocation.  It has no counterpart in the
 * reference and was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch memory of tnk_ksd_rows lives in a
 * static device array of TNK_STATIC_SCRATCH_BYTES inside the library (held by every GPU that loads it until the
 * process exits); the call enqueues the same kernels with the same arguments otherwise and calls no cudaMalloc,
 * cudaFree or allocating library routine.  Problems whose tnk_ksd_workspace_bytes exceeds the array are rejected
 * with an error naming both sizes (16 MiB holds whole evaluations up to n = 19 in TNK_PREC_FP64, n = 20 in
 * TNK_PREC_FP32 and n = 15 in TNK_PREC_TC32; larger n only for row ranges whose workspace fits).
 * All calls on one device share the array: calls that overlap on the GPU (two streams, concurrent graph replays)
 * race - serialise them.  Nothing in it carries over from one call to the next.
 *   tnk_ksd_rows_static_capacity   TNK_STATIC_SCRATCH_BYTES of the built library */
#define TNK_STATIC_SCRATCH_BYTES (16u << 20)
size_t tnk_ksd_rows_static_capacity(void);
int tnk_ksd_rows_static(int n_latent, double length_scale, int precision,
                        const void* q, int q_dtype, const void* ST, const void* Cst,
                        int64_t row_begin, int64_t row_end,
                        double* u, double* ksd_part, int* n_launch, void* stream);

/* tnk_ksd_rows with the exchange step of the row-sharded evaluation fused into its reduction
 * kernels (replaces the all_gather + all_reduce a host would otherwise issue after
 * the loop of reference ksd_vi.py:114-131 on every rank).  Rank `my_rank` of `n_peers` owns rows
 * [row_begin,row_end); the kernel that finishes a row stores it into the u buffer of EVERY rank
 * and the kernel that finishes this rank's partial of q.u stores it into slot `my_rank` of EVERY
 * rank's partial array - plain st.global on peer-mapped device addresses (NVLink P2P / symmetric
 * memory; the caller maps them, e.g. torch symmetric memory or cudaIpc).  After a cross-rank
 * barrier on the same streams each rank holds the full u and all n_peers partials, whose sum in
 * slot order is K.  No NCCL call is involved.
 *   peer_u      host array [n_peers] of device double* [2^n]   (entry my_rank = the local buffer)
 *   peer_parts  host array [n_peers] of device double* [n_peers]
 *   ksd_part    device double [1] local copy of this rank's partial (out)
 * The caller must not reuse the same sink buffers for the next evaluation until another barrier
 * has passed (double-buffer them). */
int tnk_ksd_rows_peers(int n_latent, double length_scale, int precision,
                       const void* q, int q_dtype, const void* ST, const void* Cst,
                       int64_t row_begin, int64_t row_end,
                       int n_peers, int my_rank, void* const* peer_u, void* const* peer_parts,
                       double* ksd_part,
                       void* workspace, size_t workspace_bytes, int* n_launch, void* stream);

/* The same u = K_p q and partial of q.u through the Kronecker (mode-product) factorisation of the
 * Hamming base kernel (reference stein_utils.py:30-55: k = prod_p a^(z_p xor z'_p)), O(n^2 2^n)
 * instead of O(4^n); fp64 throughout.  Replaces the same reference lines as tnk_ksd_rows
 * (ksd_vi.py:110-131,145).  Work is split by score component: the call handles bit positions
 * [bit_begin,bit_end) of [0,n) and writes that partial of u for ALL 2^n rows; partials of
 * disjoint bit ranges add up to u and K (sum-allreduce across ranks).
 *   S_cols   device double (or float, per `storage`) [2][bit_end-bit_begin][2^n]: column-major score components of the bit
 *            range, once in state order (read by the first pass) and once in the order the last
 *            pass walks (a private, plan-dependent permutation of the states);
 *   r_sum    device double [2^n] = sum over the range of (2 - s_p(z)), in the last-pass order;
 *            both written once per table by tnk_modeproduct_tables from
 *   ST64     device double [n_pad][np]: the packed table tnk_score emits for TNK_PREC_FP64
 *   u        device double [2^n] (out, partial), ksd_part device double [1] (out, partial)
 *   workspace: tnk_ksd_modeproduct_workspace_bytes(n, bit_begin, bit_end) bytes, 256-byte aligned */
#define TNK_MP_STORE_FP64 0   /* score columns and transformed vectors kept in fp64 (results to ~1e-15 of the fp64 pair kernel) */
#define TNK_MP_STORE_FP32 1   /* kept in fp32 in HBM (half the traffic); arithmetic, r_sum, u and K stay fp64 */
size_t tnk_ksd_modeproduct_workspace_bytes(int n_latent, int bit_begin, int bit_end, int storage);
/* Plan knob (tests): at most max_row_bits (1..11, default 11) high bits per strided pass, so that the
 * three-and-more-pass plans of n > 24 can be exercised at small n.  Returns the previous value; tables and
 * workspaces made under one setting must not be used under another. */
int tnk_modeproduct_plan_limit(int max_row_bits);
int tnk_modeproduct_tables(int n_latent, const double* ST64, int bit_begin, int bit_end, int storage,
                           void* S_cols, double* r_sum, void* stream);
int tnk_ksd_modeproduct(int n_latent, double length_scale, const void* q, int q_dtype, int storage,
                        const void* S_cols, const double* r_sum, int bit_begin, int bit_end,
                        double* u, double* ksd_part,
                        void* workspace, size_t workspace_bytes, int* n_launch, void* stream);

/* Classical Born machine over 2^n logits on the GPU (one unconditional distribution, fp32 like the
 * reference's parameters): q = softmax(logits) and H = -sum q log clamp(q, 1e-10) in two streaming
 * passes, and the backward pass of both in two more.  Replaces the tensor expressions of reference
 * born_machine_classical_sim.py:74-99 (get_probabilities) and :177-181 (entropy) and their autograd
 * graph.  Block partials are fp64 and folded in a fixed order.
 *   tnk_born_blocks(N)   number of block partials (entropy_partial, scratch) for N states
 *   max_logit            device float [1]: max over the logits (the shift of :97)
 *   q                    device float [N] (out);  entropy_partial device double [blocks] (out): H = their sum
 *   sum_exp              device double [1] (out): sum exp(logit - max)
 *   grad_q               device double/float [N] or NULL (grad_dtype 0/1): dL/dq
 *   grad_entropy         device double [1] or NULL: dL/dH
 *   grad_logits          device float [N] (out): q (g - sum q g), g = grad_q + grad_entropy dH/dq */
int tnk_born_blocks(int64_t num_states);
int tnk_born_forward(const float* logits, int64_t num_states, const float* max_logit,
                     float* q, double* entropy_partial, double* sum_exp, double* scratch, void* stream);
int tnk_born_backward(const float* q, const void* grad_q, int grad_dtype, const double* grad_entropy,
                      int64_t num_states, float* grad_logits, double* scratch, void* stream);

/* Introspection of the tensor-core variant (TNK_PREC_TC32): padded score length np, number of
 * tf32 chunks (4 floats) per state k32, total 16-byte chunks per state of the column table kb
 * (tf32 chunks followed by bf16 chunks), the tf32 MMA K-step list - step k multiplies row chunks
 * {s[4k], s[4k]+s[4k+1]} with column chunks {s[4k+2], s[4k+2]+s[4k+3]} - and the number of bf16
 * K-steps (K = 16; step k uses bf16 chunks 2k, 2k+1 on both sides).  steps: host int32 [4*32] or
 * NULL.  Used by the CPU tests to check that the steps cover the bracket exactly. */
int tnk_tc_plan(int n_latent, int* np, int* k32, int* kb, int* nsteps, int* nsteps16, int32_t* steps);

/* Symmetric tensor-core evaluation of the same quadratic form (reference ksd_vi.py:110-131, whose k_p is
 * symmetric, stein_utils.py:161-196): the pair space is cut into 128 x 128-state tiles and every
 * UNORDERED pair of tiles is evaluated once - its values feed the rows of one tile and, mirrored, the
 * rows of the other - so N(N+128)/2 of the 4^n ordered pairs are executed.  fp32-grade
 * arithmetic (1e-4 mode); 7 <= n <= 24.
 *   split      operand split of the tensor-core products:
 *              TNK_SYM_SPLIT_TF32   tf32 high parts + bf16 cross terms (as TNK_PREC_TC32; ~2^-19 per product)
 *              TNK_SYM_SPLIT_BF16X2 every real factor as two bf16 pieces (~2^-17 per product, 10 instead of 13
 *                                   MMA K-steps per tile at n = 20)
 *   q          device double/float [N];  ST, Cst: the fp32 tables tnk_score emits for TNK_PREC_TC32
 *   tile_offset, tile_stride   this call owns row tiles tile_offset + k*tile_stride (k >= 0): rank g of G
 *              passes (g, G); (0, 1) is the whole evaluation
 *   reuse_table  non-zero: the workspace still holds the q-independent column table an earlier call built
 *              for the same n, length_scale, split, ST, Cst (skips the packing pass)
 *   u_partial  device double [N] (out): contributions of the owned tiles to u = K_p q, rows and mirrored
 *              columns; the sum over all offsets of one stride is u
 *   ksd_part   device double [1] (out): sum_z q_z u_partial[z]; the sum over all offsets is KSD^2
 * The mirrored sums are accumulated with fp64 atomics: the last bits of u_partial depend on their order. */
#define TNK_SYM_SPLIT_TF32 0
#define TNK_SYM_SPLIT_BF16X2 1
#define TNK_SYM_SPLIT_TF32_TILE 2   /* as TF32, with the T1/T2 terms of the bit positions >= 7 (constant over a 128-state tile
                                       on one side) taken off the MMAs and folded into the row / column sums through four
                                       128-point Kronecker transforms per tile: 9 instead of 13 MMA K-steps at n = 20, 10
                                       instead of 16 at n = 24; same accuracy class */
#define TNK_SYM_SPLIT_TF32_DOT 3    /* only the score dot product s(z).s(z') stays in the pair kernel (6 K-steps for every
                                       n <= 24); C(z) + C(z') and all T1/T2 terms go through n + 2 fp64 Kronecker transforms */
size_t tnk_ksd_sym_workspace_bytes(int n_latent, int split, int tile_offset, int tile_stride);
int tnk_ksd_sym(int n_latent, double length_scale, int split, const void* q, int q_dtype,
                const void* ST, const void* Cst, int tile_offset, int tile_stride, int reuse_table,
                double* u_partial, double* ksd_part,
                void* workspace, size_t workspace_bytes, int* n_launch, void* stream);

/* Symmetric fp64 evaluation of the same quadratic form (reference ksd_vi.py:110-131; k_p symmetric, stein_utils.py:161-196):
 * the 1e-6 mode at N(N+128)/2 instead of 4^n pair evaluations.  fp64 throughout (score dot product as fp64 tensor-core
 * MMAs, everything else DFMA/DADD); the result is bit-reproducible run to run (fixed summation order, no floating-point
 * atomics between different threads).  7 <= n <= 24.
 *   ST, Cst   the fp64 tables tnk_score emits for TNK_PREC_FP64
 *   unit_offset, unit_stride   the pair space is cut into tnk_ksd_sym64_units(n) work units (pairs of tile panels); this
 *             call owns units unit_offset + k*unit_stride: rank g of G passes (g, G); (0, 1) is the whole evaluation
 *   u_partial device double [N] (out): contributions of the owned units to u = K_p q; the sum over all offsets of one
 *             stride is u.  ksd_part device double [1] (out): sum_z q_z u_partial[z]. */
#define TNK_SYM64_MIN_LATENT 7
#define TNK_SYM64_MAX_LATENT 24
size_t tnk_ksd_sym64_workspace_bytes(int n_latent, int unit_offset, int unit_stride);
int tnk_ksd_sym64_units(int n_latent);
int tnk_ksd_sym64(int n_latent, double length_scale, const void* q, int q_dtype,
                  const double* ST, const double* Cst, int unit_offset, int unit_stride,
                  double* u_partial, double* ksd_part,
                  void* workspace, size_t workspace_bytes, int* n_launch, void* stream);

/* ---- exchange step of the sharded evaluations: one process per GPU of one NVSwitch box ----
 * Replaces what a host would otherwise issue after the sharded loop of reference ksd_vi.py:114-131 on every rank:
 * the sum-all-reduce of the partial u = K_p q vectors and of the partial KSD^2 scalars.  No NCCL: every rank owns an
 * exchange buffer (send region, recv region, barrier flags) exported as a CUDA IPC handle and mapped by all peers
 * over NVLink; the all-reduce is barrier -> one reduce-scatter + all-gather kernel (slice g of every rank's send
 * region summed in rank order by rank g and stored into every rank's recv region) -> barrier, all on `stream`.
 * All ranks obtain bit-identical sums.
 *   tnk_comm_create    allocates this rank's buffer on the current device (cudaMalloc: the one allocation the library
 *                      makes) for vectors of up to max_doubles doubles and returns its IPC handle (64 bytes)
 *   tnk_comm_connect   all_handles: the world x 64 handle bytes in rank order, all-gathered by the host (any transport)
 *   tnk_comm_sendbuf   device double* [max_doubles]: the evaluation calls write their partial results here
 *                      (pass it as u_partial / ksd_part), so nothing is copied before the exchange
 *   tnk_comm_allreduce_sum   sums the first `count` doubles of all send regions; the result is in tnk_comm_recvbuf
 *                      and, if out != NULL, copied to `out` (device).  The send region may be rewritten after the call
 *                      is enqueued (stream order); the recv region is valid until the next all-reduce on this stream. */
#define TNK_COMM_HANDLE_BYTES 64
typedef struct tnk_comm tnk_comm;
int tnk_comm_create(int rank, int world, int64_t max_doubles, tnk_comm** comm, unsigned char* handle_out);
int tnk_comm_connect(tnk_comm* comm, const unsigned char* all_handles);
int tnk_comm_destroy(tnk_comm* comm);
int tnk_comm_rank(const tnk_comm* comm);
int tnk_comm_world(const tnk_comm* comm);
double* tnk_comm_sendbuf(tnk_comm* comm);
double* tnk_comm_recvbuf(tnk_comm* comm);
int tnk_comm_allreduce_sum(tnk_comm* comm, int64_t count, double* out, int* n_launch, void* stream);

/* tnk_ksd_sym on rank tnk_comm_rank(comm) of tnk_comm_world(comm) (row tiles dealt round-robin) followed by the
 * exchange: every rank receives the full u = K_p q [N] and KSD^2 [1].  The communicator must have been created with
 * max_doubles >= 2^n + 1.  workspace: tnk_ksd_sym_workspace_bytes(n, split, rank, world). */
int tnk_ksd_sym_allreduce(int n_latent, double length_scale, int split, const void* q, int q_dtype,
                          const void* ST, const void* Cst, int reuse_table, tnk_comm* comm,
                          double* u, double* ksd,
                          void* workspace, size_t workspace_bytes, int* n_launch, void* stream);

/* Measurement hook: when both are non-NULL cudaEvent_t handles, every following tnk_ksd_rows,
 * tnk_ksd_sym or tnk_ksd_sym64 call on this host thread records them on its stream immediately before and after the pair kernel
 * (K3) alone, so a benchmark can time the dominant kernel live.  Pass NULL, NULL to switch off. */
int tnk_ksd_timing_events(void* start_event, void* stop_event);

/* ALU roofline probe: runs a dependent-chain FMA kernel (precision 0: DFMA, 1: FFMA, 3: packed FFMA2) that
 * fills the chip, times it with CUDA events on `stream` (synchronises) and returns the
 * achieved lane-ops/s (one FMA on one lane = 1).  scratch: device double [>= 1]. */
int tnk_alu_peak(int precision, int iters, double* scratch, double* lane_ops_per_s, float* ms, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TNK_H */
