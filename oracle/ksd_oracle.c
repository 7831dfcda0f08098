/* CPU oracle, C restatement of the reference's exact-KSD path.  TEST INFRASTRUCTURE ONLY:
 * used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
 * never by the product path.  Pinned to the reference through tests/test_oracle_golden.py
 * (same fixtures as oracle/ksd_oracle.py; the joint and score tables must be bit-identical).
 *
 * Each function cites the reference lines it follows (paths relative to the reference repo).
 * Build: make -C oracle   (gcc -O2 -pthread -ffp-contract=off; no fast-math, so IEEE order is kept)
 */
#include <math.h>
#include <stdint.h>
#include <pthread.h>
#include <stdlib.h>
#include <unistd.h>

#define NODE_LATENT 0
#define NODE_CONST 1
#define NODE_HIDDEN 2

int orc_max_threads(void) {
    long c = sysconf(_SC_NPROCESSORS_ONLN);
    return c > 0 ? (int)c : 1;
}

/* compute_prob_joint_xz for every latent state (stein_utils.py:58-112): evidence clamped,
 * latent bits from the state index (MSB-first, utils.py:77-91), hidden nodes summed from 0.0 in
 * generate_all_binary_outcomes order; each term is get_joint_probability's sequential product
 * from 1.0 in bn.nodes order (bayesian_network.py:129-144). */
void orc_joint_table(int n_latent, int n_hidden, int n_nodes, const int32_t* kind, const int32_t* arg,
                     const int32_t* pa_off, const int32_t* pa_idx, const int32_t* cpt_off,
                     const double* cpt, double* P) {
    const int64_t N = (int64_t)1 << n_latent;
    for (int64_t idx = 0; idx < N; ++idx) {
        double total = 0.0;
        for (uint32_t h = 0; h < (1u << n_hidden); ++h) {
            uint64_t assign = 0;
            for (int v = 0; v < n_nodes; ++v) {
                uint64_t bit;
                if (kind[v] == NODE_LATENT) bit = (uint64_t)(idx >> arg[v]) & 1u;
                else if (kind[v] == NODE_CONST) bit = (uint64_t)(arg[v] & 1);
                else bit = (h >> arg[v]) & 1u;
                assign |= bit << v;
            }
            double prob = 1.0;
            for (int v = 0; v < n_nodes; ++v) {
                int cfg = 0;
                for (int t = pa_off[v]; t < pa_off[v + 1]; ++t) cfg = (cfg << 1) | (int)((assign >> pa_idx[t]) & 1u);
                prob *= cpt[2 * ((int64_t)cpt_off[v] + cfg) + (int)((assign >> v) & 1u)];
            }
            if (n_hidden == 0) total = prob; else total += prob;
        }
        P[idx] = total;
    }
}

/* get_score_function_sp_for_z for every state (stein_utils.py:115-136); S is [N][n] in latent
 * order; flip_bit(z,k) (stein_utils.py:7-11) is idx ^ (1 << (n-1-k)). */
void orc_score_table(int n, const double* P, double* S) {
    const int64_t N = (int64_t)1 << n;
    for (int64_t i = 0; i < N; ++i) {
        const double p = P[i];
        for (int k = 0; k < n; ++k) {
            if (fabs(p) < 1e-12) S[i * n + k] = 0.0;                     /* :126-128 */
            else S[i * n + k] = 1.0 - (P[i ^ ((int64_t)1 << (n - 1 - k))] / p); /* :134 */
        }
    }
}

static inline double base_kernel(const double* lut, int64_t z1, int64_t z2) {
    /* base_hamming_kernel_torch (stein_utils.py:30-55): exp(-hamming/(n*l)); the n+1 possible
     * values are tabulated once by the caller, each as exp(-h/(n*l)). */
    return lut[__builtin_popcountll((unsigned long long)(z1 ^ z2))];
}

/* get_stein_kernel_kp_value (stein_utils.py:138-197): the four Eq.-13 terms with their
 * 5n+1 base-kernel evaluations, in the reference's order. */
double orc_stein_kernel_value(int n, const double* lut, int64_t z1, int64_t z2,
                              const double* sp1, const double* sp2) {
    if (n == 0) return 0.0;                                              /* :151-152 */
    const double k12 = base_kernel(lut, z1, z2);
    double dot = 0.0;
    for (int i = 0; i < n; ++i) dot += sp1[i] * sp2[i];
    const double term1 = dot * k12;                                      /* :161-164 */
    double t2 = 0.0, t3 = 0.0, t4 = 0.0;
    for (int j = 0; j < n; ++j) {                                        /* :166-171 */
        const int64_t f = (int64_t)1 << (n - 1 - j);
        t2 += sp1[j] * (k12 - base_kernel(lut, z1, z2 ^ f));
    }
    for (int i = 0; i < n; ++i) {                                        /* :173-178 */
        const int64_t f = (int64_t)1 << (n - 1 - i);
        t3 += (k12 - base_kernel(lut, z1 ^ f, z2)) * sp2[i];
    }
    for (int i = 0; i < n; ++i) {                                        /* :180-194 */
        const int64_t f = (int64_t)1 << (n - 1 - i);
        t4 += k12 - base_kernel(lut, z1, z2 ^ f) - base_kernel(lut, z1 ^ f, z2) + base_kernel(lut, z1 ^ f, z2 ^ f);
    }
    return term1 - t2 - t3 + t4;
}

/* Rows [r0,r1) of the double loop of ksd_vi.py:114-131: u_i = sum_j k_p(z_i,z_j) q_j (so that
 * K = sum_i q_i u_i and the autograd gradient of ksd_vi.py:145 is 2u).  Returns the partial K.
 * `nthreads` pthreads take interleaved rows (the reference itself is single-threaded Python). */
typedef struct {
    int n, tid, nthreads;
    const double *lut, *S, *q;
    int64_t r0, r1;
    double* u;
} row_job;

static void* row_worker(void* arg) {
    row_job* job = (row_job*)arg;
    const int n = job->n;
    const int64_t N = (int64_t)1 << n;
    for (int64_t i = job->r0 + job->tid; i < job->r1; i += job->nthreads) {
        double acc = 0.0;
        for (int64_t j = 0; j < N; ++j)
            acc += job->q[j] * orc_stein_kernel_value(n, job->lut, i, j, job->S + i * n, job->S + j * n);
        job->u[i - job->r0] = acc;
    }
    return NULL;
}

double orc_ksd_rows(int n, double length_scale, const double* S, const double* q,
                    int64_t r0, int64_t r1, double* u, int nthreads) {
    double lut[64];
    for (int h = 0; h <= n; ++h) lut[h] = (n > 0) ? exp(-(double)h / ((double)n * length_scale)) : 1.0;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    row_job jobs[256];
    for (int t = 0; t < nthreads; ++t) {
        row_job jb = {n, t, nthreads, lut, S, q, r0, r1, u};
        jobs[t] = jb;
        if (t > 0) pthread_create(&th[t], NULL, row_worker, &jobs[t]);
    }
    row_worker(&jobs[0]);
    for (int t = 1; t < nthreads; ++t) pthread_join(th[t], NULL);
    double total = 0.0;
    for (int64_t i = r0; i < r1; ++i) total += q[i] * u[i - r0];
    return total;
}

/* ------------------------------------------------------------------------------------------------
 * O2 in C: the same K = q^T K_p q and u = K_p q through the Kronecker factorisation of the Hamming base
 * kernel (stein_utils.py:49-55: exp(-h/(n l)) = prod_p a^(z_p xor z'_p), i.e. K = M^{(x)n}, M = [[1,a],[a,1]]).
 * Restates oracle/ksd_oracle.py:ksd_modeproduct (which is validated against the term-by-term O1 form and the
 * reference goldens) at the sizes numpy cannot reach in test time (n = 20, 24).  With K^(i) = K with axis i's
 * factor replaced by (1-a)[[1,-1],[-1,1]] = (I - X_i) K (X_i = flip of latent variable i), w = K q and
 * W_i = K (q .* s_i):
 *     u = sum_i [ s_i .* W_i  +  r_i .* (w - X_i w)  +  (I - X_i)(w - W_i) ],      r_i = 1 - s_i,
 * the three sums of the numpy form term for term.  Scores come straight from the joint table by the rule of
 * stein_utils.py:115-136 (zero row under the 1e-12 threshold), so no [N][n] table is held (n = 24: 3.2 GB).
 * Pinned in tests/test_oracle_golden.py against the numpy O2/O1 forms and against orc_ksd_rows rows. */
typedef struct {
    int n, tid, nthreads, phase, var;
    double a;
    const double *P, *q, *w, *Wi;
    double *v, *u;
} mp_job;

static inline double mp_score(const double* P, int64_t z, int64_t flip) {
    const double p = P[z];
    if (fabs(p) < 1e-12) return 0.0;                                     /* stein_utils.py:126-128 */
    return 1.0 - (P[z ^ flip] / p);                                      /* :134 */
}

#define MP_LB 12          /* low bits transformed inside one contiguous 32 KB block */
#define MP_CW 8           /* doubles per strided column chunk (one cache line) */

static void mp_butterflies(double* x, int64_t len, double a) {
    /* in-place (x)_bits [[1,a],[a,1]] over `len` (a power of two) contiguous elements */
    for (int64_t half = 1; half < len; half <<= 1)
        for (int64_t base = 0; base < len; base += 2 * half)
            for (int64_t k = base; k < base + half; ++k) {
                const double lo = x[k], hi = x[k + half];
                x[k] = lo + a * hi;
                x[k + half] = a * lo + hi;
            }
}

static void* mp_worker(void* arg) {
    mp_job* jb = (mp_job*)arg;
    const int n = jb->n;
    const int64_t N = (int64_t)1 << n;
    const int lb = n < MP_LB ? n : MP_LB;
    const int64_t L = (int64_t)1 << lb, H = N >> lb;
    if (jb->phase == 0) {               /* v = q .* s_var (var >= 0) or q (var < 0) */
        const int64_t flip = jb->var >= 0 ? (int64_t)1 << (n - 1 - jb->var) : 0;
        for (int64_t z = jb->tid; z < N; z += jb->nthreads)
            jb->v[z] = jb->var >= 0 ? jb->q[z] * mp_score(jb->P, z, flip) : jb->q[z];
    } else if (jb->phase == 1) {        /* low bits: contiguous blocks */
        for (int64_t b = jb->tid; b < H; b += jb->nthreads) mp_butterflies(jb->v + b * L, L, jb->a);
    } else if (jb->phase == 2) {        /* high bits: column chunks of MP_CW, gathered into a local buffer */
        if (H == 1) return NULL;
        const int64_t cw = L < MP_CW ? L : MP_CW;
        double* buf = (double*)malloc((size_t)H * cw * sizeof(double));
        for (int64_t c = jb->tid * cw; c < L; c += jb->nthreads * cw) {
            for (int64_t h = 0; h < H; ++h)
                for (int64_t e = 0; e < cw; ++e) buf[h * cw + e] = jb->v[h * L + c + e];
            for (int64_t half = 1; half < H; half <<= 1)
                for (int64_t base = 0; base < H; base += 2 * half)
                    for (int64_t k = base; k < base + half; ++k)
                        for (int64_t e = 0; e < cw; ++e) {
                            const double lo = buf[k * cw + e], hi = buf[(k + half) * cw + e];
                            buf[k * cw + e] = lo + jb->a * hi;
                            buf[(k + half) * cw + e] = jb->a * lo + hi;
                        }
            for (int64_t h = 0; h < H; ++h)
                for (int64_t e = 0; e < cw; ++e) jb->v[h * L + c + e] = buf[h * cw + e];
        }
        free(buf);
    } else {                            /* phase 3: fold variable `var` into u */
        const int64_t flip = (int64_t)1 << (n - 1 - jb->var);
        const double *w = jb->w, *W = jb->Wi;
        for (int64_t z = jb->tid; z < N; z += jb->nthreads) {
            const double s = mp_score(jb->P, z, flip), r = 1.0 - s;
            const int64_t zf = z ^ flip;
            jb->u[z] += s * W[z] + r * (w[z] - w[zf]) + ((w[z] - W[z]) - (w[zf] - W[zf]));
        }
    }
    return NULL;
}

static void mp_run(mp_job* proto, int phase, int nthreads) {
    pthread_t th[256];
    mp_job jobs[256];
    for (int t = 0; t < nthreads; ++t) {
        jobs[t] = *proto;
        jobs[t].tid = t; jobs[t].nthreads = nthreads; jobs[t].phase = phase;
        if (t > 0) pthread_create(&th[t], NULL, mp_worker, &jobs[t]);
    }
    mp_worker(&jobs[0]);
    for (int t = 1; t < nthreads; ++t) pthread_join(th[t], NULL);
}

/* u[N] = K_p q (the gradient of ksd_vi.py:145 is 2u) and the return value K = q.u, from the joint table P. */
double orc_ksd_modeproduct(int n, double length_scale, const double* P, const double* q, double* u, int nthreads) {
    const int64_t N = (int64_t)1 << n;
    for (int64_t z = 0; z < N; ++z) u[z] = 0.0;
    if (n == 0) return 0.0;                                              /* stein_utils.py:151-152 */
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    double* w = (double*)malloc((size_t)N * sizeof(double));
    double* W = (double*)malloc((size_t)N * sizeof(double));
    mp_job jb = {n, 0, nthreads, 0, -1, exp(-1.0 / ((double)n * length_scale)), P, q, w, W, w, u};
    mp_run(&jb, 0, nthreads); mp_run(&jb, 1, nthreads); mp_run(&jb, 2, nthreads);       /* w = K q */
    jb.v = W;
    for (int i = 0; i < n; ++i) {
        jb.var = i;
        mp_run(&jb, 0, nthreads); mp_run(&jb, 1, nthreads); mp_run(&jb, 2, nthreads);   /* W_i = K (q .* s_i) */
        mp_run(&jb, 3, nthreads);
    }
    free(w); free(W);
    double total = 0.0;
    for (int64_t z = 0; z < N; ++z) total += q[z] * u[z];
    return total;
}
