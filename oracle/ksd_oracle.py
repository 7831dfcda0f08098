"""CPU oracle for the exact-KSD hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product path (tensornetworks_b200/) never does.

It restates, in numpy fp64, what the reference computes on the path
(file:line relative to the reference checkout):

  enumerate_states      utils.py:77-91         MSB-first enumeration order
  joint_table           stein_utils.py:58-112  p(x,z): evidence clamped, hidden summed
                        bayesian_network.py:111-146  CPT product in bn.nodes order from 1.0
  score_table           stein_utils.py:115-136 s_i = 1 - p(x,flip_i z)/p(x,z), 0 if |p|<1e-12
  base_kernel_lut       stein_utils.py:30-55   k = exp(-h/(n*l))
  stein_matrix_dense    stein_utils.py:138-197 the four Eq.-13 terms, term by term (O1)
  ksd_dense             ksd_vi.py:110-131,145  K = sum_ij q_i q_j k_p(z_i,z_j), dK/dq = 2 K_p q
  ksd_modeproduct       same quantity through the Kronecker factorisation K = M^{(x)n} (O2)

Parity pinning: checked in tests/test_oracle_golden.py against (a) the eight known-answer
asserts of stein_utils.py:205-251 and (b) fixtures under tests/golden/ produced by
tests/golden/make_golden.py, which imports and runs the reference's own functions.
"""
import numpy as np

SCORE_ZERO_THRESHOLD = 1e-12  # stein_utils.py:126


def enumerate_states(n):
    """[2^n, n] int8 array; row i is bin(i) zero-filled to n digits (utils.py:77-91)."""
    if n == 0:
        return np.zeros((1, 0), dtype=np.int8)
    idx = np.arange(1 << n, dtype=np.int64)
    shifts = np.arange(n - 1, -1, -1, dtype=np.int64)
    return ((idx[:, None] >> shifts[None, :]) & 1).astype(np.int8)


def _cpt_lookup(cpt_entry, parent_values):
    """bayesian_network.py:135-141: callable or dict CPT -> {0:p0, 1:p1}."""
    if callable(cpt_entry):
        return cpt_entry(parent_values)
    return cpt_entry.get(parent_values)


def joint_table(bn, x_dict, latent_vars_names):
    """p(x,z) for every latent state z, in enumeration order (fp64 [2^n]).

    Follows compute_prob_joint_xz (stein_utils.py:58-112): assignment = observed, then
    latent on top (62-64); nodes of bn.nodes in neither set are summed out over
    generate_all_binary_outcomes(m) starting from 0.0 (95-111); each term is the
    sequential CPT product of get_joint_probability (bayesian_network.py:129-144).
    """
    n = len(latent_vars_names)
    states = enumerate_states(n)  # [N, n]
    N = states.shape[0]
    assigned = dict(x_dict) if x_dict else {}
    latent_pos = {}
    for k, name in enumerate(latent_vars_names):
        assigned[name] = ('latent', k)
        latent_pos[name] = k
    hidden = [node for node in bn.nodes if node not in assigned]
    hidden_states = enumerate_states(len(hidden))  # [2^m, m]

    def column(name, hrow):
        v = assigned.get(name)
        if isinstance(v, tuple):
            return states[:, v[1]].astype(np.int64)
        if v is None:
            return np.full(N, int(hrow[hidden.index(name)]), dtype=np.int64)
        return np.full(N, int(v), dtype=np.int64)

    total = np.zeros(N, dtype=np.float64)
    for hrow in hidden_states:
        prob = np.ones(N, dtype=np.float64)
        for node in bn.nodes:
            val = column(node, hrow)
            pas = list(bn.parents[node]) if node in bn.parents else []
            pa_cols = [column(p, hrow) for p in pas]
            k = len(pas)
            cfg = np.zeros(N, dtype=np.int64)
            for c in pa_cols:
                cfg = (cfg << 1) | c
            # tabulate both outcomes for each parent configuration that occurs
            tab = np.zeros((1 << k, 2), dtype=np.float64)
            for c in np.unique(cfg):
                key = tuple(int((c >> (k - 1 - t)) & 1) for t in range(k))
                pd = _cpt_lookup(bn.cpts[node], key)
                if pd is None:
                    raise ValueError(f"CPT entry for node {node} with parent values {key} not found.")
                tab[c, 0] = pd[0]
                tab[c, 1] = pd[1]
            prob = prob * tab[cfg, val]
        if len(hidden) == 0:
            return prob
        total = total + prob
    return total


def score_table(P, n):
    """S[N, n] with S[z,i] = 1 - P[flip_i z]/P[z]; all-zero row where |P[z]|<1e-12.

    stein_utils.py:115-136.  flip_bit(z,i) (stein_utils.py:7-11) is index ^ (1 << (n-1-i)).
    """
    N = P.shape[0]
    idx = np.arange(N, dtype=np.int64)
    S = np.zeros((N, n), dtype=np.float64)
    live = ~(np.abs(P) < SCORE_ZERO_THRESHOLD)
    with np.errstate(divide='ignore', invalid='ignore'):
        for i in range(n):
            nb = P[idx ^ (1 << (n - 1 - i))]
            S[:, i] = np.where(live, 1.0 - nb / P, 0.0)
    return S


def base_kernel_lut(n, length_scale=1.0):
    """w[h] = exp(-h/(n*l)), h = 0..n (stein_utils.py:49-55)."""
    h = np.arange(n + 1, dtype=np.float64)
    return np.exp(-h / (float(n) * float(length_scale)))


def _popcount(x):
    x = x.astype(np.uint64)
    c = np.zeros(x.shape, dtype=np.int64)
    while np.any(x):
        c += (x & np.uint64(1)).astype(np.int64)
        x = x >> np.uint64(1)
    return c


def stein_matrix_dense(S, n, length_scale=1.0, rows=None):
    """K_p[rows, :] as a dense fp64 block, term by term as stein_utils.py:161-196 (O1).

    For a pair (z,z') with h = hamming(z,z') the reference evaluates the base kernel at
    (z,z'), (z,flip_j z'), (flip_i z,z') and (flip_i z,flip_i z').  Flipping one bit of one
    argument moves h by +1 where the two states agree on that bit and by -1 where they
    differ, so each of those evaluations is a LUT entry w[h+-1]; flipping the same bit
    in both leaves h unchanged.
    """
    N = S.shape[0]
    if n == 0:
        return np.zeros((N if rows is None else len(rows), N))
    w = base_kernel_lut(n, length_scale)
    idx = np.arange(N, dtype=np.int64)
    ri = idx if rows is None else np.asarray(rows, dtype=np.int64)
    X = ri[:, None] ^ idx[None, :]
    h = _popcount(X)
    k = w[h]                                                    # k(z,z')
    wpad = np.concatenate([[0.0], w, [0.0]])                    # w[h-1], w[h+1] safely
    k_minus = wpad[h]                                           # w[h-1]
    k_plus = wpad[h + 2]                                        # w[h+1]
    Si = S[ri]
    term1 = (Si @ S.T) * k                                      # 161-164
    term2 = np.zeros_like(k)
    term3 = np.zeros_like(k)
    term4 = np.zeros_like(k)
    for j in range(n):
        d = ((X >> (n - 1 - j)) & 1).astype(bool)
        k_flip = np.where(d, k_minus, k_plus)                   # one argument flipped at bit j
        delta = k - k_flip
        term2 -= Si[:, j][:, None] * delta                      # 166-171
        term3 -= delta * S[:, j][None, :]                       # 173-178
        term4 += k - k_flip - k_flip + k                        # 180-194
    return term1 + term2 + term3 + term4


def ksd_dense(q, S, n, length_scale=1.0, block=2048):
    """(K, dK/dq) with K = sum_ij q_i q_j k_p(z_i,z_j) (ksd_vi.py:114-131) and the autograd
    gradient 2 K_p q (ksd_vi.py:145), via dense row blocks of stein_matrix_dense."""
    q = np.asarray(q, dtype=np.float64)
    N = q.shape[0]
    u = np.zeros(N)
    for r0 in range(0, N, block):
        rows = np.arange(r0, min(N, r0 + block))
        u[rows] = stein_matrix_dense(S, n, length_scale, rows) @ q
    return float(q @ u), 2.0 * u


def _kron_apply(v, n, mats):
    """Apply (x)_p mats[p] to v viewed as a (2,)*n tensor; axis p = latent variable p (MSB first)."""
    t = v.reshape((2,) * n)
    for p in range(n):
        t = np.moveaxis(np.tensordot(mats[p], t, axes=([1], [p])), 0, p)
    return t.reshape(-1)


def ksd_modeproduct(q, S, n, length_scale=1.0):
    """Same (K, dK/dq) in O(n^2 2^n) through K = M^{(x)n}, M=[[1,a],[a,1]] (O2).

    u = sum_i [ s_i .* K(q .* s_i) + r_i .* K^(i) q + K^(i)(q .* r_i) ],  r = 1 - s,
    where K^(i) replaces axis i's factor by (1-a)[[1,-1],[-1,1]].  Derivation in DESIGN.md;
    validated against ksd_dense and the reference loop in tests/.
    """
    q = np.asarray(q, dtype=np.float64)
    if n == 0:
        return 0.0, np.zeros_like(q)
    a = float(np.exp(-1.0 / (float(n) * float(length_scale))))
    M = np.array([[1.0, a], [a, 1.0]])
    C = (1.0 - a) * np.array([[1.0, -1.0], [-1.0, 1.0]])
    base = [M] * n
    u = np.zeros_like(q)
    r = 1.0 - S
    for i in range(n):
        mats_i = list(base)
        mats_i[i] = C
        u += S[:, i] * _kron_apply(q * S[:, i], n, base)
        u += r[:, i] * _kron_apply(q, n, mats_i)
        u += _kron_apply(q * r[:, i], n, mats_i)
    return float(q @ u), 2.0 * u
