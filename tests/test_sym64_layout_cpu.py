"""CPU check of the decomposition behind the symmetric fp64 kernel (csrc/tnk_sym64.cu, "K3-S64") and the operand formats
tf32_tile / tf32_dot of the symmetric tcgen05 kernel (csrc/tnk_tc.cu):

  * the pair kernel keeps  a^popc(i^j) * s(i).s(j)  with the Hamming weight factorised over the bit fields of the
    accumulator fragment:  LUT[tile and block bits] * thread constant * a^popc(a^b) * a^((r&1)^e);
  * everything else of k_p (reference stein_utils.py:138-197), summed over all columns, is Kronecker-structured and is
    what tnk_hibits_pass (csrc/tnk_modeprod.cu) adds:  C(i) K(q)_i + K(C.q)_i + sum_p [ g_p(i) K(b_p.q)_i + b_p(i) K(g_p.q)_i ],
    with K(b_p.q) obtained elementwise from w = K q.

Numpy restatements of both pieces must reproduce the oracle's dense K_p; the unit enumeration must be a bijection."""
import numpy as np
import pytest

from oracle import ksd_oracle as orc
from tensornetworks_b200.synthetic import illconditioned_binary_bn, random_binary_bn, random_q


def _popc(x):
    return np.array([bin(int(v)).count("1") for v in np.ravel(x)]).reshape(np.shape(x))


def _setup(n, gen):
    bn, lat, obs, x = gen(n, seed=n)
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    N = 1 << n
    a_ = np.exp(-1.0 / n)
    alpha, gamma = 1.0 - a_, a_ - 1.0 / a_
    z = np.arange(N)
    s = S[:, ::-1]                                               # component p = bit position p
    bit = (z[:, None] >> np.arange(n)[None, :]) & 1
    g = gamma * (1.0 - s) * (1 - 2 * bit)                        # g_p(z)
    C = alpha * (1.0 - s).sum(1) + gamma * ((1.0 - s) * bit).sum(1)
    return S, N, a_, z, s, bit, g, C


@pytest.mark.parametrize("n,gen", [(8, random_binary_bn), (9, illconditioned_binary_bn)])
def test_weight_factorisation_and_split_of_the_bracket(n, gen):
    S, N, a_, z, s, bit, g, C = _setup(n, gen)
    Kp = orc.stein_matrix_dense(S, n)
    wl = np.exp(-np.arange(n + 1) / n)
    i, j = z[:, None], z[None, :]
    I, J = i >> 7, j >> 7
    wr, a, r = (i >> 5) & 3, (i >> 3) & 3, i & 7
    cb, b, k4, e = (j >> 5) & 3, (j >> 3) & 3, (j >> 1) & 3, j & 1
    weight = (wl[_popc(I ^ J) + _popc(wr ^ cb)] * a_ ** _popc(((r >> 1) & 3) ^ k4)
              * a_ ** _popc(a ^ b) * a_ ** ((r & 1) ^ e))
    assert np.max(np.abs(weight - wl[_popc(i ^ j)])) <= 1e-15
    rest = C[:, None] + C[None, :] + g @ bit.T + bit @ g.T       # C(i) + C(j) + T1 + T2 (gamma inside g)
    got = weight * (s @ s.T + rest)
    assert np.max(np.abs(got - Kp)) <= 1e-12 * np.max(np.abs(Kp))


@pytest.mark.parametrize("n,gen,p_lo,with_c", [(8, random_binary_bn, 0, True), (9, illconditioned_binary_bn, 0, True),
                                               (9, random_binary_bn, 7, False), (10, random_binary_bn, 7, False)])
def test_kronecker_pass_identity(n, gen, p_lo, with_c):
    """What hib_build / mp_transform / hib_fold compute: the terms of bit positions >= p_lo (and, with_c, the C terms)
    for all rows at once, K(b_p.q) taken elementwise from w = K q."""
    S, N, a_, z, s, bit, g, C = _setup(n, gen)
    q = random_q(N, seed=n).numpy()
    wl = np.exp(-np.arange(n + 1) / n)
    Kb = wl[_popc(z[:, None] ^ z[None, :])]                      # the base kernel matrix K = M^{(x)n}
    T = sum(np.outer(g[:, p], bit[:, p]) + np.outer(bit[:, p], g[:, p]) for p in range(p_lo, n))
    if with_c:
        T = T + C[:, None] + C[None, :]
    want = (Kb * T) @ q
    w = Kb @ q
    got = np.zeros(N)
    for p in range(p_lo, n):
        on = bit[:, p].astype(bool)
        A = np.where(on, 1.0, a_) * (w[z | (1 << p)] - a_ * w[z & ~(1 << p)]) / (1.0 - a_ * a_)
        assert np.max(np.abs(A - Kb @ (bit[:, p] * q))) <= 1e-12 * np.max(np.abs(w))
        got += g[:, p] * A + np.where(on, Kb @ (g[:, p] * q), 0.0)
    if with_c:
        got += C * w + Kb @ (C * q)
    assert np.max(np.abs(got - want)) <= 1e-12 * np.max(np.abs(want))


def test_unit_enumeration_is_a_bijection():
    for B in (1, 2, 5, 32, 128):
        seen = set()
        for k in range(B * (B + 1) // 2):
            a, kk = 0, k
            while kk >= B - a:
                kk -= B - a
                a += 1
            pa, pb = a, a + kk
            assert 0 <= pa <= pb < B
            assert pa * B - pa * (pa - 1) // 2 + (pb - pa) == k
            seen.add((pa, pb))
        assert len(seen) == B * (B + 1) // 2


@pytest.mark.parametrize("n,G", [(9, 2), (10, 3), (9, 8), (8, 16)])
def test_side_pass_shares_add_up(n, G):
    """The Kronecker side pass is dealt over the G calls of a stride (tnk_hibits_pass: share g takes the bit positions
    g, g + G, ...; share 0 also the C terms; shares beyond the last bit are empty): the shares must add up to the whole."""
    S, N, a_, z, s, bit, g, C = _setup(n, random_binary_bn)
    q = random_q(N, seed=n).numpy()
    wl = np.exp(-np.arange(n + 1) / n)
    Kb = wl[_popc(z[:, None] ^ z[None, :])]
    T = sum(np.outer(g[:, p], bit[:, p]) + np.outer(bit[:, p], g[:, p]) for p in range(n)) + C[:, None] + C[None, :]
    want = (Kb * T) @ q
    w = Kb @ q
    total = np.zeros(N)
    owned = []
    for share in range(G):
        bits = list(range(share, n, G))
        owned += bits
        part = np.zeros(N)
        for p in bits:
            on = bit[:, p].astype(bool)
            A = np.where(on, 1.0, a_) * (w[z | (1 << p)] - a_ * w[z & ~(1 << p)]) / (1.0 - a_ * a_)
            part += g[:, p] * A + np.where(on, Kb @ (g[:, p] * q), 0.0)
        if share == 0:
            part += C * w + Kb @ (C * q)
        total += part
    assert sorted(owned) == list(range(n))
    assert np.max(np.abs(total - want)) <= 1e-12 * np.max(np.abs(want))
