"""CPU check of the decomposition the symmetric fp64 kernel uses (csrc/tnk_sym64.cu, "K3-S64"): per pair the bracket
is   DMMA part (score dot product)  +  row seed  +  column seed  +  the T1/T2 terms of bit positions 3,4 (held by the
row-block / column-group indices a, b of the accumulator fragment: a Gray walk on the row side, g_3, g_4 of the column
on the other),
and the Hamming weight is  LUT[tile and block bits] * thread constant * a^popc(a^b) * a^((r&1)^e).
A numpy restatement of those pieces, indexed by the same bit fields of i and j the kernel uses, must reproduce the
oracle's dense K_p (reference stein_utils.py:138-197); the unit enumeration must be a bijection."""
import numpy as np
import pytest

from oracle import ksd_oracle as orc
from tensornetworks_b200.synthetic import illconditioned_binary_bn, random_binary_bn


def _popc(x):
    return np.array([bin(int(v)).count("1") for v in np.ravel(x)]).reshape(np.shape(x))


@pytest.mark.parametrize("n,gen", [(8, random_binary_bn), (9, illconditioned_binary_bn)])
def test_seed_and_weight_factorisation(n, gen):
    bn, lat, obs, x = gen(n, seed=n)
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    Kp = orc.stein_matrix_dense(S, n)
    N = 1 << n
    a_ = np.exp(-1.0 / n)
    alpha, gamma = 1.0 - a_, a_ - 1.0 / a_
    z = np.arange(N)
    s = S[:, ::-1]                                               # component p = bit position p
    bit = (z[:, None] >> np.arange(n)[None, :]) & 1
    g = gamma * (1.0 - s) * (1 - 2 * bit)                        # g_p(z)
    C = alpha * (1.0 - s).sum(1) + gamma * ((1.0 - s) * bit).sum(1)
    wl = np.exp(-np.arange(n + 1) / n)

    i, j = z[:, None], z[None, :]
    I, J = i >> 7, j >> 7
    wr, a, r = (i >> 5) & 3, (i >> 3) & 3, i & 7
    cb, b, k4, e = (j >> 5) & 3, (j >> 3) & 3, (j >> 1) & 3, j & 1
    wc = cb >> 1
    gi = lambda p: g[:, p][:, None]
    gj = lambda p: g[:, p][None, :]
    hi_bits = range(7, n)
    # row seed: C(i) + column bits 1,2 (k4), 6 (wc), >= 7 (J), 5 (cb & 1), 0 (e)
    row = (C[:, None] + (k4 & 1) * gi(1) + (k4 >> 1) * gi(2) + wc * gi(6)
           + sum(((J >> (p - 7)) & 1) * gi(p) for p in hi_bits) + (cb & 1) * gi(5) + e * gi(0))
    # column seed: colA[wr][c] + colB[r][c]
    col = (C[None, :] + sum(((I >> (p - 7)) & 1) * gj(p) for p in hi_bits) + (wr & 1) * gj(5) + (wr >> 1) * gj(6)
           + (r & 1) * gj(0) + ((r >> 1) & 1) * gj(1) + (r >> 2) * gj(2))
    # dot product (DMMA) + the bits-3,4 terms: g3(i) b3(j) + g4(i) b4(j) on the row seed, b3(i) g3(j) + b4(i) g4(j) on the column seed
    mma = s @ s.T + gi(3) * (b & 1) + gi(4) * (b >> 1) + (a & 1) * gj(3) + (a >> 1) * gj(4)
    weight = (wl[_popc(I ^ J) + _popc(wr ^ cb)] * a_ ** _popc(((r >> 1) & 3) ^ k4)
              * a_ ** _popc(a ^ b) * a_ ** ((r & 1) ^ e))
    got = weight * (mma + row + col)
    assert np.max(np.abs(got - Kp)) <= 1e-12 * np.max(np.abs(Kp))


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
