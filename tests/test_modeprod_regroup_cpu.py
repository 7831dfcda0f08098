"""CPU check of the algebra behind the mode-product kernels (csrc/tnk_modeprod.cu header): with flips folded
in front of the transform for the "pre" bits and taken from the transformed tile for the "post" bits,

    u = - sum_p r_p .* K(q.*s_p + X_p q) + w .* (n + sum_p r_p) + K(-sum_pre X_p (q.*r_p))
        + sum_post [ X_p K(q.*s_p + X_p q) - X_p w - w ]

must equal the oracle's u = K_p q for every split between pre and post bits and for bit-range partials."""
import numpy as np
import pytest

from oracle import ksd_oracle as orc
from tensornetworks_b200.synthetic import random_binary_bn, random_q


def _regrouped(q, S, n, length_scale, split, bit_begin=0, bit_end=None):
    bit_end = n if bit_end is None else bit_end
    N = 1 << n
    a = np.exp(-1.0 / (n * length_scale))
    z = np.arange(N)

    def K(v):
        for p in range(n):
            v = v + a * v[z ^ (1 << p)]
        return v

    w = K(q)
    bits = range(bit_begin, bit_end)
    s = {p: S[:, n - 1 - p] for p in bits}              # device tables are in bit order
    r = {p: 1.0 - s[p] for p in bits}
    n_post = sum(1 for p in bits if p >= split)
    Y = np.zeros(N)
    for p in bits:
        if p < split:
            Y -= (q * r[p])[z ^ (1 << p)]
    u = w * (len(bits) + sum(r[p] for p in bits) - n_post) + K(Y)
    for p in bits:
        Wp = K(q * s[p] + q[z ^ (1 << p)])
        u -= r[p] * Wp
        if p >= split:
            u += Wp[z ^ (1 << p)] - w[z ^ (1 << p)]
    return u


@pytest.mark.parametrize("n", [3, 5, 8])
@pytest.mark.parametrize("length_scale", [1.0, 0.6])
def test_regrouped_sum_equals_oracle(n, length_scale):
    bn, lat, obs, x = random_binary_bn(n, seed=n)
    P = orc.joint_table(bn, x, lat)
    S = orc.score_table(P, n)
    q = random_q(1 << n, seed=n).numpy()
    Kref, gref = orc.ksd_dense(q, S, n, length_scale)
    for split in (0, n // 2, n):
        u = _regrouped(q, S, n, length_scale, split)
        assert np.max(np.abs(2.0 * u - gref)) <= 1e-12 * np.max(np.abs(gref)), split
        assert abs(q @ u - Kref) <= 1e-12 * abs(Kref), split
    # bit-range partials (the multi-GPU split) add up, whatever side of the split they are on
    cut = max(1, n // 3)
    ua = _regrouped(q, S, n, length_scale, n // 2, 0, cut)
    ub = _regrouped(q, S, n, length_scale, n // 2, cut, n)
    assert np.max(np.abs(2.0 * (ua + ub) - gref)) <= 1e-12 * np.max(np.abs(gref))
