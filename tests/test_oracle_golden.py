"""Pins the CPU oracle (oracle/ksd_oracle.py) to the reference: the known-answer asserts of
reference stein_utils.py:205-251 and the fixtures made by running the reference
(tests/golden/make_golden.py).  Joint/score tables must be bit-identical."""
import numpy as np
import pytest

from oracle import ksd_oracle as orc
from tensornetworks_b200.bayesian_network import BayesianNetwork, get_sprinkler_network
from tensornetworks_b200.synthetic import random_binary_bn

RANDOM_CASES = [(n, n, 0) for n in range(2, 7)] + [(4, 104, 2)]
TABLE_CASES = [(10, 0, 0), (10, 1, 2), (8, 7, 1)]


def ab_network():
    bn = BayesianNetwork()
    bn.add_node('A', cpt={(): {0: 0.8, 1: 0.2}})
    bn.add_node('B', cpt={(0,): {0: 0.7, 1: 0.3}, (1,): {0: 0.4, 1: 0.6}}, parent_names=['A'])
    return bn


def test_enumeration_order_msb_first():
    # reference utils.py:77-91 and the flip_bit assert stein_utils.py:205
    st = orc.enumerate_states(3)
    assert [tuple(r) for r in st] == [(0, 0, 0), (0, 0, 1), (0, 1, 0), (0, 1, 1),
                                      (1, 0, 0), (1, 0, 1), (1, 1, 0), (1, 1, 1)]
    assert orc.enumerate_states(0).shape == (1, 0)
    assert [tuple(r) for r in orc.enumerate_states(1)] == [(0,), (1,)]


def test_known_answers_ab(golden):
    # stein_utils.py:229-251
    P = orc.joint_table(ab_network(), {'B': 1}, ['A'])
    S = orc.score_table(P, 1)
    assert np.isclose(P[1], 0.12)
    assert np.isclose(S[1, 0], -1.0) and np.isclose(S[0, 0], 0.5)
    assert np.array_equal(P, golden["ab_P"]) and np.array_equal(S, golden["ab_S"])
    Kp = orc.stein_matrix_dense(S, 1)
    assert np.isclose(Kp[0, 1], 2 * np.exp(-1.0) - 2.5) and np.isclose(Kp[0, 0], 1.25 - np.exp(-1.0))
    assert np.isclose(Kp[0, 1], golden["ab_kp01"], rtol=1e-14)
    assert np.isclose(Kp[0, 0], golden["ab_kp00"], rtol=1e-14)
    # stein_utils.py:208-217: hamming([0,0,1,1],[1,0,0,1]) = 2, kernel exp(-2/4)
    assert orc._popcount(np.array([0b0011 ^ 0b1001]))[0] == 2
    assert np.isclose(orc.base_kernel_lut(4)[2], np.exp(-0.5))


def test_sprinkler_tables_and_ksd(golden):
    bn = get_sprinkler_network(False)
    P = orc.joint_table(bn, {'W': 1}, ['C', 'S', 'R'])
    S = orc.score_table(P, 3)
    assert np.array_equal(P, golden["spr_P"])
    assert np.array_equal(S, golden["spr_S"])
    assert np.isclose(P.sum(), 0.65) and np.allclose(P / P.sum(), golden["spr_post"], rtol=1e-15)
    assert np.isclose(golden["spr_K_uniform"], 2364.8706394412725, rtol=1e-13)
    for name in ("uniform", "ramp", "post"):
        q = golden[f"spr_q_{name}"]
        for fn in (orc.ksd_dense, orc.ksd_modeproduct):
            K, g = fn(q, S, 3)
            gref = golden[f"spr_g_{name}"]
            if name == "post":  # K ~ 0: compare absolutely against the scale of K_p
                assert abs(K - golden[f"spr_K_{name}"]) < 1e-11
            else:
                assert np.isclose(K, golden[f"spr_K_{name}"], rtol=1e-12)
                assert np.max(np.abs(g - gref)) <= 1e-12 * np.max(np.abs(gref))
    for ls in (0.5, 2.0):
        for fn in (orc.ksd_dense, orc.ksd_modeproduct):
            K, g = fn(golden["spr_q_ramp"], S, 3, length_scale=ls)
            assert np.isclose(K, golden[f"spr_K_ramp_ls{ls}"], rtol=1e-12)
            gref = golden[f"spr_g_ramp_ls{ls}"]
            assert np.max(np.abs(g - gref)) <= 1e-12 * np.max(np.abs(gref))


@pytest.mark.parametrize("n,seed,nh", RANDOM_CASES)
def test_random_bn_against_reference_loop(golden, n, seed, nh):
    bn, lat, obs, x = random_binary_bn(n, seed=seed, n_hidden=nh)
    tag = f"rnd_n{n}_s{seed}_h{nh}"
    P = orc.joint_table(bn, x, lat)
    S = orc.score_table(P, n)
    assert np.array_equal(P, golden[tag + "_P"])
    assert np.array_equal(S, golden[tag + "_S"])
    for fn in (orc.ksd_dense, orc.ksd_modeproduct):
        K, g = fn(golden[tag + "_q"], S, n)
        assert np.isclose(K, golden[tag + "_K"], rtol=1e-11)
        gref = golden[tag + "_g"]
        assert np.max(np.abs(g - gref)) <= 1e-11 * np.max(np.abs(gref))


@pytest.mark.parametrize("n,seed,nh", TABLE_CASES)
def test_enumeration_tables_bit_exact(golden, n, seed, nh):
    bn, lat, obs, x = random_binary_bn(n, seed=seed, n_hidden=nh)
    tag = f"tab_n{n}_s{seed}_h{nh}"
    P = orc.joint_table(bn, x, lat)
    assert np.array_equal(P, golden[tag + "_P"])
    assert np.array_equal(orc.score_table(P, n), golden[tag + "_S"])


def test_dense_vs_modeproduct_n11():
    bn, lat, obs, x = random_binary_bn(11, seed=3)
    S = orc.score_table(orc.joint_table(bn, x, lat), 11)
    q = np.random.default_rng(0).dirichlet(np.ones(1 << 11))
    K1, g1 = orc.ksd_dense(q, S, 11)
    K2, g2 = orc.ksd_modeproduct(q, S, 11)
    assert np.isclose(K1, K2, rtol=1e-10)
    assert np.max(np.abs(g1 - g2)) <= 1e-10 * np.max(np.abs(g1))


# ---- the C mode-product oracle (oracle/ksd_oracle.c:orc_ksd_modeproduct), the checker of the n = 20 / 24 GPU tests
@pytest.mark.parametrize("n,seed,nh", RANDOM_CASES)
def test_c_modeproduct_against_reference_loop(golden, n, seed, nh):
    from oracle import c_oracle
    tag = f"rnd_n{n}_s{seed}_h{nh}"
    K, g = c_oracle.ksd_modeproduct(golden[tag + "_q"], golden[tag + "_P"], n)
    assert np.isclose(K, golden[tag + "_K"], rtol=1e-11)
    gref = golden[tag + "_g"]
    assert np.max(np.abs(g - gref)) <= 1e-11 * np.max(np.abs(gref))


def test_c_modeproduct_sprinkler_and_length_scales(golden):
    from oracle import c_oracle
    P = golden["spr_P"]
    for name in ("uniform", "ramp"):
        K, g = c_oracle.ksd_modeproduct(golden[f"spr_q_{name}"], P, 3)
        assert np.isclose(K, golden[f"spr_K_{name}"], rtol=1e-12)
        assert np.max(np.abs(g - golden[f"spr_g_{name}"])) <= 1e-12 * np.max(np.abs(golden[f"spr_g_{name}"]))
    for ls in (0.5, 2.0):
        K, g = c_oracle.ksd_modeproduct(golden["spr_q_ramp"], P, 3, length_scale=ls)
        assert np.isclose(K, golden[f"spr_K_ramp_ls{ls}"], rtol=1e-12)
    K, g = c_oracle.ksd_modeproduct(np.ones(1), np.ones(1), 0)
    assert K == 0.0 and g[0] == 0.0                     # n = 0: k_p = 0 (stein_utils.py:151-152)


@pytest.mark.parametrize("n,threads", [(9, 1), (12, 3), (13, 8), (15, 5)])
def test_c_modeproduct_vs_numpy_forms_and_literal_rows(n, threads):
    """Sizes on both sides of the cache-blocking threshold (2^12 states), odd thread counts, dead rows."""
    from oracle import c_oracle
    from tensornetworks_b200.flatten import flatten_bn
    from tensornetworks_b200.synthetic import illconditioned_binary_bn, random_q
    gen = illconditioned_binary_bn if n % 2 else random_binary_bn
    bn, lat, obs, x = gen(n, seed=n)
    P = c_oracle.joint_table(flatten_bn(bn, x, lat))
    assert np.array_equal(P, orc.joint_table(bn, x, lat))
    S = orc.score_table(P, n)
    q = random_q(1 << n, seed=n).numpy()
    K0, g0 = orc.ksd_dense(q, S, n) if n <= 12 else orc.ksd_modeproduct(q, S, n)
    K1, g1 = c_oracle.ksd_modeproduct(q, P, n, nthreads=threads)
    assert np.isclose(K0, K1, rtol=1e-11)
    assert np.max(np.abs(g0 - g1)) <= 1e-11 * np.max(np.abs(g0))
    rows = (1 << n) - 5
    u, _ = c_oracle.ksd_rows(q, S, n, rows, rows + 5)
    assert np.max(np.abs(2 * u - g1[rows:rows + 5])) <= 1e-11 * np.max(np.abs(g1))
