"""CPU-only tests: C oracle vs goldens, C-ABI exports, BN flattening, enumeration helpers,
and the world_size-2 exchange step over gloo."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import c_oracle, ksd_oracle as orc
from tensornetworks_b200 import _lib
from tensornetworks_b200.bayesian_network import BayesianNetwork, get_sprinkler_network
from tensornetworks_b200.flatten import flatten_bn, NODE_CONST, NODE_HIDDEN, NODE_LATENT
from tensornetworks_b200.synthetic import random_binary_bn
from tensornetworks_b200.utils import calculate_tvd, generate_all_binary_outcomes, get_binary_key, get_outcome_tuple

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_c_abi_exports_every_declared_symbol():
    header = open(os.path.join(REPO, "include", "tnk.h")).read()
    declared = set(re.findall(r"\b(tnk_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.EXPORTS)
    lib = ctypes.CDLL(_lib.build())
    for name in declared:
        assert hasattr(lib, name), name
    assert _lib.load().tnk_version() == _lib.ABI_VERSION
    assert _lib.layout(20, _lib.PREC_FP32) == (20, 1 << 20, 4)
    assert _lib.layout(3, _lib.PREC_FP64) == (4, 1024, 8)


def test_no_cpu_fallback():
    from tensornetworks_b200 import ops
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    bn = get_sprinkler_network(False)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.joint_table(bn, {'W': 1}, ['C', 'S', 'R'], device="cuda")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.joint_table(bn, {'W': 1}, ['C', 'S', 'R'], device="cpu")


@pytest.mark.parametrize("n,seed,nh", [(n, n, 0) for n in range(2, 7)] + [(4, 104, 2)])
def test_c_oracle_against_reference_goldens(golden, n, seed, nh):
    bn, lat, obs, x = random_binary_bn(n, seed=seed, n_hidden=nh)
    tag = f"rnd_n{n}_s{seed}_h{nh}"
    P = c_oracle.joint_table(flatten_bn(bn, x, lat))
    S = c_oracle.score_table(P, n)
    assert np.array_equal(P, golden[tag + "_P"]) and np.array_equal(S, golden[tag + "_S"])
    u, K = c_oracle.ksd_rows(golden[tag + "_q"], S, n, 0, 1 << n)
    assert np.isclose(K, golden[tag + "_K"], rtol=1e-11)
    gref = golden[tag + "_g"]
    assert np.max(np.abs(2 * u - gref)) <= 1e-11 * np.max(np.abs(gref))


@pytest.mark.parametrize("n,seed,nh", [(10, 0, 0), (10, 1, 2), (8, 7, 1)])
def test_c_oracle_tables_bit_exact(golden, n, seed, nh):
    bn, lat, obs, x = random_binary_bn(n, seed=seed, n_hidden=nh)
    tag = f"tab_n{n}_s{seed}_h{nh}"
    P = c_oracle.joint_table(flatten_bn(bn, x, lat))
    assert np.array_equal(P, golden[tag + "_P"])
    assert np.array_equal(c_oracle.score_table(P, n), golden[tag + "_S"])


def test_flatten_semantics():
    bn = get_sprinkler_network(False)
    flat = flatten_bn(bn, {'W': 1}, ['C', 'S', 'R'])
    assert list(flat.node_kind) == [NODE_LATENT] * 3 + [NODE_CONST]
    assert list(flat.node_arg) == [2, 1, 0, 1]          # C is the most significant bit
    assert list(flat.pa_off) == [0, 0, 1, 2, 4] and list(flat.pa_idx) == [0, 0, 1, 2]
    assert flat.cpt.shape == (1 + 2 + 2 + 4, 2)
    # a latent name wins a clash with the evidence; un-mentioned nodes are hidden (stein_utils.py:62-67)
    flat = flatten_bn(bn, {'W': 1, 'R': 0}, ['C', 'R'])
    assert list(flat.node_kind) == [NODE_LATENT, NODE_HIDDEN, NODE_LATENT, NODE_CONST]
    assert flat.n_hidden == 1
    # callable CPTs are tabulated
    bn2 = BayesianNetwork()
    bn2.add_node('A', cpt=lambda pa: {0: 0.25, 1: 0.75})
    bn2.add_node('B', cpt=lambda pa: {0: 0.9 - 0.5 * pa[0], 1: 0.1 + 0.5 * pa[0]}, parent_names=['A'])
    flat = flatten_bn(bn2, {'B': 1}, ['A'])
    assert np.allclose(flat.cpt, [[0.25, 0.75], [0.9, 0.1], [0.4, 0.6]])
    assert np.allclose(c_oracle.joint_table(flat), [0.25 * 0.1, 0.75 * 0.6])
    with pytest.raises(ValueError):
        flatten_bn(bn, {'W': 2}, ['C', 'S', 'R'])


def test_utils_drop_in():
    outs = generate_all_binary_outcomes(3)
    assert len(outs) == 8 and outs[5] == (1, 0, 1) and outs.index((1, 1, 0)) == 6
    assert list(generate_all_binary_outcomes(0)) == [()]
    assert list(generate_all_binary_outcomes(1)) == [(0,), (1,)]
    assert outs == [tuple(int(c) for c in bin(i)[2:].zfill(3)) for i in range(8)]
    assert len(generate_all_binary_outcomes(24)) == 1 << 24          # lazy: no 16.7 M tuples
    assert get_binary_key((0, 1, 1)) == "011" and get_outcome_tuple("011") == (0, 1, 1)
    p1 = {'00': 0.25, '01': 0.25, '10': 0.25, '11': 0.25}
    p2 = {'00': 0.5, '01': 0.1, '10': 0.1, '11': 0.3}
    assert np.isclose(calculate_tvd(p1, p2), 0.3)                     # reference utils.py:98
    assert np.isclose(calculate_tvd(np.array(list(p1.values())), np.array(list(p2.values()))), 0.3)
    with pytest.raises(TypeError):
        calculate_tvd(p1, np.zeros(4))


def test_bayesian_network_drop_in(golden):
    bn = get_sprinkler_network(False)
    post, pw = bn.get_true_posterior(['C', 'S', 'R'], {'W': 1})
    assert np.isclose(pw, golden["spr_evidence"], rtol=1e-15)
    assert np.allclose([post[z] for z in generate_all_binary_outcomes(3)], golden["spr_post"], rtol=1e-14)
    prior = bn.get_prior_distribution(['C', 'W'])
    assert np.isclose(sum(prior.values()), 1.0)
    with pytest.raises(ValueError):
        bn.add_node('C', cpt={})
    with pytest.raises(ValueError):
        bn.add_node('Q', cpt={}, parent_names=['nope'])
    with pytest.raises(ValueError):
        bn.get_joint_probability((0, 1))
    np.random.seed(0)
    dicts, tuples = get_sprinkler_network(False).sample_forward(5)
    assert len(dicts) == 5 and all(len(t) == 4 for t in tuples)


def _gloo_worker(rank, world, port, n, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from tensornetworks_b200 import ops
    bn, lat, obs, x = random_binary_bn(n, seed=2)
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q = np.random.default_rng(0).dirichlet(np.ones(1 << n))
    for N_eff in (1 << n, (1 << n) - 3):  # even split -> all_gather path; ragged -> all_reduce path
        r0, r1 = ops.shard_rows(N_eff, rank, world)
        u_rows, part = c_oracle.ksd_rows(q, S, n, r0, r1)   # oracle stands in for the CUDA rows in this CPU test
        u, K = ops.combine_row_shards(torch.from_numpy(u_rows), torch.tensor([part], dtype=torch.float64), N_eff, rank, world)
        ret[(rank, N_eff)] = (u.numpy().copy(), float(K.item()))
    dist.destroy_process_group()


def test_row_shard_exchange_gloo_world2():
    n, world, port = 7, 2, 29731
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gloo_worker, args=(world, port, n, ret), nprocs=world, join=True)
    bn, lat, obs, x = random_binary_bn(n, seed=2)
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q = np.random.default_rng(0).dirichlet(np.ones(1 << n))
    K, g = orc.ksd_dense(q, S, n)
    for rank in range(world):
        u, Kr = ret[(rank, 1 << n)]
        assert np.allclose(2 * u, g, rtol=1e-12) and np.isclose(Kr, K, rtol=1e-12)
        u, Kr = ret[(rank, (1 << n) - 3)]
        assert np.allclose(2 * u, g[:-3], rtol=1e-12)
        assert np.isclose(Kr, float(q[:-3] @ (g[:-3] / 2)), rtol=1e-12)


def _gloo_sym_worker(rank, world, port, n, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from tensornetworks_b200 import ops
    bn, lat, obs, x = random_binary_bn(n, seed=2)
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q = np.random.default_rng(0).dirichlet(np.ones(1 << n))
    N, T = 1 << n, (1 << n) // 128
    offset, stride = ops.sym_tile_stripe(rank, world)
    u = np.zeros(N)                                   # what tnk_ksd_sym(tile_offset, tile_stride) returns: the oracle's
    for I in range(offset, T, stride):                # dense blocks stand in for the CUDA tiles in this CPU test
        rows = np.arange(128 * I, 128 * I + 128)
        Kp = orc.stein_matrix_dense(S, n, 1.0, rows)
        for J in range(I, T):
            cols = slice(128 * J, 128 * J + 128)
            u[rows] += Kp[:, cols] @ q[cols]
            if J != I:
                u[cols] += Kp[:, cols].T @ q[rows]
    uu, K = ops.combine_partial_sums(torch.from_numpy(u), torch.tensor([float(q @ u)], dtype=torch.float64))
    ret[rank] = (uu.numpy().copy(), float(K.item()))
    dist.destroy_process_group()


def test_symmetric_stripe_exchange_gloo_world2():
    """N > 1 path of algorithm "pairs_sym": row tiles dealt round-robin, full-length partials, one sum-all-reduce."""
    n, world, port = 9, 2, 29741
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gloo_sym_worker, args=(world, port, n, ret), nprocs=world, join=True)
    bn, lat, obs, x = random_binary_bn(n, seed=2)
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q = np.random.default_rng(0).dirichlet(np.ones(1 << n))
    K, g = orc.ksd_dense(q, S, n)
    for rank in range(world):
        u, Kr = ret[rank]
        assert np.allclose(2 * u, g, rtol=1e-10, atol=1e-10 * np.max(np.abs(g))) and np.isclose(Kr, K, rtol=1e-10)
