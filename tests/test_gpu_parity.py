"""GPU parity tests: the CUDA path (through the C-ABI) against the oracle and the
reference-generated goldens.  Tolerances are BASELINE.json's: enumeration bit-exact;
KSD^2 and gradient (normwise) 1e-6 in fp64 mode, 1e-4 in fp32 mode."""
import numpy as np
import pytest
import torch

from oracle import ksd_oracle as orc
from tensornetworks_b200.bayesian_network import BayesianNetwork, get_sprinkler_network
from tensornetworks_b200.synthetic import random_binary_bn, random_q

pytestmark = pytest.mark.gpu

TOL = {"fp64": 1e-6, "fp32": 1e-4, "fp64_simple": 1e-6, "tc32": 1e-4}
RANDOM_CASES = [(n, n, 0) for n in range(2, 7)] + [(4, 104, 2)]
TABLE_CASES = [(10, 0, 0), (10, 1, 2), (8, 7, 1)]


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from tensornetworks_b200 import ops as _ops
    return _ops


def _tables(ops, bn, x, lat, precision, length_scale=1.0):
    return ops.SteinTables.from_network(bn, x, lat, length_scale=length_scale, precision=precision, keep_scores=True)


def _check(K, g, Kref, gref, tol, Kscale=None):
    scale = abs(Kref) if Kscale is None else Kscale
    assert abs(K - Kref) <= tol * scale, (K, Kref)
    assert np.max(np.abs(g - gref)) <= tol * np.max(np.abs(gref)), np.max(np.abs(g - gref)) / np.max(np.abs(gref))


def _eval(ops, tabs, q):
    qd = torch.as_tensor(q, dtype=torch.float64, device="cuda").requires_grad_(True)
    K = ops.ksd_quadratic_form(qd, tabs)
    K.backward()
    return float(K.item()), qd.grad.cpu().numpy()


def test_known_answers_ab(ops, golden):
    bn = BayesianNetwork()
    bn.add_node('A', cpt={(): {0: 0.8, 1: 0.2}})
    bn.add_node('B', cpt={(0,): {0: 0.7, 1: 0.3}, (1,): {0: 0.4, 1: 0.6}}, parent_names=['A'])
    tabs = _tables(ops, bn, {'B': 1}, ['A'], "fp64")
    assert np.array_equal(tabs.P.cpu().numpy(), golden["ab_P"])
    assert np.array_equal(tabs.S.cpu().numpy(), golden["ab_S"])
    # k_p(0,1) and k_p(0,0) of stein_utils.py:245-251 through one-hot q
    e0 = torch.tensor([1.0, 0.0], dtype=torch.float64, device="cuda")
    u, _ = tabs.ksd_rows(e0)
    torch.cuda.synchronize()
    assert np.isclose(u[0].item(), 1.25 - np.exp(-1.0), rtol=1e-12)
    assert np.isclose(u[1].item(), 2 * np.exp(-1.0) - 2.5, rtol=1e-12)


@pytest.mark.parametrize("precision", ["fp64", "fp32", "fp64_simple", "tc32"])
def test_sprinkler(ops, golden, precision):
    bn = get_sprinkler_network(False)
    tabs = _tables(ops, bn, {'W': 1}, ['C', 'S', 'R'], precision)
    assert np.array_equal(tabs.P.cpu().numpy(), golden["spr_P"])
    assert np.array_equal(tabs.S.cpu().numpy(), golden["spr_S"])
    for name in ("uniform", "ramp"):
        K, g = _eval(ops, tabs, golden[f"spr_q_{name}"])
        _check(K, g, float(golden[f"spr_K_{name}"]), golden[f"spr_g_{name}"], TOL[precision])
    # q = exact posterior: K ~ 0, compare against the scale of the uniform-q value
    K, g = _eval(ops, tabs, golden["spr_q_post"])
    assert abs(K - float(golden["spr_K_post"])) <= TOL[precision] * float(golden["spr_K_uniform"])
    for ls in (0.5, 2.0):
        t2 = _tables(ops, bn, {'W': 1}, ['C', 'S', 'R'], precision, length_scale=ls)
        K, g = _eval(ops, t2, golden["spr_q_ramp"])
        _check(K, g, float(golden[f"spr_K_ramp_ls{ls}"]), golden[f"spr_g_ramp_ls{ls}"], TOL[precision])


@pytest.mark.parametrize("precision", ["fp64", "fp32", "fp64_simple"])
@pytest.mark.parametrize("n,seed,nh", RANDOM_CASES)
def test_random_bn_vs_reference_loop(ops, golden, n, seed, nh, precision):
    bn, lat, obs, x = random_binary_bn(n, seed=seed, n_hidden=nh)
    tag = f"rnd_n{n}_s{seed}_h{nh}"
    tabs = _tables(ops, bn, x, lat, precision)
    assert np.array_equal(tabs.P.cpu().numpy(), golden[tag + "_P"])
    assert np.array_equal(tabs.S.cpu().numpy(), golden[tag + "_S"])
    K, g = _eval(ops, tabs, golden[tag + "_q"])
    _check(K, g, float(golden[tag + "_K"]), golden[tag + "_g"], TOL[precision])


@pytest.mark.parametrize("n,seed,nh", TABLE_CASES)
def test_enumeration_tables_bit_exact(ops, golden, n, seed, nh):
    bn, lat, obs, x = random_binary_bn(n, seed=seed, n_hidden=nh)
    tag = f"tab_n{n}_s{seed}_h{nh}"
    tabs = _tables(ops, bn, x, lat, "fp64")
    assert np.array_equal(tabs.P.cpu().numpy(), golden[tag + "_P"])
    assert np.array_equal(tabs.S.cpu().numpy(), golden[tag + "_S"])


@pytest.mark.parametrize("precision", ["fp64", "fp32", "tc32"])
@pytest.mark.parametrize("n,seed", [(7, 0), (9, 1), (10, 2), (11, 3), (12, 0)])
def test_vs_dense_oracle(ops, n, seed, precision):
    bn, lat, obs, x = random_binary_bn(n, seed=seed)
    tabs = _tables(ops, bn, x, lat, precision)
    P = orc.joint_table(bn, x, lat)
    S = orc.score_table(P, n)
    assert np.array_equal(tabs.P.cpu().numpy(), P)
    assert np.array_equal(tabs.S.cpu().numpy(), S)
    q = random_q(1 << n, seed=seed).numpy()
    Kref, gref = orc.ksd_dense(q, S, n)
    K, g = _eval(ops, tabs, q)
    _check(K, g, Kref, gref, TOL[precision])


@pytest.mark.parametrize("precision", ["fp64", "fp32", "tc32"])
@pytest.mark.parametrize("n,seed", [(13, 1), (14, 2), (16, 0)])
def test_vs_modeproduct_oracle(ops, n, seed, precision):
    bn, lat, obs, x = random_binary_bn(n, seed=seed)
    tabs = _tables(ops, bn, x, lat, precision)
    P = orc.joint_table(bn, x, lat)
    S = orc.score_table(P, n)
    assert np.array_equal(tabs.P.cpu().numpy(), P)
    assert np.array_equal((np.abs(P) < 1e-12), (tabs.S.cpu().numpy() == 0).all(axis=1) & (np.abs(P) < 1e-12))
    assert np.array_equal(tabs.S.cpu().numpy(), S)
    q = random_q(1 << n, seed=seed).numpy()
    Kref, gref = orc.ksd_modeproduct(q, S, n)
    K, g = _eval(ops, tabs, q)
    _check(K, g, Kref, gref, TOL[precision])


def _deterministic_bn(n):
    """Chain Z0 -> ... -> Z(n-1) in which Z3 copies Z2 and Z6 is the negation of Z5 (CPT entries exactly 0 and 1):
    three quarters of the states have p(x,z) = 0 (zero score rows, stein_utils.py:126-128) and live states have
    dead neighbours (ratio 0, score 1)."""
    bn = BayesianNetwork()
    bn.add_node('Z0', cpt={(): {0: 0.35, 1: 0.65}})
    for k in range(1, n):
        if k == 3:
            cpt = {(0,): {0: 1.0, 1: 0.0}, (1,): {0: 0.0, 1: 1.0}}
        elif k == 6:
            cpt = {(0,): {0: 0.0, 1: 1.0}, (1,): {0: 1.0, 1: 0.0}}
        else:
            cpt = {(0,): {0: 0.2 + 0.05 * k, 1: 0.8 - 0.05 * k}, (1,): {0: 0.6 - 0.03 * k, 1: 0.4 + 0.03 * k}}
        bn.add_node(f'Z{k}', cpt=cpt, parent_names=[f'Z{k-1}'])
    bn.add_node('X', cpt={(0, 0): {0: 0.9, 1: 0.1}, (0, 1): {0: 0.5, 1: 0.5}, (1, 0): {0: 0.3, 1: 0.7},
                          (1, 1): {0: 0.05, 1: 0.95}}, parent_names=['Z0', f'Z{n-1}'])
    return bn, [f'Z{k}' for k in range(n)], {'X': 1}


@pytest.mark.parametrize("precision", ["fp64", "fp32", "tc32"])
def test_zero_probability_states(ops, precision):
    n = 9
    bn, lat, x = _deterministic_bn(n)
    tabs = _tables(ops, bn, x, lat, precision)
    P = orc.joint_table(bn, x, lat)
    S = orc.score_table(P, n)
    assert (P == 0).sum() == 3 * (1 << n) // 4 and np.all(S[P == 0] == 0) and np.any(S == 1.0)
    assert np.array_equal(tabs.P.cpu().numpy(), P)
    assert np.array_equal(tabs.S.cpu().numpy(), S)
    q = random_q(1 << n, seed=4).numpy()
    Kref, gref = orc.ksd_dense(q, S, n)
    K, g = _eval(ops, tabs, q)
    _check(K, g, Kref, gref, TOL[precision])
    Km, gm = _eval_mp(ops, tabs, q)                      # mode-product, storage following the table precision
    _check(Km, gm, Kref, gref, 1e-9 if precision == "fp64" else 1e-5)


def test_row_ranges_add_up(ops):
    """Sharding property: disjoint row ranges reproduce the full u and K (incl. ranges that are
    not aligned to the kernel's row block)."""
    n = 12
    bn, lat, obs, x = random_binary_bn(n, seed=5)
    tabs = _tables(ops, bn, x, lat, "fp64")
    q = random_q(1 << n, seed=5).cuda()
    u_full, k_full = tabs.ksd_rows(q)
    u = torch.zeros_like(u_full)
    parts = []
    cuts = [0, 1000, 1024, 2500, 4096]
    for a, b in zip(cuts[:-1], cuts[1:]):
        _, p = tabs.ksd_rows(q, a, b, u=u)
        parts.append(p.clone())
    torch.cuda.synchronize()
    assert torch.allclose(u, u_full, rtol=1e-13, atol=0)
    assert np.isclose(sum(p.item() for p in parts), k_full.item(), rtol=1e-12)


def test_q_dtypes_and_cpu_q(ops, golden):
    """q may be fp32 or fp64, on CPU (PennyLane) or CUDA; grad comes back in q's dtype/device."""
    bn = get_sprinkler_network(False)
    tabs = _tables(ops, bn, {'W': 1}, ['C', 'S', 'R'], "fp64")
    q32 = torch.tensor(golden["spr_q_ramp"], dtype=torch.float32, requires_grad=True)  # CPU fp32
    K = ops.ksd_quadratic_form(q32, tabs)
    K.backward()
    assert K.device.type == "cpu" and K.dtype == torch.float64
    assert q32.grad.dtype == torch.float32 and q32.grad.device.type == "cpu"
    Kref, gref = orc.ksd_dense(q32.detach().double().numpy(), golden["spr_S"], 3)
    assert np.isclose(K.item(), Kref, rtol=1e-9)
    assert np.allclose(q32.grad.numpy(), gref, rtol=1e-5)
    # chain rule through sqrt(clamp(K)) as in ksd_vi.py:133-134
    q = torch.tensor(golden["spr_q_ramp"], dtype=torch.float64, device="cuda", requires_grad=True)
    loss = torch.sqrt(ops.ksd_quadratic_form(q, tabs).clamp(min=1e-12))
    loss.backward()
    assert np.allclose(q.grad.cpu().numpy(), golden["spr_g_ramp"] / (2 * np.sqrt(golden["spr_K_ramp"])), rtol=1e-9)


def test_cpu_q_gradient_arrives_with_the_value(ops, golden):
    """A host q that wants a gradient gets dK/dq downloaded behind K (one synchronisation); scaled upstream gradients,
    a second backward through a retained graph and q that needs no gradient all give the reference's numbers."""
    bn = get_sprinkler_network(False)
    tabs = _tables(ops, bn, {'W': 1}, ['C', 'S', 'R'], "fp64")
    gref, Kref = golden["spr_g_ramp"], float(golden["spr_K_ramp"])
    q = torch.tensor(golden["spr_q_ramp"], dtype=torch.float64, requires_grad=True)          # CPU fp64
    loss = torch.sqrt(ops.ksd_quadratic_form(q, tabs).clamp(min=1e-12))                       # upstream gradient != 1
    loss.backward()
    assert np.allclose(q.grad.numpy(), gref / (2 * np.sqrt(Kref)), rtol=1e-9)
    q.grad = None
    K = ops.ksd_quadratic_form(q, tabs)
    K.backward(retain_graph=True)
    assert np.allclose(q.grad.numpy(), gref, rtol=1e-9)
    K.backward()                                                                              # second pass: accumulates
    assert np.allclose(q.grad.numpy(), 2 * gref, rtol=1e-9)
    with torch.no_grad():
        assert np.isclose(ops.ksd_quadratic_form(q, tabs).item(), Kref, rtol=1e-9)
    assert np.isclose(ops.ksd_quadratic_form(q.detach(), tabs).item(), Kref, rtol=1e-9)


# ---------------------------------------------------------------- mode-product (Kronecker) path
def _eval_mp(ops, tabs, q):
    qd = torch.as_tensor(q, dtype=torch.float64, device="cuda").requires_grad_(True)
    K = ops.ksd_quadratic_form(qd, tabs, algorithm="modeproduct")
    K.backward()
    return float(K.item()), qd.grad.cpu().numpy()


def test_modeproduct_sprinkler_golden(ops, golden):
    bn = get_sprinkler_network(False)
    tabs = _tables(ops, bn, {'W': 1}, ['C', 'S', 'R'], "fp64")
    for name in ("uniform", "ramp"):
        K, g = _eval_mp(ops, tabs, golden[f"spr_q_{name}"])
        _check(K, g, float(golden[f"spr_K_{name}"]), golden[f"spr_g_{name}"], 1e-9)
    t2 = _tables(ops, bn, {'W': 1}, ['C', 'S', 'R'], "fp32", length_scale=0.5)   # fp32 tables build ST64 lazily
    K, g = _eval_mp(ops, t2, golden["spr_q_ramp"])                               # and default to fp32 storage
    _check(K, g, float(golden["spr_K_ramp_ls0.5"]), golden["spr_g_ramp_ls0.5"], 1e-5)
    u, part = t2.ksd_modeproduct(torch.as_tensor(golden["spr_q_ramp"], dtype=torch.float64, device="cuda"), storage="fp64")
    _check(float(part), 2.0 * u.cpu().numpy(), float(golden["spr_K_ramp_ls0.5"]), golden["spr_g_ramp_ls0.5"], 1e-9)


@pytest.mark.parametrize("n,seed,nh", RANDOM_CASES)
def test_modeproduct_vs_reference_loop(ops, golden, n, seed, nh):
    bn, lat, obs, x = random_binary_bn(n, seed=seed, n_hidden=nh)
    tag = f"rnd_n{n}_s{seed}_h{nh}"
    tabs = _tables(ops, bn, x, lat, "fp64")
    K, g = _eval_mp(ops, tabs, golden[tag + "_q"])
    _check(K, g, float(golden[tag + "_K"]), golden[tag + "_g"], 1e-9)


@pytest.mark.parametrize("n,seed", [(1, 0), (7, 0), (11, 3), (12, 0), (13, 1), (14, 2), (16, 0), (18, 1)])
def test_modeproduct_vs_oracle(ops, n, seed):
    """Covers every pass plan: single tile (n<=12), one strided pass (13..21) incl. 1-bit and 6-bit ones."""
    bn, lat, obs, x = random_binary_bn(n, seed=seed)
    tabs = _tables(ops, bn, x, lat, "fp64")
    S = tabs.S.cpu().numpy()
    q = random_q(1 << n, seed=seed).numpy()
    Kref, gref = (orc.ksd_dense if n <= 12 else orc.ksd_modeproduct)(q, S, n)
    K, g = _eval_mp(ops, tabs, q)
    _check(K, g, Kref, gref, 1e-9)


@pytest.mark.parametrize("n,limit", [(15, 1), (17, 2), (18, 3), (19, 2), (20, 4)])
def test_modeproduct_multi_pass_plans(ops, n, limit):
    """The plans of n > 24 (in-place middle passes, natural V order, unblocked last pass) reached at small n by
    limiting the row bits per strided pass; compared with the fp64 pair kernel and with bit-range partials."""
    from tensornetworks_b200 import _lib
    lib = _lib.load()
    bn, lat, obs, x = random_binary_bn(n, seed=n)
    q = random_q(1 << n, seed=n).cuda()
    ref = _tables(ops, bn, x, lat, "fp64")
    u0, K0 = ref.ksd_rows(q)
    prev = lib.tnk_modeproduct_plan_limit(limit)
    try:
        tabs = _tables(ops, bn, x, lat, "fp64")          # fresh tables: the column copies follow the plan
        u1, K1 = tabs.ksd_modeproduct(q)
        ua, Ka = tabs.ksd_modeproduct(q, 0, n // 2)
        ub, Kb = tabs.ksd_modeproduct(q, n // 2, n)
        torch.cuda.synchronize()
    finally:
        lib.tnk_modeproduct_plan_limit(prev)
    den = float(u0.abs().max())
    assert float((u1 - u0).abs().max()) <= 1e-12 * den
    assert float((ua + ub - u0).abs().max()) <= 1e-12 * den
    assert abs(float(K1) - float(K0)) <= 1e-12 * abs(float(K0))
    assert abs(float(Ka + Kb) - float(K0)) <= 1e-12 * abs(float(K0))


@pytest.mark.parametrize("n,seed", [(5, 1), (12, 0), (13, 2), (15, 0), (18, 1), (20, 0)])
def test_modeproduct_fp32_storage(ops, n, seed):
    """storage="fp32" (score columns and transformed vectors in fp32 in HBM, fp64 arithmetic) against the fp64
    storage: the fp32-mode tolerance with a wide margin, for fp64 and fp32 q, full and split bit ranges."""
    bn, lat, obs, x = random_binary_bn(n, seed=seed)
    tabs = _tables(ops, bn, x, lat, "fp64")
    q = random_q(1 << n, seed=seed).cuda()
    u64, k64 = tabs.ksd_modeproduct(q)                      # fp64 tables default to fp64 storage
    u32, k32 = tabs.ksd_modeproduct(q, storage="fp32")
    ua, ka = tabs.ksd_modeproduct(q.float(), 0, n // 2, storage="fp32")
    ub, kb = tabs.ksd_modeproduct(q.float(), n // 2, n, storage="fp32")
    torch.cuda.synchronize()
    den = u64.abs().max().item()
    assert (u32 - u64).abs().max().item() <= 2e-6 * den
    assert abs(k32.item() - k64.item()) <= 2e-6 * abs(k64.item())
    assert (ua + ub - u64).abs().max().item() <= 2e-6 * den
    t32 = _tables(ops, bn, x, lat, "fp32")                  # fp32 tables default to fp32 storage
    ud, kd = t32.ksd_modeproduct(q)
    assert (ud - u64).abs().max().item() <= 2e-6 * den
    with pytest.raises(ValueError):
        tabs.ksd_modeproduct(q, storage="bf16")


def test_modeproduct_matches_pair_kernel_n20(ops):
    """Full size of BASELINE.json's metric: the two CUDA algorithms agree (no CPU oracle at this size
    in test time), plus the bit-range partials add up (the multi-GPU split)."""
    n = 20
    bn, lat, obs, x = random_binary_bn(n, seed=0)
    tabs = _tables(ops, bn, x, lat, "fp64")
    q = random_q(1 << n, seed=0).cuda()
    u_mp, k_mp = tabs.ksd_modeproduct(q)
    ua, ka = tabs.ksd_modeproduct(q, 0, 7)
    ub, kb = tabs.ksd_modeproduct(q, 7, n)
    torch.cuda.synchronize()
    assert torch.allclose(ua + ub, u_mp, rtol=1e-12, atol=1e-9 * u_mp.abs().max().item())
    assert np.isclose(ka.item() + kb.item(), k_mp.item(), rtol=1e-12)
    u_pair, k_pair = tabs.ksd_rows(q)
    torch.cuda.synchronize()
    assert abs(k_pair.item() - k_mp.item()) <= 1e-6 * abs(k_mp.item())
    assert (u_pair - u_mp).abs().max().item() <= 1e-6 * u_mp.abs().max().item()


def test_tensor_core_row_ranges_and_splits(ops):
    """tc32: disjoint row ranges (not aligned to the 128-row tile) reproduce the full u and K, and
    the tensor-core result agrees with the fp64 pair kernel within the fp32-mode tolerance."""
    n = 12
    bn, lat, obs, x = random_binary_bn(n, seed=5)
    tabs = _tables(ops, bn, x, lat, "tc32")
    ref = _tables(ops, bn, x, lat, "fp64")
    q = random_q(1 << n, seed=5).cuda()
    u_full, k_full = tabs.ksd_rows(q)
    u_ref, k_ref = ref.ksd_rows(q)
    u = torch.zeros_like(u_full)
    parts = []
    cuts = [0, 100, 128, 1000, 2500, 4096]
    for a, b in zip(cuts[:-1], cuts[1:]):
        _, p = tabs.ksd_rows(q, a, b, u=u)
        parts.append(p.clone())
    torch.cuda.synchronize()
    assert torch.equal(u, u_full)          # same tiles, same order of operations per row
    assert np.isclose(sum(p.item() for p in parts), k_full.item(), rtol=1e-12)
    assert abs(k_full.item() - k_ref.item()) <= 1e-4 * abs(k_ref.item())
    assert (u_full - u_ref).abs().max().item() <= 1e-4 * u_ref.abs().max().item()


def test_tensor_core_q_dtypes_and_posterior(ops):
    """tc32 with fp32 q (the classical Born machine's dtype) and at q = posterior (KSD^2 ~ 0: the
    gradient must still agree normwise with the fp64 kernel; K is compared on the scale of |q|^T|K_p||q|)."""
    n = 10
    bn, lat, obs, x = random_binary_bn(n, seed=3)
    tabs = _tables(ops, bn, x, lat, "tc32")
    ref = _tables(ops, bn, x, lat, "fp64")
    q32 = random_q(1 << n, seed=3).to(torch.float32).cuda()
    u1, k1 = tabs.ksd_rows(q32)
    u0, k0 = ref.ksd_rows(q32)
    assert abs(k1.item() - k0.item()) <= 1e-4 * abs(k0.item())
    assert (u1 - u0).abs().max().item() <= 1e-4 * u0.abs().max().item()
    post = ref.posterior()
    u1, k1 = tabs.ksd_rows(post)
    u0, k0 = ref.ksd_rows(post)
    uni = torch.full_like(post, 1.0 / post.numel())
    _, kscale = ref.ksd_rows(uni)
    assert abs(k1.item() - k0.item()) <= 1e-4 * abs(kscale.item())


def test_tensor_core_matches_modeproduct_n20(ops):
    """Full size of BASELINE.json's metric through the tensor-core variant."""
    n = 20
    bn, lat, obs, x = random_binary_bn(n, seed=0)
    tabs = _tables(ops, bn, x, lat, "tc32")
    q = random_q(1 << n, seed=0).cuda()
    u_mp, k_mp = tabs.ksd_modeproduct(q)
    u_tc, k_tc = tabs.ksd_rows(q)
    torch.cuda.synchronize()
    assert abs(k_tc.item() - k_mp.item()) <= 1e-4 * abs(k_mp.item())
    assert (u_tc - u_mp).abs().max().item() <= 1e-4 * u_mp.abs().max().item()


def test_fused_peer_exchange_world2(ops):
    """tnk_ksd_rows_peers on 2 GPUs of the box: same u and K as the NCCL path on every rank."""
    import os, subprocess, sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs on one box")
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29531",
           os.path.join(repo, "scripts", "check_fused_exchange.py"), "12", "fp64"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    assert '"fused_exchange": "ok"' in res.stdout


def test_tnk_comm_world2(ops):
    """tnk_comm (include/tnk.h) on 2 GPUs of the box: bit-identical to the NCCL all-reduce of the same partial vectors,
    and the sharded pairs_sym / pairs_sym64 / modeproduct evaluations through it against the single-rank results."""
    import os, subprocess, sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs on one box")
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29533",
           os.path.join(repo, "scripts", "check_comm.py"), "12"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    assert '"tnk_comm": "ok"' in res.stdout
