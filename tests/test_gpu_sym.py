"""GPU parity tests of the symmetric tensor-core evaluation (tnk_ksd_sym / algorithm "pairs_sym"): every
unordered pair of 128-state tiles evaluated once, mirrored half from k_p(z,z') = k_p(z',z)
(reference stein_utils.py:161-196 is symmetric in its two arguments).  Tolerance: BASELINE.json's fp32 mode, 1e-4."""
import numpy as np
import pytest
import torch

from oracle import ksd_oracle as orc
from tensornetworks_b200.synthetic import random_binary_bn, random_q
from test_gpu_parity import _deterministic_bn

pytestmark = pytest.mark.gpu
TOL = 1e-4


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from tensornetworks_b200 import ops as _ops
    return _ops


def _eval(ops, tabs, q):
    qd = torch.as_tensor(q, dtype=torch.float64, device="cuda").requires_grad_(True)
    K = ops.ksd_quadratic_form(qd, tabs, algorithm="pairs_sym")
    K.backward()
    return float(K.item()), qd.grad.cpu().numpy()


def _check(K, g, Kref, gref, tol=TOL):
    assert abs(K - Kref) <= tol * abs(Kref), (K, Kref)
    assert np.max(np.abs(g - gref)) <= tol * np.max(np.abs(gref)), np.max(np.abs(g - gref)) / np.max(np.abs(gref))


@pytest.mark.parametrize("n,seed", [(7, 0), (8, 5), (9, 1), (10, 2), (11, 3), (12, 0)])
def test_sym_vs_dense_oracle(ops, n, seed):
    bn, lat, obs, x = random_binary_bn(n, seed=seed)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q = random_q(1 << n, seed=seed).numpy()
    Kref, gref = orc.ksd_dense(q, S, n)
    K, g = _eval(ops, tabs, q)
    _check(K, g, Kref, gref)


@pytest.mark.parametrize("n,seed", [(13, 1), (14, 2), (16, 0), (17, 3)])
def test_sym_vs_modeproduct_oracle(ops, n, seed):
    """n = 13..16 cover 4 chunks per field, n = 17 five (the n = 20 instance)."""
    bn, lat, obs, x = random_binary_bn(n, seed=seed)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q = random_q(1 << n, seed=seed).numpy()
    Kref, gref = orc.ksd_modeproduct(q, S, n)
    K, g = _eval(ops, tabs, q)
    _check(K, g, Kref, gref)


def test_sym_length_scale_and_hidden_nodes(ops):
    n = 9
    bn, lat, obs, x = random_binary_bn(n, seed=11, n_hidden=2)
    for ls in (0.5, 2.0):
        tabs = ops.SteinTables.from_network(bn, x, lat, length_scale=ls, precision="tc32")
        S = orc.score_table(orc.joint_table(bn, x, lat), n)
        q = random_q(1 << n, seed=11).numpy()
        Kref, gref = orc.ksd_dense(q, S, n, length_scale=ls)
        K, g = _eval(ops, tabs, q)
        _check(K, g, Kref, gref)


def test_sym_zero_probability_states(ops):
    n = 9
    bn, lat, x = _deterministic_bn(n)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q = random_q(1 << n, seed=4).numpy()
    Kref, gref = orc.ksd_dense(q, S, n)
    K, g = _eval(ops, tabs, q)
    _check(K, g, Kref, gref)


@pytest.mark.parametrize("stride", [2, 3, 8])
def test_sym_row_tile_stripes_add_up(ops, stride):
    """Row tiles dealt round-robin (what rank g of G evaluates) add up to the whole u and K; an offset beyond
    the last tile contributes zeros; a second call reuses the column table in the workspace."""
    n = 12
    bn, lat, obs, x = random_binary_bn(n, seed=5)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    ref = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
    q = random_q(1 << n, seed=5).cuda()
    u_ref, k_ref = ref.ksd_rows(q)
    u_full, k_full = tabs.ksd_sym(q)
    u_again, k_again = tabs.ksd_sym(q)
    parts = [tabs.ksd_sym(q, g, stride) for g in range(stride)]
    torch.cuda.synchronize()
    u_sum = sum(p[0] for p in parts)
    k_sum = sum(p[1].item() for p in parts)
    scale = u_ref.abs().max().item()
    assert (u_again - u_full).abs().max().item() <= 1e-12 * scale and abs(k_again.item() - k_full.item()) <= 1e-12 * abs(k_ref.item())
    assert (u_sum - u_full).abs().max().item() <= 1e-6 * scale      # fp32 chunk sums: the stripes round differently
    assert abs(k_sum - k_full.item()) <= 1e-6 * abs(k_ref.item())
    assert (u_full - u_ref).abs().max().item() <= TOL * scale
    assert abs(k_full.item() - k_ref.item()) <= TOL * abs(k_ref.item())
    u_none, k_none = tabs.ksd_sym(q, 1 << 20, 1 << 20)       # owns no row tile
    assert float(u_none.abs().max()) == 0.0 and float(k_none) == 0.0


def test_sym_q_dtypes_and_posterior(ops):
    n = 10
    bn, lat, obs, x = random_binary_bn(n, seed=3)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    ref = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
    q32 = random_q(1 << n, seed=3).to(torch.float32).cuda()
    u1, k1 = tabs.ksd_sym(q32)
    u0, k0 = ref.ksd_rows(q32)
    assert abs(k1.item() - k0.item()) <= TOL * abs(k0.item())
    assert (u1 - u0).abs().max().item() <= TOL * u0.abs().max().item()
    post = ref.posterior()
    u1, k1 = tabs.ksd_sym(post)
    u0, k0 = ref.ksd_rows(post)
    uni = torch.full_like(post, 1.0 / post.numel())
    _, kscale = ref.ksd_rows(uni)
    assert abs(k1.item() - k0.item()) <= TOL * abs(kscale.item())
    # CPU q (PennyLane's probabilities live on the host) through the public op
    q_cpu = random_q(1 << n, seed=3).requires_grad_(True)
    K = ops.ksd_quadratic_form(q_cpu, tabs, algorithm="pairs_sym")
    K.backward()
    assert K.device.type == "cpu" and q_cpu.grad.device.type == "cpu"
    u64, k64 = ref.ksd_rows(q_cpu.detach().cuda())
    assert abs(K.item() - k64.item()) <= TOL * abs(k64.item())
    assert (q_cpu.grad.cuda() - 2 * u64).abs().max().item() <= TOL * (2 * u64).abs().max().item()


def test_sym_rejects_unsupported(ops):
    bn, lat, obs, x = random_binary_bn(5, seed=0)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    with pytest.raises(RuntimeError):
        tabs.ksd_sym(random_q(32, seed=0).cuda())              # n < 7: no tile
    bn, lat, obs, x = random_binary_bn(8, seed=0)
    tabs64 = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
    with pytest.raises(ValueError):
        tabs64.ksd_sym(random_q(256, seed=0).cuda())           # needs the fp32 tables


def test_sym_matches_modeproduct_n20(ops):
    """Full size of BASELINE.json's metric: the symmetric tile kernel against the Kronecker kernels (two
    independent algorithms) and against the tile kernel that executes all 4^n ordered pairs."""
    n = 20
    bn, lat, obs, x = random_binary_bn(n, seed=0)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    q = random_q(1 << n, seed=0).cuda()
    u_mp, k_mp = tabs.ksd_modeproduct(q, storage="fp64")
    u_sy, k_sy = tabs.ksd_sym(q)
    u_tc, k_tc = tabs.ksd_rows(q)
    torch.cuda.synchronize()
    scale = u_mp.abs().max().item()
    assert abs(k_sy.item() - k_mp.item()) <= TOL * abs(k_mp.item())
    assert (u_sy - u_mp).abs().max().item() <= TOL * scale
    assert (u_sy - u_tc).abs().max().item() <= TOL * scale


def test_trainer_epochs_pairs_sym_match_fp64(ops):
    """KSDVariationalInference.train with algorithm="pairs_sym" (eager and CUDA-graph epochs) follows the fp64
    pair-kernel trajectory within the fp32-mode tolerance."""
    from tensornetworks_b200.ksd_vi import KSDVariationalInference
    n = 9
    bn, lat, obs, x = random_binary_bn(n, seed=5)
    hist = {}
    for name, kw in (("fp64", dict(precision="fp64", algorithm="pairs")),
                     ("sym", dict(precision="tc32", algorithm="pairs_sym")),
                     ("sym_graph", dict(precision="tc32", algorithm="pairs_sym", cuda_graph=True))):
        torch.manual_seed(1)
        vi = KSDVariationalInference(bn, lat, obs, {"use_logits": True, "conditioning_dim": 0}, device="cuda",
                                     compute_device="cuda", **kw)
        hist[name] = vi.train(x, num_epochs=10, lr_born_machine=0.05, verbose=False)
    for name in ("sym", "sym_graph"):
        for key in ("loss_ksd", "entropy", "grad_norm"):
            a, b = np.array(hist["fp64"][key], dtype=float), np.array(hist[name][key], dtype=float)
            assert a.shape == b.shape == (10,), key
            assert np.max(np.abs(a - b) / np.maximum(np.abs(a), 1e-12)) < 2e-3, (name, key, a, b)


@pytest.mark.parametrize("n,seed", [(7, 0), (10, 2), (12, 0), (13, 1), (16, 0), (17, 3)])
def test_sym_bf16x2_operands(ops, n, seed):
    """split="bf16x2": every real factor as two bf16 pieces (fewer MMA K-steps); errors ~1e-5, inside the 1e-4 mode."""
    bn, lat, obs, x = random_binary_bn(n, seed=seed)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    ref = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
    q = random_q(1 << n, seed=seed).cuda()
    u0, k0 = ref.ksd_rows(q) if n <= 14 else ref.ksd_modeproduct(q)
    u1, k1 = tabs.ksd_sym(q, split="bf16x2")
    ua, ka = tabs.ksd_sym(q, 0, 2, split="bf16x2")
    ub, kb = tabs.ksd_sym(q, 1, 2, split="bf16x2")
    torch.cuda.synchronize()
    scale = u0.abs().max().item()
    assert abs(k1.item() - k0.item()) <= TOL * abs(k0.item())
    assert (u1 - u0).abs().max().item() <= TOL * scale
    assert (ua + ub - u1).abs().max().item() <= 1e-6 * scale        # fp32 chunk sums: the stripes round differently
    with pytest.raises(ValueError):
        tabs.ksd_sym(q, split="fp8")


@pytest.mark.parametrize("split", ["tf32_tile", "tf32_dot"])
@pytest.mark.parametrize("n,seed", [(7, 0), (8, 1), (9, 2), (10, 2), (12, 0), (13, 1), (14, 4), (16, 0), (17, 3), (18, 1)])
def test_sym_tile_format(ops, n, seed, split):
    """split="tf32_tile": the T1/T2 terms of the bit positions >= 7 leave the pair kernel (9 instead of 13 K-steps at
    n = 20) and are added for all rows at once by a pass of fp64 Kronecker transforms; split="tf32_dot": only the score dot
    product stays (6 K-steps).  Same 1e-4 mode, stripes add up, results within 5e-6 of the 13-step kernel's."""
    bn, lat, obs, x = random_binary_bn(n, seed=seed)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    ref = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
    q = random_q(1 << n, seed=seed).cuda()
    u0, k0 = ref.ksd_rows(q) if n <= 14 else ref.ksd_modeproduct(q)
    u1, k1 = tabs.ksd_sym(q, split=split)
    u2, k2 = tabs.ksd_sym(q, split="tf32")
    parts = [tabs.ksd_sym(q, g, 3, split=split) for g in range(3)]
    torch.cuda.synchronize()
    scale = u0.abs().max().item()
    assert abs(k1.item() - k0.item()) <= TOL * abs(k0.item())
    assert (u1 - u0).abs().max().item() <= TOL * scale
    assert (u1 - u2).abs().max().item() <= 5e-6 * scale
    assert (sum(p[0] for p in parts) - u1).abs().max().item() <= 1e-6 * scale
    assert abs(sum(p[1].item() for p in parts) - k1.item()) <= 1e-6 * abs(k1.item())
    q32 = q.float()
    u3, k3 = tabs.ksd_sym(q32, split=split)
    assert (u3 - u1).abs().max().item() <= 1e-5 * scale


@pytest.mark.parametrize("split", ["tf32", "tf32_tile", "tf32_dot"])
@pytest.mark.parametrize("rowtiles,splits", [("1", "1"), ("2", "4"), ("4", "2"), ("4", "16")])
def test_row_tile_instances_and_column_splits_agree(ops, monkeypatch, split, rowtiles, splits):
    """Every instance of the tile kernel (TNK_TCS_ROWTILES = 1: one row tile per CTA, 2 / 4: tcs_tilesN_kernel) and any
    number of column splits gives the oracle's result, on stripes with a ragged last CTA (22 and 21 row tiles) and on the
    whole problem."""
    n = 13
    monkeypatch.setenv("TNK_TCS_ROWTILES", rowtiles)
    monkeypatch.setenv("TNK_SYM_SPLITS", splits)
    bn, lat, obs, x = random_binary_bn(n, seed=4)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    S = orc.score_table(orc.joint_table(bn, x, lat), n)
    q = random_q(1 << n, seed=4)
    Kref, gref = orc.ksd_modeproduct(q.numpy(), S, n)
    qd = q.cuda()
    u = torch.zeros(1 << n, dtype=torch.float64, device="cuda")
    K = 0.0
    for off in range(3):                                   # 64 tiles dealt to 3 "ranks": 22, 21, 21 row tiles
        up, part = tabs.ksd_sym(qd, off, 3, split=split)
        u += up
        K += float(part.item())
    _check(K, 2.0 * u.cpu().numpy(), Kref, gref)
    up, part = tabs.ksd_sym(qd, split=split)               # 64 row tiles, one call
    _check(float(part.item()), 2.0 * up.cpu().numpy(), Kref, gref)
