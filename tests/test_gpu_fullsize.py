"""GPU parity at the sizes the metric and the configs are quoted on, against the CPU oracle (not against another
CUDA kernel):

  n = 20, seeds {0,1,2}   BASELINE.json configs[3] - every evaluation path of the library against the C mode-product
                          oracle (oracle/ksd_oracle.c:orc_ksd_modeproduct, itself checked here against rows of the
                          literal four-term loop of reference stein_utils.py:161-196)
  n = 24, seed 0          BASELINE.json configs[4] - the symmetric tile kernel over the whole pair space, the Kronecker
                          kernels, and the fp64 / fp32 / tcgen05 all-pairs kernels on row ranges taken from the first,
                          a middle and the last row tile
  ill-conditioned family  joint tables straddling the 1e-12 score threshold of stein_utils.py:126-128 (dead rows next
                          to live ones, scores up to 1e8), n = 12 and 16, plus q = posterior

Tolerances are BASELINE.json's: joint table bit-exact; KSD^2 relative and gradient normwise 1e-6 in fp64 mode, 1e-4 in
the fp32-grade modes (fp32, tc32, pairs_sym)."""
import numpy as np
import pytest
import torch

from oracle import c_oracle
from oracle import ksd_oracle as orc
from tensornetworks_b200.flatten import flatten_bn
from tensornetworks_b200.synthetic import illconditioned_binary_bn, random_binary_bn, random_q

pytestmark = pytest.mark.gpu

# (tables precision, algorithm, keyword arguments of the evaluation, tolerance)
PATHS = {
    "fp64-pairs": ("fp64", "rows", {}, 1e-6),
    "fp64-pairs_sym64": ("fp64", "sym64", {}, 1e-6),
    "fp32-pairs": ("fp32", "rows", {}, 1e-4),
    "tc32-pairs": ("tc32", "rows", {}, 1e-4),
    "tc32-pairs_sym-tf32": ("tc32", "sym", {"split": "tf32"}, 1e-4),
    "tc32-pairs_sym-bf16x2": ("tc32", "sym", {"split": "bf16x2"}, 1e-4),
    "tc32-pairs_sym-tf32_tile": ("tc32", "sym", {"split": "tf32_tile"}, 1e-4),
    "tc32-pairs_sym-tf32_dot": ("tc32", "sym", {"split": "tf32_dot"}, 1e-4),
    "modeproduct-fp64": ("fp64", "mp", {"storage": "fp64"}, 1e-6),
    "modeproduct-fp32": ("tc32", "mp", {"storage": "fp32"}, 1e-4),
}


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from tensornetworks_b200 import ops as _ops
    return _ops


def _run(tabs, kind, q, kw):
    if kind == "rows":
        u, k = tabs.ksd_rows(q)
    elif kind == "sym":
        u, k = tabs.ksd_sym(q, **kw)
    elif kind == "sym64":
        u, k = tabs.ksd_sym64(q)
    else:
        u, k = tabs.ksd_modeproduct(q, **kw)
    torch.cuda.synchronize()
    return float(k.item()), u.cpu().numpy()


def _assert_close(K, u, Kref, uref, tol, Kscale=None, uscale=None):
    Kscale = abs(Kref) if Kscale is None else Kscale
    uscale = np.max(np.abs(uref)) if uscale is None else uscale
    assert abs(K - Kref) <= tol * Kscale, (K, Kref, abs(K - Kref) / Kscale)
    err = np.max(np.abs(u - uref)) / uscale
    assert err <= tol, err


_CACHE = {}


def _case(n, seed, gen=random_binary_bn):
    """Network, oracle joint table, q and the oracle's (K, u) for one (family, n, seed), computed once per module."""
    key = (gen.__name__, n, seed)
    if key not in _CACHE:
        bn, lat, obs, x = gen(n, seed=seed)
        P = c_oracle.joint_table(flatten_bn(bn, x, lat))
        q = random_q(1 << n, seed=seed).numpy()
        K, g = c_oracle.ksd_modeproduct(q, P, n)
        if len(_CACHE) >= 2:                       # keep at most two cases alive (n = 24: 0.5 GB each)
            _CACHE.pop(next(iter(_CACHE)))
        _CACHE[key] = (bn, lat, x, P, q, K, g / 2.0)
    return _CACHE[key]


# ------------------------------------------------------------------------------------------ n = 20, three seeds
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_n20_oracle_is_pinned_to_the_literal_loop(seed):
    """The C mode-product oracle at n = 20 against rows of the literal four-term pair loop (first, middle, last tile)."""
    n = 20
    bn, lat, x, P, q, Kref, uref = _case(n, seed)
    S = c_oracle.score_table(P, n)
    for r0 in (0, (1 << n) // 2 - 4, (1 << n) - 8):
        u, _ = c_oracle.ksd_rows(q, S, n, r0, r0 + 8)
        assert np.max(np.abs(u - uref[r0:r0 + 8])) <= 1e-12 * np.max(np.abs(uref))


@pytest.mark.parametrize("path", list(PATHS))
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_n20_every_path_vs_cpu_oracle(ops, seed, path):
    n = 20
    precision, kind, kw, tol = PATHS[path]
    bn, lat, x, P, q, Kref, uref = _case(n, seed)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision=precision)
    assert np.array_equal(tabs.P.cpu().numpy(), P)
    K, u = _run(tabs, kind, torch.from_numpy(q).cuda(), kw)
    _assert_close(K, u, Kref, uref, tol)


def test_n20_public_op_vs_cpu_oracle(ops):
    """The drop-in call itself (autograd Function, host q in, host gradient out) at the metric's size."""
    n = 20
    bn, lat, x, P, q, Kref, uref = _case(n, 0)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    qh = torch.from_numpy(q).requires_grad_(True)
    K = ops.ksd_quadratic_form(qh, tabs, algorithm="pairs_sym")
    K.backward()
    _assert_close(float(K.item()), qh.grad.numpy() / 2.0, Kref, uref, 1e-4)


# ------------------------------------------------------------------------------------------ n = 24, one seed
N24_RANGES = lambda N: [(0, 1024), (N // 2 - 512, N // 2 + 512), (N - 1024, N)]   # first, middle, last row tiles


@pytest.fixture(scope="module")
def case24():
    return _case(24, 0)


def test_n24_oracle_is_pinned_to_the_literal_loop(case24):
    n = 24
    bn, lat, x, P, q, Kref, uref = case24
    N = 1 << n
    rows = [0, 127, N // 2 - 1, N // 2, N - 128, N - 1]        # both ends of the first / last tile and both triangle halves
    for r in rows:
        u, _ = c_oracle.ksd_rows(q, _score_rows(P, n), n, r, r + 1)
        assert abs(u[0] - uref[r]) <= 1e-12 * np.max(np.abs(uref))


_S24 = {}


def _score_rows(P, n):
    if "S" not in _S24:
        _S24["S"] = c_oracle.score_table(P, n)
    return _S24["S"]


def test_n24_symmetric_tiles_and_modeproduct_vs_cpu_oracle(ops, case24):
    """The whole 2^24 x 2^24 pair space through the symmetric tile kernel (the kernel of config #5's epochs) and
    through the Kronecker kernels, every row compared with the oracle."""
    n = 24
    bn, lat, x, P, q, Kref, uref = case24
    _S24.clear()
    tabs = ops.SteinTables.from_network(bn, x, lat, precision="tc32")
    assert np.array_equal(tabs.P.cpu().numpy(), P)
    qd = torch.from_numpy(q).cuda()
    K, u = _run(tabs, "sym", qd, {"split": "tf32"})
    _assert_close(K, u, Kref, uref, 1e-4)
    tabs._ws.clear()
    tabs._sym_tables = set()
    for split in ("tf32_tile", "tf32_dot"):
        K, u = _run(tabs, "sym", qd, {"split": split})
        _assert_close(K, u, Kref, uref, 1e-4)
        tabs._ws.clear()
        tabs._sym_tables = set()
    K, u = _run(tabs, "mp", qd, {"storage": "fp32"})
    _assert_close(K, u, Kref, uref, 1e-4)
    tabs._ws.clear()
    del tabs
    t64 = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
    K, u = _run(t64, "mp", qd, {"storage": "fp64"})
    _assert_close(K, u, Kref, uref, 1e-6)


@pytest.mark.parametrize("precision,tol", [("fp64", 1e-6), ("fp32", 1e-4), ("tc32", 1e-4)])
def test_n24_pair_kernels_on_row_ranges_vs_cpu_oracle(ops, case24, precision, tol):
    n = 24
    bn, lat, x, P, q, Kref, uref = case24
    tabs = ops.SteinTables.from_network(bn, x, lat, precision=precision)
    qd = torch.from_numpy(q).cuda()
    scale = np.max(np.abs(uref))
    for r0, r1 in N24_RANGES(1 << n):
        u, part = tabs.ksd_rows(qd, r0, r1)
        torch.cuda.synchronize()
        got = u[r0:r1].cpu().numpy()
        assert np.max(np.abs(got - uref[r0:r1])) <= tol * scale
        pref = float(np.dot(q[r0:r1], uref[r0:r1]))
        assert abs(part.item() - pref) <= tol * max(abs(pref), abs(Kref) * (r1 - r0) / (1 << n))
        tabs._ws.clear()


# ------------------------------------------------------------------------------------------ ill-conditioned family
ILL_PATHS = ["fp64-pairs", "fp64-pairs_sym64", "fp32-pairs", "tc32-pairs", "tc32-pairs_sym-tf32", "tc32-pairs_sym-tf32_tile",
             "tc32-pairs_sym-tf32_dot", "modeproduct-fp64"]


@pytest.mark.parametrize("path", ILL_PATHS)
@pytest.mark.parametrize("n,seed", [(12, 0), (12, 1), (12, 2), (16, 0), (16, 1), (16, 2)])
def test_near_threshold_states(ops, n, seed, path):
    """>= 7 % of the states within [1e-13, 1e-10] of the 1e-12 threshold, ~80-90 % dead rows, scores up to 1e8."""
    precision, kind, kw, tol = PATHS[path]
    bn, lat, x, P, q, Kref, uref = _case(n, seed, illconditioned_binary_bn)
    near = np.mean((P >= 1e-13) & (P <= 1e-10))
    assert near >= 0.05, near
    tabs = ops.SteinTables.from_network(bn, x, lat, precision=precision, keep_scores=True)
    assert np.array_equal(tabs.P.cpu().numpy(), P)
    assert np.array_equal(tabs.S.cpu().numpy(), c_oracle.score_table(P, n))
    K, u = _run(tabs, kind, torch.from_numpy(q).cuda(), kw)
    _assert_close(K, u, Kref, uref, tol)
    N = 1 << n
    quni = np.full(N, 1.0 / N)
    Kuni, guni = c_oracle.ksd_modeproduct(quni, P, n)
    K, u = _run(tabs, kind, torch.from_numpy(quni).cuda(), kw)
    _assert_close(K, u, Kuni, guni / 2.0, tol)


@pytest.mark.parametrize("path", ILL_PATHS)
def test_near_threshold_states_at_the_posterior(ops, path):
    """q = p(z|x) on the ill-conditioned family (n = 12): KSD^2 and u cancel almost completely; the error is bounded on
    the scale of the terms that cancel, sum_ij q_i q_j |k_p(i,j)| and max_i sum_j |k_p(i,j)| q_j (dense oracle)."""
    n, seed = 12, 1
    precision, kind, kw, tol = PATHS[path]
    bn, lat, x, P, q, _, _ = _case(n, seed, illconditioned_binary_bn)
    post = P / P.sum()
    S = c_oracle.score_table(P, n)
    Kp = orc.stein_matrix_dense(S, n)
    uref = Kp @ post
    Kref = float(post @ uref)
    uscale = np.max(np.abs(Kp) @ post)
    Kscale = float(post @ (np.abs(Kp) @ post))
    tabs = ops.SteinTables.from_network(bn, x, lat, precision=precision)
    K, u = _run(tabs, kind, torch.from_numpy(post).cuda(), kw)
    _assert_close(K, u, Kref, uref, tol, Kscale=Kscale, uscale=uscale)


@pytest.mark.parametrize("path", ["fp64-pairs", "fp64-pairs_sym64", "fp32-pairs", "tc32-pairs", "tc32-pairs_sym-tf32",
                                  "tc32-pairs_sym-tf32_tile", "tc32-pairs_sym-tf32_dot"])
def test_posterior_n16(ops, path):
    """q = exact posterior at n = 16 (random family): K ~ 0 and u ~ 0 (Stein identity); compared on the scale of the
    uniform-q values, as the reference-golden Sprinkler test does."""
    n, seed = 16, 0
    precision, kind, kw, tol = PATHS[path]
    bn, lat, x, P, q, _, _ = _case(n, seed)
    N = 1 << n
    post = P / P.sum()
    Kref, gref = c_oracle.ksd_modeproduct(post, P, n)
    Kuni, guni = c_oracle.ksd_modeproduct(np.full(N, 1.0 / N), P, n)
    tabs = ops.SteinTables.from_network(bn, x, lat, precision=precision)
    K, u = _run(tabs, kind, torch.from_numpy(post).cuda(), kw)
    _assert_close(K, u, Kref, gref / 2.0, tol, Kscale=abs(Kuni), uscale=np.max(np.abs(guni / 2.0)))
