"""Streaming Born-machine kernels (tnk_born_forward / tnk_born_backward) against the tensor expressions of
reference born_machine_classical_sim.py:74-99 (softmax(raw - max)) and :177-181 (entropy), and the trainer
with and without them."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from tensornetworks_b200 import ops as _ops
    return _ops


def _reference(logits, w, c):
    l = logits.detach().clone().double().requires_grad_(True)
    q = torch.softmax(l - l.max(), dim=0)
    H = -(q * torch.log(q.clamp(min=1e-10))).sum()
    loss = (w.double() * q).sum() + c * H
    loss.backward()
    return q.detach(), H.detach(), l.grad


@pytest.mark.parametrize("N", [1, 5, 2048, 65536, 100003, 1 << 20])
@pytest.mark.parametrize("scale", [1.0, 30.0])
def test_softmax_entropy_forward_backward(ops, N, scale):
    g = torch.Generator().manual_seed(N)
    logits = (scale * torch.randn(N, generator=g)).cuda().requires_grad_(True)
    w = torch.randn(N, generator=g).cuda()
    q, H = ops.born_softmax_entropy(logits)
    assert q.dtype == torch.float32 and q.shape == (N,) and H.dim() == 0
    loss = (w * q).sum() + 0.3 * H
    loss.backward()
    q_ref, H_ref, g_ref = _reference(logits, w, 0.3)
    assert float((q.detach().double() - q_ref).abs().max() / q_ref.abs().max()) < 2e-6      # fp32 storage
    assert abs(float(H.detach()) - float(H_ref)) <= 2e-6 * max(1.0, abs(float(H_ref)))
    # normwise 2e-5, or - for a nearly one-hot q, where q (g - q.g) cancels below fp32 resolution in the
    # reference's own fp32 graph as well - the fp32 rounding of the largest term q_max * |g_in|
    g_in = float(w.abs().max()) + 0.3 * (float(torch.log(q_ref.clamp(min=1e-10)).abs().max()) + 1.0)
    tol = max(2e-5 * float(g_ref.abs().max()), 5e-6 * float(q_ref.max()) * g_in)
    assert float((logits.grad.double() - g_ref).abs().max()) < tol
    assert abs(float(q.detach().double().sum()) - 1.0) < 1e-5


def test_single_output_gradients(ops):
    N = 1 << 17
    g = torch.Generator().manual_seed(7)
    base = torch.randn(N, generator=g).cuda()
    w = torch.randn(N, generator=g).cuda()
    for use_q, use_H, gdtype in ((True, False, torch.float64), (False, True, None), (True, True, torch.float32)):
        logits = base.clone().requires_grad_(True)
        q, H = ops.born_softmax_entropy(logits)
        loss = 0.0
        if use_q:
            loss = loss + (w.to(gdtype) * q.to(gdtype)).sum()        # fp64 incoming gradient when gdtype is double
        if use_H:
            loss = loss + 2.0 * H
        loss.backward()
        l = base.double().requires_grad_(True)
        qr = torch.softmax(l - l.max(), dim=0)
        Hr = -(qr * torch.log(qr.clamp(min=1e-10))).sum()
        ref = (w.double() * qr).sum() * (1.0 if use_q else 0.0) + 2.0 * Hr * (1.0 if use_H else 0.0)
        ref.backward()
        assert float((logits.grad.double() - l.grad).abs().max() / l.grad.abs().max()) < 2e-5


def test_born_machine_streaming_matches_tensor_path():
    from tensornetworks_b200.born_machine_classical_sim import ClassicalBornMachine
    torch.manual_seed(0)
    bm = ClassicalBornMachine(16, use_logits=True).cuda()
    p = bm.get_probabilities()
    H = bm.entropy()
    assert p.shape == (1, 1 << 16)
    raw = bm.params.detach().unsqueeze(0)
    p_ref = torch.softmax(raw - raw.max(dim=-1, keepdim=True)[0], dim=-1)
    H_ref = -(p_ref.squeeze() * torch.log(p_ref.squeeze().clamp(min=1e-10))).sum()
    assert float((p.detach() - p_ref).abs().max() / p_ref.max()) < 2e-6
    assert abs(float(H.detach()) - float(H_ref)) < 1e-4
    # q and H of one step share a forward pass; both feed one backward
    loss = (p.squeeze() * torch.arange(1 << 16, device="cuda", dtype=torch.float32)).sum() * 1e-4 - 0.01 * H
    loss.backward()
    g1 = bm.params.grad.clone()
    # fp64 evaluation of the reference's expressions (its fp32 graph adds the rounding of sum(grad) to the
    # arg-max logit through the max shift, so fp32 against fp32 would compare two roundings)
    l = bm.params.detach().double().requires_grad_(True)
    p2 = torch.softmax(l - l.max(), dim=0)
    H2 = -(p2 * torch.log(p2.clamp(min=1e-10))).sum()
    ((p2 * torch.arange(1 << 16, device="cuda", dtype=torch.float64)).sum() * 1e-4 - 0.01 * H2).backward()
    assert float((g1.double() - l.grad).abs().max() / l.grad.abs().max()) < 2e-5
    import copy
    bm.get_probabilities()
    twin = copy.deepcopy(bm)                  # the cached pair (with its graph) is not copied
    assert twin._pair_cache is None and torch.equal(twin.params, bm.params)
    # asking twice in one step never reuses a consumed graph
    a = bm.get_probabilities(); b = bm.get_probabilities()
    a.sum().backward(); b.sum().backward()


def test_separate_backward_passes_for_q_and_entropy():
    """q and H of one step come from one autograd node; a backward pass through q alone must not leave a cached H with
    a freed graph behind (separate backward calls for the two loss terms, gradient accumulation)."""
    from tensornetworks_b200.born_machine_classical_sim import ClassicalBornMachine
    torch.manual_seed(1)
    bm = ClassicalBornMachine(16, use_logits=True).cuda()
    w = torch.linspace(0.0, 1.0, 1 << 16, device="cuda")
    (bm.get_probabilities().squeeze() * w).sum().backward()          # backward through q only
    g_q = bm.params.grad.clone()
    bm.params.grad = None
    (-0.01 * bm.entropy()).backward()                                 # no parameter update in between
    g_h = bm.params.grad.clone()
    bm.params.grad = None
    p, H = bm.get_probabilities(), bm.entropy()
    ((p.squeeze() * w).sum() - 0.01 * H).backward()
    assert float((bm.params.grad - (g_q + g_h)).abs().max()) <= 1e-6 * float((g_q + g_h).abs().max()) + 1e-12
    # the other order, and twice each
    bm.params.grad = None
    bm.entropy().backward()
    bm.get_probabilities().sum().backward()
    bm.entropy().backward()


def test_trainer_epochs_with_and_without_streaming_kernels(monkeypatch):
    from tensornetworks_b200.born_machine_classical_sim import ClassicalBornMachine
    from tensornetworks_b200.ksd_vi import KSDVariationalInference
    from tensornetworks_b200.synthetic import random_binary_bn
    n = 16
    bn, lat, obs, x = random_binary_bn(n, seed=3)
    hist = []
    for streaming in (True, False):
        if not streaming:
            monkeypatch.setattr(ClassicalBornMachine, "_streaming", lambda self, raw: False)
        torch.manual_seed(0)
        vi = KSDVariationalInference(bn, lat, obs, {"use_logits": True, "conditioning_dim": 0}, device="cuda",
                                     precision="fp64", compute_device="cuda", algorithm="modeproduct")
        hist.append(vi.train(x, num_epochs=6, lr_born_machine=0.05, verbose=False))
    for key in ("loss_ksd", "entropy", "grad_norm"):
        a, b = np.array(hist[0][key], dtype=float), np.array(hist[1][key], dtype=float)
        assert np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-12)) < 1e-4, key


@pytest.mark.parametrize("n,precision,algorithm", [(10, "fp64", "pairs"), (12, "tc32", "pairs"), (16, "fp64", "modeproduct")])
def test_cuda_graph_epochs_match_eager(n, precision, algorithm):
    """cuda_graph=True replays forward+backward+clipping and the Adam step as two CUDA graphs per epoch; the history
    (with TVD tracking, cosine schedule, best-q restore) must be the eager one up to rounding."""
    from tensornetworks_b200 import ops
    from tensornetworks_b200.ksd_vi import KSDVariationalInference
    from tensornetworks_b200.synthetic import random_binary_bn
    bn, lat, obs, x = random_binary_bn(n, seed=5)
    hist, final = [], []
    for graph in (False, True):
        torch.manual_seed(1)
        vi = KSDVariationalInference(bn, lat, obs, {"use_logits": True, "conditioning_dim": 0}, device="cuda",
                                     precision=precision, compute_device="cuda", algorithm=algorithm, cuda_graph=graph)
        p_true = ops.SteinTables.from_network(bn, x, lat, precision="fp64").posterior()
        hist.append(vi.train(x, num_epochs=12, lr_born_machine=0.05, verbose=False, true_posterior_for_tvd=p_true))
        final.append(vi.born_machine.get_probabilities().squeeze().detach().double().cpu().numpy())
        assert all(p.grad is not None for p in vi.born_machine.parameters())
    tol = 1e-5 if precision == "fp64" else 2e-3
    for key in ("loss_ksd", "entropy", "grad_norm", "tvd"):
        a, b = np.array(hist[0][key], dtype=float), np.array(hist[1][key], dtype=float)
        assert a.shape == b.shape == (12,), key
        assert np.max(np.abs(a - b) / np.maximum(np.abs(a), 1e-12)) < tol, (key, a, b)
    assert np.max(np.abs(final[0] - final[1])) <= tol * np.max(final[0])
