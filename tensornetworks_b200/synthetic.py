"""Seeded synthetic workloads for the parity tests and the benchmark.

`random_binary_bn` builds its network only through `add_node`, so the very same generator
drives the reference's own `BayesianNetwork` (in tests/golden/make_golden.py) and ours.
"""
import numpy as np


def random_binary_bn(n_latent, n_obs=None, max_in_degree=3, seed=0, n_hidden=0, bn_cls=None):
    """Random binary BN: nodes Z0..Z{n-1}, then H0..H{n_hidden-1} (neither latent nor
    observed -> summed out), then X0..X{n_obs-1}, added parents-first.  Node idx draws
    k ~ U{0..min(max_in_degree, idx)} parents among earlier nodes; each CPT row
    p1 ~ U(0.01, 0.99) (the range of the reference's random Sprinkler,
    bayesian_network.py:323), stored as {0: 1-p1, 1: p1}.

    Returns (bn, latent_names, observed_names, evidence) with evidence X_j = 1 if j even else 0.
    """
    if bn_cls is None:
        from .bayesian_network import BayesianNetwork as bn_cls
    if n_obs is None:
        n_obs = max(1, n_latent // 3)
    rng = np.random.default_rng(seed)
    bn = bn_cls()
    latent = [f"Z{i}" for i in range(n_latent)]
    hidden = [f"H{i}" for i in range(n_hidden)]
    observed = [f"X{i}" for i in range(n_obs)]
    names = latent + hidden + observed
    for idx, name in enumerate(names):
        k = int(rng.integers(0, min(max_in_degree, idx) + 1))
        pa_idx = sorted(int(v) for v in rng.choice(idx, size=k, replace=False)) if k else []
        cpt = {}
        for cfg in range(1 << k):
            key = tuple((cfg >> (k - 1 - t)) & 1 for t in range(k))
            p1 = float(rng.uniform(0.01, 0.99))
            cpt[key] = {0: 1 - p1, 1: p1}
        bn.add_node(name, cpt=cpt, parent_names=[names[p] for p in pa_idx] or None)
    evidence = {name: (1 if j % 2 == 0 else 0) for j, name in enumerate(observed)}
    return bn, latent, observed, evidence


def random_q(num_states, seed=0, dtype="float64"):
    """q = softmax(randn) drawn in fp32 and cast up, as the classical Born machine hands it to
    the KSD loop (reference ksd_vi.py:105,112).  Returns a CPU torch tensor."""
    import torch
    g = torch.Generator().manual_seed(int(seed))
    q = torch.softmax(torch.randn(num_states, generator=g, dtype=torch.float32), dim=0)
    return q.to(getattr(torch, dtype))
