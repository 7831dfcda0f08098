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


def illconditioned_binary_bn(n_latent, n_obs=None, seed=0, n_sharp=9, bn_cls=None):
    """Random binary BN whose joint table straddles the score threshold of reference stein_utils.py:126-128
    (|p| < 1e-12 -> zero score row): the latent nodes form a chain Z0 -> Z1 -> ... whose first `n_sharp` links
    are sharp - Z_{k+1} copies Z_k except with a probability eps ~ logU(1e-4, 1e-2), drawn independently per
    CPT row -, every third latent node additionally draws one earlier random parent, and all remaining CPT rows
    (later chain links, observed nodes with 1..3 latent parents) are p1 ~ U(1e-4, 1 - 1e-4).  States with
    three to four broken sharp links land around 1e-12: a sizeable share of the states lies within
    [1e-13, 1e-10], dead rows sit next to live ones and the live scores reach 1e4...1e8 - the regime where a
    reduced-precision operand split would lose bits.  Same return convention as random_binary_bn."""
    if bn_cls is None:
        from .bayesian_network import BayesianNetwork as bn_cls
    if n_obs is None:
        n_obs = max(1, n_latent // 3)
    rng = np.random.default_rng(seed)
    bn = bn_cls()
    latent = [f"Z{i}" for i in range(n_latent)]
    observed = [f"X{i}" for i in range(n_obs)]
    names = latent + observed
    for idx, name in enumerate(names):
        if idx == 0:
            pa_idx = []
        elif idx < n_latent:
            pa_idx = [idx - 1]
            if idx % 3 == 2 and idx >= 2:
                pa_idx = sorted({int(rng.integers(0, idx - 1)), idx - 1})
        else:
            k = int(rng.integers(1, 4))
            pa_idx = sorted(int(v) for v in rng.choice(n_latent, size=min(k, n_latent), replace=False))
        k = len(pa_idx)
        cpt = {}
        for cfg in range(1 << k):
            key = tuple((cfg >> (k - 1 - t)) & 1 for t in range(k))
            if 0 < idx <= min(n_sharp, n_latent - 1):
                prev = key[pa_idx.index(idx - 1)]
                eps = float(np.exp(rng.uniform(np.log(1e-4), np.log(1e-2))))
                p1 = 1.0 - eps if prev == 1 else eps
            else:
                p1 = float(rng.uniform(1e-4, 1.0 - 1e-4))
            cpt[key] = {0: 1 - p1, 1: p1}
        bn.add_node(name, cpt=cpt, parent_names=[names[p] for p in pa_idx] or None)
    evidence = {name: (1 if j % 2 == 0 else 0) for j, name in enumerate(observed)}
    return bn, latent, observed, evidence
