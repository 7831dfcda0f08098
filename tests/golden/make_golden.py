"""Generate tests/golden/ref_golden.npz by RUNNING THE REFERENCE (/root/reference).

Run in the build container only (the GPU box has no /root/reference):
    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

What it records, all from the reference's own functions (no oracle, no product code on
the value path; our seeded generator only *builds* the networks through the reference's
BayesianNetwork.add_node):
  * the known answers asserted in stein_utils.py:205-251;
  * Sprinkler (bayesian_network.py:358-381), latent C,S,R | W=1: p(x,z), score table,
    posterior, and KSD^2 + autograd gradient from the ksd_vi.py:114-131 loop for three q;
  * seeded random BNs n=2..6 (+ one with 2 hidden nodes): same quantities;
  * p(x,z) and score tables for every state of n=10 networks (seeds 0,1) incl. hidden nodes.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"

# utils.py:3 imports matplotlib.pyplot, which is not installed: stub it.
mpl = types.ModuleType("matplotlib")
plt = types.ModuleType("matplotlib.pyplot")
plt.show = lambda *a, **k: None
mpl.pyplot = plt
sys.modules["matplotlib"] = mpl
sys.modules["matplotlib.pyplot"] = plt
sys.dont_write_bytecode = True
sys.path.insert(0, REF)

import stein_utils as ref_su                      # noqa: E402
import bayesian_network as ref_bn                 # noqa: E402
from utils import generate_all_binary_outcomes    # noqa: E402
from functools import partial                     # noqa: E402

sys.path.insert(0, REPO)
from tensornetworks_b200.synthetic import random_binary_bn, random_q  # noqa: E402


def ref_tables(bn, x, latent, observed):
    n = len(latent)
    states = generate_all_binary_outcomes(n)
    P = np.array([ref_su.compute_prob_joint_xz(bn, x, z, latent, observed) for z in states])
    S = np.stack([ref_su.get_score_function_sp_for_z(bn, x, z, latent, observed).numpy() for z in states])
    return P, S.reshape(len(states), n)


def ref_ksd(bn, x, latent, observed, S, q64, length_scale=1.0):
    """The hot loop of ksd_vi.py:110-131 followed by backward (ksd_vi.py:145)."""
    n = len(latent)
    states = generate_all_binary_outcomes(n)
    base = partial(ref_su.base_hamming_kernel_torch, num_vars=n, length_scale=length_scale)
    q = q64.clone().requires_grad_(True)
    sp = [torch.from_numpy(S[i].copy()) for i in range(len(states))]
    total = torch.tensor(0.0, dtype=torch.float64)
    for i, zi in enumerate(states):
        for j, zj in enumerate(states):
            kp = ref_su.get_stein_kernel_kp_value(zi, zj, x, bn, latent, observed, base, sp[i], sp[j]).to(torch.float64)
            total = total + q[i] * q[j] * kp
    total.backward()
    return float(total.item()), q.grad.numpy().copy()


def main():
    out = {}
    # --- known answers of stein_utils.py:221-251 (A -> B, evidence B=1)
    ab = ref_bn.BayesianNetwork()
    ab.add_node('A', cpt={(): {0: 0.8, 1: 0.2}})
    ab.add_node('B', cpt={(0,): {0: 0.7, 1: 0.3}, (1,): {0: 0.4, 1: 0.6}}, parent_names=['A'])
    P, S = ref_tables(ab, {'B': 1}, ['A'], ['B'])
    out["ab_P"], out["ab_S"] = P, S
    base1 = partial(ref_su.base_hamming_kernel_torch, num_vars=1, length_scale=1.0)
    s0, s1 = torch.from_numpy(S[0].copy()), torch.from_numpy(S[1].copy())
    out["ab_kp01"] = ref_su.get_stein_kernel_kp_value((0,), (1,), {'B': 1}, ab, ['A'], ['B'], base1, s0, s1).item()
    out["ab_kp00"] = ref_su.get_stein_kernel_kp_value((0,), (0,), {'B': 1}, ab, ['A'], ['B'], base1, s0, s0).item()

    # --- Sprinkler
    sp = ref_bn.get_sprinkler_network(False)
    lat, obs, x = ['C', 'S', 'R'], ['W'], {'W': 1}
    P, S = ref_tables(sp, x, lat, obs)
    post, pw = sp.get_true_posterior(lat, x)
    out["spr_P"], out["spr_S"], out["spr_evidence"] = P, S, pw
    out["spr_post"] = np.array([post[z] for z in generate_all_binary_outcomes(3)])
    qs = {"uniform": torch.full((8,), 1.0 / 8, dtype=torch.float64),
          "ramp": torch.arange(1, 9, dtype=torch.float64) / 36.0,
          "post": torch.from_numpy(out["spr_post"].copy())}
    for name, q in qs.items():
        K, g = ref_ksd(sp, x, lat, obs, S, q)
        out[f"spr_q_{name}"], out[f"spr_K_{name}"], out[f"spr_g_{name}"] = q.numpy(), K, g
    for ls in (0.5, 2.0):
        K, g = ref_ksd(sp, x, lat, obs, S, qs["ramp"], length_scale=ls)
        out[f"spr_K_ramp_ls{ls}"], out[f"spr_g_ramp_ls{ls}"] = K, g

    # --- seeded random BNs (generator drives the reference's BayesianNetwork.add_node)
    cases = [(n, n, 0) for n in range(2, 7)] + [(4, 104, 2)]
    for n, seed, nh in cases:
        bn, lat, obs, x = random_binary_bn(n, seed=seed, n_hidden=nh, bn_cls=ref_bn.BayesianNetwork)
        P, S = ref_tables(bn, x, lat, obs)
        q = random_q(1 << n, seed=seed)
        K, g = ref_ksd(bn, x, lat, obs, S, q)
        tag = f"rnd_n{n}_s{seed}_h{nh}"
        out[tag + "_P"], out[tag + "_S"], out[tag + "_q"] = P, S, q.numpy()
        out[tag + "_K"], out[tag + "_g"] = K, g
        print(tag, "K =", K, flush=True)

    # --- enumeration tables only, larger n
    for n, seed, nh in [(10, 0, 0), (10, 1, 2), (8, 7, 1)]:
        bn, lat, obs, x = random_binary_bn(n, seed=seed, n_hidden=nh, bn_cls=ref_bn.BayesianNetwork)
        P, S = ref_tables(bn, x, lat, obs)
        tag = f"tab_n{n}_s{seed}_h{nh}"
        out[tag + "_P"], out[tag + "_S"] = P, S
        print(tag, "zero-score states:", int((np.abs(P) < 1e-12).sum()), flush=True)

    np.savez_compressed(os.path.join(HERE, "ref_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "ref_golden.npz"), len(out), "arrays")


if __name__ == "__main__":
    main()
