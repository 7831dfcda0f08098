"""CPU tests of the drop-in modules' host side: the reference's own asserts
(stein_utils.py:205-251) through our scalar API, the Born machines, and the import shims."""
import os
import subprocess
import sys
from functools import partial

import numpy as np
import pytest
import torch

from oracle import ksd_oracle as orc
from tensornetworks_b200 import stein_utils as su
from tensornetworks_b200.bayesian_network import BayesianNetwork, get_sprinkler_network
from tensornetworks_b200.born_machine_classical_sim import ClassicalBornMachine
from tensornetworks_b200.quantum_born_machine import QuantumBornMachine
from tensornetworks_b200.synthetic import random_binary_bn
from tensornetworks_b200.utils import generate_all_binary_outcomes

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
F64 = torch.float64


def test_reference_self_checks():
    """Every assert of reference stein_utils.py:205-251, run against our module."""
    assert su.flip_bit((0, 0, 0), 0) == (1, 0, 0)
    z1 = torch.tensor([0, 0, 1, 1], dtype=F64)
    z2 = torch.tensor([1, 0, 0, 1], dtype=F64)
    assert su.hamming_distance_torch(z1, z2).item() == 2.0
    assert torch.isclose(su.base_hamming_kernel_torch(z1, z2, 4, length_scale=1.0), torch.tensor(np.exp(-0.5), dtype=F64))
    bn = BayesianNetwork()
    bn.add_node('A', cpt={(): {0: 0.8, 1: 0.2}})
    bn.add_node('B', cpt={(0,): {0: 0.7, 1: 0.3}, (1,): {0: 0.4, 1: 0.6}}, parent_names=['A'])
    x = {'B': 1}
    assert np.isclose(su.compute_prob_joint_xz(bn, x, (1,), ['A'], ['B']), 0.12)
    s1 = su.get_score_function_sp_for_z(bn, x, (1,), ['A'], ['B'])
    s0 = su.get_score_function_sp_for_z(bn, x, (0,), ['A'], ['B'])
    assert s1.dtype == F64 and torch.isclose(s1[0], torch.tensor(-1.0, dtype=F64))
    assert torch.isclose(s0[0], torch.tensor(0.5, dtype=F64))
    bk = partial(su.base_hamming_kernel_torch, num_vars=1, length_scale=1.0)
    kp01 = su.get_stein_kernel_kp_value((0,), (1,), x, bn, ['A'], ['B'], bk, s0, s1)
    kp00 = su.get_stein_kernel_kp_value((0,), (0,), x, bn, ['A'], ['B'], bk, s0, s0)
    assert torch.isclose(kp01, torch.tensor(2 * np.exp(-1.0) - 2.5, dtype=F64))
    assert torch.isclose(kp00, torch.tensor(1.25 - np.exp(-1.0), dtype=F64))


def test_base_kernel_edge_cases():
    z = torch.tensor([1.0, 0.0])
    assert su.base_hamming_kernel_torch(z, z, 0).item() == 1.0                  # num_vars == 0 (:36-40)
    assert su.base_hamming_kernel_torch(z, z, 2, length_scale=0.0).item() == 1.0  # zero denominator (:52-54)
    assert su.base_hamming_kernel_torch(z, 1 - z, 2, length_scale=0.0).item() == 0.0
    out = su.base_hamming_kernel_torch(z, (1 - z).to(F64), 2)                   # dtype promotion (:19-26)
    assert out.dtype == F64 and np.isclose(out.item(), np.exp(-1.0))
    assert su.get_stein_kernel_kp_value((), (), {}, None, [], [], None, torch.zeros(0), torch.zeros(0)).item() == 0.0


@pytest.mark.parametrize("n,seed,nh", [(3, 3, 0), (4, 104, 2)])
def test_scalar_api_matches_goldens(golden, n, seed, nh):
    bn, lat, obs, x = random_binary_bn(n, seed=seed, n_hidden=nh)
    tag = f"rnd_n{n}_s{seed}_h{nh}"
    states = generate_all_binary_outcomes(n)
    P = np.array([su.compute_prob_joint_xz(bn, x, z, lat, obs) for z in states])
    S = np.stack([su.get_score_function_sp_for_z(bn, x, z, lat, obs).numpy() for z in states])
    assert np.array_equal(P, golden[tag + "_P"]) and np.array_equal(S, golden[tag + "_S"])
    bk = partial(su.base_hamming_kernel_torch, num_vars=n, length_scale=1.0)
    Kp = orc.stein_matrix_dense(S, n)
    for i, j in [(0, 0), (1, 2), ((1 << n) - 1, 3)]:
        v = su.get_stein_kernel_kp_value(states[i], states[j], x, bn, lat, obs, bk,
                                         torch.from_numpy(S[i]), torch.from_numpy(S[j]))
        assert np.isclose(v.item(), Kp[i, j], rtol=1e-12)


def test_classical_born_machine_api():
    torch.manual_seed(0)
    bm = ClassicalBornMachine(3)
    p = bm.get_probabilities()
    assert p.shape == (1, 8) and p.dtype == torch.float32 and np.isclose(p.sum().item(), 1.0, atol=1e-6)
    assert [n for n, _ in bm.named_parameters()] == ['params']
    d = bm.get_prob_dict()
    assert list(d.keys()) == list(generate_all_binary_outcomes(3)) and np.isclose(sum(d.values()), 1.0, atol=1e-6)
    ent = bm.entropy()
    assert np.isclose(ent.item(), -(p * p.clamp(min=1e-10).log()).sum().item())
    s = bm.sample(5)
    assert s.shape == (5, 3) and s.dtype == torch.float32
    lq = bm.get_log_q_z_x(s)
    idx = (s.long() * torch.tensor([4, 2, 1])).sum(-1)
    assert torch.allclose(lq, p[0, idx].log())
    with pytest.raises(ValueError):
        bm.get_probabilities(x_condition=torch.zeros(1))
    with pytest.raises(ValueError):
        bm.get_log_q_z_x(torch.tensor([[0., 2., 1.]]))
    bm.set_fixed_probs(torch.full((8,), 0.125))
    assert torch.equal(bm.get_probabilities(), torch.full((1, 8), 0.125))
    bm.clear_fixed_probs()
    cond = ClassicalBornMachine(3, conditioning_dim=1)
    assert [type(m).__name__ for m in cond.param_generator_net] == ['Linear', 'ReLU', 'Dropout', 'Linear', 'ReLU', 'Dropout', 'Linear']
    assert cond.param_generator_net[0].out_features == 64 and cond.param_generator_net[3].out_features == 32
    assert cond.get_probabilities(torch.tensor([1.0])).shape == (1, 8)
    with pytest.raises(ValueError):
        cond.get_probabilities()
    assert cond.sample(4, torch.tensor([[1.0], [0.0]])).shape == (2, 4, 3)


def test_quantum_born_machine_simulator():
    torch.manual_seed(1)
    for kind, per in (("hardware_efficient", 3), ("all_to_all", 3), ("basic", 2)):
        qbm = QuantumBornMachine(3, ansatz_layers=4, ansatz_type=kind, backend="torch")
        assert qbm.theta.numel() == 4 * per * 3 and qbm.theta.dtype == torch.float32   # reference :31-38
        p = qbm.get_probabilities()
        assert p.shape == (8,) and np.isclose(p.sum().item(), 1.0)
        assert list(qbm.get_prob_dict().keys()) == list(generate_all_binary_outcomes(3))
    # zero parameters: H layer then identity rotations; CNOT/CZ permute/phase a uniform state -> uniform q
    qbm = QuantumBornMachine(3, ansatz_layers=2, init_method="zero", backend="torch")
    assert torch.allclose(qbm.get_probabilities(), torch.full((8,), 0.125, dtype=F64))
    # basic ansatz with zero angles leaves |000>
    qbm = QuantumBornMachine(3, ansatz_layers=1, ansatz_type="basic", init_method="zero", backend="torch")
    assert np.isclose(qbm.get_probabilities()[0].item(), 1.0)
    # autograd gradient == parameter-shift gradient (what PennyLane would return)
    qbm = QuantumBornMachine(3, ansatz_layers=2, init_method="random", backend="torch")
    w = torch.randn(8, dtype=F64)
    (qbm.get_probabilities() * w).sum().backward()
    from tensornetworks_b200.quantum_born_machine import simulate_probs
    th = qbm.theta.detach().to(F64)
    for k in (0, 5, 17):
        e = torch.zeros_like(th); e[k] = np.pi / 2
        shift = 0.5 * ((simulate_probs(th + e, 3, 2, "hardware_efficient") - simulate_probs(th - e, 3, 2, "hardware_efficient")) * w).sum()
        assert np.isclose(qbm.theta.grad[k].item(), shift.item(), rtol=1e-5, atol=1e-7)


def test_flat_module_import_shims():
    code = ("import sys; sys.path.insert(0, %r);"
            "import utils, bayesian_network, stein_utils, born_machine_classical_sim, quantum_born_machine;"
            "from ksd_vi import KSDVariationalInference as A;"
            "from ksd_vi_quantum import KSDVariationalInference as B;"
            "from stein_utils import flip_bit, hamming_distance_torch, base_hamming_kernel_torch, compute_prob_joint_xz,"
            " get_score_function_sp_for_z, get_stein_kernel_kp_value;"
            "from utils import calculate_tvd, generate_all_binary_outcomes;"
            "from bayesian_network import BayesianNetwork, get_sprinkler_network;"
            "assert A is not B; print('ok')") % os.path.join(REPO, "dropin")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.strip() == "ok", out.stderr


def test_trainer_argument_errors():
    from tensornetworks_b200.ksd_vi import KSDVariationalInference, posterior_vector
    bn = get_sprinkler_network(False)
    vi = KSDVariationalInference(bn, ['C', 'S', 'R'], ['W'], {'use_logits': True, 'conditioning_dim': 1, 'init_method': 'uniform'})
    assert vi.num_latent_vars == 3 and vi.num_possible_latent_states == 8 and len(vi._score_function_cache) == 0
    assert vi.all_latent_states_tuples[6] == (1, 1, 0)
    with pytest.raises(ValueError, match="Keys in x_observation_dict"):
        vi.train({'Q': 1}, 1, 0.01)                                    # ksd_vi.py:68-69
    vi2 = KSDVariationalInference(bn, ['C', 'S', 'R'], [], {'conditioning_dim': 1})
    with pytest.raises(ValueError, match="no observed vars"):
        vi2.train({}, 1, 0.01)                                         # :76-77
    vi3 = KSDVariationalInference(bn, ['C', 'S'], ['R', 'W'], {'conditioning_dim': 1})
    with pytest.raises(ValueError, match="conditioning_dim must match"):
        vi3.train({'R': 1, 'W': 1}, 1, 0.01)                           # :78-79
    vec, stray = posterior_vector({(0, 0): 0.25, (1, 1): 0.5, 'junk': 0.25}, 2, 'cpu')
    assert vec.tolist() == [0.25, 0, 0, 0.5] and stray == 0.25
