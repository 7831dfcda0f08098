"""Pins the torch state-vector simulator of tensornetworks_b200/quantum_born_machine.py to an independent dense
restatement of the reference's three circuits (oracle/quantum_oracle.py; reference quantum_born_machine.py:57-128):
explicit 2^n x 2^n unitaries built gate by gate with np.kron, wire 0 = MSB."""
import numpy as np
import pytest
import torch

from oracle import quantum_oracle as qorc
from tensornetworks_b200.quantum_born_machine import QuantumBornMachine, simulate_probs

ANSATZ = ["hardware_efficient", "all_to_all", "basic"]


def _nparams(n, layers, kind):
    return layers * (2 if kind == "basic" else 3) * n


@pytest.mark.parametrize("kind", ANSATZ)
@pytest.mark.parametrize("layers", [1, 2, 3, 4])
@pytest.mark.parametrize("n", [1, 2, 3, 4, 5])
def test_simulator_matches_dense_unitaries(n, layers, kind):
    rng = np.random.default_rng(1000 * n + 10 * layers + ANSATZ.index(kind))
    w = rng.uniform(-np.pi, np.pi, size=_nparams(n, layers, kind))
    ref = qorc.circuit_probs(w, n, layers, kind)
    got = simulate_probs(torch.tensor(w, dtype=torch.float64), n, layers, kind).numpy()
    assert np.isclose(ref.sum(), 1.0, atol=1e-12)
    assert np.max(np.abs(got - ref)) <= 1e-12


def test_unitaries_are_unitary_and_msb_first():
    U = qorc.circuit_unitary(np.linspace(0.1, 1.7, 18), 3, 2, "hardware_efficient")
    assert np.allclose(U.conj().T @ U, np.eye(8), atol=1e-12)
    # wire 0 is the most significant bit: an X on wire 0 alone moves |000> to index 4 (utils.py:77-91 order)
    flip0 = qorc._embed({0: qorc.X}, 3)
    assert np.argmax(np.abs(flip0[:, 0])) == 4
    # CNOT(control 2, target 0) on |001> gives |101>
    cn = qorc._controlled(2, 0, qorc.X, 3)
    assert np.argmax(np.abs(cn[:, 1])) == 5


@pytest.mark.parametrize("kind", ANSATZ)
def test_module_get_probabilities_matches_oracle(kind):
    torch.manual_seed(3)
    m = QuantumBornMachine(4, ansatz_layers=3, ansatz_type=kind, init_method="random", backend="torch")
    ref = qorc.circuit_probs(m.theta.detach().double().numpy(), 4, 3, kind)
    got = m.get_probabilities().detach().numpy()
    assert np.max(np.abs(got - ref)) <= 1e-12
    d = m.get_prob_dict()
    assert abs(d[(0, 1, 1, 0)] - ref[0b0110]) <= 1e-12
