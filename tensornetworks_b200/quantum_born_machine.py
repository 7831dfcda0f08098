"""Quantum Born machine q(z) = |<z|U(theta)|0>|^2 - the q supplier of the quantum KSD runs.

Drop-in for the reference's `quantum_born_machine.QuantumBornMachine`
(reference quantum_born_machine.py:7-200): same constructor, `theta` parameter, ansatz
layouts, `get_probabilities`, `get_prob_dict`.  The reference evaluates the circuit with a
PennyLane `default.qubit` QNode and parameter-shift gradients.  PennyLane is used here too
when it is importable; otherwise the same circuits run on a small torch state-vector
simulator (wire 0 = most significant bit = the enumeration order of utils.py:77-91), whose
autograd gradient equals the parameter-shift gradient analytically.  Either way the KSD
itself goes through the CUDA pair kernel.
"""
import math

import torch
import torch.nn as nn

from .utils import generate_all_binary_outcomes

try:  # optional, exactly as the reference uses it
    import pennylane as qml  # noqa: F401
    HAVE_PENNYLANE = True
except Exception:  # pragma: no cover - not installed in this image
    qml = None
    HAVE_PENNYLANE = False


# ---------------------------------------------------------------- torch state-vector backend
def _apply_1q(state, gate, wire):
    state = torch.tensordot(gate, state, dims=([1], [wire]))
    return torch.movedim(state, 0, wire)


def _rx(t):
    c, s = torch.cos(t / 2), torch.sin(t / 2)
    return torch.stack([torch.stack([c + 0j, -1j * s]), torch.stack([-1j * s, c + 0j])])


def _ry(t):
    c, s = torch.cos(t / 2), torch.sin(t / 2)
    return torch.stack([torch.stack([c + 0j, -s + 0j]), torch.stack([s + 0j, c + 0j])])


def _rz(t):
    e = torch.exp(-0.5j * t)
    zero = torch.zeros((), dtype=e.dtype, device=e.device)
    return torch.stack([torch.stack([e, zero]), torch.stack([zero, torch.conj(e)])])


def _cnot(state, control, target):
    idx = [slice(None)] * state.dim()
    idx[control] = 1
    sub = state[tuple(idx)]
    t_axis = target if target < control else target - 1
    flipped = torch.flip(sub, dims=[t_axis])
    parts = [state.select(control, 0), flipped]
    return torch.stack(parts, dim=control)


def _cz(state, a, b):
    n = state.dim()
    shape = [1] * n
    shape[a] = 2
    ia = torch.arange(2, device=state.device).reshape(shape)
    shape = [1] * n
    shape[b] = 2
    ib = torch.arange(2, device=state.device).reshape(shape)
    return state * (1.0 - 2.0 * (ia * ib)).to(state.dtype)


def simulate_probs(weights, num_qubits, layers, ansatz_type):
    """Outcome probabilities [2^n] (float64) of the reference's ansaetze
    (quantum_born_machine.py:57-128)."""
    n = num_qubits
    if n == 0:
        return torch.ones(1, dtype=torch.float64, device=weights.device)
    w = weights.to(torch.float64)
    state = torch.zeros((2,) * n, dtype=torch.complex128, device=weights.device)
    state[(0,) * n] = 1.0
    had = torch.tensor([[1, 1], [1, -1]], dtype=torch.complex128, device=weights.device) / math.sqrt(2.0)
    k = 0
    if ansatz_type in ("hardware_efficient", "all_to_all"):
        for i in range(n):
            state = _apply_1q(state, had, i)
    for layer in range(layers):
        for i in range(n):
            if ansatz_type in ("hardware_efficient", "all_to_all"):
                state = _apply_1q(state, _rx(w[k]), i); k += 1
            state = _apply_1q(state, _ry(w[k]), i); k += 1
            state = _apply_1q(state, _rz(w[k]), i); k += 1
        if n > 1:
            if ansatz_type == "all_to_all":
                for i in range(n):
                    for j in range(i + 1, n):
                        state = _cz(state, i, j)
            else:
                for i in range(n - 1):
                    state = _cnot(state, i, i + 1)
                if n > 2:
                    state = _cnot(state, n - 1, 0)
                if ansatz_type == "hardware_efficient" and layer % 2 == 0 and n > 2:
                    for i in range(0, n - 2, 2):
                        state = _cz(state, i, i + 2)
    return (state.real ** 2 + state.imag ** 2).reshape(-1)


class QuantumBornMachine(nn.Module):
    def __init__(self, num_latent_vars, ansatz_layers=1, conditioning_dim=0,
                 device_name="default.qubit", ansatz_type="hardware_efficient",
                 init_method="small_random", backend=None):
        super().__init__()
        self.num_latent_vars = num_latent_vars
        self.conditioning_dim = conditioning_dim
        self.ansatz_type = ansatz_type
        self.ansatz_layers = ansatz_layers
        per_qubit = 3 if ansatz_type in ("hardware_efficient", "all_to_all") else 2
        self.num_ansatz_params = ansatz_layers * per_qubit * num_latent_vars   # reference :31-38
        if init_method == "zero":
            theta = torch.zeros(self.num_ansatz_params, dtype=torch.float32)
        elif init_method == "small_random":
            theta = 0.1 * torch.randn(self.num_ansatz_params, dtype=torch.float32)
        else:
            theta = torch.rand(self.num_ansatz_params, dtype=torch.float32) * 2 * torch.pi
        self.theta = nn.Parameter(theta)
        self.all_outcomes_tuples = generate_all_binary_outcomes(num_latent_vars)
        self.backend = backend or ("pennylane" if HAVE_PENNYLANE else "torch")
        self.pqc = self._build_pennylane(device_name) if self.backend == "pennylane" else self._torch_pqc

    def _torch_pqc(self, weights, x_inputs=None):
        return simulate_probs(weights, self.num_latent_vars, self.ansatz_layers, self.ansatz_type)

    def _build_pennylane(self, device_name):
        if not HAVE_PENNYLANE:
            raise ImportError("backend='pennylane' requested but pennylane is not installed")
        n, layers, kind = self.num_latent_vars, self.ansatz_layers, self.ansatz_type
        dev = qml.device(device_name, wires=n, shots=None)

        @qml.qnode(dev, interface="torch", diff_method="parameter-shift")
        def pqc(weights, x_inputs=None):
            k = 0
            if kind in ("hardware_efficient", "all_to_all"):
                for i in range(n):
                    qml.Hadamard(wires=i)
            for layer in range(layers):
                for i in range(n):
                    if kind in ("hardware_efficient", "all_to_all"):
                        qml.RX(weights[k], wires=i); k += 1
                    qml.RY(weights[k], wires=i); k += 1
                    qml.RZ(weights[k], wires=i); k += 1
                if n > 1:
                    if kind == "all_to_all":
                        for i in range(n):
                            for j in range(i + 1, n):
                                qml.CZ(wires=[i, j])
                    else:
                        for i in range(n - 1):
                            qml.CNOT(wires=[i, i + 1])
                        if n > 2:
                            qml.CNOT(wires=[n - 1, 0])
                        if kind == "hardware_efficient" and layer % 2 == 0 and n > 2:
                            for i in range(0, n - 2, 2):
                                qml.CZ(wires=[i, i + 2])
            return qml.probs(wires=range(n))

        return pqc

    def get_probabilities(self, x_condition=None):
        """q_theta(z) over all 2^n outcomes, enumeration order (reference :132-137)."""
        if self.conditioning_dim > 0 and x_condition is not None:
            print("Warning: Conditioning with x_condition not fully implemented in PQC ansatz yet.")
        return self.pqc(weights=self.theta)

    def get_prob_dict(self, x_condition=None):
        """{outcome tuple: probability} (reference :139-151)."""
        probs = self.get_probabilities(x_condition).detach().cpu().numpy()
        return {outcome: probs[i] for i, outcome in enumerate(self.all_outcomes_tuples)}

    def sample(self, num_samples=1, x_condition=None):
        """float32 samples [num_samples, n] drawn from q_theta."""
        probs = self.get_probabilities(x_condition).detach().to(torch.float64)
        idx = torch.multinomial(probs / probs.sum(), num_samples, replacement=True)
        shifts = torch.arange(self.num_latent_vars - 1, -1, -1, device=idx.device)
        return ((idx.unsqueeze(-1) >> shifts) & 1).to(torch.float32)
