"""Dense restatement of the three ansatz circuits of the reference's QuantumBornMachine.  TEST INFRASTRUCTURE ONLY.

Reference quantum_born_machine.py:57-128 builds the circuits from PennyLane gates (Hadamard, RX, RY, RZ, CNOT, CZ on
`default.qubit`, probabilities over wires 0..n-1 with wire 0 the most significant bit of the outcome index).
PennyLane is not installed in this image or on the GPU box, so its published gate matrices are restated here and the
circuit is multiplied out as explicit 2^n x 2^n unitaries, gate by gate with np.kron - no state-vector tricks, no
shared code with tensornetworks_b200/quantum_born_machine.py.  Parity note: without PennyLane the gate conventions
themselves (RX = exp(-i t X/2), RY = exp(-i t Y/2), RZ = exp(-i t Z/2) = diag(e^{-it/2}, e^{it/2}); CNOT(wires=[c,t]))
are pinned only to PennyLane's documentation, not to a run of it.
"""
import numpy as np

I2 = np.eye(2, dtype=np.complex128)
X = np.array([[0, 1], [1, 0]], dtype=np.complex128)
Y = np.array([[0, -1j], [1j, 0]], dtype=np.complex128)
Z = np.array([[1, 0], [0, -1]], dtype=np.complex128)
H = np.array([[1, 1], [1, -1]], dtype=np.complex128) / np.sqrt(2.0)
P0 = np.array([[1, 0], [0, 0]], dtype=np.complex128)
P1 = np.array([[0, 0], [0, 1]], dtype=np.complex128)


def _rot(pauli, t):
    return np.cos(t / 2.0) * I2 - 1j * np.sin(t / 2.0) * pauli


def _embed(ops, n):
    """kron of single-qubit operators; ops: {wire: 2x2}, identity elsewhere; wire 0 is the leftmost (most significant) factor."""
    out = np.ones((1, 1), dtype=np.complex128)
    for w in range(n):
        out = np.kron(out, ops.get(w, I2))
    return out


def _controlled(control, target, gate, n):
    return _embed({control: P0}, n) + _embed({control: P1, target: gate}, n)


def circuit_unitary(weights, n, layers, ansatz_type):
    """U such that the circuit's output state is U|0...0>, following quantum_born_machine.py:57-128 line by line."""
    w = np.asarray(weights, dtype=np.float64)
    U = np.eye(1 << n, dtype=np.complex128)
    k = 0

    def apply(G):
        nonlocal U
        U = G @ U

    if ansatz_type in ("hardware_efficient", "all_to_all"):
        for i in range(n):                                        # :63-64, :97-98
            apply(_embed({i: H}, n))
        for layer in range(layers):
            for i in range(n):                                    # :68-71, :102-105
                apply(_embed({i: _rot(X, w[k])}, n)); k += 1
                apply(_embed({i: _rot(Y, w[k])}, n)); k += 1
                apply(_embed({i: _rot(Z, w[k])}, n)); k += 1
            if n > 1:
                if ansatz_type == "hardware_efficient":
                    for i in range(n - 1):                        # :76-77
                        apply(_controlled(i, i + 1, X, n))
                    if n > 2:                                     # :79-80
                        apply(_controlled(n - 1, 0, X, n))
                    if layer % 2 == 0 and n > 2:                  # :83-85
                        for i in range(0, n - 2, 2):
                            apply(_controlled(i, i + 2, Z, n))
                else:
                    for i in range(n):                            # :108-111
                        for j in range(i + 1, n):
                            apply(_controlled(i, j, Z, n))
    else:                                                         # basic, :116-126
        for layer in range(layers):
            for i in range(n):
                apply(_embed({i: _rot(Y, w[k])}, n)); k += 1
                apply(_embed({i: _rot(Z, w[k])}, n)); k += 1
            if n > 1:
                for i in range(n - 1):
                    apply(_controlled(i, i + 1, X, n))
                if n > 2:
                    apply(_controlled(n - 1, 0, X, n))
    assert k == len(w), (k, len(w))
    return U


def circuit_probs(weights, n, layers, ansatz_type):
    """qml.probs(wires=range(n)) of the circuit: |U|0>|^2 in outcome-index order (wire 0 = MSB), :87/:113/:128."""
    if n == 0:
        return np.ones(1)
    amp = circuit_unitary(weights, n, layers, ansatz_type)[:, 0]
    return (amp.real ** 2 + amp.imag ** 2)
