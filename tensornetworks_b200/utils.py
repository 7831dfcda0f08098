"""Enumeration order, TVD and key helpers (drop-in for the reference's `utils` module).

The state order is the bit-exact contract of the whole path: outcome i is the n-digit
binary expansion of i, most significant digit first (reference utils.py:77-91), so latent
variable k is bit (n-1-k) of the packed index.  `generate_all_binary_outcomes` returns a
lazy, tensor-free sequence so that n = 24 does not build 16.7 M Python tuples up front.
"""
from collections.abc import Sequence

import numpy as np


class BinaryOutcomes(Sequence):
    """Lazy list of all n-bit outcomes as tuples, MSB first (reference utils.py:77-91).

    Behaves like the reference's list for len/indexing/iteration/`index`/`==`, but holds
    no storage; `packed(t)` gives the integer index of an outcome tuple.
    """

    def __init__(self, num_vars):
        self.num_vars = int(num_vars)

    def __len__(self):
        return 1 << self.num_vars

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        i = int(i)
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError("outcome index out of range")
        n = self.num_vars
        return tuple((i >> (n - 1 - k)) & 1 for k in range(n))

    def __iter__(self):
        for i in range(len(self)):
            yield self[i]

    def __contains__(self, t):
        return isinstance(t, tuple) and len(t) == self.num_vars and all(v in (0, 1) for v in t)

    def packed(self, outcome_tuple):
        idx = 0
        for v in outcome_tuple:
            idx = (idx << 1) | int(v)
        return idx

    def index(self, outcome_tuple, *args):
        if tuple(outcome_tuple) not in self:
            raise ValueError(f"{outcome_tuple} is not in list")
        return self.packed(outcome_tuple)

    def __eq__(self, other):
        if isinstance(other, BinaryOutcomes):
            return other.num_vars == self.num_vars
        if isinstance(other, (list, tuple)):
            return len(other) == len(self) and all(a == b for a, b in zip(self, other))
        return NotImplemented

    def __repr__(self):
        if self.num_vars <= 4:
            return repr(list(self))
        return f"BinaryOutcomes(num_vars={self.num_vars})"


def generate_all_binary_outcomes(num_vars):
    """All 2^num_vars outcomes, MSB first; num_vars=0 -> [()] (reference utils.py:77-91)."""
    return BinaryOutcomes(num_vars)


def get_binary_key(outcome_tuple):
    """(0,1,1) -> '011' (reference utils.py:69-71)."""
    return "".join(str(v) for v in outcome_tuple)


def get_outcome_tuple(binary_key):
    """'011' -> (0,1,1) (reference utils.py:73-75)."""
    return tuple(int(c) for c in binary_key)


def calculate_tvd(p_true, p_approx):
    """Total variation distance of two dicts or two equal-shape arrays (reference utils.py:6-36)."""
    if isinstance(p_true, dict) and isinstance(p_approx, dict):
        keys = set(p_true) | set(p_approx)
        return 0.5 * sum(np.abs(p_true.get(k, 0.0) - p_approx.get(k, 0.0)) for k in keys)
    if isinstance(p_true, np.ndarray) and isinstance(p_approx, np.ndarray):
        if p_true.shape != p_approx.shape:
            raise ValueError("Probability arrays must have the same shape for simple TVD calculation.")
        return 0.5 * np.sum(np.abs(p_true - p_approx))
    raise TypeError("Inputs p_true and p_approx must be both dicts or both np.arrays.")


def plot_training_results(results_dict, title="Training Results"):
    """Headless stand-in for reference utils.py:38-67: no display stack on the B200 boxes,
    so the curves are summarised on stdout instead of drawn."""
    if not results_dict:
        print("No results to plot.")
        return
    print(title)
    for name, values in results_dict.items():
        vals = np.asarray([float(v) for v in values], dtype=np.float64)
        if vals.size:
            print(f"  {name}: first={vals[0]:.6g} last={vals[-1]:.6g} min={np.nanmin(vals):.6g}")
