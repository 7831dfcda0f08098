"""Flatten a BayesianNetwork (+ evidence, latent names) into the arrays tnk_bn_joint reads.

Mirrors how compute_prob_joint_xz builds its assignment (reference stein_utils.py:62-67):
observed values first, latent values written on top (a latent name wins a clash), every
other node of `bn.nodes` is hidden and summed out in `bn.nodes` order, most significant
hidden node first (generate_all_binary_outcomes order, stein_utils.py:96).  Callable CPTs
(bayesian_network.py:136-137) are tabulated over all parent configurations here, once.
"""
from dataclasses import dataclass

import numpy as np

NODE_LATENT, NODE_CONST, NODE_HIDDEN = 0, 1, 2
MAX_NODES = 64
MAX_LATENT = 30


@dataclass
class FlatBN:
    n_latent: int
    n_hidden: int
    node_kind: np.ndarray   # int32 [V]
    node_arg: np.ndarray    # int32 [V]
    pa_off: np.ndarray      # int32 [V+1]
    pa_idx: np.ndarray      # int32 [sum k_v]
    cpt_off: np.ndarray     # int32 [V]
    cpt: np.ndarray         # float64 [rows, 2]

    @property
    def n_nodes(self):
        return int(self.node_kind.shape[0])


def flatten_bn(bn, x_dict, latent_vars_names, n_latent=None):
    """Build the descriptor.  `n_latent` defaults to len(latent_vars_names); latent names
    that are not nodes of `bn` simply do not influence p(x,z), as in the reference."""
    n = len(latent_vars_names) if n_latent is None else int(n_latent)
    if n > MAX_LATENT:
        raise ValueError(f"{n} latent variables exceed the supported maximum of {MAX_LATENT}")
    nodes = list(bn.nodes)
    if len(nodes) > MAX_NODES:
        raise ValueError(f"{len(nodes)} nodes exceed the supported maximum of {MAX_NODES}")
    x_dict = dict(x_dict) if x_dict else {}
    latent_index = {}
    for k, name in enumerate(latent_vars_names):
        latent_index[name] = k  # a repeated name keeps its last position, like dict assignment
    hidden = [name for name in nodes if name not in latent_index and name not in x_dict]
    m = len(hidden)
    if m > 20:
        raise ValueError(f"{m} hidden nodes to sum out exceed the supported maximum of 20")
    node_pos = {name: v for v, name in enumerate(nodes)}

    kind = np.zeros(len(nodes), dtype=np.int32)
    arg = np.zeros(len(nodes), dtype=np.int32)
    pa_off = np.zeros(len(nodes) + 1, dtype=np.int32)
    pa_idx = []
    cpt_off = np.zeros(len(nodes), dtype=np.int32)
    rows = []
    for v, name in enumerate(nodes):
        if name in latent_index:
            kind[v], arg[v] = NODE_LATENT, n - 1 - latent_index[name]   # MSB-first enumeration
        elif name in x_dict:
            value = int(x_dict[name])
            if value not in (0, 1):
                raise ValueError(f"evidence for {name} must be 0 or 1, got {x_dict[name]!r}")
            kind[v], arg[v] = NODE_CONST, value
        else:
            kind[v], arg[v] = NODE_HIDDEN, m - 1 - hidden.index(name)
        parents = list(bn.parents[name]) if name in bn.parents else []
        for p in parents:
            pa_idx.append(node_pos[p])
        pa_off[v + 1] = len(pa_idx)
        cpt_off[v] = len(rows)
        entry = bn.cpts[name]
        k = len(parents)
        for cfg in range(1 << k):
            key = tuple((cfg >> (k - 1 - t)) & 1 for t in range(k))
            row = entry(key) if callable(entry) else entry.get(key)
            if row is None:
                # the reference only fails if this row is actually looked up
                # (bayesian_network.py:141-142): poison it, joint_table() raises if it is hit
                rows.append((float("nan"), float("nan")))
            else:
                rows.append((float(row[0]), float(row[1])))
    cpt = np.asarray(rows, dtype=np.float64).reshape(-1, 2)
    return FlatBN(n, m, kind, arg, pa_off, np.asarray(pa_idx, dtype=np.int32), cpt_off, cpt)
