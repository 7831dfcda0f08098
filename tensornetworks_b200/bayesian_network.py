"""Binary Bayesian-network container (drop-in for the reference's `bayesian_network` module).

Same public surface as reference bayesian_network.py:6-383 (`nodes`, `parents`, `cpts`,
`node_to_index`, `add_node`, `sample_forward`, `get_joint_probability`, `get_true_posterior`,
`get_prior_distribution`, `get_sprinkler_network`).  Only `nodes`/`parents`/`cpts` and
`get_joint_probability` lie on the KSD hot path; the GPU path reads them through
`tensornetworks_b200.flatten.flatten_bn`.
"""
from collections import defaultdict

import numpy as np

from .utils import generate_all_binary_outcomes


class BayesianNetwork:
    """Ordered binary nodes with dict or callable CPTs (reference bayesian_network.py:6-16)."""

    def __init__(self):
        self.nodes = []
        self.parents = defaultdict(list)
        self.cpts = {}
        self.node_to_index = {}

    def add_node(self, name, cpt, parent_names=None):
        """Append a node; parents must already exist, so `nodes` is a topological order
        (reference bayesian_network.py:18-51).  CPT keys are parent-value tuples in
        `parent_names` order; values are {0: p0, 1: p1}; a callable CPT maps the key to that dict."""
        if name in self.nodes:
            raise ValueError(f"Node {name} already exists.")
        self.nodes.append(name)
        self.node_to_index[name] = len(self.nodes) - 1
        if parent_names:
            for parent in parent_names:
                if parent not in self.nodes:
                    raise ValueError(f"Parent node {parent} for {name} not found. Add parents first.")
            self.parents[name] = list(parent_names)
        self.cpts[name] = cpt

    def _cpt_row(self, node_name, parent_values_tuple):
        entry = self.cpts[node_name]
        return entry(parent_values_tuple) if callable(entry) else entry.get(parent_values_tuple)

    def _get_prob_node_given_parents(self, node_name, parent_values_tuple):
        """P(node=1 | parents) with the normalisation check (reference bayesian_network.py:53-74)."""
        entry = self.cpts[node_name]
        if not callable(entry) and not isinstance(entry, dict):
            raise TypeError(f"CPT for node {node_name} has an invalid type.")
        row = self._cpt_row(node_name, parent_values_tuple)
        if row is None:
            raise ValueError(f"CPT entry for node {node_name} with parent values {parent_values_tuple} not found.")
        if not isinstance(row, dict) or 0 not in row or 1 not in row:
            raise ValueError(f"CPT for {node_name} with parent values {parent_values_tuple} must return a dict {{0: p0, 1: p1}}")
        if not np.isclose(row[0] + row[1], 1.0):
            raise ValueError(f"Probabilities for node {node_name} given parents {parent_values_tuple} do not sum to 1: {row}")
        return row[1]

    def sample_forward(self, num_samples=1):
        """Ancestral sampling in `nodes` order (reference bayesian_network.py:76-109)."""
        dicts, tuples = [], []
        for _ in range(num_samples):
            sample = {}
            for name in self.nodes:
                pa = tuple(sample[p] for p in self.parents[name])
                sample[name] = 1 if np.random.rand() < self._get_prob_node_given_parents(name, pa) else 0
            dicts.append(sample)
            tuples.append(tuple(sample[name] for name in self.nodes))
        return dicts, tuples

    def get_joint_probability(self, full_assignment_tuple):
        """Product over `nodes` order, starting at 1.0, of CPT[parents][value]
        (reference bayesian_network.py:111-146).  The CUDA joint kernel reproduces exactly
        this multiplication order so the 1e-12 score threshold falls on the same states."""
        if len(full_assignment_tuple) != len(self.nodes):
            raise ValueError("Full assignment tuple length must match the number of nodes.")
        value = dict(zip(self.nodes, full_assignment_tuple))
        prob = 1.0
        for name in self.nodes:
            pa = tuple(value[p] for p in self.parents[name])
            row = self._cpt_row(name, pa)
            if row is None:
                raise ValueError(f"CPT entry for node {name} with parent values {pa} not found.")
            prob *= row[value[name]]
        return prob

    def _marginal_over_rest(self, fixed):
        """Sum of the joint over every node not in `fixed` (dict name -> value)."""
        rest = [name for name in self.nodes if name not in fixed]
        total = 0.0
        for other in generate_all_binary_outcomes(len(rest)):
            full = dict(fixed)
            full.update(zip(rest, other))
            total += self.get_joint_probability(tuple(full[name] for name in self.nodes))
        return total

    def get_true_posterior(self, latent_vars_names, observed_vars_dict):
        """P(latent | observed) by enumeration -> (dict, P(observed))
        (reference bayesian_network.py:148-253)."""
        if not all(name in self.nodes for name in latent_vars_names):
            raise ValueError("One or more latent variable names not in the network.")
        if not all(name in self.nodes for name in observed_vars_dict.keys()):
            raise ValueError("One or more observed variable names not in the network.")
        if set(latent_vars_names) & set(observed_vars_dict.keys()):
            raise ValueError("Latent and observed variables must be disjoint.")
        unnorm = {}
        for z in generate_all_binary_outcomes(len(latent_vars_names)):
            fixed = dict(observed_vars_dict)
            fixed.update(zip(latent_vars_names, z))
            unnorm[z] = self._marginal_over_rest(fixed)
        evidence = sum(unnorm.values())
        if evidence == 0:
            print(f"Warning: P(Observed) is zero for evidence {observed_vars_dict}. Posterior is ill-defined.")
            return {k: 0.0 for k in unnorm}, evidence
        return {k: v / evidence for k, v in unnorm.items()}, evidence

    def get_prior_distribution(self, var_names_ordered):
        """P(var_names_ordered), other nodes summed out (reference bayesian_network.py:255-306)."""
        prior = {}
        for t in generate_all_binary_outcomes(len(var_names_ordered)):
            prior[t] = self._marginal_over_rest(dict(zip(var_names_ordered, t)))
        if not np.isclose(sum(prior.values()), 1.0):
            print(f"Warning: Prior probabilities for {var_names_ordered} sum to {sum(prior.values())}, not 1.0.")
        return prior


def _row(p1):
    return {0: 1 - p1, 1: p1}


def get_sprinkler_network(random_cpts=False):
    """Cloudy -> {Sprinkler, Rain} -> GrassWet, node order C,S,R,W
    (reference bayesian_network.py:312-383; textbook CPTs 358-381, random CPTs ~U(0.01,0.99) 320-355)."""
    bn = BayesianNetwork()
    if random_cpts:
        draw = lambda: np.random.uniform(0.01, 0.99)
        bn.add_node('C', cpt={(): _row(draw())})
        s0, s1 = draw(), draw()
        bn.add_node('S', cpt={(0,): _row(s0), (1,): _row(s1)}, parent_names=['C'])
        r0, r1 = draw(), draw()
        bn.add_node('R', cpt={(0,): _row(r0), (1,): _row(r1)}, parent_names=['C'])
        w = [draw() for _ in range(4)]
        bn.add_node('W', cpt={(0, 0): _row(w[0]), (0, 1): _row(w[1]), (1, 0): _row(w[2]), (1, 1): _row(w[3])},
                    parent_names=['S', 'R'])
        return bn
    bn.add_node('C', cpt={(): {0: 0.5, 1: 0.5}})
    bn.add_node('S', cpt={(0,): {0: 0.5, 1: 0.5}, (1,): {0: 0.9, 1: 0.1}}, parent_names=['C'])
    bn.add_node('R', cpt={(0,): {0: 0.8, 1: 0.2}, (1,): {0: 0.2, 1: 0.8}}, parent_names=['C'])
    bn.add_node('W', cpt={(0, 0): {0: 0.99, 1: 0.01}, (0, 1): {0: 0.1, 1: 0.9},
                          (1, 0): {0: 0.1, 1: 0.9}, (1, 1): {0: 0.01, 1: 0.99}}, parent_names=['S', 'R'])
    return bn
