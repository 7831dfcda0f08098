"""Drop-in for the reference's `stein_utils` module (same names, arguments and errors).

The scalar entry points keep working exactly as in the reference - they are its unit of
test (stein_utils.py:199-253) - and are thin host-side code.  The 2^n / 4^n work the
reference performs by calling them in Python loops goes through the batched CUDA entry
points instead: `precompute_tables` (joint + score for all states) and
`tensornetworks_b200.ops.ksd_quadratic_form` (the pair space).
"""
import numpy as np
import torch

from .utils import generate_all_binary_outcomes

SCORE_ZERO_THRESHOLD = 1e-12  # reference stein_utils.py:126


def flip_bit(z_tuple, index):
    """Flip one entry of a binary tuple (reference stein_utils.py:7-11)."""
    return tuple(1 - v if i == index else v for i, v in enumerate(z_tuple))


def _common(z1_tensor, z2_tensor):
    if z1_tensor.device != z2_tensor.device:
        z2_tensor = z2_tensor.to(z1_tensor.device)
    if z1_tensor.dtype != z2_tensor.dtype:
        dtype = torch.float64 if torch.float64 in (z1_tensor.dtype, z2_tensor.dtype) else torch.float32
        z1_tensor, z2_tensor = z1_tensor.to(dtype), z2_tensor.to(dtype)
    return z1_tensor, z2_tensor


def hamming_distance_torch(z1_tensor, z2_tensor):
    """L1 distance of binary vectors along the last axis, with the reference's device/dtype
    promotion (reference stein_utils.py:13-28).  On packed states this is popc(i ^ j)."""
    z1_tensor, z2_tensor = _common(z1_tensor, z2_tensor)
    return (z1_tensor - z2_tensor).abs().sum(dim=-1)


def base_hamming_kernel_torch(z1_tensor, z2_tensor, num_vars, length_scale=1.0):
    """k(z1,z2) = exp(-||z1-z2||_1 / (num_vars*length_scale)); 1.0 when num_vars == 0; the
    zero-denominator branch of the reference is kept (reference stein_utils.py:30-55)."""
    if num_vars == 0:
        ref = z1_tensor if hasattr(z1_tensor, 'dtype') else z2_tensor
        dtype = ref.dtype if hasattr(ref, 'dtype') else torch.float64
        device = ref.device if hasattr(ref, 'device') else 'cpu'
        return torch.tensor(1.0, device=device, dtype=dtype)
    dtype = torch.float64 if torch.float64 in (z1_tensor.dtype, z2_tensor.dtype) else torch.float32
    if z1_tensor.device != z2_tensor.device:
        z2_tensor = z2_tensor.to(z1_tensor.device)
    distance = hamming_distance_torch(z1_tensor.to(dtype), z2_tensor.to(dtype))
    denominator = float(num_vars) * float(length_scale)
    if denominator == 0:
        return torch.ones_like(distance) if torch.all(distance == 0) else torch.zeros_like(distance)
    return torch.exp(-distance / torch.tensor(denominator, device=distance.device, dtype=distance.dtype))


def compute_prob_joint_xz(bn, x_dict, z_tuple, latent_vars_names, observed_vars_names, device='cpu'):
    """p(x, z) for ONE latent assignment as a Python float: evidence clamped, latent values on
    top, every other node summed out (reference stein_utils.py:58-112).  For all 2^n states at
    once use `precompute_tables` / ops.joint_table (bit-identical values)."""
    assignment = dict(x_dict) if x_dict else {}
    for idx, name in enumerate(latent_vars_names):
        assignment[name] = z_tuple[idx]
    others = [node for node in bn.nodes if node not in assignment]
    if not others:
        return float(bn.get_joint_probability(tuple(assignment[node] for node in bn.nodes)))
    total = 0.0
    for other_values in generate_all_binary_outcomes(len(others)):
        full = dict(assignment)
        full.update(zip(others, other_values))
        total += bn.get_joint_probability(tuple(full[node] for node in bn.nodes))
    return float(total)


def get_score_function_sp_for_z(bn, x_dict, z_tuple, latent_vars_names, observed_vars_names, device='cpu'):
    """s_p(x,z)_i = 1 - p(x, flip_i z)/p(x,z) as a float64 tensor [n]; all zeros when
    |p(x,z)| < 1e-12 (reference stein_utils.py:115-136)."""
    n = len(latent_vars_names)
    s_p = torch.zeros(n, device=device, dtype=torch.float64)
    p = compute_prob_joint_xz(bn, x_dict, z_tuple, latent_vars_names, observed_vars_names, device=device)
    if abs(p) < SCORE_ZERO_THRESHOLD:
        return s_p
    for i in range(n):
        p_flip = compute_prob_joint_xz(bn, x_dict, flip_bit(z_tuple, i), latent_vars_names, observed_vars_names, device=device)
        s_p[i] = 1.0 - (p_flip / p)
    return s_p


def get_stein_kernel_kp_value(z1_tuple, z2_tuple, x_dict, bn, latent_vars_names, observed_vars_names,
                              base_kernel_func, sp_at_z1, sp_at_z2, device='cpu'):
    """One Stein-kernel value k_p(z1,z2|x), float64 0-dim tensor (reference stein_utils.py:138-197).

    Same four Eq.-13 terms for ANY base kernel function; the reference's 5n+1 base-kernel calls
    become 3n+1 because term 4 re-uses the flipped evaluations of terms 2 and 3.
    `x_dict`, `bn`, `observed_vars_names` are accepted and unused, as in the reference."""
    n = len(latent_vars_names)
    dtype = torch.float64
    if n == 0:
        return torch.tensor(0.0, device=device, dtype=dtype)
    as_vec = lambda t: torch.tensor(t, dtype=dtype, device=device)
    z1, z2 = as_vec(z1_tuple), as_vec(z2_tuple)
    sp1 = sp_at_z1.to(dtype=dtype, device=device)
    sp2 = sp_at_z2.to(dtype=dtype, device=device)
    k = base_kernel_func(z1, z2).to(dtype)
    k_z2_flip = torch.zeros(n, device=device, dtype=dtype)   # k(z1, flip_j z2)
    k_z1_flip = torch.zeros(n, device=device, dtype=dtype)   # k(flip_i z1, z2)
    k_both = torch.zeros(n, device=device, dtype=dtype)      # k(flip_i z1, flip_i z2)
    for j in range(n):
        z1_j, z2_j = as_vec(flip_bit(z1_tuple, j)), as_vec(flip_bit(z2_tuple, j))
        k_z2_flip[j] = base_kernel_func(z1, z2_j).to(dtype)
        k_z1_flip[j] = base_kernel_func(z1_j, z2).to(dtype)
        k_both[j] = base_kernel_func(z1_j, z2_j).to(dtype)
    term1 = torch.dot(sp1, sp2) * k                          # reference :161-164
    term2 = -torch.dot(sp1, k - k_z2_flip)                   # :166-171
    term3 = -torch.dot(k - k_z1_flip, sp2)                   # :173-178
    term4 = (k - k_z2_flip - k_z1_flip + k_both).sum()       # :180-194
    return term1 + term2 + term3 + term4


def precompute_tables(bn, x_dict, latent_vars_names, observed_vars_names=None, length_scale=1.0,
                      precision="fp32", device="cuda", keep_scores=False, n_latent=None):
    """Batched replacement of the `_precompute_all_s_p` loop (reference ksd_vi.py:55-60): joint
    table and score table of all 2^n states on the GPU (ops.SteinTables)."""
    from . import ops
    return ops.SteinTables.from_network(bn, x_dict, latent_vars_names, length_scale=length_scale,
                                        precision=precision, device=device, keep_scores=keep_scores,
                                        n_latent=n_latent)
