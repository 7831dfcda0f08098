"""KSD variational inference with a quantum Born machine (PennyLane- or simulator-supplied q).

Drop-in for the reference's `ksd_vi_quantum.KSDVariationalInference`
(reference ksd_vi_quantum.py:18-192): same constructor (note `num_latent_vars` comes from
`qbm_num_latent_vars`, :35), same `train` signature, history keys `loss_ksd, tvd, grad_norm`,
no entropy term and no early stopping.  The hot loop (:122-142) is the CUDA pair kernel;
`loss.backward()` hands dK/dq = 2 K_p q back to whatever produced q (PennyLane's
parameter-shift vjp, or autograd through the torch state-vector simulator).
"""
import copy
from functools import partial

import numpy as np
import torch
import torch.nn.utils as nn_utils
import torch.optim as optim

from . import ops
from .ksd_vi import ScoreCache, posterior_vector
from .quantum_born_machine import QuantumBornMachine
from .stein_utils import base_hamming_kernel_torch
from .utils import generate_all_binary_outcomes


class KSDVariationalInference:
    def __init__(self, bayesian_network, latent_vars_names, observed_vars_names,
                 qbm_num_latent_vars, qbm_ansatz_layers=1, qbm_conditioning_dim=0,
                 qbm_pennylane_device_name="default.qubit", qbm_ansatz_type="hardware_efficient",
                 qbm_init_method="small_random", base_kernel_length_scale=1.0, pytorch_device='cpu',
                 *, precision="fp64", compute_device="cuda", algorithm="pairs", group=None):
        self.bn = bayesian_network
        self.latent_vars_names = latent_vars_names
        self.observed_vars_names = observed_vars_names
        self.num_latent_vars = qbm_num_latent_vars
        self.num_observed_vars = len(observed_vars_names)
        self.pytorch_device = pytorch_device
        self.born_machine = QuantumBornMachine(
            num_latent_vars=self.num_latent_vars, ansatz_layers=qbm_ansatz_layers,
            conditioning_dim=qbm_conditioning_dim, device_name=qbm_pennylane_device_name,
            ansatz_type=qbm_ansatz_type, init_method=qbm_init_method).to(pytorch_device)
        self.all_latent_states_tuples = generate_all_binary_outcomes(self.num_latent_vars)
        self.num_possible_latent_states = len(self.all_latent_states_tuples)
        self.base_kernel_length_scale = base_kernel_length_scale
        self.base_kernel_func = partial(base_hamming_kernel_torch, num_vars=self.num_latent_vars,
                                        length_scale=base_kernel_length_scale)
        self.precision, self.compute_device, self.algorithm = precision, compute_device, algorithm
        # opt-in sharding of the pair space over a torch.distributed group whose ranks all train the same circuit
        # (parameters broadcast from the group's rank 0 at the start of train())
        self.group = group
        self.tables = None
        self._score_function_cache = ScoreCache()

    def _get_precomputed_s_p(self, z_tuple, x_dict):
        if z_tuple not in self._score_function_cache:
            self._precompute_all_s_p(x_dict)
        return self._score_function_cache[z_tuple]

    def _precompute_all_s_p(self, x_dict):
        """Reference ksd_vi_quantum.py:70-75, as two kernel launches."""
        if len(self.latent_vars_names) != self.num_latent_vars:
            # the reference would index z_tuple[idx] past its end here (stein_utils.py:63-64)
            raise IndexError("qbm_num_latent_vars does not match len(latent_vars_names)")
        self._score_function_cache.clear()
        print("Precomputing score functions s_p(x,z)...")
        self.tables = ops.SteinTables.from_network(
            self.bn, x_dict, self.latent_vars_names, length_scale=self.base_kernel_length_scale,
            precision=self.precision, device=self.compute_device)
        self._score_function_cache.bind(self.tables, self.all_latent_states_tuples, self.pytorch_device)
        print("Score functions precomputed.")

    def train(self, x_observation_dict, num_epochs, lr_born_machine,
              verbose=True, true_posterior_for_tvd=None,
              use_lr_scheduler=True, gradient_clip_norm=10.0,
              optimizer_type="adam", adam_betas=(0.9, 0.999)):
        if self.num_observed_vars > 0 and set(x_observation_dict.keys()) != set(self.observed_vars_names):
            raise ValueError("Keys in x_observation_dict must match self.observed_vars_names.")
        x_condition = None
        if self.num_observed_vars > 0 and self.born_machine.conditioning_dim > 0:
            x_condition = torch.tensor([x_observation_dict[name] for name in self.observed_vars_names],
                                       dtype=torch.float32, device=self.pytorch_device)
        self._precompute_all_s_p(x_observation_dict)

        params = list(self.born_machine.parameters())
        if self.group is not None:
            import torch.distributed as dist
            src = dist.get_global_rank(self.group, 0)
            with torch.no_grad():
                for p in params:
                    buf = p.data.to(self.tables.device) if dist.get_backend(self.group) == "nccl" and not p.is_cuda else p.data
                    dist.broadcast(buf, src=src, group=self.group)
                    if buf is not p.data:
                        p.data.copy_(buf)
        if optimizer_type == "adam":
            optimizer = optim.Adam(params, lr=lr_born_machine, betas=adam_betas)
        elif optimizer_type == "sgd":
            optimizer = optim.SGD(params, lr=lr_born_machine, momentum=0.9)
        else:
            optimizer = optim.Adam(params, lr=lr_born_machine)
        scheduler = None
        if use_lr_scheduler:
            scheduler = optim.lr_scheduler.CosineAnnealingLR(optimizer, T_max=num_epochs, eta_min=lr_born_machine / 10)

        history = {'loss_ksd': [], 'tvd': [], 'grad_norm': []}
        p_true, stray_mass = (None, 0.0)
        if true_posterior_for_tvd is not None:
            p_true, stray_mass = posterior_vector(true_posterior_for_tvd, self.num_latent_vars, self.pytorch_device)
        best_tvd, best_params = float('inf'), None
        last_grad_norm = 0.0
        log_every = num_epochs // 10 if num_epochs >= 10 else 1

        for epoch in range(num_epochs):
            optimizer.zero_grad()
            q = self.born_machine.get_probabilities(x_condition=x_condition).squeeze().to(torch.float64)
            if verbose and epoch % log_every == 0:
                print(f"  Epoch {epoch+1} Q Probs (first 4): {q.detach().cpu().numpy()[:4]}")
            if q.shape[0] != self.num_possible_latent_states:
                raise ValueError("Probabilities from Born machine have unexpected shape")

            K = ops.ksd_quadratic_form(q, self.tables, group=self.group, algorithm=self.algorithm)   # ksd_vi_quantum.py:122-142
            loss = torch.sqrt(K.clamp(min=1e-12))                    # :144-145
            if torch.isnan(loss) or torch.isinf(loss):
                print(f"Warning: NaN or Inf KSD loss: {loss.item()}. Skipping update.")
            else:
                loss.backward()
                last_grad_norm = nn_utils.clip_grad_norm_(params, gradient_clip_norm)
                if verbose and epoch % log_every == 0:
                    print(f"  Epoch {epoch+1} Grad Norm (after clipping): {last_grad_norm:.4f}")
                optimizer.step()
                if scheduler is not None:
                    scheduler.step()
            history['loss_ksd'].append(loss.item())
            history['grad_norm'].append(last_grad_norm)   # a tensor, as in the reference (:164)

            if p_true is not None:
                with torch.no_grad():
                    q_now = self.born_machine.get_probabilities(x_condition=x_condition).squeeze().to(torch.float64)
                    tvd = float(0.5 * ((p_true - q_now).abs().sum().item() + stray_mass))
                history['tvd'].append(tvd)
                if tvd < best_tvd:
                    best_tvd = tvd
                    # the reference keeps an aliased state_dict() here (:175), so its "best" is in
                    # fact the final parameters; a real snapshot is what the comment there intends
                    best_params = copy.deepcopy(self.born_machine.state_dict())
            else:
                history['tvd'].append(np.nan)

            if verbose and (epoch % max(1, num_epochs // 20) == 0 or epoch == num_epochs - 1):
                msg = f"Epoch {epoch+1}/{num_epochs} | KSD: {loss.item():.6f}"
                if scheduler is not None:
                    msg += f" | LR: {scheduler.get_last_lr()[0]:.6f}"
                if p_true is not None and not np.isnan(history['tvd'][-1]):
                    msg += f" | TVD: {history['tvd'][-1]:.6f}"
                print(msg)

        if best_params is not None and verbose:   # restore only when verbose, as the reference (:188-190)
            print(f"\nRestoring best parameters (TVD: {best_tvd:.6f})")
            self.born_machine.load_state_dict(best_params)
        return history
