"""KSD variational inference with a classical Born machine.

Drop-in for the reference's `ksd_vi.KSDVariationalInference` (reference ksd_vi.py:19-215):
same constructor, attributes, `train` signature, history keys, errors and quirks.  What
changes is where the work happens:

  _precompute_all_s_p   one joint + one score kernel over all 2^n states (was 2^n*(n+1)
                        Python calls, ksd_vi.py:55-60)
  the O(4^n) hot loop   ops.ksd_quadratic_form: one tiled CUDA pass that returns K and keeps
                        u = K_p q for the backward (was ksd_vi.py:110-131 + autograd graph)
  TVD tracking          tensor arithmetic in enumeration order (was a 2^n-entry dict per epoch)

Extra keyword-only arguments (defaults keep reference behaviour): `precision` ("fp64" |
"fp32" | "tc32") selects the pair-kernel arithmetic, `algorithm` ("pairs" = all 4^n pairs on the
fly, "pairs_sym" = every unordered pair of 128-state tiles once on the tensor cores (precision "tc32",
7 <= n <= 24), "modeproduct" = the Kronecker form, O(n^2 2^n)), `compute_device` the CUDA device of the KSD
kernels (the Born machine itself stays on `device`, as in the reference).
"""
import math
from functools import partial

import numpy as np
import torch
import torch.nn.utils as nn_utils
import torch.optim as optim

from . import ops
from .born_machine_classical_sim import ClassicalBornMachine
from .stein_utils import base_hamming_kernel_torch
from .utils import generate_all_binary_outcomes


class ScoreCache:
    """Dict-like view `z_tuple -> float64 score vector` over the device score table; stands in
    for the reference's `_score_function_cache` dict (ksd_vi.py:41-53) without 2^n tensors."""

    def __init__(self):
        self._tables = None
        self._outcomes = None

    def bind(self, tables, outcomes, device):
        self._tables, self._outcomes, self._device = tables, outcomes, device

    def clear(self):
        self._tables = None

    def __len__(self):
        return 0 if self._tables is None else self._tables.N

    def __contains__(self, z_tuple):
        return self._tables is not None and z_tuple in self._outcomes

    def __getitem__(self, z_tuple):
        if z_tuple not in self:
            raise KeyError(z_tuple)
        return self._tables.scores()[self._outcomes.packed(z_tuple)].to(self._device)


def posterior_vector(true_posterior, num_latent_vars, device):
    """Enumeration-ordered fp64 tensor of a posterior given as dict {tuple: p}, array or tensor,
    plus the probability mass sitting on keys that are not n-bit outcomes (they still count in
    calculate_tvd's union of keys, reference utils.py:19-26)."""
    N = 1 << num_latent_vars
    if isinstance(true_posterior, dict):
        vec = np.zeros(N, dtype=np.float64)
        outcomes = generate_all_binary_outcomes(num_latent_vars)
        stray = 0.0
        for key, p in true_posterior.items():
            if key in outcomes:
                vec[outcomes.packed(key)] = p
            else:
                stray += abs(p)
        return torch.from_numpy(vec).to(device), stray
    vec = torch.as_tensor(true_posterior, dtype=torch.float64).reshape(-1)
    if vec.numel() != N:
        raise ValueError("posterior vector must have 2^n entries")
    return vec.to(device), 0.0


class KSDVariationalInference:
    def __init__(self, bayesian_network, latent_vars_names, observed_vars_names,
                 born_machine_config, base_kernel_length_scale=1.0, device='cpu',
                 *, precision="fp64", compute_device="cuda", algorithm="pairs", cuda_graph=False,
                 distributed=False, group=None):
        self.bn = bayesian_network
        self.latent_vars_names = latent_vars_names
        self.observed_vars_names = observed_vars_names
        self.num_latent_vars = len(latent_vars_names)
        self.num_observed_vars = len(observed_vars_names)
        self.device = device
        self.cuda_graph = cuda_graph
        # the reference overrides the caller's init_method (ksd_vi.py:30)
        born_machine_config = {**born_machine_config, 'init_method': 'small_random'}
        self.born_machine = ClassicalBornMachine(num_latent_vars=self.num_latent_vars,
                                                 **born_machine_config).to(device)
        self.all_latent_states_tuples = generate_all_binary_outcomes(self.num_latent_vars)
        self.num_possible_latent_states = len(self.all_latent_states_tuples)
        self.base_kernel_length_scale = base_kernel_length_scale
        self.base_kernel_func = partial(base_hamming_kernel_torch, num_vars=self.num_latent_vars,
                                        length_scale=base_kernel_length_scale)
        self.precision = precision
        self.compute_device = compute_device
        self.algorithm = algorithm
        # Sharding the pair space over the ranks of a torch.distributed group is opt-in (every rank must train the same
        # Born machine on the same evidence): distributed=True uses `group` (default: the WORLD group).
        self.distributed = bool(distributed) or group is not None
        self.group = group
        self._exchange = None
        self.tables = None
        self._score_function_cache = ScoreCache()

    def _get_precomputed_s_p(self, z_tuple, x_dict):
        if z_tuple not in self._score_function_cache:
            self._precompute_all_s_p(x_dict)
        return self._score_function_cache[z_tuple]

    def _precompute_all_s_p(self, x_dict):
        """All 2^n score vectors in two kernel launches (reference ksd_vi.py:55-60)."""
        self._score_function_cache.clear()
        print("Precomputing score functions s_p(x,z)...")
        self.tables = ops.SteinTables.from_network(
            self.bn, x_dict, self.latent_vars_names, length_scale=self.base_kernel_length_scale,
            precision=self.precision, device=self.compute_device)
        self._score_function_cache.bind(self.tables, self.all_latent_states_tuples, self.device)
        print("Score functions precomputed.")

    def ksd_squared(self, q_probs_all_z):
        """K = sum_ij q_i q_j k_p(z_i,z_j) as a differentiable fp64 scalar (the hot loop)."""
        group, exchange = self._sharding()
        return ops.ksd_quadratic_form(q_probs_all_z, self.tables, group=group, algorithm=self.algorithm,
                                      exchange=exchange)

    def _world(self):
        import torch.distributed as dist
        if not self.distributed:
            return 1
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("distributed=True needs an initialised torch.distributed process group")
        return dist.get_world_size(self.group)

    def _sharding(self):
        """(group, exchange) of the sharded evaluation, or (None, None) on a single process.  The exchange object is
        ops.PeerComm (tnk_comm of include/tnk.h: our all-reduce kernels over peer-mapped buffers); its constructor is
        collective and either succeeds on every rank or raises on every rank, and the fallback to the plain
        NCCL/gloo collective is therefore taken by all ranks together."""
        import torch.distributed as dist
        if self._world() == 1:
            return None, None
        group = dist.group.WORLD if self.group is None else self.group
        if self._exchange is None:
            try:
                self._exchange = ops.PeerComm(self.tables.N + 1, self.tables.device, group=group)
            except Exception as exc:
                print(f"tnk_comm unavailable ({exc}); the exchange goes through torch.distributed collectives")
                self._exchange = False
        return group, (self._exchange or None)

    def _sync_replicas(self):
        """Sharded training assumes identical replicas: parameters are broadcast from rank 0 of the group at the start
        of train(); ksd_squared's inputs are then identical on every rank as long as the optimiser is deterministic."""
        import torch.distributed as dist
        if self._world() == 1:
            return
        group = dist.group.WORLD if self.group is None else self.group
        src = dist.get_global_rank(group, 0)
        backend = dist.get_backend(group)
        with torch.no_grad():
            for p in self.born_machine.parameters():
                buf = p.data if (p.is_cuda or backend != "nccl") else p.data.to(self.tables.device)
                dist.broadcast(buf, src=src, group=group)
                if buf is not p.data:
                    p.data.copy_(buf)

    def _graph_supported(self, params, optimizer_type):
        """The epoch can be replayed as CUDA graphs when everything it launches is on one CUDA device and
        fixed from epoch to epoch: unconditional Born machine (no dropout), Adam, a single process."""
        import torch.distributed as dist
        if not (self.cuda_graph and params and optimizer_type == "adam" and self.born_machine.conditioning_dim == 0):
            return False
        if self.distributed and dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            return False
        return all(p.is_cuda and p.device == self.tables.device for p in params)

    def _capture_epoch(self, params, optimizer, gradient_clip_norm, entropy_weight):
        """Two CUDA graphs per epoch: A = forward, loss, backward, gradient clipping; B = optimizer step.
        The host reads the three scalars between them (the reference's NaN/Inf guard decides whether B runs).
        Returns (graph_a, graph_b, scalars [ksd, entropy, loss] fp64, grad_norm)."""
        dev = params[0].device
        scal = torch.zeros(3, dtype=torch.float64, device=dev)
        gnorm = torch.zeros((), dtype=torch.float32, device=dev)

        def forward_backward():
            optimizer.zero_grad(set_to_none=True)
            q = self.born_machine.get_probabilities().squeeze()
            ksd_loss = torch.sqrt(self.ksd_squared(q).clamp(min=1e-12))
            entropy = self.born_machine.entropy()
            loss = ksd_loss - entropy_weight * entropy.to(ksd_loss.dtype)
            scal.copy_(torch.stack([ksd_loss.detach().double(), entropy.detach().double(), loss.detach().double()]))
            loss.backward()
            gnorm.copy_(nn_utils.clip_grad_norm_(params, gradient_clip_norm).detach().float())

        # warm-up on a side stream (workspaces, optimizer state), then put parameters and state back
        snapshot = [p.detach().clone() for p in params]
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(2):
                self.born_machine._pair_cache = None
                forward_backward()
                optimizer.step()
        torch.cuda.current_stream(dev).wait_stream(side)
        with torch.no_grad():
            for p, keep in zip(params, snapshot):
                p.copy_(keep)
            for state in optimizer.state.values():
                for value in state.values():
                    if torch.is_tensor(value):
                        value.zero_()
        self.born_machine._pair_cache = None
        graph_a, graph_b = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph_a, capture_error_mode="thread_local"):
            forward_backward()
        with torch.cuda.graph(graph_b, pool=graph_a.pool(), capture_error_mode="thread_local"):
            optimizer.step()
        self.born_machine._pair_cache = None
        return graph_a, graph_b, scal, gnorm

    def train(self, x_observation_dict, num_epochs, lr_born_machine,
              verbose=True, true_posterior_for_tvd=None,
              use_lr_scheduler=True, gradient_clip_norm=10.0,
              optimizer_type="adam", adam_betas=(0.9, 0.999),
              entropy_weight=0.01, patience=200):
        """Reference ksd_vi.py:62-215: loss = sqrt(clamp(K,1e-12)) - entropy_weight*H(q); Adam or
        SGD(momentum .9); cosine LR to lr/10; gradient clipping; NaN/Inf-skip; TVD tracking with
        best-q snapshot, early stop (patience, epoch > 300) and restore via set_fixed_probs.
        Returns history with keys loss_ksd, tvd, grad_norm, entropy."""
        if self.num_observed_vars > 0 and set(x_observation_dict.keys()) != set(self.observed_vars_names):
            raise ValueError("Keys in x_observation_dict must match self.observed_vars_names.")
        x_obs = [x_observation_dict[name] for name in self.observed_vars_names] if self.num_observed_vars > 0 else []
        x_obs_tensor = torch.tensor(x_obs, dtype=torch.float32, device=self.device)
        x_condition = None
        if self.born_machine.conditioning_dim > 0:
            if self.num_observed_vars == 0:
                raise ValueError("Born machine is conditional but no observed vars specified.")
            if self.born_machine.conditioning_dim != self.num_observed_vars:
                raise ValueError("Born machine conditioning_dim must match num_observed_vars.")
            x_condition = x_obs_tensor

        self._precompute_all_s_p(x_observation_dict)
        self._sync_replicas()

        params = list(self.born_machine.parameters())
        graphs = None
        if self._graph_supported(params, optimizer_type):
            # CUDA-graph epochs: the learning rate lives in a device scalar that the host
            # refreshes from an ordinary scheduler on a stand-in optimizer, so the schedule is the reference's.
            lr_device = torch.tensor(float(lr_born_machine), dtype=torch.float32, device=params[0].device)
            optimizer = optim.Adam(params, lr=lr_device, betas=adam_betas, capturable=True, fused=True)
            schedule_owner = optim.SGD([torch.zeros(1, requires_grad=True)], lr=lr_born_machine)
            graphs = self._capture_epoch(params, optimizer, gradient_clip_norm, entropy_weight)
        elif optimizer_type == "adam":
            # same update; on CUDA parameters one fused kernel instead of ~10 foreach passes over 2^n floats
            fused = bool(params) and all(p.is_cuda for p in params)
            optimizer = optim.Adam(params, lr=lr_born_machine, betas=adam_betas, fused=True if fused else None)
            schedule_owner = optimizer
        else:
            optimizer = optim.SGD(params, lr=lr_born_machine, momentum=0.9)
            schedule_owner = optimizer
        scheduler = None
        if use_lr_scheduler:
            scheduler = optim.lr_scheduler.CosineAnnealingLR(schedule_owner, T_max=num_epochs, eta_min=lr_born_machine / 10)

        history = {'loss_ksd': [], 'tvd': [], 'grad_norm': [], 'entropy': []}
        p_true, stray_mass = (None, 0.0)
        if true_posterior_for_tvd is not None:
            p_true, stray_mass = posterior_vector(true_posterior_for_tvd, self.num_latent_vars, self.device)

        def current_tvd():
            with torch.no_grad():
                # NB: like the reference (ksd_vi.py:161), this is a fresh forward pass, so an active
                # Dropout in the conditional Born machine draws a new mask here.
                q_now = self.born_machine.get_probabilities(x_condition=x_condition).squeeze()
                return float(0.5 * ((p_true - q_now.to(torch.float64)).abs().sum().item() + stray_mass))

        best_tvd, best_epoch, best_probs = float('inf'), -1, None
        epochs_without_improvement = 0
        last_grad_norm, last_entropy = 0.0, 0.0

        for epoch in range(num_epochs):
            if graphs is not None:
                graph_a, graph_b, scalars, grad_norm_dev = graphs
                graph_a.replay()                                      # forward, loss, backward, clipping
                ksd_value, last_entropy, loss_value = scalars.tolist()
                if math.isnan(loss_value) or math.isinf(loss_value):
                    print(f"Warning: NaN or Inf loss: {loss_value}. Skipping update.")
                else:
                    last_grad_norm = grad_norm_dev.clone()
                    graph_b.replay()                                  # optimizer step
                    if scheduler is not None:
                        schedule_owner.step()                         # stand-in optimizer: keeps the scheduler's call order
                        scheduler.step()
                        lr_device.fill_(scheduler.get_last_lr()[0])
                self.born_machine._pair_cache = None                  # replays change the parameters in place
            else:
                optimizer.zero_grad()
                q = self.born_machine.get_probabilities(x_condition=x_condition).squeeze()
                if q.shape[0] != self.num_possible_latent_states:
                    raise ValueError(f"Probabilities shape mismatch: {q.shape}")

                K = self.ksd_squared(q)                                   # ksd_vi.py:110-131
                ksd_loss = torch.sqrt(K.clamp(min=1e-12))                 # :133-134
                entropy = self.born_machine.entropy(x_condition=x_condition)
                loss = ksd_loss - entropy_weight * entropy.to(ksd_loss.dtype)
                # one device->host read for the three scalars the reference reads with three .item() calls
                ksd_value, last_entropy, loss_value = torch.stack(
                    [ksd_loss.detach().double(), entropy.detach().double(), loss.detach().double()]).tolist()

                if math.isnan(loss_value) or math.isinf(loss_value):
                    print(f"Warning: NaN or Inf loss: {loss_value}. Skipping update.")
                else:
                    loss.backward()
                    last_grad_norm = nn_utils.clip_grad_norm_(params, gradient_clip_norm).detach()   # read after the loop
                    optimizer.step()
                    if scheduler is not None:
                        scheduler.step()

            history['loss_ksd'].append(ksd_value)
            history['grad_norm'].append(last_grad_norm)   # after a skipped step the previous value is re-logged (:156)
            history['entropy'].append(last_entropy)

            if p_true is not None:
                tvd = current_tvd()
                history['tvd'].append(tvd)
                if tvd < best_tvd:
                    best_tvd, best_epoch, epochs_without_improvement = tvd, epoch, 0
                    with torch.no_grad():
                        best_probs = self.born_machine.get_probabilities(x_condition=x_condition).squeeze().clone()
                    if verbose and tvd < 0.05:
                        print(f"  -> New best TVD: {tvd:.6f} at epoch {epoch+1}")
                else:
                    epochs_without_improvement += 1
                if epochs_without_improvement > patience and epoch > 300:
                    if verbose:
                        print(f"\nEarly stopping at epoch {epoch+1} (no improvement for {patience} epochs)")
                    break
            else:
                history['tvd'].append(np.nan)

            if verbose and (epoch % max(1, num_epochs // 20) == 0 or epoch == num_epochs - 1):
                msg = f"Epoch {epoch+1}/{num_epochs} | KSD: {ksd_value:.6f}"
                if scheduler is not None:
                    msg += f" | LR: {scheduler.get_last_lr()[0]:.6f}"
                msg += f" | Entropy: {last_entropy:.4f}"
                if p_true is not None and not np.isnan(history['tvd'][-1]):
                    msg += f" | TVD: {history['tvd'][-1]:.6f}"
                print(msg)

        if graphs is not None:
            for p in params:                    # gradients live in the graphs' memory pool: hand out copies
                if p.grad is not None:
                    p.grad = p.grad.detach().clone()
            graphs = None

        # gradient norms were kept on the device during the loop: one read for all epochs
        pending = [i for i, g in enumerate(history['grad_norm']) if torch.is_tensor(g)]
        if pending:
            values = torch.stack([history['grad_norm'][i].double().reshape(()) for i in pending]).tolist()
            for i, v in zip(pending, values):
                history['grad_norm'][i] = v

        if best_probs is not None:
            if verbose:
                print(f"\nRestoring best probabilities (TVD: {best_tvd:.6f} from epoch {best_epoch+1})")
            self.born_machine.set_fixed_probs(best_probs)
            final_tvd = current_tvd()
            if abs(final_tvd - best_tvd) > 1e-6:
                print(f"WARNING: Still have restoration issue! Expected TVD: {best_tvd:.6f}, Got: {final_tvd:.6f}")
            elif verbose:
                print(f"Successfully restored best probabilities! Final TVD: {final_tvd:.6f}")
        return history
