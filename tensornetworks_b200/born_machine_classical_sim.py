"""Classical Born machine q(z|x) = softmax(logits) over the 2^n outcomes.

Drop-in for the reference's `born_machine_classical_sim.ClassicalBornMachine`
(reference born_machine_classical_sim.py:7-181): same constructor, parameter names
(`params` / `param_generator_net`), fixed-probability mode, entropy, sampling and
dictionary view.  It stays plain PyTorch (it runs on the GPU as ordinary torch ops); the
only change is that `all_outcome_tuples` is a lazy sequence, so n = 24 does not build
16.7 M tuples, and the sample/log-prob paths pack bits instead of searching that list.
"""
import torch
import torch.nn as nn

from .utils import generate_all_binary_outcomes


class ClassicalBornMachine(nn.Module):
    def __init__(self, num_latent_vars, use_logits=True, conditioning_dim=0,
                 init_method='small_random', hidden_dims=None, use_layer_norm=False):
        super().__init__()
        self.num_latent_vars = num_latent_vars
        self.num_outcomes = 2 ** num_latent_vars
        self.use_logits = use_logits
        self.conditioning_dim = conditioning_dim
        self.use_layer_norm = use_layer_norm
        self._fixed_probs = None
        self._use_fixed_probs = False
        self._pair_cache = None

        if conditioning_dim > 0:
            # MLP x -> logits (reference :24-48): Linear/[LayerNorm]/ReLU/Dropout(0.1) blocks,
            # Xavier weights, zero biases.
            if hidden_dims is None:
                hidden_dims = [max(conditioning_dim * 4, 64), max(conditioning_dim * 2, 32)]
            blocks, width = [], conditioning_dim
            for h in hidden_dims:
                blocks.append(nn.Linear(width, h))
                if use_layer_norm:
                    blocks.append(nn.LayerNorm(h))
                blocks += [nn.ReLU(), nn.Dropout(0.1)]
                width = h
            blocks.append(nn.Linear(width, self.num_outcomes))
            self.param_generator_net = nn.Sequential(*blocks)
            for module in self.param_generator_net.modules():
                if isinstance(module, nn.Linear):
                    nn.init.xavier_uniform_(module.weight)
                    nn.init.zeros_(module.bias)
        else:
            # free logits (reference :49-60)
            if init_method == 'zero':
                start = torch.zeros(self.num_outcomes)
            elif init_method == 'small_random':
                start = 0.1 * torch.randn(self.num_outcomes)
            elif init_method == 'uniform':
                start = torch.log(torch.ones(self.num_outcomes) / self.num_outcomes) + 0.01 * torch.randn(self.num_outcomes)
            else:
                start = torch.randn(self.num_outcomes)
            self.params = nn.Parameter(start)

        self.all_outcome_tuples = generate_all_binary_outcomes(num_latent_vars)

    # ---- fixed-probability mode (reference :64-72)
    def set_fixed_probs(self, prob_tensor):
        self._fixed_probs = prob_tensor.detach().clone()
        self._use_fixed_probs = True

    def clear_fixed_probs(self):
        self._fixed_probs = None
        self._use_fixed_probs = False

    def get_probabilities(self, x_condition=None):
        """[1, 2^n] (or [B, 2^n]) probabilities in enumeration order (reference :74-99)."""
        if self._use_fixed_probs and self._fixed_probs is not None:
            return self._fixed_probs.unsqueeze(0) if self._fixed_probs.ndim == 1 else self._fixed_probs
        if self.conditioning_dim > 0:
            if x_condition is None:
                raise ValueError("x_condition must be provided for conditional Born machine.")
            raw = self.param_generator_net(x_condition.unsqueeze(0) if x_condition.ndim == 1 else x_condition)
        else:
            if x_condition is not None:
                raise ValueError("x_condition provided but conditioning_dim is 0.")
            raw = self.params.unsqueeze(0)
        if self.use_logits:
            if self._streaming(raw):
                return self._probs_and_entropy("q_taken")[0].unsqueeze(0)
            shifted = raw - raw.max(dim=-1, keepdim=True)[0]
            if raw.is_cuda and raw.shape[0] == 1 and raw.shape[-1] >= (1 << 16):
                e = torch.exp(shifted)              # conditional net, one long row: same function from
                return e / e.sum(dim=-1, keepdim=True)   # grid-wide ops (torch.softmax: one block per row)
            return torch.softmax(shifted, dim=-1)
        mag = raw.abs()
        return mag / mag.sum(dim=-1, keepdim=True)

    def __getstate__(self):
        state = self.__dict__.copy()
        state["_pair_cache"] = None          # cached (q, H) carry an autograd graph: never copied or pickled
        return state

    def _streaming(self, raw):
        """One unconditional row of >= 2^16 fp32 logits on the GPU goes through the streaming kernels
        (ops.born_softmax_entropy); torch.softmax would run that single row on one SM."""
        return (self.conditioning_dim == 0 and raw.is_cuda and raw.dtype == torch.float32
                and raw.shape[-1] >= (1 << 16))

    def _probs_and_entropy(self, which):
        """(q, H) of the current parameters.  get_probabilities() and entropy() of the same training step
        share one forward pass: the pair is cached per parameter version and grad mode, and each half is
        handed out once (a second request recomputes, so no autograd graph is ever used twice)."""
        from . import ops
        key = (self.params._version, self.params.data_ptr(), torch.is_grad_enabled())
        c = self._pair_cache
        if c is None or c["key"] != key or c[which]:
            q, H = ops.born_softmax_entropy(self.params)
            c = self._pair_cache = {"key": key, "q": q, "H": H, "q_taken": False, "H_taken": False}
            if q.requires_grad:
                # q and H are two outputs of ONE autograd node: once a backward pass has run through either of them
                # the node's buffers are freed, so the other half must not be handed out any more (separate backward
                # calls for the KSD and the entropy term, gradient accumulation) - drop the pair when that happens
                def _invalidate(_grad, pair=c):
                    if self._pair_cache is pair:
                        self._pair_cache = None
                q.register_hook(_invalidate)
                H.register_hook(_invalidate)
        c[which] = True
        return c["q"], c["H"]

    def _bits(self, indices, device):
        n = self.num_latent_vars
        shifts = torch.arange(n - 1, -1, -1, device=indices.device)
        return ((indices.unsqueeze(-1) >> shifts) & 1).to(torch.float32).to(device)

    def sample(self, num_samples=1, x_condition=None):
        """float32 samples [num_samples, n] (or [B, num_samples, n]) (reference :101-124)."""
        probs = self.get_probabilities(x_condition) + 1e-10
        probs = probs / probs.sum(dim=-1, keepdim=True)
        batched = probs.shape[0] > 1 or (x_condition is not None and x_condition.ndim > 1)
        if batched:
            return torch.stack([self._bits(torch.multinomial(p, num_samples, replacement=True), probs.device)
                                for p in probs])
        return self._bits(torch.multinomial(probs.squeeze(0), num_samples, replacement=True), probs.device)

    def get_prob_dict(self, x_condition=None):
        """{outcome tuple: probability} for a single distribution (reference :126-138)."""
        probs = self.get_probabilities(x_condition)
        if probs.shape[0] > 1 and not (probs.ndim == 1 and self.conditioning_dim == 0):
            raise ValueError("get_prob_dict is for a single distribution.")
        flat = probs.squeeze().detach().cpu().numpy()
        return {outcome: flat[i] for i, outcome in enumerate(self.all_outcome_tuples)}

    def get_log_q_z_x(self, z_samples, x_condition=None):
        """log q(z|x) of given samples (reference :140-175)."""
        if self.conditioning_dim > 0 and x_condition is None:
            raise ValueError("x_condition must be provided for conditional Born machine.")
        if self.conditioning_dim == 0 and x_condition is not None:
            raise ValueError("x_condition provided but Born machine is not conditional.")
        log_probs = torch.log(self.get_probabilities(x_condition).clamp(min=1e-10))
        bits = z_samples.long()
        if not bool(((bits == 0) | (bits == 1)).all()):
            bad = tuple(z_samples[0].cpu().long().tolist())
            raise ValueError(f"Sample {bad} is not a valid outcome.")
        n = self.num_latent_vars
        weights = (1 << torch.arange(n - 1, -1, -1, device=bits.device)).long()
        idx = (bits * weights).sum(dim=-1).to(log_probs.device)
        if log_probs.shape[0] == 1:
            return log_probs[0, idx]
        if log_probs.shape[0] == z_samples.shape[0]:
            return log_probs[torch.arange(idx.shape[0], device=idx.device), idx]
        raise ValueError(f"Batch size mismatch: x_condition ({log_probs.shape[0]}) vs z_samples ({z_samples.shape[0]}).")

    def entropy(self, x_condition=None):
        """H(q) = -sum q log clamp(q, 1e-10) (reference :177-181)."""
        if (self.use_logits and self.conditioning_dim == 0 and x_condition is None
                and not (self._use_fixed_probs and self._fixed_probs is not None) and self._streaming(self.params)):
            return self._probs_and_entropy("H_taken")[1]
        probs = self.get_probabilities(x_condition).squeeze()
        return -(probs * torch.log(probs.clamp(min=1e-10))).sum()
