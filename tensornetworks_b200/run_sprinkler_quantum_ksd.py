"""BASELINE.json config #2: Sprinkler KSD with q from a quantum Born machine, GPU KSD drop-in.
Headless counterpart of the reference's run_sprinkler_quantum_ksd.py (hyper-parameters of
:35-48,77-87: 4 hardware-efficient layers, lr 0.005, 1000 epochs, clip 10).  q comes from
PennyLane default.qubit when it is installed, else from the torch state-vector simulator."""
import argparse

import numpy as np
import torch

from .bayesian_network import get_sprinkler_network
from .ksd_vi_quantum import KSDVariationalInference
from .utils import calculate_tvd


def run_sprinkler_quantum_ksd_experiment(num_epochs=1000, precision="fp64", seed=0, verbose=True):
    torch.manual_seed(seed)
    np.random.seed(seed)
    bn = get_sprinkler_network(random_cpts=False)
    latent, observed, evidence = ['C', 'S', 'R'], ['W'], {'W': 1}
    posterior, _ = bn.get_true_posterior(latent, evidence)
    model = KSDVariationalInference(
        bayesian_network=bn, latent_vars_names=latent, observed_vars_names=observed,
        qbm_num_latent_vars=3, qbm_ansatz_layers=4, qbm_conditioning_dim=0,
        qbm_pennylane_device_name="default.qubit", qbm_ansatz_type="hardware_efficient",
        qbm_init_method="small_random", base_kernel_length_scale=1.0, pytorch_device='cpu', precision=precision)
    if verbose:
        print(f"quantum Born machine backend: {model.born_machine.backend}, {model.born_machine.num_ansatz_params} parameters")
    history = model.train(x_observation_dict=evidence, num_epochs=num_epochs, lr_born_machine=0.005,
                          verbose=verbose, true_posterior_for_tvd=posterior, use_lr_scheduler=True,
                          gradient_clip_norm=10.0, optimizer_type="adam")
    final_tvd = calculate_tvd(posterior, model.born_machine.get_prob_dict())
    if verbose:
        print(f"Final TVD: {final_tvd:.6f}  (best during training: {min(history['tvd']):.6f})")
    return history, final_tvd


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=1000)
    ap.add_argument("--precision", default="fp64")
    args = ap.parse_args()
    run_sprinkler_quantum_ksd_experiment(num_epochs=args.epochs, precision=args.precision)
