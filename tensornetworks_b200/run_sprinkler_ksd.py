"""BASELINE.json config #1: Sprinkler BN, latent C,S,R | W=1, classical Born machine, exact KSD.
Headless counterpart of the reference's run_sprinkler_ksd.py (hyper-parameters of :32-38,53-74);
the matplotlib figures of the reference are replaced by a printed summary."""
import argparse

import numpy as np
import torch

from .bayesian_network import get_sprinkler_network
from .ksd_vi import KSDVariationalInference
from .utils import calculate_tvd, plot_training_results


def run_sprinkler_ksd_experiment(num_epochs=2000, device=None, precision="fp64", seed=0, verbose=True):
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    torch.manual_seed(seed)
    np.random.seed(seed)
    bn = get_sprinkler_network(random_cpts=False)
    latent, observed, evidence = ['C', 'S', 'R'], ['W'], {'W': 1}
    posterior, p_evidence = bn.get_true_posterior(latent, evidence)
    if verbose:
        print(f"P(W=1) = {p_evidence:.6f}")
    model = KSDVariationalInference(
        bayesian_network=bn, latent_vars_names=latent, observed_vars_names=observed,
        born_machine_config={'use_logits': True, 'conditioning_dim': len(observed), 'init_method': 'uniform'},
        base_kernel_length_scale=1.0, device=device, precision=precision)
    history = model.train(
        x_observation_dict=evidence, num_epochs=num_epochs, lr_born_machine=0.003, verbose=verbose,
        true_posterior_for_tvd=posterior, use_lr_scheduler=True, gradient_clip_norm=5.0,
        optimizer_type="adam", entropy_weight=0.001, patience=200)
    x_cond = torch.tensor([1.0], device=device)
    final = model.born_machine.get_prob_dict(x_condition=x_cond)
    final_tvd = calculate_tvd(posterior, final)
    if verbose:
        for z in model.all_latent_states_tuples:
            print(f"  z={z}  true={posterior[z]:.4f}  q={float(final[z]):.4f}")
        print(f"Final TVD: {final_tvd:.6f}  (best during training: {min(history['tvd']):.6f})")
        plot_training_results({k: v for k, v in history.items() if len(v)}, "KSD VI on Sprinkler")
    return history, final_tvd


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=2000)
    ap.add_argument("--precision", default="fp64")
    args = ap.parse_args()
    run_sprinkler_ksd_experiment(num_epochs=args.epochs, precision=args.precision)
