"""Generate tests/golden/ref_train_golden.npz by RUNNING THE REFERENCE's trainer (/root/reference).

Run in the build container only:  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_train_golden.py

BASELINE.json config #1 (run_sprinkler_ksd.py): Sprinkler BN, latent C,S,R | W=1, classical Born machine,
exact KSD over the 8 states.  Two seeded runs of reference ksd_vi.KSDVariationalInference.train are recorded
epoch by epoch (loss_ksd, tvd, grad_norm, entropy) together with the final Born-machine probabilities:
  uncond  the unconditional Born machine (a logits vector; deterministic)
  cond    the conditional MLP Born machine of run_sprinkler_ksd.py:32-38 with that script's hyper-parameters
          (Dropout(0.1) is active, so the trajectory also pins the order and number of forward passes)
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
mpl = types.ModuleType("matplotlib")
plt = types.ModuleType("matplotlib.pyplot")
plt.show = lambda *a, **k: None
mpl.pyplot = plt
sys.modules["matplotlib"] = mpl
sys.modules["matplotlib.pyplot"] = plt
sys.dont_write_bytecode = True
sys.path.insert(0, REF)

from bayesian_network import get_sprinkler_network     # noqa: E402
from ksd_vi import KSDVariationalInference              # noqa: E402

EPOCHS = 25
SEED = 1234
RUNS = {
    "uncond": dict(config={"use_logits": True, "conditioning_dim": 0},
                   train=dict(lr_born_machine=0.05, gradient_clip_norm=10.0, entropy_weight=0.01)),
    "cond": dict(config={"use_logits": True, "conditioning_dim": 1, "init_method": "uniform", "hidden_dims": None,
                         "use_layer_norm": False},
                 train=dict(lr_born_machine=0.003, gradient_clip_norm=5.0, entropy_weight=0.001, patience=200)),
}


def main():
    out = {"epochs": np.int64(EPOCHS), "seed": np.int64(SEED)}
    bn = get_sprinkler_network(random_cpts=False)
    latent, observed, x = ["C", "S", "R"], ["W"], {"W": 1}
    posterior, _ = bn.get_true_posterior(latent, x)
    for name, run in RUNS.items():
        torch.manual_seed(SEED)
        vi = KSDVariationalInference(bn, latent, observed, run["config"], base_kernel_length_scale=1.0, device="cpu")
        hist = vi.train(x, num_epochs=EPOCHS, verbose=False, true_posterior_for_tvd=posterior, **run["train"])
        for key in ("loss_ksd", "tvd", "grad_norm", "entropy"):
            out[f"{name}_{key}"] = np.asarray([float(v) for v in hist[key]], dtype=np.float64)
        vi.born_machine.eval()
        with torch.no_grad():
            xc = torch.tensor([1.0]) if run["config"]["conditioning_dim"] else None
            out[f"{name}_final_q"] = vi.born_machine.get_probabilities(x_condition=xc).squeeze().double().numpy()
        print(name, out[f"{name}_loss_ksd"][[0, -1]], out[f"{name}_tvd"][[0, -1]])
    np.savez(os.path.join(HERE, "ref_train_golden.npz"), **out)


if __name__ == "__main__":
    main()
