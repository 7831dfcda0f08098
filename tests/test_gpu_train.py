"""Drop-in parity of the trainer (BASELINE.json config #1): our KSDVariationalInference.train, with the hot
loop on the GPU, must retrace the REFERENCE's own training trajectory epoch by epoch
(tests/golden/ref_train_golden.npz, produced by tests/golden/make_train_golden.py from /root/reference)."""
import os

import numpy as np
import pytest
import torch

from tensornetworks_b200.bayesian_network import get_sprinkler_network
from tensornetworks_b200.ksd_vi import KSDVariationalInference

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

RUNS = {
    "uncond": dict(config={"use_logits": True, "conditioning_dim": 0},
                   train=dict(lr_born_machine=0.05, gradient_clip_norm=10.0, entropy_weight=0.01)),
    "cond": dict(config={"use_logits": True, "conditioning_dim": 1, "init_method": "uniform", "hidden_dims": None,
                         "use_layer_norm": False},
                 train=dict(lr_born_machine=0.003, gradient_clip_norm=5.0, entropy_weight=0.001, patience=200)),
}


@pytest.fixture(scope="module")
def train_golden():
    return dict(np.load(os.path.join(REPO, "tests", "golden", "ref_train_golden.npz")))


@pytest.mark.parametrize("precision,rtol", [("fp64", 2e-5), ("fp32", 2e-3), ("tc32", 2e-3)])
@pytest.mark.parametrize("name", ["uncond", "cond"])
def test_training_trajectory_matches_reference(train_golden, name, precision, rtol):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    g = train_golden
    run = RUNS[name]
    bn = get_sprinkler_network(random_cpts=False)
    latent, observed, x = ["C", "S", "R"], ["W"], {"W": 1}
    posterior, _ = bn.get_true_posterior(latent, x)
    torch.manual_seed(int(g["seed"]))
    vi = KSDVariationalInference(bn, latent, observed, run["config"], base_kernel_length_scale=1.0, device="cpu",
                                 precision=precision, compute_device="cuda")
    hist = vi.train(x, num_epochs=int(g["epochs"]), verbose=False, true_posterior_for_tvd=posterior, **run["train"])
    assert sorted(hist) == ["entropy", "grad_norm", "loss_ksd", "tvd"]
    for key in ("loss_ksd", "tvd", "grad_norm", "entropy"):
        ours = np.asarray([float(v) for v in hist[key]], dtype=np.float64)
        ref = g[f"{name}_{key}"]
        assert ours.shape == ref.shape, key
        scale = np.max(np.abs(ref))
        assert np.max(np.abs(ours - ref)) <= rtol * scale, (key, np.max(np.abs(ours - ref)) / scale)
    vi.born_machine.eval()
    with torch.no_grad():
        xc = torch.tensor([1.0]) if run["config"]["conditioning_dim"] else None
        q = vi.born_machine.get_probabilities(x_condition=xc).squeeze().double().numpy()
    assert np.max(np.abs(q - g[f"{name}_final_q"])) <= rtol
