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


def test_quantum_trainer_runs_config2():
    """BASELINE.json config #2 (run_sprinkler_quantum_ksd.py): Sprinkler KSD with q from a quantum Born
    machine; the KSD part is the GPU drop-in.  PennyLane is not installed in this image, so the q supplier
    is the torch state-vector simulator of the same ansatz; the first epoch's loss must equal the oracle's
    sqrt(KSD^2) for the simulator's q, the history must have the reference's keys and the loss must fall."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from oracle import ksd_oracle as orc
    from tensornetworks_b200.ksd_vi_quantum import KSDVariationalInference as QuantumKSD
    bn = get_sprinkler_network(random_cpts=False)
    latent, observed, x = ["C", "S", "R"], ["W"], {"W": 1}
    posterior, _ = bn.get_true_posterior(latent, x)
    torch.manual_seed(7)
    vi = QuantumKSD(bn, latent, observed, qbm_num_latent_vars=3, qbm_ansatz_layers=3, pytorch_device="cpu",
                    precision="fp64", compute_device="cuda")
    q0 = vi.born_machine.get_probabilities().detach().double().numpy()
    P = orc.joint_table(bn, x, latent)
    K0, _ = orc.ksd_dense(q0, orc.score_table(P, 3), 3)
    hist = vi.train(x, num_epochs=40, lr_born_machine=0.05, verbose=False, true_posterior_for_tvd=posterior)
    assert sorted(hist) == ["grad_norm", "loss_ksd", "tvd"]
    assert len(hist["loss_ksd"]) == 40 and np.all(np.isfinite(hist["loss_ksd"]))
    assert abs(hist["loss_ksd"][0] - np.sqrt(K0)) <= 1e-9 * np.sqrt(K0)
    assert min(hist["loss_ksd"][-5:]) < 0.8 * hist["loss_ksd"][0]
    assert hist["tvd"][-1] < hist["tvd"][0]
