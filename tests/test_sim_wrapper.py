"""Host-side mirrors of the reference's simulator glue against golden outputs of the reference itself
(tests/golden/sim_7x7_seed5.npz, made by tests/make_golden_sim.py from meshnet/model_utils.py, cloth_network.py
and train_meshnet_sim.py:322-359).  CPU part: Normalizer and update_prediction (pure torch, no GNN call).
GPU part (-m gpu): ClothMeshSimulator.predict_acceleration / predict_velocity / the 3-step training loss and its
gradients through the fused GNN forward + backward, within bf16 tolerance of the reference's fp32 numbers."""
import os

import numpy as np
import pytest
import torch

import conftest

FX = os.path.join(conftest.ROOT, "tests", "golden", "sim_7x7_seed5.npz")


def _fx():
    return np.load(FX)


def test_normalizer_matches_reference():
    from meshnet.model_utils import Normalizer

    fx = _fx()
    nz = Normalizer(size=8, name="node_normalizer", device="cpu")
    g = torch.Generator().manual_seed(11)
    for i in range(3):
        last = nz(torch.randn((50, 8), generator=g) * (i + 1) + 0.5 * i, True)
    assert np.allclose(last.numpy(), fx["norm_last"], rtol=1e-6, atol=1e-6)
    assert np.allclose(nz._mean().numpy(), fx["norm_mean"], rtol=1e-6, atol=1e-7)
    assert np.allclose(nz._std_with_epsilon().numpy(), fx["norm_std"], rtol=1e-6, atol=1e-7)
    assert np.allclose(nz.inverse(last).numpy(), fx["norm_inverse"], rtol=1e-6, atol=1e-6)
    assert set(nz.get_variable()) == {"_max_accumulations", "_std_epsilon", "_acc_count", "_num_accumulations",
                                      "_acc_sum", "_acc_sum_squared", "name"}
    frozen = nz._acc_count.clone()
    nz(torch.ones(4, 8), False)
    assert torch.equal(frozen, nz._acc_count)


def test_update_prediction_matches_reference_bit_for_bit():
    from meshnet.rollout import compute_edge_features, update_prediction

    fx = _fx()
    t = {k[3:]: torch.from_numpy(fx[k]) for k in fx.files if k.startswith("in_")}
    v0 = t["velocity"].clone()
    v2, e2, p2 = update_prediction(t["noise"], t["velocity"], torch.from_numpy(fx["up_acc"]), t["positions"],
                                   t["edge_index"], t["actions"][:, 0], t["actions"][:, 1])
    assert torch.equal(t["velocity"], v0)  # functional: the caller's history is not modified
    assert v2.dtype == torch.float64 and p2.dtype == torch.float32 and e2.dtype == torch.float32
    assert np.array_equal(v2.numpy(), fx["up_velocity"])
    assert np.array_equal(p2.numpy(), fx["up_positions"])
    assert np.array_equal(e2.numpy(), fx["up_edge_features"])
    assert np.allclose(compute_edge_features(t["positions"], t["edge_index"]).numpy(), fx["in_edge_features"], atol=1e-7)


def _simulator():
    from meshnet import ClothMeshSimulator

    torch.manual_seed(5)
    return ClothMeshSimulator(3, 8, 4, 128, 15, 2, 128, 2, 2, normalize=True, device="cuda")


@pytest.mark.gpu
def test_simulator_predictions_and_multi_step_loss_vs_reference():
    from meshnet.rollout import multi_step_loss

    fx = _fx()
    t = {k[3:]: torch.from_numpy(fx[k]).cuda() for k in fx.files if k.startswith("in_")}
    sim = _simulator()
    sim.train()
    g = torch.Generator().manual_seed(11)
    for i in range(3):
        sim._node_normalizer((torch.randn((50, 8), generator=g) * (i + 1) + 0.5 * i).cuda(), True)
    pred, tgt = sim.predict_acceleration(t["velocity"], t["node_types"], t["edge_index"], t["edge_features"],
                                         t["target"][:, 0], velocity_noise=t["noise"])
    sc = float(np.abs(fx["pred_acc"]).max())
    assert float(np.abs(pred.detach().cpu().numpy() - fx["pred_acc"]).max()) < 5e-2 * sc
    assert np.allclose(tgt.cpu().numpy(), fx["target_acc"], rtol=1e-5, atol=1e-6)
    sim.eval()
    with torch.no_grad():
        pv = sim.predict_velocity(t["velocity"].float(), t["node_types"], t["edge_index"], t["edge_features"])
    assert float(np.abs(pv.cpu().numpy() - fx["pred_velocity"]).max()) < 5e-2 * float(np.abs(fx["pred_velocity"]).max())
    # three unrolled predictions, gradients through dL/d edge_features of the later ones
    loss = multi_step_loss(sim, t["velocity"], t["node_types"], t["edge_index"], t["edge_features"], t["target"],
                           t["actions"], t["positions"], 3, velocity_noise=t["noise"])
    loss.backward()
    assert abs(float(loss) - float(fx["loss"])) < 2e-2 * float(fx["loss"])
    sd = dict(sim.named_parameters())
    for k in fx.files:
        if k.startswith("grad:"):
            a, b = sd[k[5:]].grad.double().cpu().flatten(), torch.from_numpy(fx[k]).double().flatten()
            assert float((a - b).norm() / b.norm()) < 0.25, k  # bf16 forward: see test_gnn_backward_gpu.py
            assert float(a @ b / (a.norm() * b.norm())) > 0.97, k
    total = np.sqrt(sum(float(p.grad.double().pow(2).sum()) for p in sim.parameters()))
    assert abs(total - float(fx["grad_norm_total"])) < 0.1 * float(fx["grad_norm_total"])


@pytest.mark.gpu
def test_checkpoint_roundtrip(tmp_path):
    sim = _simulator()
    sim.train()
    sim._node_normalizer(torch.randn(30, 8).cuda(), True)
    sim.save(str(tmp_path / "model-12.pt"))
    sim.save(str(tmp_path / "model-3.pt"))
    other = _simulator()
    with torch.no_grad():
        for p in other.parameters():
            p.add_(1.0)
    other.load(str(tmp_path), "latest")
    for (ka, a), (kb, b) in zip(sim.state_dict().items(), other.state_dict().items()):
        assert ka == kb and torch.equal(a, b)
    assert torch.equal(other._node_normalizer._acc_sum, sim._node_normalizer._acc_sum)
