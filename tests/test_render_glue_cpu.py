"""`gaussian_renderer.render`, `SinusoidalEncoder` and `ResidualMeshSimulator` against the reference's OWN functions
(tests/make_golden_render.py lifts them from gaussian_renderer/__init__.py:39-206 and meshnet/meshnet_network.py:256-380;
tests/golden/render_glue.npz).  Both sides call the same CPU stand-in rasterizer (tests/fake_scene.StandInRasterizer), so
every difference is the glue's: argument order, settings fields, time-index rounding, projections, the gradient paths."""
import os

import numpy as np
import pytest
import torch

import fake_scene
from helpers import same_as_golden
from make_golden_render import CASES, run_case

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "render_glue.npz"))


@pytest.fixture()
def mirror_render(monkeypatch):
    import gaussian_renderer

    monkeypatch.setattr(gaussian_renderer, "GaussianRasterizer", fake_scene.StandInRasterizer)
    return gaussian_renderer.render


def test_render_results_have_the_reference_fields():
    from gaussian_renderer import RenderResults

    assert list(RenderResults._fields) == G["fields"].tolist()
    assert str(G["ref_commit"]).startswith("KTH-RPL/cloth-splatting ")


@pytest.mark.parametrize("name", list(CASES))
def test_render_matches_the_reference_function(mirror_render, name):
    out = run_case(mirror_render, None, name, reference=False)
    keys = {k[len(name) + 1:] for k in G.files if k.startswith(name + "/")}
    assert keys == set(out), keys ^ set(out)
    for k, v in out.items():
        # the rotation oracle's SVD and torch's CPU exp / sin agree to a few ulp between hosts; gradients pass through
        # sums over pixels and Gaussians
        tol = dict(rtol=2e-4, atol=2e-6) if k.startswith("grad/") else dict(rtol=1e-5, atol=1e-6)
        assert same_as_golden(v, G[f"{name}/{k}"], **tol), (name, k, np.abs(np.asarray(v, np.float64) - G[f"{name}/{k}"]).max())


def test_render_raises_past_the_last_mesh_prediction_like_the_reference(mirror_render):
    m, sim, cams, bg = fake_scene.render_scene()
    pipe = type("Pipe", (), {"compute_cov3D_python": False})()
    assert str(G["past_end"]) == "ValueError"
    with pytest.raises(ValueError):
        mirror_render(cams[5], m, sim, pipe, bg)


def test_sinusoidal_encoder_matches_the_reference_class():
    from meshnet.meshnet_network import SinusoidalEncoder

    for nf, ident, scale in ((6, True, 1.0), (4, False, 0.5), (0, True, 1.0)):
        e = SinusoidalEncoder(input_dim=1, num_freqs=nf, scale=scale, use_identity=ident)
        tag = f"enc/{nf}_{int(ident)}"
        assert e.output_dim == int(G[f"{tag}/output_dim"])
        assert same_as_golden(e(torch.from_numpy(G[f"{tag}/x"])).numpy(), G[f"{tag}/y"], rtol=1e-6, atol=1e-6), tag


@pytest.mark.parametrize("n_times", [1, 2, 5])
def test_residual_simulator_matches_the_reference_class(n_times):
    """Time-index rounding (round-half-even of t / dt in fp32), the n_times == 1 case, the out-of-bounds error; through the
    reference-compatible `forward(time_vector)` (device-side check) and the host-checked path render() uses."""
    from meshnet.meshnet_network import ResidualMeshSimulator

    s = ResidualMeshSimulator(torch.from_numpy(G[f"sim/{n_times}/preds"]), device="cpu")
    assert s.time_delta == float(G[f"sim/{n_times}/time_delta"])
    torch.manual_seed(n_times)
    with torch.no_grad():
        for p in s.parameters():
            p.copy_(torch.randn(p.shape) * 0.1)
    assert np.allclose([float(v.double().sum()) for v in s.state_dict().values()], G[f"sim/{n_times}/state_sum"], rtol=1e-9)
    assert list(s.state_dict()) == ["encoder.freqs", "input.weight", "input.bias", "hidden.weight", "hidden.bias", "output.weight", "output.bias"]
    for t in G[f"sim/{n_times}/times"].tolist():
        want = G[f"sim/{n_times}/t{t}"]
        tv = torch.tensor(t).repeat(7, 1)
        if want.dtype.kind == "U":
            assert str(want) == "ValueError"
            with pytest.raises(ValueError):
                s(tv)
            with pytest.raises(ValueError):
                s.check_time(t)
        else:
            assert same_as_golden(s(tv).detach().numpy(), want, rtol=1e-5, atol=1e-6), t
            s.check_time(t)
            assert same_as_golden(s(tv, checked=True).detach().numpy(), want, rtol=1e-5, atol=1e-6), t
