"""Host-side pieces of the refinement step added around the kernels (scene_reconstruction/train_utils.py, gaussian_mesh.py,
meshnet_network.py): they change how often work is done, never its result."""
import numpy as np
import torch

import conftest  # noqa: F401
import fake_scene


def test_forward_times_equals_the_per_view_forward():
    """ResidualMeshSimulator.forward_times(times)[v] == forward(time_vector filled with times[v], checked=True)
    (meshnet/meshnet_network.py:361-373 per view; batched over the views of a step)."""
    _, sim, _, _ = fake_scene.render_scene()
    times = torch.tensor([0.0, 0.26, 0.5, 0.74, 1.0])
    batched = sim.forward_times(times)
    assert batched.shape == (5, 9, 3)
    for v, t in enumerate(times.tolist()):
        single = sim(torch.full((9, 1), t), checked=True)
        assert torch.allclose(batched[v], single, rtol=1e-5, atol=1e-6), t
    # gradients reach the same parameters
    batched.square().sum().backward()
    g_b = {n: p.grad.clone() for n, p in sim.named_parameters()}
    sim.zero_grad()
    sum(sim(torch.full((9, 1), t), checked=True).square().sum() for t in times.tolist()).backward()
    for n, p in sim.named_parameters():
        assert torch.allclose(g_b[n], p.grad, rtol=1e-4, atol=1e-6), n


def test_shared_activations_are_evaluated_once_and_give_the_same_gradients():
    m, _, _, _ = fake_scene.render_scene()
    calls = {"n": 0}
    real_sigmoid = torch.sigmoid

    def counting_sigmoid(x):
        calls["n"] += 1
        return real_sigmoid(x)

    torch.sigmoid = counting_sigmoid
    try:
        with m.shared_activations():
            a, b = m.get_opacity, m.get_opacity
            s1, f1 = m.get_scaling, m.get_features
            assert a is b and m.get_scaling is s1 and m.get_features is f1
        assert calls["n"] == 1
        c = m.get_opacity  # outside the scope: evaluated per access, like the reference's property
        assert c is not a and calls["n"] == 2 and torch.equal(a, c)
    finally:
        torch.sigmoid = real_sigmoid
    # two "views" sharing the node accumulate the same gradient as two separate evaluations
    with m.shared_activations():
        (m.get_opacity.sum() * 2 + m.get_opacity.square().sum()).backward()
    g_shared = m._opacity.grad.clone()
    m._opacity.grad = None
    (m.get_opacity.sum() * 2 + m.get_opacity.square().sum()).backward()
    assert torch.allclose(g_shared, m._opacity.grad)


def test_resident_planes_cache_by_object_and_version_with_lru_eviction():
    from scene_reconstruction.train_utils import ResidentPlanes

    rp = ResidentPlanes()
    rp.budget = 3 * 4 * 100  # room for three 100-float planes
    planes = [torch.full((100,), float(i)) for i in range(4)]
    d0 = rp.fetch(planes[0], "cpu")
    assert rp.misses == 1 and rp.fetch(planes[0], "cpu") is d0 and rp.hits == 1
    planes[0].add_(1)                                  # edited in place: uploaded again
    d0b = rp.fetch(planes[0], "cpu")
    assert rp.misses == 2 and torch.equal(d0b, planes[0])
    for p in planes[1:]:
        rp.fetch(p, "cpu")
    assert rp.used <= rp.budget and len(rp.entries) == 3 and id(planes[0]) not in rp.entries   # the oldest was evicted
    into = torch.empty(100)
    assert rp.fetch(planes[3], "cpu", into=into) is into and torch.equal(into, planes[3])
    clone_same_content = planes[3].clone()             # another object: its own entry
    rp.fetch(clone_same_content, "cpu")
    assert id(clone_same_content) in rp.entries
    rp.budget = 0                                      # switched off: plain copies, nothing kept
    rp.clear()
    assert torch.equal(rp.fetch(planes[1], "cpu"), planes[1]) and not rp.entries


def test_precomputed_transform_slices_give_the_same_render(monkeypatch):
    """render(..., precomputed=(vertices, xyz, rot)) == render() computing them itself (CPU stand-in rasterizer)."""
    import gaussian_renderer

    monkeypatch.setattr(gaussian_renderer, "GaussianRasterizer", fake_scene.StandInRasterizer)
    m, sim, cams, bg = fake_scene.render_scene()
    pipe = type("Pipe", (), {"compute_cov3D_python": False})()
    cam = cams[2]
    a = gaussian_renderer.render(cam, m, sim, pipe, bg)
    verts = sim.forward_times(torch.tensor([cam.time]))[0]
    b = gaussian_renderer.render(cam, m, sim, pipe, bg, precomputed=(verts, m.get_xyz(verts), m.get_rotation(verts)))
    for f in ("render", "depth", "radii", "means3D_deform", "vertice_deform", "rotations", "projections"):
        assert np.allclose(getattr(a, f).detach().numpy(), getattr(b, f).detach().numpy(), rtol=1e-5, atol=1e-6), f
