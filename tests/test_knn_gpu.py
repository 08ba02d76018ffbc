"""distCUDA2 parity: bit-exact against the CPU oracle (exhaustive 3-NN) and the reference CUDA."""
import numpy as np
import pytest
import torch

import conftest  # noqa: F401
import csb_synth as S

pytestmark = pytest.mark.gpu


def _dist(points_np):
    from simple_knn._C import distCUDA2

    return distCUDA2(torch.from_numpy(points_np).cuda()).cpu().numpy()


@pytest.mark.parametrize("P", [1, 2, 3, 4, 5, 33, 1000, 10108])
def test_random_cloud_bit_exact_vs_oracle(P):
    import oracle

    rng = np.random.default_rng(P)
    pts = rng.standard_normal((P, 3)).astype(np.float32)
    got, want = _dist(pts), oracle.dist2(pts)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_cloth_gaussians_bit_exact_vs_oracle():
    import oracle

    g = S.mesh_gaussians(20, 14, 1)  # planar cloud: most grid cells are empty
    got, want = _dist(g["means3D"]), oracle.dist2(g["means3D"])
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_degenerate_clouds():
    import oracle

    same = np.tile(np.array([[0.3, -1.0, 2.0]], np.float32), (50, 1))
    assert np.array_equal(_dist(same), np.zeros(50, np.float32))
    line = np.zeros((300, 3), np.float32)
    line[:, 0] = np.linspace(0, 1, 300)
    assert np.array_equal(_dist(line).view(np.uint32), oracle.dist2(line).view(np.uint32))
    # a tight cluster plus far outliers: ring search has to grow
    rng = np.random.default_rng(0)
    pts = np.concatenate([rng.standard_normal((2000, 3)) * 1e-3, rng.standard_normal((5, 3)) * 50]).astype(np.float32)
    assert np.array_equal(_dist(pts).view(np.uint32), oracle.dist2(pts).view(np.uint32))
    # duplicates count as neighbours at distance 0
    dup = np.concatenate([pts[:100], pts[:100]]).astype(np.float32)
    assert np.array_equal(_dist(dup).view(np.uint32), oracle.dist2(dup).view(np.uint32))


def test_lattice_known_answer():
    h = 0.125
    ax = np.arange(20) * h
    pts = np.stack(np.meshgrid(ax, ax, ax, indexing="ij"), -1).reshape(-1, 3).astype(np.float32)
    assert np.allclose(_dist(pts), h * h, rtol=1e-6)


def test_vs_reference_cuda():
    from oracle import ref_cuda

    if not ref_cuda.available():
        pytest.skip("oracle/_ref/libref_cuda.so not built")
    for P, seed in ((10108, 1), (200_000, 2)):
        rng = np.random.default_rng(seed)
        pts = torch.from_numpy(rng.random((P, 3), dtype=np.float32) * 2 - 1).cuda()
        from simple_knn._C import distCUDA2

        got = distCUDA2(pts)
        want = ref_cuda.dist_cuda2(pts)
        assert torch.equal(got.view(torch.int32), want.view(torch.int32))


def test_million_points_properties():
    rng = np.random.default_rng(9)
    pts = (rng.random((1_000_000, 3), dtype=np.float32) * 2 - 1).astype(np.float32)
    d = _dist(pts)
    assert np.isfinite(d).all() and d.min() >= 0
    # spot-check 64 points against brute force
    idx = rng.integers(0, pts.shape[0], 64)
    for i in idx:
        diff = pts - pts[i]
        d2 = np.sort((diff[:, 0] * diff[:, 0] + diff[:, 1] * diff[:, 1] + diff[:, 2] * diff[:, 2]))[1:4]
        assert abs(d[i] - d2.mean()) <= 1e-5 * max(d2.mean(), 1e-12) + 1e-12
