"""The visit-culling rule of the compositing kernels (render_lanes.cuh::quad_mask: bounding box of the alpha >= 1/255 ellipse,
then the exact ellipse-vs-quad test), restated in fp32 numpy by tools/cull_census.py, on the CPU oracle's tile lists: no
8x8 quad that holds a blending pixel may be dropped, and the exact test must actually remove visits.  (The CUDA function
itself is checked on the GPU by tests/test_render_variants_gpu.py: bit-identical images with the test on and off.)"""
import importlib.util
import os

import numpy as np

import conftest
import csb_synth as S
import helpers as Hh

spec = importlib.util.spec_from_file_location("cull_census", os.path.join(conftest.ROOT, "tools", "cull_census.py"))
CC = importlib.util.module_from_spec(spec)
spec.loader.exec_module(CC)


def test_no_blending_quad_is_dropped_small_scene(capsys):
    g, cam = Hh.small_scene(seed=11)
    assert CC.census("small", g, cam) == 0


def test_no_blending_quad_is_dropped_needles(capsys):
    # axis ratios up to 1 : 200 at every orientation, opacities from below 1/255 to 1: the conics whose terms cancel most
    g = S.random_gaussians(4000, seed=77, log_scale=(0.001, 0.2))
    g["opacities"] = np.random.default_rng(78).uniform(0.004, 1.0, g["opacities"].shape).astype(np.float32)
    cam = S.front_camera(320, 180, 1.2)
    assert CC.census("needles", g, cam) == 0
    out = capsys.readouterr().out
    kept = float(out.split("with the exact test")[1].split("(")[1].split(")")[0])
    assert kept < 0.8, out          # the exact test removes more than a fifth of the bounding box's visits here
