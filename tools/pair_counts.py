"""Counts the (pixel, Gaussian) pairs of the bench workloads with the CPU oracle (TEST INFRASTRUCTURE: imports oracle/).

  python tools/pair_counts.py [c2] [c3] [c4]   ->  profiles/pair_counts.json

Per workload and view: instances R, pairs the reference's per-pixel loop evaluates (forward.cu:309-367, every list entry a
pixel walks before it is done), pairs that pass power <= 0 (one exp each) and pairs that are blended.  bench.py turns them
into the algorithmic FLOPs of the compositing kernels (forward 10 / evaluated + 9 / blended, backward
10 / evaluated + 60 / blended) for its FP32 roofline; the counts depend on the seeded inputs only, not on the GPU.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "cloth-splatting_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
import csb_synth as S  # noqa: E402
import oracle  # noqa: E402

PARAMS = ("means3D", "scales", "rotations", "opacities", "shs")


def workload(name):
    if name == "c2":
        g, cams = S.config_c2(4, 800)
    elif name == "c3":
        g, cams = S.config_c3(1_000_000, 1920, 1080)
    elif name == "c4":
        g, cams = S.config_c4(100_000, 64, 800)
    else:
        raise SystemExit(name)
    return {k: g[k] for k in PARAMS}, cams


def main():
    out_path = os.path.join(ROOT, "profiles", "pair_counts.json")
    out = json.load(open(out_path)) if os.path.exists(out_path) else {}
    for name in (sys.argv[1:] or ["c2", "c3", "c4"]):
        g, cams = workload(name)
        views = []
        t0 = time.time()
        for cam in cams:
            s = oracle.rasterize_forward(g["means3D"], g["opacities"], cam.world_view_transform, cam.full_proj_transform,
                                         cam.camera_center, np.ones(3, np.float32), cam.tanfovx, cam.tanfovy,
                                         cam.image_height, cam.image_width, shs=g["shs"], scales=g["scales"],
                                         rotations=g["rotations"], sh_degree=3)
            views.append({"instances": int(s.num_rendered), "evaluated": int(s.counts[0]), "exp": int(s.counts[1]),
                          "blended": int(s.counts[2])})
            print(name, len(views), views[-1], "%.1fs" % (time.time() - t0), flush=True)
        out[name] = {"gaussians": int(g["means3D"].shape[0]), "width": cams[0].image_width,
                     "height": cams[0].image_height, "views": views,
                     "how": "oracle/rast_oracle.c oracle_render counts[] on csb_synth's seeded workload"}
        json.dump(out, open(out_path, "w"), indent=1)


if __name__ == "__main__":
    main()
