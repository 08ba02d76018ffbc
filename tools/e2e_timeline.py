"""Device timeline of one end-to-end C2 step (host buffers): python tools/e2e_timeline.py [staged=1]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import torch  # noqa: E402
import diff_gaussian_rasterization as dgr  # noqa: E402

use_staged = (sys.argv[1] if len(sys.argv) > 1 else "1") == "1"
dev = torch.device("cuda", 0)
g_np, cams, dl_np, desc = bench.make_workload("c2", 0, 1)
sl = bench.settings_list(cams, dev)
P = bench.PARAM_NAMES
host = {k: torch.from_numpy(g_np[k]).pin_memory() for k in P}
dl_host = torch.from_numpy(dl_np).pin_memory()
leaves = {k: torch.empty(g_np[k].shape, device=dev).requires_grad_(True) for k in P}
dl_dev = torch.empty(dl_np.shape, device=dev)
grads_host = {k: torch.empty(g_np[k].shape).pin_memory() for k in P}
stream = torch.cuda.current_stream()
copy_stream = torch.cuda.Stream()
staged = dgr.StagedPixelGrads(dl_dev, copy_stream)
whole = torch.cuda.Event(enable_timing=True)


def ev():
    return torch.cuda.Event(enable_timing=True)


for it in range(6):
    marks = {}
    t0 = ev(); t0.record(stream)
    copy_stream.wait_stream(stream)
    with torch.no_grad():
        for k in P:
            leaves[k].copy_(host[k], non_blocking=True)
            leaves[k].grad = None
    marks["params up"] = ev(); marks["params up"].record(stream)
    if use_staged:
        with torch.cuda.stream(copy_stream):
            for v in range(4):
                dl_dev[v].copy_(dl_host[v], non_blocking=True)
                staged.events[v].record(copy_stream)
                marks[f"dl view {v} up"] = ev(); marks[f"dl view {v} up"].record(copy_stream)
    else:
        with torch.cuda.stream(copy_stream):
            dl_dev.copy_(dl_host, non_blocking=True)
            marks["dl up"] = ev(); marks["dl up"].record(copy_stream)
    color, radii, depth = dgr.rasterize_views(sl, leaves["means3D"], leaves["opacities"], shs=leaves["shs"],
                                              scales=leaves["scales"], rotations=leaves["rotations"])
    marks["forward"] = ev(); marks["forward"].record(stream)
    if use_staged:
        dgr.backward_staged(color, staged)
    else:
        stream.wait_stream(copy_stream)
        color.backward(dl_dev)
    marks["backward"] = ev(); marks["backward"].record(stream)
    for k in P:
        grads_host[k].copy_(leaves[k].grad, non_blocking=True)
    marks["grads down"] = ev(); marks["grads down"].record(stream)
    torch.cuda.synchronize()
    if it >= 4:
        print("staged" if use_staged else "plain", {k: round(t0.elapsed_time(m) * 1e3) for k, m in marks.items()}, "us")
