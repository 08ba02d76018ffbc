"""Per-kernel device time of one C5 training step on this GPU: `python tests/gnn_profile_train.py [graphs]`."""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench  # noqa: E402
import torch  # noqa: E402

graphs = int(sys.argv[1]) if len(sys.argv) > 1 else 128
r = bench.gnn_train_bench(torch.device("cuda", 0), "b200", {}, graphs, iters=2)
print("ms_per_step", round(r["ms_per_step"], 2), "loss", r["loss"])
for k, v in sorted(r["kernels"].items(), key=lambda kv: -kv[1]["ms_per_step"]):
    print(f"  {k:18s} {v['launches']:6d} launches {v['ms_per_step']:9.3f} ms  {v['ms_per_step'] / v['launches'] * 1e3:9.1f} us/launch")
