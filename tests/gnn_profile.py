"""Short GNN workload for ncu / per-kernel timing: `python tests/gnn_profile.py [graphs] [n] [iters]`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import conftest  # noqa
import numpy as np
import torch

import csb_cabi
import csb_synth as S
from meshnet import EncodeProcessDecode

graphs = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
x, ei, ef, _, _ = S.gnn_inputs(n, 4)
N = x.shape[0]
xs, efs = np.tile(x, (graphs, 1)), np.tile(ef, (graphs, 1))
eis = np.ascontiguousarray(np.concatenate([ei + g * N for g in range(graphs)], 1))
torch.manual_seed(0)
m = EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).eval().cuda()
xt, eit, eft = torch.from_numpy(xs).cuda(), torch.from_numpy(eis).cuda(), torch.from_numpy(efs).cuda()
lib = csb_cabi.lib()
with torch.no_grad():
    m(xt, eit, eft)
    torch.cuda.synchronize()
    lib.csb_profile_enable(1)
    for _ in range(iters):
        m(xt, eit, eft)
    torch.cuda.synchronize()
prof = csb_cabi.profile_read()
lib.csb_profile_enable(0)
tot = sum(v[0] for v in prof.values())
print(f"graphs={graphs} N={xs.shape[0]} E={eis.shape[1]} total {tot / iters:.3f} ms/forward")
for k, (ms, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    print(f"  {k:16s} {cnt / iters:5.1f} launches/fwd  {ms / cnt * 1e3:9.1f} us/launch  {ms / iters:8.3f} ms/fwd")
