"""Per-tensor error report of the GNN backward against torch autograd over the oracle (run on the GPU box)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import conftest  # noqa
import torch

from test_gnn_backward_gpu import _rel, run_case

cases = [(0, 5), (1, 9), (3, 13), (15, 20)] if len(sys.argv) < 2 else [(int(sys.argv[1]), int(sys.argv[2]))]
for steps, n in cases:
    for emulate in (False, True):
        got, ref, y, yr = run_case(steps, n, 3, emulate)
        print(f"== steps={steps} n={n} emulate_bf16={emulate}  y err {float((y - yr).abs().max()):.3e} / {float(yr.abs().max()):.3e}", flush=True)
        worst = []
        for k, r in ref.items():
            if r is None:
                continue
            worst.append((_rel(got[k], r), k, float(r.norm()), float(got[k].norm())))
        worst.sort(reverse=True)
        for e, k, rn, gn in (worst if "-v" in sys.argv else worst[:12]):
            print(f"   {e:9.3e}  {k:55s} |ref| {rn:.3e} |got| {gn:.3e}")

# ---- noise floor: the bf16-emulating oracle against ITSELF under a permuted edge order (same maths, other
# summation order of the scatter-add -> other bf16 roundings downstream), and against the fp32 oracle
import numpy as np
import csb_synth as S
from oracle import gnn_oracle
from test_gnn_backward_gpu import _model

def oracle_grads(m, x, ei, ef, w, emulate, perm=None):
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    xr = torch.from_numpy(x).cuda().requires_grad_(True)
    er = torch.from_numpy(ef).cuda().requires_grad_(True)
    eit = torch.from_numpy(ei).cuda()
    e_in = er
    if perm is not None:
        pt = torch.from_numpy(perm).cuda()
        eit, e_in = eit[:, pt], er[pt]
    yr = gnn_oracle.forward(sd, xr, eit, e_in, emulate_bf16=emulate, device="cuda", differentiable=True)
    (yr * w).sum().backward()
    g = {k: v.grad for k, v in sd.items()}
    g["x"], g["edge_features"] = xr.grad, er.grad
    return g

for steps, n in cases:
    m = _model(3, steps)
    x, ei, ef, _, _ = S.gnn_inputs(n, 3)
    w = torch.from_numpy(np.random.default_rng(103).standard_normal((x.shape[0], 3)).astype(np.float32)).cuda()
    a = oracle_grads(m, x, ei, ef, w, True)
    b = oracle_grads(m, x, ei, ef, w, True, np.random.default_rng(0).permutation(ei.shape[1]))
    c = oracle_grads(m, x, ei, ef, w, False)
    e1 = max(_rel(b[k], a[k]) for k in a if a[k] is not None)
    e2 = max(_rel(a[k], c[k]) for k in a if a[k] is not None)
    print(f"noise floor steps={steps} n={n}: emulated oracle vs itself (permuted edges) max rel {e1:.3e}; emulated vs fp32 oracle {e2:.3e}", flush=True)
