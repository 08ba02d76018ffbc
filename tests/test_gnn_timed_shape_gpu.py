"""GNN parity at the shape bench.py times: 40x40 cloth graphs batched as one disjoint union (BASELINE.json configs[4]).
8 graphs = 12 800 nodes, 74 256 edges = 581 edge tiles > 148 SMs, so the edge role runs as gnn_edge2_kernel (two tiles in
flight per CTA) and every role walks several tiles per CTA -- the paths the 7x7 .. 20x20 cases never reach.

Forward against the fp32 oracle (= the reference's op sequence, oracle/gnn_oracle.py pinned on the reference module) and
against the bf16-emulating oracle; one backward against torch autograd over both oracles.  Tolerances as in
test_gnn_gpu.py / test_gnn_backward_gpu.py (north_star: "GNN outputs within bf16 tolerance of the reference fp32 path")."""
import numpy as np
import pytest
import torch

import conftest  # noqa: F401
import csb_synth as S
from oracle import gnn_oracle

pytestmark = pytest.mark.gpu


def _batch(n, graphs, seed):
    xs, eis, efs, off = [], [], [], 0
    for g in range(graphs):
        x, ei, ef, _, _ = S.gnn_inputs(n, seed + g)  # every graph has its own velocities
        xs.append(x)
        eis.append(ei + off)
        efs.append(ef)
        off += x.shape[0]
    return np.concatenate(xs), np.ascontiguousarray(np.concatenate(eis, 1)), np.concatenate(efs)


def _rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / max(float(b.norm()), 1e-30))


def test_forward_and_backward_40x40_x8():
    from meshnet import EncodeProcessDecode

    torch.manual_seed(0)
    m = EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).cuda()
    x, ei, ef = _batch(40, 8, 4)
    assert ei.shape[1] // 128 > 148
    rng = np.random.default_rng(7)
    w = torch.from_numpy(rng.standard_normal((x.shape[0], 3)).astype(np.float32)).cuda()
    xt = torch.from_numpy(x).cuda().requires_grad_(True)
    eft = torch.from_numpy(ef).cuda().requires_grad_(True)
    eit = torch.from_numpy(ei).cuda()
    with torch.no_grad():
        y_inf = m(xt.detach(), eit, eft.detach())  # inference path (csb_gnn_forward)
    y = m(xt, eit, eft)                            # training path (csb_gnn_forward_train)
    (y * w).sum().backward()
    torch.cuda.synchronize()
    assert torch.equal(y.detach(), y_inf), "the training forward and the inference forward disagree"
    got = {k: p.grad.detach().clone() for k, p in m.named_parameters()}
    got["x"], got["edge_features"] = xt.grad.clone(), eft.grad.clone()

    refs, outs = [], []
    for emu in (False, True):
        sd = {k: v.detach().clone().requires_grad_(True) for k, v in m.state_dict().items()}
        xr = torch.from_numpy(x).cuda().requires_grad_(True)
        er = torch.from_numpy(ef).cuda().requires_grad_(True)
        yr = gnn_oracle.forward(sd, xr, eit, er, emulate_bf16=emu, device="cuda", differentiable=True)
        (yr * w).sum().backward()
        r = {k: v.grad for k, v in sd.items()}
        r["x"], r["edge_features"] = xr.grad, er.grad
        refs.append(r)
        outs.append(yr.detach())
    scale = float(outs[0].abs().max())
    e32 = float((y.detach() - outs[0]).abs().max()) / scale
    e16 = float((y.detach() - outs[1]).abs().max()) / scale
    assert e32 < 5e-2, f"forward vs the fp32 oracle: {e32} of the output scale"
    assert e16 < 3e-2, f"forward vs the bf16-emulating oracle: {e16} of the output scale"
    bad = {}
    for k, r in refs[0].items():
        if r is None or float(r.norm()) == 0.0:
            continue
        e, floor = _rel(got[k], r), _rel(refs[1][k], r)
        # measured (tools/diag_gnn_grad.py): worst tensor 0.29 against a floor of 0.21; the kernel's TOTAL gradient is
        # closer to fp32 (0.113) than the emulating oracle's (0.136)
        if not e < 1.5 * floor + 6e-2:
            bad[k] = (e, floor)
    assert not bad, bad
    keys = [k for k in refs[0] if k not in ("x", "edge_features") and refs[0][k] is not None]
    a = torch.cat([got[k].flatten() for k in keys]).double()
    b = torch.cat([refs[0][k].flatten() for k in keys]).double()
    assert float(a @ b / (a.norm() * b.norm())) > 0.98
    c = torch.cat([refs[1][k].flatten() for k in keys]).double()
    # the whole parameter gradient is no further from fp32 than the bf16-emulating oracle's is (+ 2 % of slack)
    assert float((a - b).norm() / b.norm()) < float((c - b).norm() / b.norm()) + 2e-2
