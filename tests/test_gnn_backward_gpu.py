"""GPU parity of the meshnet GNN backward (csb_gnn_forward_train + csb_gnn_backward behind the drop-in
EncodeProcessDecode's autograd) against torch autograd over the oracle (oracle/gnn_oracle.py with
differentiable=True): the gradients the reference's loss.backward() produces in train_meshnet_sim.py:529.

Tolerances.  The CUDA path rounds the operands of every tensor-core contraction (activations, dZ_l, weights) to
bf16 and accumulates in fp32.  Gradients of a ReLU network are discontinuous in the forward values (a hidden
unit whose pre-activation changes sign under a rounding flips its whole gradient path), so:
  * where the bf16-emulating oracle reproduces the kernel's forward bit for bit (graphs of a few dozen rows), every
    gradient tensor must agree to 2e-2 relative L2 (measured 5e-3 .. 8e-3: the bf16 rounding of dZ_l);
  * on larger graphs a gradient tensor may deviate from the fp32 oracle by at most 1.5x what the bf16-emulating
    oracle itself deviates from the fp32 oracle on that case, + 3e-2 (measured on B200, 15 steps / 20x20:
    kernel 0.15, emulating oracle 0.14; the emulating oracle against itself under a permuted edge order: 0.02-0.08);
  * the whole flat parameter gradient must have cosine similarity > 0.98 with the fp32 oracle's."""
import numpy as np
import pytest
import torch

import conftest  # noqa: F401
import csb_synth as S
from oracle import gnn_oracle

pytestmark = pytest.mark.gpu


def _model(seed, steps, nin=8, ein=4, nout=3):
    from meshnet import EncodeProcessDecode

    torch.manual_seed(seed)
    return EncodeProcessDecode(nin, nout, ein, 128, steps, 2, 128).cuda()


def _rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / max(float(b.norm()), 1e-30))


def run_case(steps, n, seed, emulate=False):
    m = _model(seed, steps)
    x, ei, ef, _, _ = S.gnn_inputs(n, seed)
    rng = np.random.default_rng(seed + 100)
    w = torch.from_numpy(rng.standard_normal((x.shape[0], 3)).astype(np.float32)).cuda()
    xt = torch.from_numpy(x).cuda().requires_grad_(True)
    eft = torch.from_numpy(ef).cuda().requires_grad_(True)
    eit = torch.from_numpy(ei).cuda()
    y = m(xt, eit, eft)
    (y * w).sum().backward()
    torch.cuda.synchronize()
    got = {k: p.grad.detach().clone() for k, p in m.named_parameters()}
    got["x"], got["edge_features"] = xt.grad.clone(), eft.grad.clone()

    sd = {k: v.detach().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    xr = torch.from_numpy(x).cuda().requires_grad_(True)
    er = torch.from_numpy(ef).cuda().requires_grad_(True)
    yr = gnn_oracle.forward(sd, xr, eit, er, emulate_bf16=emulate, device="cuda", differentiable=True)
    (yr * w).sum().backward()
    ref = {k: v.grad for k, v in sd.items()}
    ref["x"], ref["edge_features"] = xr.grad, er.grad
    return got, ref, y.detach(), yr.detach()


@pytest.mark.parametrize("steps,n", [(0, 5), (3, 3), (3, 4), (1, 4)])
def test_gradients_tight_where_forward_is_reproduced(steps, n):
    got, ref, y, yr = run_case(steps, n, 3, emulate=True)
    if float((y - yr).abs().max()) > 1e-6 * max(float(yr.abs().max()), 1e-6):
        # The order of the float atomics of the scatter-add is the hardware's: on some runs a bf16 rounding of `agg` falls the
        # other way and the emulating oracle no longer reproduces this forward bit for bit (graphs with more than one warp
        # of edges per node tile).  The tight bound then does not apply -- ReLU masks may differ -- and the case is held
        # to the bound of the larger graphs instead: within 1.5x the oracle's own bf16-vs-fp32 deviation + 3e-2.
        _, ref32, _, _ = run_case(steps, n, 3, emulate=False)
        _assert_within_floor(got, ref32, ref, f"steps {steps}, n {n} (forward not reproduced bit for bit)")
        return
    bad = {k: _rel(got[k], r) for k, r in ref.items() if r is not None and not _rel(got[k], r) < 2e-2}
    assert not bad, bad


@pytest.mark.parametrize("steps,n", [(1, 9), (3, 13), (15, 20)])
def test_gradients_within_bf16_deviation_of_fp32_oracle(steps, n):
    got, ref32, y, yr = run_case(steps, n, 3, emulate=False)
    _, ref16, _, _ = run_case(steps, n, 3, emulate=True)
    assert float((y - yr).abs().max()) < 5e-2 * max(float(yr.abs().max()), 1e-6)
    bad = {}
    for k, r in ref32.items():
        if r is None:
            continue
        e, floor = _rel(got[k], r), _rel(ref16[k], r)
        if not e < 1.5 * floor + 3e-2:
            bad[k] = (e, floor)
    assert not bad, bad
    keys = [k for k in ref32 if k not in ("x", "edge_features") and ref32[k] is not None]
    a = torch.cat([got[k].flatten() for k in keys]).double()
    b = torch.cat([ref32[k].flatten() for k in keys]).double()
    assert float(a @ b / (a.norm() * b.norm())) > 0.98


def _assert_within_floor(got, ref32, ref16, tag=""):
    bad = {}
    for k, r in ref32.items():
        if r is None or float(r.norm()) == 0.0:
            continue
        e, floor = _rel(got[k], r), _rel(ref16[k], r)
        if not e < 1.5 * floor + 3e-2:
            bad[k] = (e, floor)
    assert not bad, (tag, bad)


def test_ragged_and_no_edges():
    for n in (2, 12):  # N = 4 (one partial tile); N = 144 (two node tiles), E not a multiple of 128
        got, ref32, _, _ = run_case(2, n, 4, emulate=False)
        _, ref16, _, _ = run_case(2, n, 4, emulate=True)
        _assert_within_floor(got, ref32, ref16, n)
    m = _model(1, 2)
    x, ei, ef, _, _ = S.gnn_inputs(3, 0)
    e0, f0 = torch.from_numpy(ei[:, :0]).cuda(), torch.from_numpy(ef[:0]).cuda()
    xt = torch.from_numpy(x).cuda().requires_grad_(True)
    m(xt, e0, f0).sum().backward()
    got = {k: p.grad for k, p in m.named_parameters()}
    got["x"] = xt.grad
    refs = []
    for emu in (False, True):
        sd = {k: v.detach().clone().requires_grad_(True) for k, v in m.state_dict().items()}
        xr = torch.from_numpy(x).cuda().requires_grad_(True)
        gnn_oracle.forward(sd, xr, e0, f0, emulate_bf16=emu, device="cuda", differentiable=True).sum().backward()
        r = {k: v.grad for k, v in sd.items()}
        r["x"] = xr.grad
        refs.append(r)
    _assert_within_floor(got, refs[0], refs[1], "no edges")
    # parameters no row ever reaches (edge encoder, edge MLPs) get exact zeros, as autograd's None/zero
    for k, p in m.named_parameters():
        if "edge_fn" in k:
            assert float(p.grad.abs().max()) == 0.0, k


def test_batched_graphs_equal_sum_of_single_graph_gradients():
    m = _model(0, 2)
    parts = [S.gnn_inputs(n, s) for n, s in ((6, 0), (9, 1))]
    tot = None
    for p in parts:
        m.zero_grad()
        y = m(torch.from_numpy(p[0]).cuda(), torch.from_numpy(p[1]).cuda(), torch.from_numpy(p[2]).cuda())
        y.square().sum().backward()
        g = torch.cat([q.grad.flatten() for q in m.parameters()])
        tot = g if tot is None else tot + g
    m.zero_grad()
    off, xs, eis, efs = 0, [], [], []
    for p in parts:
        xs.append(p[0]); eis.append(p[1] + off); efs.append(p[2])
        off += p[0].shape[0]
    y = m(torch.from_numpy(np.concatenate(xs)).cuda(), torch.from_numpy(np.ascontiguousarray(np.concatenate(eis, 1))).cuda(),
          torch.from_numpy(np.concatenate(efs)).cuda())
    y.square().sum().backward()
    g = torch.cat([q.grad.flatten() for q in m.parameters()])
    assert _rel(g, tot) < 5e-2


def test_two_step_rollout_gradient_flows_through_inputs():
    """pred -> next inputs -> pred (train_meshnet_sim.py:491-532): gradients reach step-0 parameters via dL/dx."""
    m = _model(2, 2)
    x, ei, ef, _, _ = S.gnn_inputs(7, 2)
    xt, eit, eft = torch.from_numpy(x).cuda(), torch.from_numpy(ei).cuda(), torch.from_numpy(ef).cuda()

    def rollout(f):
        a1 = f(xt, eft)
        x2 = torch.cat([xt[:, 3:6], xt[:, 3:6] + a1, xt[:, 6:]], 1)
        e2 = eft + 0.1 * torch.cat([a1[eit[0]] - a1[eit[1]], (a1[eit[0]] - a1[eit[1]]).norm(dim=1, keepdim=True)], 1)
        a2 = f(x2, e2)
        return a1.square().mean() + a2.square().mean()

    rollout(lambda a, b: m(a, eit, b)).backward()
    got = {k: p.grad for k, p in m.named_parameters()}
    refs = []
    for emu in (False, True):
        sd = {k: v.detach().clone().requires_grad_(True) for k, v in m.state_dict().items()}
        rollout(lambda a, b: gnn_oracle.forward(sd, a, eit, b, emulate_bf16=emu, device="cuda",
                                                differentiable=True)).backward()
        refs.append({k: v.grad for k, v in sd.items()})
    _assert_within_floor(got, refs[0], refs[1], "rollout")
