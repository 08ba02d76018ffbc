"""Headline benchmark: rasterizer fwd+bwd Mpix/s on BASELINE.json configs[1] (C2).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c2|c3|c4]

One "step" = forward + backward of every view of this rank's view batch (C2: 4 views at 800x800 of the
10 108 mesh-anchored Gaussians of the 20x20 cloth).  With N > 1 (torchrun) every rank renders its own 4
ring views of the replicated Gaussians and the parameter gradients are summed with ONE flat NCCL
all-reduce per step ("scaling": "weak").

  value   device-resident inputs, C-ABI calls back to back, CUDA events, L2 flushed between steps
  e2e     the package's public API (rasterize_views + autograd) with pinned HOST buffers: H2D of the
          Gaussian parameters and dL/dpixel, D2H of the parameter gradients inside the timed region
  --impl reference: the reference's own CUDA rasterizer (oracle/_ref/libref_cuda.so, compiled unmodified
          from /root/reference) driven view by view the way its Python loop does, on the same inputs.
          The reference has no CPU implementation of the rasterizer; if the .so is absent the CPU
          oracle port is timed instead on a bounded sample.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "cloth-splatting_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import csb_synth as S  # noqa: E402

PARAM_NAMES = ("means3D", "scales", "rotations", "opacities", "shs")


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def pair_counts(workload, views):
    """(pixel, Gaussian) pair counts of the seeded workload, per view (profiles/pair_counts.json, made by
    tools/pair_counts.py with the CPU oracle; they depend on the inputs only)."""
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "pair_counts.json")))[workload]["views"]
        sel = [rec[v] for v in views]
        return {k: sum(r[k] for r in sel) for k in ("instances", "evaluated", "blended")}
    except Exception:
        return None


def fp32_peaks(dev):
    """FP32 FMA and SFU peaks of this GPU, measured now by the library's micro-benchmark (csb_measure_fp32_peaks)."""
    import ctypes

    import csb_cabi

    lib = csb_cabi.lib()
    scratch = torch.empty(148 * 8 * 256 * 4 * 2, dtype=torch.uint8, device=dev)
    a, b, c = ctypes.c_double(0), ctypes.c_double(0), ctypes.c_double(0)
    rc = lib.csb_measure_fp32_peaks(scratch.data_ptr(), scratch.numel(), ctypes.byref(a), ctypes.byref(b), ctypes.byref(c),
                                    torch.cuda.current_stream().cuda_stream)
    csb_cabi.check(rc, "csb_measure_fp32_peaks")
    return {"ffma_tflops": 2 * a.value / 1e12, "ffma2_tflops": 2 * b.value / 1e12, "mufu_tops": c.value / 1e12,
            "how": "csb_measure_fp32_peaks: dependent-chain-free fma.rn.f32 / fma.rn.f32x2 / ex2.approx loops, 8 blocks of "
                   "256 threads per SM, best of 3, CUDA events, this run"}


# ------------------------------------------------------------------------------------------
# workloads
# ------------------------------------------------------------------------------------------
def make_workload(name, rank, world):
    if name == "c2":
        g = S.mesh_gaussians(20, 14, 1)
        center = g["verts"].mean(0)
        cams = S.ring_cameras(4 * world, 800, 800, 0.6911, center)[rank::world]
        desc = "C2: 20x20 cloth, 10108 mesh-anchored Gaussians, SH deg 3, 4 ring views 800x800 per GPU, fwd+bwd"
    elif name == "c3":
        g, cams = S.config_c3(1_000_000, 1920, 1080)
        desc = "C3: 1M random Gaussians, SH deg 3, one 1920x1080 view per GPU, fwd+bwd"
    elif name == "c4":
        g, cams = S.config_c4(100_000, 64, 800)
        per = max(1, 64 // world)
        cams = cams[rank * per:(rank + 1) * per]
        desc = "C4: 100k Gaussians x 64 views 800x800 split over the GPUs, fwd+bwd + gradient all-reduce"
    else:
        raise SystemExit(f"unknown workload {name}")
    g = {k: g[k] for k in PARAM_NAMES}
    H, W = cams[0].image_height, cams[0].image_width
    dl = S.pixel_grad(len(cams), H, W, seed=5 + rank)
    return g, cams, dl, desc


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def l2_flusher(dev):
    buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def flush():
        buf.fill_(1)
    return flush


def settings_list(cams, dev, sh_degree=3):
    from diff_gaussian_rasterization import GaussianRasterizationSettings

    out = []
    for c in cams:
        out.append(GaussianRasterizationSettings(
            image_height=c.image_height, image_width=c.image_width, tanfovx=c.tanfovx, tanfovy=c.tanfovy,
            bg=torch.ones(3, device=dev), scale_modifier=1.0,
            viewmatrix=torch.from_numpy(c.world_view_transform).to(dev),
            projmatrix=torch.from_numpy(c.full_proj_transform).to(dev), sh_degree=sh_degree,
            campos=torch.from_numpy(c.camera_center).to(dev), prefiltered=False, debug=False))
    return out


def batches(n, size=16):
    return [list(range(i, min(n, i + size))) for i in range(0, n, size)]



# ------------------------------------------------------------------------------------------
# GNN leg (BASELINE.json metric part 2: "GNN graphs/s per GPU")
# ------------------------------------------------------------------------------------------
def gnn_flops(N, E, steps=15, nin=8, ein=4, nout=3):
    return steps * (163840.0 * E + 131072.0 * N) + 2.0 * N * (nin * 128 + 2 * 128 * 128) + \
        2.0 * E * (ein * 128 + 2 * 128 * 128) + 2.0 * N * (2 * 128 * 128 + nout * 128)


def gnn_batch_inputs(n, graphs, seed=4):
    """`graphs` copies of the n x n cloth as one disjoint-union batch (PyG Batch semantics)."""
    x, ei, ef, _, _ = S.gnn_inputs(n, seed)
    N = x.shape[0]
    xs = np.tile(x, (graphs, 1))
    efs = np.tile(ef, (graphs, 1))
    eis = np.concatenate([ei + g * N for g in range(graphs)], 1)
    return xs, np.ascontiguousarray(eis), efs


def _time_cuda(fn, iters, warm, flush=None):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ms = []
    st = torch.cuda.current_stream()
    for _ in range(iters):
        if flush:
            flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        fn()
        b.record(st)
        b.synchronize()
        ms.append(a.elapsed_time(b))
    return statistics.median(ms), ms


def gnn_bench(dev, impl, peaks, flush, with_cpu=True, iters=30):
    """C1 (one 20x20 graph, one-step prediction = forward) and a 128-graph 40x40 batch, forward only."""
    from oracle import gnn_oracle

    out = {"metric": "GNN forward graphs/s", "unit": "graphs/s", "dtype": "bf16 (fp32 accumulate, LN, residuals)",
           "impl": impl}
    from meshnet import EncodeProcessDecode

    torch.manual_seed(0)
    model = EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).eval().to(dev)
    sd = {k: v.detach() for k, v in model.state_dict().items()}
    ref_model = None
    if impl == "reference":
        mod = reference_gnn_module()
        if mod is not None:
            ref_model = mod.EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).eval().to(dev)
            ref_model.load_state_dict(model.state_dict())
        out["reference_impl"] = ("the reference's meshnet/graph_network.py (nn.Module) under oracle/pyg_stub" if ref_model is not None
                                 else "oracle/gnn_oracle.py (the module file did not travel: baseline/_ref absent)")
    cases = {"c1_20x20": (20, 1), "batch128_40x40": (40, 128)}
    for name, (n, graphs) in cases.items():
        x, ei, ef = gnn_batch_inputs(n, graphs)
        N, E = x.shape[0], ei.shape[1]
        xt, eit, eft = torch.from_numpy(x).to(dev), torch.from_numpy(ei).to(dev), torch.from_numpy(ef).to(dev)
        xh, efh = torch.from_numpy(x).pin_memory(), torch.from_numpy(ef).pin_memory()
        yh = torch.empty((N, 3), dtype=torch.float32).pin_memory()
        if impl == "b200":
            def step():
                with torch.no_grad():
                    return model(xt, eit, eft)

            def e2e():
                xt.copy_(xh, non_blocking=True)
                eft.copy_(efh, non_blocking=True)
                with torch.no_grad():
                    y = model(xt, eit, eft)
                yh.copy_(y, non_blocking=True)
                torch.cuda.current_stream().synchronize()
        else:
            if ref_model is not None:
                def ref_fwd():
                    return ref_model(xt, eit, eft)
            else:
                def ref_fwd():
                    return gnn_oracle.forward(sd, xt, eit, eft, device=dev)

            def step():
                with torch.no_grad():
                    return ref_fwd()

            def e2e():
                xt.copy_(xh, non_blocking=True)
                eft.copy_(efh, non_blocking=True)
                with torch.no_grad():
                    y = ref_fwd()
                yh.copy_(y, non_blocking=True)
                torch.cuda.current_stream().synchronize()
        it = iters if graphs == 1 else max(5, iters // 3)
        ms, _ = _time_cuda(step, it, 3, flush)
        ms_warm, _ = _time_cuda(step, it, 3, None)
        ms_e2e, _ = _time_cuda(e2e, it, 3, flush)
        fl = gnn_flops(N, E)
        peak = float(peaks.get("bf16_tflops", 1590.0))
        out[name] = {"graphs": graphs, "nodes": N, "edges": E, "ms_per_forward": ms, "ms_l2_warm": ms_warm,
                     "graphs_per_s": graphs / (ms * 1e-3), "e2e_graphs_per_s": graphs / (ms_e2e * 1e-3),
                     "e2e_ms": ms_e2e, "h2d_bytes": int(xh.numel() * 4 + efh.numel() * 4), "d2h_bytes": int(yh.numel() * 4),
                     "tflops": fl / (ms * 1e-3) / 1e12,
                     "roofline": {"bound": "tensor", "achieved": fl / (ms * 1e-3) / 1e12, "peak": peak,
                                  "unit": "TFLOP/s", "frac": fl / (ms * 1e-3) / 1e12 / peak,
                                  "peak_source": "measured" if peaks else "fallback"}}
    if with_cpu:
        x, ei, ef = gnn_batch_inputs(20, 1)
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        sdc = {k: v.cpu() for k, v in sd.items()}
        xt, eit, eft = torch.from_numpy(x), torch.from_numpy(ei), torch.from_numpy(ef)
        n, t0 = 0, time.perf_counter()
        while True:
            with torch.no_grad():
                gnn_oracle.forward(sdc, xt, eit, eft)
            n += 1
            if time.perf_counter() - t0 > 8.0 or n >= 100:
                break
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": n / dt, "unit": "graphs/s", "cores": cores, "kind": "port",
                               "sample": f"{n} x C1 forward (20x20 graph) with the fp32 torch restatement of "
                                         "meshnet/graph_network.py (oracle/gnn_oracle.py) on the host cores"}
    return out

def c5_batch(graphs, n=40, F=10, L=2, seed=4):
    """`graphs` copies of the n x n cloth (BASELINE.json configs[4]: 1600-node graphs, 10-step rollout loss)."""
    _, ei, ef, verts, _ = S.gnn_inputs(n, seed)
    N = verts.shape[0]
    rng = np.random.default_rng(seed)
    NN = N * graphs
    d = dict(
        velocity=(rng.standard_normal((NN, 3 * L)) * 0.01).astype(np.float32),
        node_types=np.tile(np.eye(1, N, 0, dtype=np.float32).T, (graphs, 1)),
        edge_index=np.ascontiguousarray(np.concatenate([ei + g * N for g in range(graphs)], 1)),
        edge_features=np.tile(ef, (graphs, 1)),
        target=(rng.standard_normal((NN, F, 3)) * 0.01).astype(np.float32),
        actions=np.zeros((NN, F + 1, 3), np.float32),
        positions=np.tile(verts.astype(np.float32), (graphs, 1)),
        noise=(rng.standard_normal((NN, 3 * L)) * 0.003).astype(np.float32))
    d["actions"][::N] = (rng.standard_normal((graphs, F + 1, 3)) * 0.02).astype(np.float32)
    return d


def gnn_train_bench(dev, impl, peaks, graphs, F=10, iters=5, world=1, rank=0):
    """One optimisation step of the cloth simulator on this rank's share of C5 (1024 graphs of 1600 nodes over
    8 GPUs = 128 per GPU): zero_grad, 10-step rollout loss (train_meshnet_sim.py:491-527), backward, ONE flat
    all-reduce of the 2.33 M parameter gradients, fused Adam step."""
    import csb_dist
    from meshnet import ClothMeshSimulator
    from meshnet.rollout import multi_step_loss

    torch.manual_seed(0)
    sim = ClothMeshSimulator(3, 8, 4, 128, 15, 2, 128, 2, 2, normalize=True, device=dev).to(dev)
    if impl == "reference":
        from oracle import gnn_oracle

        class OracleGnn(torch.nn.Module):  # the reference's fp32 op sequence under torch autograd
            def __init__(self, inner):
                super().__init__()
                self.inner = inner

            def forward(self, x, ei, ea):
                sd = dict(self.inner.named_parameters())
                return gnn_oracle.forward(sd, x, ei, ea, device=x.device, differentiable=True)

        mod = reference_gnn_module()
        if mod is not None:  # the reference's own module under torch autograd
            ref_model = mod.EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).to(dev)
            ref_model.load_state_dict(sim._encode_process_decode.state_dict())
            sim._encode_process_decode = ref_model
        else:
            sim._encode_process_decode = OracleGnn(sim._encode_process_decode)
    sim.train()
    if world > 1:
        csb_dist.sync_normalizers(sim)  # the running statistics are those of all ranks' graphs, identical on every rank
    d = c5_batch(graphs, F=F, seed=4 + rank)  # every rank trains on its own shard of the 1024 graphs
    t = {k: torch.from_numpy(v).to(dev) for k, v in d.items()}
    N, E = t["velocity"].shape[0], t["edge_index"].shape[1]
    flat = csb_dist.attach_flat_grads(sim.parameters())
    opt = torch.optim.Adam(sim.parameters(), lr=1e-4, fused=True)
    last = {}

    def step():
        flat.zero_()
        loss = multi_step_loss(sim, t["velocity"], t["node_types"], t["edge_index"], t["edge_features"], t["target"],
                               t["actions"], t["positions"], F, velocity_noise=t["noise"])
        loss.backward()
        csb_dist.all_reduce_flat(flat, average=True)
        opt.step()
        last["loss"] = loss.detach()

    ms, _ = _time_cuda(step, iters, 3, None)
    kern = None
    if impl == "b200":  # per-kernel device time of one more step (library CUDA events)
        import csb_cabi
        csb_cabi.lib().csb_profile_enable(1)
        step()
        torch.cuda.synchronize()
        kern = {k: {"ms_per_step": v[0], "launches": int(v[1])} for k, v in csb_cabi.profile_read().items()}
        csb_cabi.lib().csb_profile_enable(0)
    fl = 3.0 * F * gnn_flops(N, E)
    peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1590.0)))
    return {"graphs": graphs, "nodes": N, "edges": E, "rollout_steps": F, "ms_per_step": ms,
            "graphs_per_s": world * graphs / (ms * 1e-3), "loss": float(last["loss"]),
            "tflops": fl / (ms * 1e-3) / 1e12,
            "roofline": {"bound": "tensor", "achieved": fl / (ms * 1e-3) / 1e12, "peak": peak, "unit": "TFLOP/s",
                         "frac": fl / (ms * 1e-3) / 1e12 / peak, "flops": "3 x forward (SURVEY.md 8d), recompute not counted",
                         "peak_source": "measured sustained" if peaks else "fallback"},
            "inputs": "resident (the batch is larger than L2: %.0f MB of saved state per prediction)" % (
                (31.0 * N + E) * 256 / 1e6), **({"kernels": kern} if kern else {})}


# ------------------------------------------------------------------------------------------
# product arm
# ------------------------------------------------------------------------------------------
def raster_leg(workload, rank, world, dev, steps, warmup, fp32, peaks, with_e2e=True):
    """Forward + backward of this rank's views of `workload`, device-resident (C-ABI wrapper layer, no autograd) and --
    with_e2e -- through the public API with host buffers.  Returns the measurements of one workload."""
    import csb_cabi
    import csb_dist
    import diff_gaussian_rasterization as dgr
    import torch.distributed as dist

    lib = csb_cabi.lib()
    g_np, cams, dl_np, desc = make_workload(workload, rank, world)
    V, H, W = len(cams), cams[0].image_height, cams[0].image_width
    P = g_np["means3D"].shape[0]
    gd = {k: torch.from_numpy(v).to(dev) for k, v in g_np.items()}
    dl = torch.from_numpy(dl_np).to(dev)
    sl = settings_list(cams, dev)
    view_batches = batches(V)
    flush = l2_flusher(dev)
    e = torch.Tensor([])
    stream = torch.cuda.current_stream()

    def views_of(b):
        return [(sl[v].viewmatrix, sl[v].projmatrix, sl[v].campos, sl[v].bg, sl[v].tanfovx, sl[v].tanfovy) for v in b]

    # 3+3+4+1+48 floats per Gaussian, one flat buffer -> one all-reduce per step
    bucket = csb_dist.FlatGradBucket({k: gd[k].shape for k in PARAM_NAMES}, dev)
    grad_flat = bucket.flat

    def pack(grads):
        d = dict(zip(("means3D", "scales", "rotations", "opacities", "shs"), grads))
        bucket.copy_(d) if len(view_batches) == 1 else bucket.add_(d)

    bucket_views = bucket.as_dict()

    def device_step():
        """Device-resident step through the C-ABI wrapper layer (no autograd).  The instance count of a forward is
        read (and its overflow check made) after the backward has been enqueued: the host never sits between two
        kernels of the step."""
        for attempt in range(3):
            try:
                return device_step_once()
            except dgr.RenderAgain:  # binning workspace too small for this scene: hints are updated, do the step again
                if attempt == 2:
                    raise

    def device_step_once():
        if len(view_batches) > 1:
            grad_flat.zero_()
        pending = []
        for b in view_batches:
            vs = views_of(b)
            fwd = dgr._forward_views(vs, gd["means3D"], e, gd["opacities"], gd["scales"], gd["rotations"], 1.0, e, H, W,
                                     gd["shs"], 3, False, defer_count=True)
            R, color, depth, radii, geom, binning, img = fwd
            pending.append(R)
            # one launch set covers all views of the batch: its gradients go straight into the flat bucket
            gb = dgr._backward_views(vs, gd["means3D"], radii, e, gd["scales"], gd["rotations"], 1.0, e, H, W,
                                     dl[b[0]:b[-1] + 1], gd["shs"], 3, geom, binning, img,
                                     out=bucket_views if len(view_batches) == 1 else None)
            if len(view_batches) > 1:
                dm2, dcol, dop, dm3, dcov, dsh, dsc, drot = gb
                pack((dm3, dsc, drot, dop, dsh))
        # the counts (long since written) are checked BEFORE the collective is enqueued: a rank that has to render again
        # must not have issued an all-reduce its peers will not match
        R = sum(p.wait() for p in pending)
        bucket.all_reduce()
        return R

    def timed(step_fn, steps, warmup):
        for _ in range(warmup):
            step_fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        # every step is bracketed by its own event pair on the launching stream; the host is not made to wait for
        # the step (a training loop does not either): the ranks then stay in lockstep behind each all-reduce
        # instead of re-starting every step with the skew of a host wake-up
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for a, b_ in ev:
            flush()
            a.record(stream)
            step_fn()
            b_.record(stream)
        torch.cuda.synchronize()
        ms = [a.elapsed_time(b_) for a, b_ in ev]
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        total = torch.tensor([sum(ms)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(total, op=dist.ReduceOp.MAX)
        return float(total.item()) / steps, ms

    launches0 = lib.csb_launch_count()
    if bucket.peer:
        bucket.peer.stats(reset=True)
    ms_step, ms_all = timed(device_step, steps, warmup)
    allreduce = None
    if world > 1:
        allreduce = {"impl": "csb_peer_all_reduce (one kernel over NVLink peer memory)" if bucket.peer else "NCCL",
                     "floats": bucket.numel}
        if bucket.peer:  # the kernel's own clock, mean over warm-up + timed calls on rank 0
            allreduce.update({k: round(v, 2) if isinstance(v, float) else v for k, v in bucket.peer.stats().items()})
    launches = (lib.csb_launch_count() - launches0) // (steps + warmup) * steps
    R_step = device_step()
    torch.cuda.synchronize()
    out = {"workload": workload, "desc": desc, "V": V, "H": H, "W": W, "P": P, "ms_step": ms_step, "ms_all": ms_all,
           "R_step": int(R_step), "launches": int(launches), "allreduce": allreduce, "peer": bool(bucket.peer)}

    # ---- end to end through the public API with host buffers --------------------------------------
    if with_e2e:
        host = {k: torch.from_numpy(g_np[k]).pin_memory() for k in PARAM_NAMES}
        dl_host = torch.from_numpy(dl_np).pin_memory()
        leaves = {k: torch.empty_like(gd[k]).requires_grad_(True) for k in PARAM_NAMES}
        dl_dev = torch.empty_like(dl)
        grads_host = {k: torch.empty(g_np[k].shape, dtype=torch.float32).pin_memory() for k in PARAM_NAMES}
        out["h2d"] = sum(t.numel() * 4 for t in host.values()) + dl_host.numel() * 4
        out["d2h"] = sum(t.numel() * 4 for t in grads_host.values())

        copy_stream = torch.cuda.Stream(device=dev)
        # one staged upload per launch set (C4: 4 sets of 16 views): a set's backward waits for ITS views only
        staged = [dgr.StagedPixelGrads(dl_dev[b[0]:b[-1] + 1], copy_stream) for b in view_batches]

        def e2e_step():
            # the parameters (2.4 MB) are enqueued FIRST: host-to-device copies of all streams go through the copy engine
            # in submission order, and the forward only needs them.  dL/dpixel (30.7 MB, needed by the backward only)
            # follows on a side stream while the forward renders
            with torch.no_grad():
                for k in PARAM_NAMES:
                    leaves[k].copy_(host[k], non_blocking=True)
                    leaves[k].grad = None
            # ... and the side stream is held back until they are through (the L2 flush in front of the step would
            # otherwise let the big copy slip ahead of them in the copy engine)
            copy_stream.wait_stream(stream)
            # one copy + event per view: view v's compositing backward starts when ITS plane has arrived
            for st, b in zip(staged, view_batches):
                st.upload(dl_host[b[0]:b[-1] + 1])
            for st, b in zip(staged, view_batches):
                color, radii, depth = dgr.rasterize_views([sl[v] for v in b], leaves["means3D"], leaves["opacities"],
                                                          shs=leaves["shs"], scales=leaves["scales"],
                                                          rotations=leaves["rotations"])
                dgr.backward_staged(color, st)
            if world > 1:
                # the autograd gradients go into the peer-memory bucket, one NVLink all-reduce kernel sums it in place
                bucket.copy_({k: leaves[k].grad for k in PARAM_NAMES})
                bucket.all_reduce()
                for k in PARAM_NAMES:
                    grads_host[k].copy_(bucket.view(k), non_blocking=True)
            else:
                for k in PARAM_NAMES:
                    grads_host[k].copy_(leaves[k].grad, non_blocking=True)
            torch.cuda.current_stream().synchronize()

        out["ms_e2e"], _ = timed(e2e_step, steps, warmup)

    # ---- per-kernel device time of the same step (library CUDA events) -> one roofline per kernel ---
    nprof = max(3, min(steps, 10))
    lib.csb_profile_enable(1)
    for _ in range(nprof):
        flush()
        device_step()
    torch.cuda.synchronize()
    prof = csb_cabi.profile_read()
    lib.csb_profile_enable(0)
    kernels = {k: {"ms_per_launch": ms / n, "launches_per_step": n / nprof, "ms_per_step": ms / nprof}
               for k, (ms, n) in prof.items()}
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    nb = max(1, len(view_batches))
    R_launch = R_step / nb
    Vb = len(view_batches[0])
    alg = {  # algorithmic bytes per launch of the HBM-bound stages
        "preprocess": 311.0 * P * Vb,
        "emit": 20.0 * P * Vb + 12.0 * R_launch,
        "sort_depth": 16.0 * P * Vb,          # per pass: (depth key, entry) pairs of the V x P entries, read + written
        "sort_tile": 16.0 * R_launch,         # per pass: (tile id, entry) pairs of the instances, read + written
        "sort_onesweep": 16.0 * R_launch,
        "sort_histogram": 4.0 * R_launch,
        # the split keys make this stage read ONE 4-byte tile id per instance (the reference reads two 8-byte keys: 16 B)
        "tile_ranges": 4.0 * R_launch + 8.0 * tiles * Vb,
        "bin_ranges": 8.0 * tiles * Vb,
        "bin_fill": 20.0 * P * Vb + 8.0 * R_launch,
        "bin_sort": 16.0 * R_launch,
        "gaussian_backward": (236.0 + 24 + 40 * Vb + 256) * P,
    }
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    for k, b_ in alg.items():
        if k in kernels:
            ach = b_ / (kernels[k]["ms_per_launch"] * 1e-3) / 1e9
            kernels[k]["roofline"] = {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                                      "frac": ach / hbm_peak, "algorithmic_bytes_per_launch": b_}
    # compositing kernels: FP32 FMA pipe + SFU.  Pairs counted by the CPU oracle on these very inputs
    # (profiles/pair_counts.json): forward 10 flop / evaluated + 9 / blended pair and one exp per evaluated pair,
    # backward 10 / evaluated + 60 / blended
    view_ids = list(range(rank, V * world, world)) if workload == "c2" else (
        list(range(rank * V, rank * V + V)) if workload == "c4" else list(range(V)))
    pc = pair_counts(workload, view_ids) if (workload != "c2" or world == 1) else None
    if pc:
        out["pairs_per_step"] = pc
        fl = {"render_forward": 10.0 * pc["evaluated"] + 9.0 * pc["blended"],
              "render_backward": 10.0 * pc["evaluated"] + 60.0 * pc["blended"]}
        for k, f in fl.items():
            if k in kernels:
                t = kernels[k]["ms_per_step"] * 1e-3
                ach = f / t / 1e12
                kernels[k]["roofline"] = {"bound": "fp32", "achieved": ach, "peak": fp32["ffma2_tflops"], "unit": "TFLOP/s",
                                          "frac": ach / fp32["ffma2_tflops"], "flops_per_step": f,
                                          "sfu_frac": pc["evaluated"] / t / 1e12 / fp32["mufu_tops"],
                                          "peak_source": "csb_measure_fp32_peaks, this run"}
    out["kernels"] = kernels
    return out


def knn_bench(dev, impl, flush, iters=10):
    """distCUDA2 (simple_knn.cu:185-221) at the sizes BASELINE.md names: the 38x38 / C2 clouds and 1 M random points."""
    out = {"metric": "distCUDA2 Mpoints/s", "impl": impl}
    clouds = {"1444": S.mesh_gaussians(20, 2, 1)["means3D"], "10108": S.mesh_gaussians(20, 14, 1)["means3D"],
              "1000000": S.config_c3(1_000_000, 1920, 1080)[0]["means3D"]}
    if impl == "b200":
        from simple_knn._C import distCUDA2 as fn
    else:
        from oracle import ref_cuda

        fn = ref_cuda.dist_cuda2
    for name, pts in clouds.items():
        t = torch.from_numpy(np.ascontiguousarray(pts)).to(dev)
        ms, _ = _time_cuda(lambda: fn(t), iters, 3, flush)
        out[name] = {"points": int(t.shape[0]), "ms": ms, "mpoints_per_s": t.shape[0] / 1e6 / (ms * 1e-3),
                     "hbm_frac_16B_per_point": 16.0 * t.shape[0] / (ms * 1e-3) / 1e9 / 6535.0}
    return out


def run_b200(args, rank, world, dev):
    import csb_cabi
    import torch.distributed as dist

    lib = csb_cabi.lib()
    if args.render_variant >= 0:
        lib.csb_debug_render_variant(args.render_variant)
    if args.cull_exact >= 0:
        lib.csb_debug_cull_exact(args.cull_exact)
    peaks = load_peaks()
    fp32 = fp32_peaks(dev)
    flush = l2_flusher(dev)
    sampler = ClockSampler(dev.index) if rank == 0 else None
    main = raster_leg(args.workload, rank, world, dev, args.steps, args.warmup, fp32, peaks)
    clocks = sampler.stop() if sampler else None
    V, H, W, P = main["V"], main["H"], main["W"], main["P"]
    ms_step, ms_e2e, ms_all, kernels = main["ms_step"], main["ms_e2e"], main["ms_all"], main["kernels"]
    strong = args.workload == "c4"  # 64 views split over the ranks: the total work is fixed
    mpix = V * world * H * W / 1e6  # all ranks' views

    top = max(kernels, key=lambda k: kernels[k]["ms_per_step"])
    roofline = dict(kernels[top].get("roofline", {"bound": "latency"}))
    roofline["kernel"] = top
    roofline["ms_per_launch"] = kernels[top]["ms_per_launch"]
    # dram__bytes_read.sum + dram__bytes_write.sum per launch of that kernel, from the committed ncu --set full
    # capture of this same command (profiles/r2_traffic.json; not measured live: ncu replays kernels)
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))[args.workload].get(top)
    except Exception:
        pass
    roofline["traffic"] = traffic
    if traffic is not None:
        roofline["traffic_source"] = "profiles/r2_traffic.json: dram bytes read + written per launch in the committed ncu --set full capture of this command"
    roofline["fp32_peaks"] = fp32

    line = {
        "metric": "rasterizer fwd+bwd Mpix/s", "value": mpix / (ms_step * 1e-3), "unit": "Mpix/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": main["desc"], "gaussians": P, "views_per_gpu": V, "width": W, "height": H,
                   "instances_per_step": main["R_step"], "l2": "flushed between steps (256 MB write)",
                   "parallelism": f"views x{world}" + ((" + 1 flat gradient all-reduce (" + ("csb_peer_all_reduce over NVLink peer memory" if main["peer"] else "NCCL")
                                                             + ")") if world > 1 else "")},
        "e2e": {"value": mpix / (ms_e2e * 1e-3), "unit": "Mpix/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": main["h2d"], "d2h_bytes_per_step": main["d2h"],
                "api": "diff_gaussian_rasterization.rasterize_views + autograd backward"
                       + " (dL/dpixel uploaded and consumed view by view: StagedPixelGrads / backward_staged)"},
        "gpu_launches": main["launches"], "clocks": clocks, "roofline": roofline, "kernels": kernels,
        "impl": "b200", "render_variant": int(lib.csb_debug_render_variant(-1)), "cull_exact": int(lib.csb_debug_cull_exact(-1)),
        # spread of the timed steps on rank 0 (device time): what a synchronous all-reduce turns into rank skew
        "step_ms_spread": {"min": min(ms_all), "median": statistics.median(ms_all), "max": max(ms_all),
                           "std": statistics.pstdev(ms_all)},
    }
    if main.get("pairs_per_step"):
        line["config"]["pairs_per_step"] = main["pairs_per_step"]
    if main["allreduce"]:
        line["allreduce"] = main["allreduce"]
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        g_np, cams, dl_np, _ = make_workload(args.workload, rank, world)
        line["cpu_baseline"] = cpu_baseline(g_np, cams, dl_np)
    # ---- the other configurations of BASELINE.json, one sub-record each ------------------------------------
    if not args.no_sub and args.workload == "c2":
        # C4: 100k Gaussians x 64 views split over the ranks (strong scaling), gradients all-reduced once per step
        c4 = raster_leg("c4", rank, world, dev, max(3, args.steps // 6), 3, fp32, peaks, with_e2e=False)
        if rank == 0:
            line["c4"] = {"workload": c4["desc"], "scaling": "strong", "views_total": c4["V"] * world,
                          "views_per_gpu": c4["V"], "ms_per_step": c4["ms_step"],
                          "value": c4["V"] * world * c4["H"] * c4["W"] / 1e6 / (c4["ms_step"] * 1e-3), "unit": "Mpix/s",
                          "instances_per_step_per_gpu": c4["R_step"], "gpu_launches": c4["launches"],
                          "kernels": c4["kernels"], "allreduce": c4["allreduce"]}
        if rank == 0 and world == 1:
            c3 = raster_leg("c3", rank, world, dev, max(3, args.steps // 3), 3, fp32, peaks, with_e2e=False)
            line["c3"] = {"workload": c3["desc"], "ms_per_step": c3["ms_step"],
                          "value": c3["H"] * c3["W"] / 1e6 / (c3["ms_step"] * 1e-3), "unit": "Mpix/s",
                          "instances_per_step": c3["R_step"], "pairs_per_step": c3.get("pairs_per_step"),
                          "gpu_launches": c3["launches"], "kernels": c3["kernels"]}
            k3 = c3["kernels"]
            mine = sum(k3[k]["ms_per_step"] for k in ("sort_onesweep", "sort_depth", "sort_tile", "sort_histogram", "emit", "tile_ranges") if k in k3)
            line["sort"] = {"instances": c3["R_step"],
                            "b200_ms": mine,
                            "b200_stages": "depth sort (4 onesweep passes over V*P u32 keys) + emit (scan + duplicate) + "
                                           "tile sort (2 passes over the 13 tile bits) + tile_ranges, inside the C3 step",
                            "reference_stages": "InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::SortPairs(u64,u32; "
                                                "45 bits) + identifyTileRanges: see the reference arm's `stages` record"}
            line["knn"] = knn_bench(dev, "b200", flush)
        if world > 1:
            dist.barrier()
    if not args.no_gnn:
        l0 = lib.csb_launch_count()
        if rank == 0:
            line["gnn"] = gnn_bench(dev, "b200", peaks, flush, with_cpu=(world == 1 and not args.no_cpu_baseline))
        if world > 1:
            dist.barrier()
        # every rank trains on its own 128 graphs of C5 (1024 graphs over 8 GPUs), gradients all-reduced
        tr = gnn_train_bench(dev, "b200", peaks, args.gnn_graphs, world=world, rank=rank)
        if world > 1:
            tms = torch.tensor([tr["ms_per_step"]], dtype=torch.float64, device=dev)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            tr["graphs_per_s"] *= tr["ms_per_step"] / float(tms.item())
            tr["ms_per_step"] = float(tms.item())
        if rank == 0:
            line["gnn"]["train_c5"] = tr
            line["gnn"]["gpu_launches"] = int(lib.csb_launch_count() - l0)
    if rank == 0 and not args.no_loss:
        l0 = lib.csb_launch_count()
        line["refine_tail"] = loss_adam_bench(dev, "b200", peaks, flush)
        line["refine_tail"]["gpu_launches"] = int(lib.csb_launch_count() - l0)
    if rank == 0 and world == 1 and not args.no_dropin and args.workload == "c2":
        try:
            line["dropin"] = dropin_bench(dev, "b200")
        except Exception as ex:  # the sub-record must never cost the headline
            line["dropin"] = {"error": repr(ex)}
    if rank == 0:
        print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
# Photometric loss + optimiser step of the refinement loop
# ------------------------------------------------------------------------------------------
def loss_adam_bench(dev, impl, peaks, flush, V=4, H=800, W=800, P=10108, iters=20):
    """image_losses(fwd+bwd) on the C2 views and the 7-group Adam step over the C2 Gaussians.  impl 'b200': the fused
    kernels through utils.loss_utils / csb_optim; impl 'reference': the reference's own torch op sequence
    (utils/loss_utils.py via oracle/loss_oracle.py, torch.optim.Adam) on the same B200."""
    rng = np.random.default_rng(5)
    gt = rng.random((V, 3, H, W), dtype=np.float32)
    img = np.clip(gt + 0.05 * rng.standard_normal(gt.shape).astype(np.float32), 0, 1)
    x = torch.from_numpy(img).to(dev).requires_grad_(True)
    y = torch.from_numpy(gt).to(dev)
    shapes = {"face_bary": (P, 3), "face_offset": (P, 1), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1),
              "scaling": (P, 3), "rotation": (P, 4)}
    lrs = {"face_bary": 1.6e-4, "face_offset": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.05,
           "scaling": 5e-3, "rotation": 1e-3}
    params = {k: torch.nn.Parameter(torch.from_numpy(rng.standard_normal(sh).astype(np.float32)).to(dev))
              for k, sh in shapes.items()}
    for k, p_ in params.items():
        p_.grad = torch.from_numpy(rng.standard_normal(shapes[k]).astype(np.float32)).to(dev)
    groups = [{"params": [params[k]], "lr": lrs[k], "name": k} for k in shapes]
    if impl == "b200":
        from csb_optim import Adam
        from utils.loss_utils import image_loss

        opt = Adam(groups, lr=0.0, eps=1e-15)

        def loss_step():
            x.grad = None
            image_loss(x, y, 0.1)[0].backward()
    else:
        from oracle import loss_oracle

        opt = torch.optim.Adam(groups, lr=0.0, eps=1e-15)

        def loss_step():
            x.grad = None
            loss_oracle.image_losses(x, y, 0.1)[0].backward()

    ms_loss, _ = _time_cuda(loss_step, iters, 3, flush)
    ms_adam, _ = _time_cuda(opt.step, iters, 3, flush)
    n = V * 3 * H * W
    elems = sum(int(np.prod(sh)) for sh in shapes.values())
    peak = float(peaks.get("hbm_gbs", 6650.0))
    out = {"impl": impl,
           "image_loss": {"shape": [V, 3, H, W], "lambda_dssim": 0.1, "ms_fwd_bwd": ms_loss,
                          "mpix_per_s": V * H * W / 1e6 / (ms_loss * 1e-3)},
           "adam": {"tensors": len(shapes), "elements": elems, "ms_per_step": ms_adam}}
    if impl == "b200":
        import csb_cabi

        lib = csb_cabi.lib()
        lib.csb_profile_enable(1)
        for _ in range(5):
            flush()
            loss_step()
            opt.step()
        torch.cuda.synchronize()
        prof = csb_cabi.profile_read()
        lib.csb_profile_enable(0)
        alg = {"image_loss_forward": 20.0 * n, "image_loss_backward": 24.0 * n, "adam_step": 28.0 * elems}
        for k, b in alg.items():
            if k in prof:
                ms, cnt = prof[k]
                ach = b / (ms / cnt * 1e-3) / 1e9
                tgt = out["image_loss"] if k.startswith("image") else out["adam"]
                tgt.setdefault("kernels", {})[k] = {
                    "ms_per_launch": ms / cnt,
                    "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                 "peak_source": "measured" if peaks else "fallback"}}
    return out


# ------------------------------------------------------------------------------------------
# The refinement iteration as a reference caller writes it (scene_reconstruction/train_utils.py:240-321)
# ------------------------------------------------------------------------------------------
def _dropin_options():
    import types

    return types.SimpleNamespace(
        lambda_dssim=0.2, lambda_deform_mag=0.1, lambda_rigid=0.5, lambda_momentum=0.05, densify_until_iter=15000,
        densify_from_iter=500, densification_interval=100, pruning_from_iter=500, pruning_interval=300,
        opacity_reset_interval=3000, bary_cleanup=700, iterations=30000, opacity_threshold_fine_init=0.5,
        opacity_threshold_fine_after=0.4, densify_grad_threshold_fine_init=1e-4, densify_grad_threshold_after=5e-5,
        position_lr_init=1.6e-4, feature_lr=2.5e-3, opacity_lr=0.05, scaling_lr=5e-3, rotation_lr=1e-3)


def dropin_bench(dev, impl, warm=8, steps=40, W=800, H=800):
    """`train_step` with the reference's signature on a C2-sized scene (20x20 cloth, 10108 mesh-anchored Gaussians, the 3
    consecutive-time views of a training item at 800x800): simulator -> get_xyz / get_rotation -> one GaussianRasterizer call per
    view -> image_losses + regularisers -> backward -> densification statistics -> both Adam steps, the loss_dict `.item()`
    reads included.  Wall clock per iteration (the loop synchronises once per iteration, like the reference's).
      b200:      the package's train_step (CUDA graph replay between schedule events).
      reference: the same loop over the reference's own CUDA rasterizer (oracle/_ref) behind an autograd wrapper, the reference's
                 torch op sequences for the loss (oracle/loss_oracle.py), get_xyz / get_rotation (oracle/mesh_oracle.py: the
                 batched SVD roma runs), torch.optim.Adam -- what train_utils.py:240-321 executes, where it can be run here."""
    import types

    from meshnet.meshnet_network import ResidualMeshSimulator
    from scene_reconstruction.gaussian_mesh import MultiGaussianMesh

    verts, faces, ei = S.cloth_mesh(20)
    pos = torch.from_numpy(verts).to(dev)
    g = torch.Generator(device=dev).manual_seed(0)
    pc = MultiGaussianMesh.from_mesh(pos, torch.from_numpy(np.ascontiguousarray(faces.T)).to(dev), gaussian_init_factor=14,
                                     generator=g)
    pc.set_edges(torch.from_numpy(ei).to(dev))
    pc.active_sh_degree = 3
    with torch.no_grad():
        pc._opacity.add_(2.0)
        pc._features_dc.add_(torch.randn(pc._features_dc.shape, device=dev, generator=g))
    opt = _dropin_options()
    pc.training_setup(opt)
    pc.init_densification_stats(0.01)
    sim = ResidualMeshSimulator(torch.stack([pos, pos + 0.01, pos + 0.02]), device=dev)
    mesh_opt = torch.optim.Adam(sim.parameters(), lr=1e-4)
    cams = S.ring_cameras(3, W, H, 0.6911, verts.mean(0))
    for i, cam in enumerate(cams):
        for k in ("world_view_transform", "full_proj_transform", "camera_center"):
            setattr(cam, k, torch.from_numpy(getattr(cam, k)).to(dev))
        cam.time, cam.mask = i / 2, None
        cam.original_image = torch.rand((3, H, W), generator=torch.Generator().manual_seed(i)).pin_memory()
    bg = torch.ones(3, device=dev)
    pipe, uargs = types.SimpleNamespace(compute_cov3D_python=False), types.SimpleNamespace(no_shadow=True)

    if impl == "b200":
        from scene_reconstruction.train_utils import train_step

        def step(it):
            return train_step(it, cams, pc, sim, mesh_opt, pipe, opt, 5.0, bg, user_args=uargs)
    else:
        step = _reference_train_step(dev, pc, sim, mesh_opt, cams, opt, bg)

    def timed_loop(first_it):
        losses, t = [], []
        for it in range(first_it, first_it + warm + steps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            _, loss, _ = step(it)
            torch.cuda.synchronize()
            t.append((time.perf_counter() - t0) * 1e3)
            losses.append(float(loss))
        ms = sorted(t[warm:])
        return ms[len(ms) // 2], ms[0], ms[-1], losses

    out = {"what": "train_step (reference signature), 3 views 800x800, 10108 Gaussians, SH 3, one iteration", "impl": impl,
           "steps": steps, "warmup": warm, "timing": "wall clock around the call, device synchronised on both sides"}
    if impl == "b200":
        # (1) the ground truth crosses PCIe in every iteration, like the reference's `.cuda()` per view (train_utils.py:262);
        # (2) the product's default: a view's planes stay resident after their first use (train_utils.ResidentPlanes)
        from scene_reconstruction import train_utils as TU

        budget = TU.resident_planes.budget
        TU.resident_planes.budget = 0
        TU.resident_planes.clear()
        med, lo, hi, losses = timed_loop(1)
        TU.resident_planes.budget = budget
        med_r, lo_r, hi_r, _ = timed_loop(1 + warm + steps)
        out.update({"ms_per_iteration": med, "ms_min": lo, "ms_max": hi, "mpix_per_s": 3 * W * H / med / 1e3,
                    "ground_truth": "uploaded from pinned host memory in every iteration (3 x 7.68 MB), as the reference does",
                    "resident_ground_truth": {"ms_per_iteration": med_r, "ms_min": lo_r, "ms_max": hi_r, "mpix_per_s": 3 * W * H / med_r / 1e3,
                                              "what": "the product's default: a view's ground-truth planes stay in HBM after their first "
                                                      "use (ResidentPlanes, LRU over CSB_GT_CACHE_MB), so steady-state iterations upload nothing"}})
    else:
        med, lo, hi, losses = timed_loop(1)
        out.update({"ms_per_iteration": med, "ms_min": lo, "ms_max": hi, "mpix_per_s": 3 * W * H / med / 1e3,
                    "ground_truth": "uploaded from pinned host memory in every iteration (3 x 7.68 MB)"})
    out.update({"loss_first": losses[0], "loss_last": losses[-1]})
    return out


def _reference_train_step(dev, pc, sim, mesh_opt, cams, opt, bg):
    """The reference's iteration with the reference's kernels and op sequences (see dropin_bench)."""
    import math

    from oracle import loss_oracle, mesh_oracle, ref_cuda
    from scene_reconstruction.train_utils import regularization

    refs = [ref_cuda.RefRasterizer() for _ in cams]

    class RefRaster(torch.autograd.Function):  # rasterize_gaussians / _RasterizeGaussians of the reference, over its own CUDA
        @staticmethod
        def forward(ctx, v, means3D, means2D, shs, opacities, scales, rotations):
            cam = cams[v]
            R, color, radii, depth = refs[v].forward(
                means3D.detach(), opacities.detach(), cam.world_view_transform, cam.full_proj_transform, cam.camera_center, bg,
                math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), int(cam.image_height), int(cam.image_width),
                shs=shs.detach().contiguous(), scales=scales.detach(), rotations=rotations.detach().contiguous(), sh_degree=3,
                sync=False)
            ctx.v = v
            ctx.mark_non_differentiable(radii, depth)
            return color, radii, depth

        @staticmethod
        def backward(ctx, dcolor, _r, _d):
            g = refs[ctx.v].backward(dcolor.contiguous(), sync=False)
            return None, g["means3D"], g["means2D"], g["shs"], g["opacities"], g["scales"], g["rotations"]

    params = [pc.face_bary, pc._features_dc, pc._features_rest, pc._opacity, pc._scaling, pc._rotation]
    groups = [{"params": [p], "lr": lr} for p, lr in zip(params, (1.6e-4, 2.5e-3, 1.25e-4, 0.05, 5e-3, 1e-3))]
    gopt = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
    rest = pc.mesh.pos
    face, ids = pc.mesh.face, pc.face_ids

    def step(it):
        pk = []
        for v, cam in enumerate(cams):
            tvec = torch.full((rest.shape[0], 1), float(cam.time), device=dev)
            verts = sim(time_vector=tvec)  # the reference's forward, device-to-host bounds check included
            xyz = mesh_oracle.get_xyz(verts, face, ids, pc.face_bary)
            rot = mesh_oracle.get_rotation(rest, verts, face, ids, pc._rotation)
            m2 = torch.zeros_like(xyz, requires_grad=True)
            color, radii, depth = RefRaster.apply(v, xyz, m2, pc.get_features, pc.get_opacity, pc.get_scaling, rot)
            pk.append((color, radii, verts, m2))
        img = torch.stack([p[0] for p in pk])
        gt = torch.stack([c.original_image.to(dev, non_blocking=True) for c in cams])
        mse = ((img - gt) ** 2).view(img.shape[0], -1).mean(1, keepdim=True)
        ps = (20 * torch.log10(1.0 / torch.sqrt(mse))).mean().double()
        loss, l1, ssim_loss = loss_oracle.image_losses(img, gt, opt.lambda_dssim)
        d = {"l1": l1.item(), "ssim_loss": ssim_loss.item()}  # the reference's two .item() reads (train_utils.py:66-71)
        loss = loss + regularization(torch.stack([p[2] for p in pk]), pc, opt)
        loss.backward()
        with torch.no_grad():
            grad = sum(p[3].grad for p in pk)
            radii = torch.stack([p[1] for p in pk]).max(0).values
            seen = torch.stack([p[1] > 0 for p in pk]).any(0)
            pc.max_radii2D[seen] = torch.max(pc.max_radii2D[seen], radii[seen].float())  # the reference's mask indexing
            pc.pos_gradient_accum[seen] += torch.norm(grad[seen, :2], dim=-1, keepdim=True)
            pc.denom[seen] += 1
            gopt.step()
            mesh_opt.step()
            gopt.zero_grad(set_to_none=True)
            mesh_opt.zero_grad()
        return ps, loss.detach(), d

    return step


# ------------------------------------------------------------------------------------------
# CPU baseline (oracle port on the host cores) and reference arm
# ------------------------------------------------------------------------------------------
def cpu_baseline(g_np, cams, dl_np, budget_s=12.0):
    import oracle

    cores = os.cpu_count() or 1
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    cam = cams[0]
    H, W = cam.image_height, cam.image_width
    n, t0 = 0, time.perf_counter()
    while True:
        s = oracle.rasterize_forward(g_np["means3D"], g_np["opacities"], cam.world_view_transform,
                                     cam.full_proj_transform, cam.camera_center, np.ones(3, np.float32), cam.tanfovx,
                                     cam.tanfovy, H, W, shs=g_np["shs"], scales=g_np["scales"],
                                     rotations=g_np["rotations"], sh_degree=3)
        oracle.rasterize_backward(s, dl_np[0])
        n += 1
        if time.perf_counter() - t0 > budget_s or n >= 20:
            break
    dt = time.perf_counter() - t0
    return {"value": n * H * W / 1e6 / dt, "unit": "Mpix/s", "cores": cores, "kind": "port",
            "sample": f"{n} x (view 0 of the workload, {W}x{H}, fwd+bwd) with the OpenMP C oracle (oracle/rast_oracle*.c)"}


def reference_gnn_module():
    """The reference's own meshnet/graph_network.py (copied unmodified into git-ignored baseline/_ref/ by
    __graft_entry__.build(), so that it travels to the GPU box) imported under oracle/pyg_stub, the stand-in for the one
    torch_geometric class it needs (real torch_geometric is not installable in this image)."""
    import importlib.util

    path = os.path.join(ROOT, "baseline", "_ref", "meshnet", "graph_network.py")
    if not os.path.exists(path):
        return None
    stub = os.path.join(ROOT, "oracle", "pyg_stub")
    if stub not in sys.path:
        sys.path.insert(0, stub)
    spec = importlib.util.spec_from_file_location("ref_graph_network", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def ref_stage_times(fn_fwd, fn_bwd):
    """Per-kernel device time of one forward and one backward of the UNMODIFIED reference, from CUPTI through
    torch.profiler (the kernels are launched by oracle/_ref/libref_cuda.so; nothing is re-stated or instrumented)."""
    try:
        from torch.profiler import ProfilerActivity, profile
    except Exception:
        return None

    def grab(fn):
        fn()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            fn()
            torch.cuda.synchronize()
        out = {}
        for ev in prof.events():
            if getattr(ev, "device_type", None) is not None and "cuda" in str(ev.device_type).lower():
                name = ev.name
                for key, stage in (("preprocessCUDA", "preprocess"), ("computeCov2DCUDA", "computeCov2D"),
                                   ("DeviceScan", "InclusiveSum"), ("duplicateWithKeys", "duplicateWithKeys"),
                                   ("DeviceRadixSort", "SortPairs"), ("identifyTileRanges", "identifyTileRanges"),
                                   ("renderCUDA", "render"), ("Memcpy", "memcpy"), ("Memset", "memset")):
                    if key in name:
                        out[stage] = out.get(stage, 0.0) + ev.device_time / 1e3  # us -> ms
                        break
                else:
                    out["other"] = out.get("other", 0.0) + ev.device_time / 1e3
        return out

    try:
        return {"forward_ms": grab(fn_fwd), "backward_ms": grab(fn_bwd),
                "how": "torch.profiler (CUPTI) kernel durations of one warm call of the unmodified reference"}
    except Exception as ex:
        return {"error": repr(ex)}


def ref_raster_leg(workload, dev, steps, warmup, with_e2e=True, with_stages=False):
    """The reference's own kernels (compiled unmodified for sm_100a) on this B200, view by view like
    train_utils.py:259 does.  The shim adds no synchronisation (torch's default stream is the legacy stream the
    reference launches on); the reference's own blocking D2H of the instance count stays."""
    from oracle import ref_cuda

    g_np, cams, dl_np, desc = make_workload(workload, 0, 1)
    V, H, W = len(cams), cams[0].image_height, cams[0].image_width
    P = g_np["means3D"].shape[0]
    host = {k: torch.from_numpy(g_np[k]).pin_memory() for k in PARAM_NAMES}
    dl_host = torch.from_numpy(dl_np).pin_memory()
    gd = {k: torch.empty(g_np[k].shape, dtype=torch.float32, device=dev) for k in PARAM_NAMES}
    dl = torch.empty(dl_np.shape, dtype=torch.float32, device=dev)
    grads_host = {k: torch.empty(g_np[k].shape, dtype=torch.float32).pin_memory() for k in PARAM_NAMES}
    cam_t = [dict(viewmatrix=torch.from_numpy(c.world_view_transform).to(dev),
                  projmatrix=torch.from_numpy(c.full_proj_transform).to(dev),
                  campos=torch.from_numpy(c.camera_center).to(dev), bg=torch.ones(3, device=dev)) for c in cams]
    ref = ref_cuda.RefRasterizer()
    flush = l2_flusher(dev)
    stream = torch.cuda.current_stream()

    def fwd(v):
        t, c = cam_t[v], cams[v]
        return ref.forward(gd["means3D"], gd["opacities"], t["viewmatrix"], t["projmatrix"], t["campos"], t["bg"],
                           c.tanfovx, c.tanfovy, H, W, shs=gd["shs"], scales=gd["scales"], rotations=gd["rotations"],
                           sh_degree=3, sync=False)

    def render_all():
        tot = None
        for v in range(V):
            fwd(v)
            g = ref.backward(dl[v], sync=False)
            if tot is None:
                tot = {k: g[k] for k in PARAM_NAMES}
            else:
                for k in PARAM_NAMES:
                    tot[k] += g[k]
        return tot

    def e2e_step():
        for k in PARAM_NAMES:
            gd[k].copy_(host[k], non_blocking=True)
        dl.copy_(dl_host, non_blocking=True)
        tot = render_all()
        for k in PARAM_NAMES:
            grads_host[k].copy_(tot[k].view(grads_host[k].shape), non_blocking=True)
        torch.cuda.synchronize()

    def timed(fn):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for a, b_ in ev:  # same loop as the product arm: events on the launching stream, read after the last step
            flush()
            a.record(stream)
            fn()
            b_.record(stream)
        torch.cuda.synchronize()
        ms = [a.elapsed_time(b_) for a, b_ in ev]
        return sum(ms) / len(ms)

    for k in PARAM_NAMES:
        gd[k].copy_(host[k])
    dl.copy_(dl_host)
    out = {"desc": desc, "V": V, "H": H, "W": W, "P": P, "ms_step": timed(render_all)}
    if with_e2e:
        out["ms_e2e"] = timed(e2e_step)
        out["h2d"] = sum(t.numel() * 4 for t in host.values()) + dl_host.numel() * 4
        out["d2h"] = sum(t.numel() * 4 for t in grads_host.values())
    if with_stages:
        out["stages"] = ref_stage_times(lambda: fwd(0), lambda: ref.backward(dl[0], sync=False))
    out["R_view0"] = int(fwd(0)[0])
    torch.cuda.synchronize()
    return out


def run_reference(args, rank, world, dev):
    from oracle import ref_cuda

    if rank != 0:
        return
    g_np, cams, dl_np, desc = make_workload(args.workload, 0, 1)
    V, H, W = len(cams), cams[0].image_height, cams[0].image_width
    P = g_np["means3D"].shape[0]
    mpix = V * H * W / 1e6
    base = {"metric": "rasterizer fwd+bwd Mpix/s", "unit": "Mpix/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "strong" if args.workload == "c4" else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": desc, "gaussians": P, "views_per_gpu": V, "width": W, "height": H}}
    if not (ref_cuda.available() and torch.cuda.is_available()):
        cb = cpu_baseline(g_np, cams, dl_np, budget_s=20.0)
        base.update(value=cb["value"], ms_per_step=mpix / cb["value"] * 1e3, cpu_baseline=cb,
                    e2e={"value": cb["value"], "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    gpu_launches=0)
        print(json.dumps(base), flush=True)
        return
    flush = l2_flusher(dev)
    peaks = load_peaks()
    sampler = ClockSampler(dev.index)
    main = ref_raster_leg(args.workload, dev, args.steps, args.warmup, with_e2e=True, with_stages=True)
    clocks = sampler.stop()
    ms_dev, ms_e2e = main["ms_step"], main["ms_e2e"]
    base.update(
        value=mpix / (ms_dev * 1e-3), ms_per_step=ms_dev, clocks=clocks,
        e2e={"value": mpix / (ms_e2e * 1e-3), "unit": "Mpix/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": main["h2d"],
             "d2h_bytes_per_step": main["d2h"]},
        cpu_baseline={"value": mpix / (ms_dev * 1e-3), "unit": "Mpix/s", "cores": 0, "kind": "reference",
                      "sample": "the reference has no CPU rasterizer: its own CUDA kernels (forward.cu, backward.cu, "
                                "rasterizer_impl.cu recompiled for sm_100a, CUB onesweep sort) on this B200, "
                                f"{V} views sequentially per step"},
        gpu_launches=0, stages=main.get("stages"))
    base["config"]["l2"] = "flushed between steps (256 MB write)"
    if not args.no_sub and args.workload == "c2":
        c4 = ref_raster_leg("c4", dev, 3, 3, with_e2e=False)
        base["c4"] = {"workload": c4["desc"], "scaling": "strong", "views_total": c4["V"], "ms_per_step": c4["ms_step"],
                      "value": c4["V"] * c4["H"] * c4["W"] / 1e6 / (c4["ms_step"] * 1e-3), "unit": "Mpix/s",
                      "note": "one GPU renders the 64 views one after another (the reference is single-GPU)"}
        c3 = ref_raster_leg("c3", dev, max(3, args.steps // 3), 3, with_e2e=False, with_stages=True)
        base["c3"] = {"workload": c3["desc"], "ms_per_step": c3["ms_step"],
                      "value": c3["H"] * c3["W"] / 1e6 / (c3["ms_step"] * 1e-3), "unit": "Mpix/s",
                      "instances_per_step": c3["R_view0"], "stages": c3.get("stages")}
        st = (c3.get("stages") or {}).get("forward_ms") or {}
        if st:
            base["sort"] = {"instances": c3["R_view0"],
                            "reference_ms": sum(st.get(k, 0.0) for k in ("InclusiveSum", "duplicateWithKeys", "SortPairs",
                                                                         "identifyTileRanges")),
                            "SortPairs_ms": st.get("SortPairs"),
                            "what": "InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::SortPairs (u64 keys, u32 "
                                    "values, bits [0, 45)) + identifyTileRanges of the reference's C3 forward "
                                    "(rasterizer_impl.cu:276-318)"}
        base["knn"] = knn_bench(dev, "reference", flush)
    if not args.no_gnn:
        # the reference's GNN = its own meshnet/graph_network.py (nn.Module, fp32 torch ops) under the MessagePassing
        # stand-in; if the copy did not travel, its op sequence restated in oracle/gnn_oracle.py
        base["gnn"] = gnn_bench(dev, "reference", peaks, flush, with_cpu=True)
        # bounded sample: torch autograd keeps every [E, 384] / [E, 128] activation of the 150 message-passing
        # steps of a 10-step rollout (~8 GB per graph-batch of 1), 128 graphs do not fit in 180 GB
        base["gnn"]["train_c5"] = gnn_train_bench(dev, "reference", peaks, args.gnn_ref_graphs, iters=3)
        base["gnn"]["train_c5"]["sample"] = f"{args.gnn_ref_graphs} of the 128 graphs a GPU owns in C5"
    if not args.no_loss:
        base["refine_tail"] = loss_adam_bench(dev, "reference", peaks, flush)
    if not args.no_dropin and args.workload == "c2":
        try:
            base["dropin"] = dropin_bench(dev, "reference")
        except Exception as ex:
            base["dropin"] = {"error": repr(ex)}
    print(json.dumps(base), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gnn", action="store_true")
    ap.add_argument("--no-loss", action="store_true", help="skip the image-loss / Adam leg")
    ap.add_argument("--cull-exact", type=int, default=-1,
                    help="A/B timing: csb_debug_cull_exact (0 = bounding-box visit culling only, 1 = exact ellipse-vs-quad test, the default)")
    ap.add_argument("--render-variant", type=lambda v: int(v, 0), default=-1,
                    help="A/B timing: csb_debug_render_variant (low nibble forward kernel, high nibble backward kernel)")
    ap.add_argument("--no-sub", action="store_true", help="skip the C3 / C4 / kNN / sort sub-records")
    ap.add_argument("--no-dropin", action="store_true", help="skip the drop-in train_step sub-record")
    ap.add_argument("--gnn-graphs", type=int, default=128, help="graphs per GPU of the C5 training leg")
    ap.add_argument("--gnn-ref-graphs", type=int, default=4, help="graphs of the reference arm's bounded sample")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference" and rank != 0:
        return
    if not torch.cuda.is_available():
        if args.impl == "reference":
            run_reference(args, rank, world, None)
            return
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if args.impl == "reference":
        run_reference(args, rank, world, dev)
        return
    saved_stdout = None
    if world > 1:
        import torch.distributed as dist

        # NCCL prints its version banner on fd 1: keep stdout to the ONE JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        t = torch.zeros(1, device=dev)
        dist.all_reduce(t)  # communicator set-up (and its banner) happens here
        torch.cuda.synchronize()
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    try:
        run_b200(args, rank, world, dev)
    finally:
        if world > 1:
            import torch.distributed as dist

            dist.destroy_process_group()


if __name__ == "__main__":
    main()
