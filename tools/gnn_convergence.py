"""Training convergence of the cloth simulator: the same seeded task trained with the B200 kernels (bf16 tensor-core
MLPs, csb_gnn_forward_train / csb_gnn_backward) and with the fp32 oracle (the reference's op sequence under torch autograd,
oracle/gnn_oracle.py; TEST INFRASTRUCTURE), same initial weights, same Adam, same batches.

  python tools/gnn_convergence.py [--n 7] [--graphs 8] [--steps 300] [--lr 1e-3] > gpurun_out/gnn_convergence.json

The training step is the reference's (train_meshnet_sim.py:491-532): multi-step loss over `F` unrolled predictions,
backward, Adam.  Task: a damped mass-spring cloth -- the target velocity of a node is its last velocity plus a Laplacian
smoothing of its neighbours' velocities and a spring force from the edge elongations, so one message-passing step carries
the needed information and the loss can fall by orders of magnitude.  Prints one JSON object with both loss curves."""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "cloth-splatting_b200")]
import csb_synth as S  # noqa: E402


def make_task(n, graphs, F, seed, dev):
    verts, faces, ei = S.cloth_mesh(n)
    N, E = verts.shape[0], ei.shape[1]
    rng = np.random.default_rng(seed)
    NN = N * graphs
    eis = np.ascontiguousarray(np.concatenate([ei + g * N for g in range(graphs)], 1))
    pos = np.tile(verts.astype(np.float32), (graphs, 1)) + (rng.standard_normal((NN, 3)) * 2e-3).astype(np.float32)
    rest = np.tile(np.linalg.norm(verts[ei[0]] - verts[ei[1]], axis=1), graphs).astype(np.float32)
    vel = (rng.standard_normal((NN, 6)) * 0.01).astype(np.float32)
    vel[:, 3:] = vel[:, :3] + (rng.standard_normal((NN, 3)) * 0.002).astype(np.float32)
    rel = pos[eis[0]] - pos[eis[1]]
    ln = np.linalg.norm(rel, axis=1, keepdims=True)
    ef = np.concatenate([rel, ln], 1).astype(np.float32)
    # next velocity: damping + velocity Laplacian + spring force (dst = eis[1] receives from src = eis[0])
    last = vel[:, 3:]
    lap = np.zeros((NN, 3), np.float32)
    np.add.at(lap, eis[1], last[eis[0]] - last[eis[1]])
    spring = np.zeros((NN, 3), np.float32)
    np.add.at(spring, eis[1], rel / ln * (ln - rest[:, None]))
    nxt = 0.98 * last + 0.15 * lap + 0.6 * spring
    target = np.repeat(nxt[:, None, :], F, 1).astype(np.float32)
    node_types = np.tile(np.eye(1, N, 0, dtype=np.float32).T, (graphs, 1))
    d = dict(velocity=vel, node_types=node_types, edge_index=eis, edge_features=ef, target=target,
             actions=np.zeros((NN, F + 1, 3), np.float32), positions=pos,
             noise=(rng.standard_normal((NN, 6)) * 1e-4).astype(np.float32))
    return {k: torch.from_numpy(v).to(dev) for k, v in d.items()}


def train(arm, task, steps, lr, F, dev):
    """arm: 'b200_kernels' | 'oracle_fp32' | 'oracle_fp32_permuted' (the same fp32 network with its edge list permuted: equal
    in exact arithmetic, different summation order in the scatter-add -- the control that shows how far two correct fp32
    runs drift apart over the same number of steps)."""
    from meshnet import ClothMeshSimulator
    from meshnet.rollout import multi_step_loss

    torch.manual_seed(0)
    sim = ClothMeshSimulator(3, 8, 4, 128, 15, 2, 128, 2, 2, normalize=True, device=dev).to(dev)
    if arm.startswith("oracle_fp32"):
        from oracle import gnn_oracle

        class OracleGnn(torch.nn.Module):
            def __init__(self, inner):
                super().__init__()
                self.inner = inner

            def forward(self, x, ei, ea):
                return gnn_oracle.forward(dict(self.inner.named_parameters()), x, ei, ea, device=x.device,
                                          differentiable=True)

        sim._encode_process_decode = OracleGnn(sim._encode_process_decode)
    sim.train()
    opt = torch.optim.Adam(sim.parameters(), lr=lr)
    sched = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=steps, eta_min=0.0)  # the final loss is then a settled value
    ei, ef = task["edge_index"], task["edge_features"]
    if arm.endswith("permuted"):
        perm = torch.randperm(ei.shape[1], generator=torch.Generator().manual_seed(1)).to(ei.device)
        ei, ef = ei[:, perm].contiguous(), ef[perm].contiguous()
    curve = []
    for _ in range(steps):
        opt.zero_grad(set_to_none=True)
        loss = multi_step_loss(sim, task["velocity"], task["node_types"], ei, ef,
                               task["target"], task["actions"], task["positions"], F, velocity_noise=task["noise"])
        loss.backward()
        opt.step()
        sched.step()
        curve.append(loss.detach())
    return [float(x) for x in torch.stack(curve).cpu()]


def run(n=7, graphs=8, steps=300, lr=1e-3, F=2, seed=5, dev="cuda"):
    task = make_task(n, graphs, F, seed, dev)
    out = {"n": n, "graphs": graphs, "steps": steps, "lr": lr, "rollout": F}
    arms = ("b200_kernels", "oracle_fp32", "oracle_fp32_permuted")
    for arm in arms:
        out[arm] = train(arm, task, steps, lr, F, dev)
    tail = max(10, steps // 15)
    fin = {arm: float(np.mean(out[arm][-tail:])) for arm in arms}
    out["final_loss"] = dict(fin, mean_over_last=tail, kernels_over_oracle=fin["b200_kernels"] / fin["oracle_fp32"],
                             permuted_oracle_over_oracle=fin["oracle_fp32_permuted"] / fin["oracle_fp32"])
    out["initial_loss"] = {arm: out[arm][0] for arm in arms}
    out["lr_schedule"] = "Adam, cosine annealing from lr to 0 over the run"
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=7)
    ap.add_argument("--graphs", type=int, default=8)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--rollout", type=int, default=2)
    a = ap.parse_args()
    print(json.dumps(run(a.n, a.graphs, a.steps, a.lr, a.rollout)))
