"""Generates tests/golden/sim_7x7_seed5.npz from the REFERENCE's own simulator wrapper, imported in the build
container (CPU, needs /root/reference):

    python tests/make_golden_sim.py

  * meshnet/cloth_network.py (ClothMeshSimulator) + meshnet/model_utils.py (Normalizer) + meshnet/graph_network.py
    are imported as they are, under oracle/pyg_stub (stand-in for torch_geometric, see tests/make_golden_gnn.py);
  * `update_prediction` is lifted out of train_meshnet_sim.py at run time (the module itself imports wandb,
    matplotlib, ... and starts training on import): its `def` block is compiled from the reference file's lines, nothing is
    copied into this repository.  The `edge_transformer` it calls (torch_geometric Cartesian(norm=False) +
    Distance(norm=False)) is restated here from the transforms' documented behaviour:
    edge_attr = [pos[row] - pos[col], |pos[col] - pos[row]|].
dtypes as in the reference's data (update_prediction only runs with them): velocity history float64,
positions / actions / targets / noise float32.  The multi-step loss follows the loop body of train_meshnet_sim.py:491-527 with the reference's pieces."""
import copy
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
NAME = "sim_7x7_seed5"


def synth(n=7, seed=5, L=2, F=3):
    sys.path.insert(0, os.path.join(ROOT, "cloth-splatting_b200"))
    import csb_synth as S

    x, ei, ef, verts, _ = S.gnn_inputs(n, seed)
    rng = np.random.default_rng(seed)
    N = verts.shape[0]
    velocity = (rng.standard_normal((N, 3 * L)) * 0.01)  # float64, as the reference's loader hands it over
    node_types = np.zeros((N, 1), np.float32)
    node_types[0] = 1
    target = (rng.standard_normal((N, F, 3)) * 0.01).astype(np.float32)
    actions = np.zeros((N, F + 1, 3), np.float32)
    actions[0] = (rng.standard_normal((F + 1, 3)) * 0.02).astype(np.float32)
    actions[0, 1, 2] = 0.0  # one zero coordinate: the reference masks element-wise
    noise = (rng.standard_normal((N, 3 * L)) * 0.003).astype(np.float32)
    return dict(velocity=velocity, node_types=node_types, edge_index=ei, edge_features=ef, target=target,
                actions=actions, positions=verts.astype(np.float32), noise=noise)


def load_reference():
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(ROOT, "oracle", "pyg_stub"))
    from meshnet.cloth_network import ClothMeshSimulator  # the reference's

    class _Data:
        def __init__(self, **kw):
            self.__dict__.update(kw)
            self.edge_attr = None

    def edge_transformer(d):
        row, col = d.edge_index
        cart = d.pos[row] - d.pos[col]
        d.edge_attr = torch.cat([cart, torch.norm(d.pos[col] - d.pos[row], p=2, dim=-1).view(-1, 1)], dim=-1)
        return d

    # the file as a whole does not parse (merge-conflict markers): take the function's lines
    lines = open(os.path.join(REF, "train_meshnet_sim.py")).read().split("\n")
    start = next(i for i, l in enumerate(lines) if l.startswith("def update_prediction("))
    end = next(i for i in range(start + 1, len(lines)) if lines[i].startswith("def "))
    ns = {"torch": torch, "copy": copy, "Data": _Data, "edge_transformer": edge_transformer}
    exec(compile("\n".join(lines[start:end]), "train_meshnet_sim.py", "exec"), ns)
    return ClothMeshSimulator, ns["update_prediction"]


def main():
    Sim, update_prediction = load_reference()
    d = synth()
    t = {k: torch.from_numpy(v) for k, v in d.items()}
    torch.manual_seed(5)
    sim = Sim(3, 8, 4, 128, 15, 2, 128, 2, 2, normalize=True, device="cpu")
    sim.train()
    out = {}
    # -- normaliser statistics after three batches
    g = torch.Generator().manual_seed(11)
    for i in range(3):
        batch = torch.randn((50, 8), generator=g) * (i + 1) + 0.5 * i
        last = sim._node_normalizer(batch, True)
    out["norm_last"] = last.numpy()
    out["norm_mean"] = sim._node_normalizer._mean().numpy()
    out["norm_std"] = sim._node_normalizer._std_with_epsilon().numpy()
    out["norm_inverse"] = sim._node_normalizer.inverse(last).numpy()
    # -- one prediction with noise and targets (training mode: both normalisers accumulate)
    pred, tgt = sim.predict_acceleration(t["velocity"], t["node_types"], t["edge_index"], t["edge_features"],
                                         t["target"][:, 0], velocity_noise=t["noise"])
    out["pred_acc"], out["target_acc"] = pred.detach().numpy(), tgt.numpy()
    # -- update_prediction on its own (fixed inputs)
    acc = torch.from_numpy(np.random.default_rng(1).standard_normal((t["velocity"].shape[0], 3)).astype(np.float32) * 0.01)
    v2, e2, p2 = update_prediction(t["noise"], t["velocity"].clone(), acc, t["positions"], t["edge_index"],
                                   t["actions"][:, 0], t["actions"][:, 1], "cpu", 2)
    out["up_acc"], out["up_velocity"], out["up_edge_features"], out["up_positions"] = acc.numpy(), v2.numpy(), e2.numpy(), p2.numpy()
    # -- rollout step in eval mode
    sim.eval()
    with torch.no_grad():
        out["pred_velocity"] = sim.predict_velocity(t["velocity"].float(), t["node_types"], t["edge_index"], t["edge_features"]).numpy()
    # -- the three-step training loss and a few gradients (statistics frozen: eval mode, as after warm-up)
    velocity, edge_features, pos = t["velocity"].clone(), t["edge_features"], t["positions"]
    loss = 0
    F = 3
    for f in range(F):
        noise = t["noise"] if f == 0 else None
        pa, ta = sim.predict_acceleration(velocity, t["node_types"], t["edge_index"], edge_features, t["target"][:, f],
                                          velocity_noise=noise)
        loss = loss + torch.mean((pa - ta) ** 2)
        if f < F - 1:
            un = sim._output_normalizer.inverse(pa)
            velocity, edge_features, pos = update_prediction(noise, velocity, un, pos, t["edge_index"],
                                                             t["actions"][:, f], t["actions"][:, f + 1], "cpu", 2)
    loss.backward()
    out["loss"] = loss.detach().numpy()
    sd = dict(sim.named_parameters())
    for k in ("_encode_process_decode._encoder.edge_fn.0.NN-0.weight", "_encode_process_decode._decoder.node_fn.NN-2.weight",
              "_encode_process_decode._processor.gnn_stacks.7.node_fn.0.NN-1.bias"):
        out["grad:" + k] = sd[k].grad.numpy()
    out["grad_norm_total"] = np.sqrt(sum(float(p.grad.double().pow(2).sum()) for p in sim.parameters()))
    for k, v in d.items():
        out["in_" + k] = v
    out["ref_commit"] = "KTH-RPL/cloth-splatting f7acc90: meshnet/cloth_network.py, model_utils.py, graph_network.py, train_meshnet_sim.py:322-359"
    np.savez_compressed(os.path.join(HERE, "golden", NAME + ".npz"), **out)
    print({k: getattr(v, "shape", v) for k, v in out.items()})
    print("loss", float(loss), "grad norm", out["grad_norm_total"])


if __name__ == "__main__":
    main()
