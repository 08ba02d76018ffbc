"""Generates tests/golden/gnn_*.npz by importing the REFERENCE's own meshnet/graph_network.py
(KTH-RPL/cloth-splatting @ f7acc90) in the build container:

    python tests/make_golden_gnn.py            # needs /root/reference; CPU only

torch_geometric is not installable here; the reference file is loaded under oracle/pyg_stub, a 28-line
stand-in for torch_geometric.nn.conv.MessagePassing that encodes its documented semantics (x_i = x[ei[1]],
x_j = x[ei[0]], sum aggregation at ei[1], update() fed from propagate()'s original kwargs).  The
fixtures hold the weights' seed (not the weights), the output and a few intermediates; tests regenerate
weights from the seed with the same torch default init (module construction order is identical, so
torch.manual_seed(seed) reproduces them bit for bit - checked through a state_dict checksum)."""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "cloth-splatting_b200"))
import csb_synth as S  # noqa: E402

REF = "/root/reference/meshnet/graph_network.py"
CASES = {"gnn_20x20_seed0": (20, 0), "gnn_7x7_seed3": (7, 3)}


def load_reference():
    sys.path.insert(0, os.path.join(ROOT, "oracle", "pyg_stub"))
    spec = importlib.util.spec_from_file_location("ref_graph_network", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def checksum(sd):
    return float(sum(v.double().abs().sum() for v in sd.values()))


def main():
    ref = load_reference()
    out = os.path.join(HERE, "golden")
    for name, (n, seed) in CASES.items():
        x, ei, ef, _, _ = S.gnn_inputs(n, seed)
        torch.manual_seed(seed)
        m = ref.EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).eval()
        # give LayerNorm / biases non-trivial values so that the fixtures exercise them
        g = torch.Generator().manual_seed(seed + 1000)
        with torch.no_grad():
            for k, p in m.named_parameters():
                if k.endswith(".1.weight"):
                    p.add_(0.1 * torch.randn(p.shape, generator=g))
                elif k.endswith(".1.bias"):
                    p.add_(0.1 * torch.randn(p.shape, generator=g))
        with torch.no_grad():
            xt, eit, eft = torch.from_numpy(x), torch.from_numpy(ei), torch.from_numpy(ef)
            xe, ee = m._encoder(xt, eft)
            x1, e1 = m._processor.gnn_stacks[0](xe, eit, ee)
            xp, ep = m._processor(xe, eit, ee)
            y = m(xt, eit, eft)
        np.savez_compressed(os.path.join(out, name + ".npz"), n=n, seed=seed, y=y.numpy(), x_enc=xe.numpy(),
                            e_enc_sum=ee.double().sum().numpy(), x_step1=x1.numpy(),
                            e_step1_ratio=(e1 / ee).mean().numpy(), x_final=xp.numpy(),
                            e_final_ratio=(ep / ee).mean().numpy(), sd_checksum=checksum(m.state_dict()),
                            ref_commit="KTH-RPL/cloth-splatting f7acc90 meshnet/graph_network.py under oracle/pyg_stub")
        print(name, y.shape, float(y.abs().max()), float((e1 / ee).mean()), float((ep / ee).mean()))


def make_state_dict(seed, module_cls):
    """Weights of a fixture: same seed, same construction order, same perturbation as main()."""
    torch.manual_seed(seed)
    m = module_cls(8, 3, 4, 128, 15, 2, 128).eval()
    g = torch.Generator().manual_seed(seed + 1000)
    with torch.no_grad():
        for k, p in m.named_parameters():
            if k.endswith(".1.weight"):
                p.add_(0.1 * torch.randn(p.shape, generator=g))
            elif k.endswith(".1.bias"):
                p.add_(0.1 * torch.randn(p.shape, generator=g))
    return m


if __name__ == "__main__":
    main()
