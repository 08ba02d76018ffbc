"""Stage-by-stage error report of the GNN CUDA path against the oracle (run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import conftest  # noqa
import numpy as np, torch
import csb_synth as S
from oracle import gnn_oracle
from meshnet import EncodeProcessDecode

for steps in (0, 1, 2, 15):
    for n in (3, 9, 20):
        torch.manual_seed(1)
        m = EncodeProcessDecode(8, 3, 4, 128, steps, 2, 128).eval().cuda()
        x, ei, ef, _, _ = S.gnn_inputs(n, 1)
        with torch.no_grad():
            y = m(torch.from_numpy(x).cuda(), torch.from_numpy(ei).cuda(), torch.from_numpy(ef).cuda(), return_latent=True)
        torch.cuda.synchronize()
        sd = {k: v.cpu() for k, v in m.state_dict().items()}
        yb, it = gnn_oracle.forward(sd, torch.from_numpy(x), torch.from_numpy(ei), torch.from_numpy(ef), emulate_bf16=True, return_intermediates=True)
        yf = gnn_oracle.forward(sd, torch.from_numpy(x), torch.from_numpy(ei), torch.from_numpy(ef))
        lat = m.last_latent.cpu()
        print(f"steps={steps} n={n} N={x.shape[0]} E={ei.shape[1]}: latent err {float((lat-it['x_final']).abs().max()):.3e} (scale {float(it['x_final'].abs().max()):.3f}) "
              f"y err vs bf16-oracle {float((y.cpu()-yb).abs().max()):.3e} vs fp32 {float((y.cpu()-yf).abs().max()):.3e} (scale {float(yf.abs().max()):.3f})", flush=True)
        if steps == 0 and n == 3:
            print("lat row0", lat[0, :8].numpy(), "\noracle  ", it['x_final'][0, :8].numpy())
