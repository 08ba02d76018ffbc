"""Station timeline of one CTA of the GNN forward (csb_debug_gnn_trace): python tools/gnn_trace.py [n] [block] [small] [graphs]"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import conftest  # noqa: F401,E402
import torch  # noqa: E402

import csb_cabi  # noqa: E402
import csb_synth as S  # noqa: E402
from meshnet import EncodeProcessDecode  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
block = int(sys.argv[2]) if len(sys.argv) > 2 else 0
small = int(sys.argv[3]) if len(sys.argv) > 3 else 1
graphs = int(sys.argv[4]) if len(sys.argv) > 4 else 1
lib = csb_cabi.lib()
lib.csb_debug_gnn_small(small)
x, ei, ef, _, _ = S.gnn_inputs(n, 4)
if graphs > 1:  # disjoint-union batch of `graphs` copies
    import numpy as np

    N0 = x.shape[0]
    x, ef = np.tile(x, (graphs, 1)), np.tile(ef, (graphs, 1))
    ei = np.ascontiguousarray(np.concatenate([ei + g * N0 for g in range(graphs)], 1))
torch.manual_seed(0)
m = EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).eval().cuda()
xt, eit, eft = torch.from_numpy(x).cuda(), torch.from_numpy(ei).cuda(), torch.from_numpy(ef).cuda()
with torch.no_grad():
    for _ in range(3):
        m(xt, eit, eft)
    torch.cuda.synchronize()
    buf = torch.zeros(8002, dtype=torch.int64, device="cuda")
    csb_cabi.check(lib.csb_debug_gnn_trace(buf.data_ptr(), block), "trace")
    m(xt, eit, eft)
    torch.cuda.synchronize()
    csb_cabi.check(lib.csb_debug_gnn_trace(None, 0), "trace")
w = buf.cpu().tolist()
cnt = w[0]
ev = sorted(((v >> 8) & ((1 << 55) - 1), v & 255) for v in w[1:1 + min(cnt, 8000)])
names = {1: "tile start", 2: "A staged", 3: "mma issued", 4: "mma done", 5: "epilogue", 6: "layernorm", 7: "tile end", 8: "barrier passed",
         9: "arrived"}
t0 = ev[0][0] if ev else 0
prev = t0
print(f"n={n} block={block} small={small}: {cnt} stamps")
for t, s in ev[:400]:
    print(f"{(t - t0) / 1e3:9.2f} us  +{(t - prev) / 1e3:6.2f}  {names.get(s, s)}")
    prev = t
