# r2m: persistent small-graph GNN kernel: parity, timing A/B
set -x
timeout 600 python -m pytest tests/test_gnn_gpu.py tests/test_gnn_backward_gpu.py tests/test_gnn_timed_shape_gpu.py tests/test_sim_wrapper.py -q --timeout 300 -x 2>&1 | tail -30 > gpurun_out/r2m_tests.log; tail -6 gpurun_out/r2m_tests.log
timeout 300 python tests/gnn_profile.py 1 20 20 > gpurun_out/r2m_c1_small.txt 2>&1; cat gpurun_out/r2m_c1_small.txt
timeout 300 python - <<'PY' > gpurun_out/r2m_c1_ab.txt 2>&1
import sys, os
sys.path.insert(0, 'tests')
import conftest, numpy as np, torch, csb_cabi, csb_synth as S
from meshnet import EncodeProcessDecode
lib = csb_cabi.lib()
for n in (20, 30, 40):
    x, ei, ef, _, _ = S.gnn_inputs(n, 4)
    torch.manual_seed(0)
    m = EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).eval().cuda()
    xt, eit, eft = torch.from_numpy(x).cuda(), torch.from_numpy(ei).cuda(), torch.from_numpy(ef).cuda()
    for mode in (0, 1):
        lib.csb_debug_gnn_small(mode)
        with torch.no_grad():
            for _ in range(5): m(xt, eit, eft)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(50): m(xt, eit, eft)
            b.record(); torch.cuda.synchronize()
        print(f"n={n} N={x.shape[0]} E={ei.shape[1]} small={mode}: {a.elapsed_time(b)/50*1e3:.1f} us/forward")
    lib.csb_debug_gnn_small(1)
PY
cat gpurun_out/r2m_c1_ab.txt
