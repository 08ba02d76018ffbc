set -x
timeout 600 python -m pytest tests/test_gnn_gpu.py tests/test_gnn_backward_gpu.py tests/test_gnn_timed_shape_gpu.py tests/test_sim_wrapper.py tests/test_gnn_convergence_gpu.py -q --timeout 300 2>&1 | tail -3
python tests/gnn_profile.py 128 40 3 > gpurun_out/r2y_gnn128.txt 2>&1; cat gpurun_out/r2y_gnn128.txt
python tools/gnn_trace.py 20 0 1 > gpurun_out/r2y_trace_persistent_edge.txt 2>&1; sed -n 1,3p gpurun_out/r2y_trace_persistent_edge.txt
python tools/gnn_trace.py 20 0 0 > gpurun_out/r2y_trace_launches.txt 2>&1
timeout 300 python - <<'PY' > gpurun_out/r2y_c1.txt 2>&1
import sys, os, json, time
sys.path.insert(0, 'tests')
sys.argv = ['bench.py']
import conftest, numpy as np, torch, csb_cabi, csb_synth as S
import bench
from meshnet import EncodeProcessDecode
lib = csb_cabi.lib()
dev = torch.device('cuda', 0)
x, ei, ef, _, _ = S.gnn_inputs(20, 4)
torch.manual_seed(0)
m = EncodeProcessDecode(8, 3, 4, 128, 15, 2, 128).eval().cuda()
xt, eit, eft = torch.from_numpy(x).cuda(), torch.from_numpy(ei).cuda(), torch.from_numpy(ef).cuda()
for mode in (0, 1):
    lib.csb_debug_gnn_small(mode)
    with torch.no_grad():
        for _ in range(5): m(xt, eit, eft)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(200): m(xt, eit, eft)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
    print(f"small={mode}: host enqueue {(t1-t0)/200*1e6:.1f} us/forward, with drain {(t2-t0)/200*1e6:.1f} us/forward")
lib.csb_debug_gnn_small(1)
flush_buf = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)
def flush(): flush_buf.fill_(1.0)
r = bench.gnn_bench(dev, 'b200', {}, flush, with_cpu=False, iters=30)
print(json.dumps({k: {kk: r[k][kk] for kk in ('ms_per_forward', 'ms_l2_warm', 'e2e_ms')} for k in ('c1_20x20', 'batch128_40x40')}))
PY
cat gpurun_out/r2y_c1.txt
