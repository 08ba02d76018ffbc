set -x
timeout 600 python -m pytest tests/test_gnn_gpu.py tests/test_gnn_backward_gpu.py tests/test_gnn_timed_shape_gpu.py tests/test_sim_wrapper.py -q --timeout 300 -x 2>&1 | tail -12 > gpurun_out/r2p_tests.log; tail -4 gpurun_out/r2p_tests.log
python tools/gnn_trace.py 20 0 0 > gpurun_out/r2p_trace_multi.txt 2>&1; sed -n 20,64p gpurun_out/r2p_trace_multi.txt
python tools/gnn_trace.py 20 0 1 > gpurun_out/r2p_trace_edge_small.txt 2>&1
python tools/gnn_trace.py 20 18 1 > gpurun_out/r2p_trace_node_small.txt 2>&1
timeout 300 python - <<'PY' > gpurun_out/r2p_c1_ab.txt 2>&1
import sys, os, json
sys.path.insert(0, 'tests')
sys.argv = ['bench.py']
import conftest, numpy as np, torch, csb_cabi, csb_synth as S
import bench
lib = csb_cabi.lib()
dev = torch.device('cuda', 0)
flush_buf = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)
def flush(): flush_buf.fill_(1.0)
for mode in (0, 1):
    lib.csb_debug_gnn_small(mode)
    r = bench.gnn_bench(dev, 'b200', {}, flush, with_cpu=False, iters=30)
    print('small =', mode, json.dumps({k: {kk: r[k][kk] for kk in ('ms_per_forward', 'ms_l2_warm', 'e2e_ms')} for k in ('c1_20x20', 'batch128_40x40')}))
PY
cat gpurun_out/r2p_c1_ab.txt
