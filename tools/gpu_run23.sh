set -x
python tools/gnn_trace.py 40 0 1 128 > gpurun_out/r2w_trace_edge2.txt 2>&1; sed -n 1,5p gpurun_out/r2w_trace_edge2.txt
python tests/gnn_profile.py 128 40 3 > gpurun_out/r2w_gnn128.txt 2>&1; cat gpurun_out/r2w_gnn128.txt
