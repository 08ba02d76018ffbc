set -x
python tools/gnn_trace.py 20 0 1 > gpurun_out/r2n_trace_edge_small.txt 2>&1; head -60 gpurun_out/r2n_trace_edge_small.txt
python tools/gnn_trace.py 20 18 1 > gpurun_out/r2n_trace_node_small.txt 2>&1; sed -n 1,50p gpurun_out/r2n_trace_node_small.txt
python tools/gnn_trace.py 20 0 0 > gpurun_out/r2n_trace_multi.txt 2>&1; sed -n 1,60p gpurun_out/r2n_trace_multi.txt
