set -x
timeout 600 python -m pytest tests/test_gnn_gpu.py tests/test_gnn_timed_shape_gpu.py -q --timeout 300 2>&1 | tail -3
python tests/gnn_profile.py 128 40 3 > gpurun_out/r2x_gnn128.txt 2>&1; cat gpurun_out/r2x_gnn128.txt
