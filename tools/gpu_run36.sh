set -x
( time python -m pytest tests/test_render_variants_gpu.py -x -m gpu -q --timeout 900 ) 2>&1 | tail -25 > gpurun_out/r3g_tests.log; tail -5 gpurun_out/r3g_tests.log
