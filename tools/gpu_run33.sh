# The no-allocation entry point's test alone in a fresh process, then the full single-GPU evidence run (which runs it again, last, after the other tests).
set -x
( time python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 600 ) > gpurun_out/r2z_no_alloc_alone.log 2>&1; tail -5 gpurun_out/r2z_no_alloc_alone.log
bash tools/gpu_run19.sh
grep -n "zz_no_alloc\|passed\|failed" gpurun_out/r2z_tests.log | tail -5
