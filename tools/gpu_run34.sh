# Final GPU suite of the round on the final code (the no-allocation test alone first, then everything with -x as the round's end runs it).
set -x
( time python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 600 ) > gpurun_out/r2z_no_alloc_alone.log 2>&1; tail -3 gpurun_out/r2z_no_alloc_alone.log
( time python -m pytest tests -x -m gpu -q --timeout 900 ) 2>&1 | tail -40 > gpurun_out/r2z_tests.log; tail -6 gpurun_out/r2z_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2z_smoke.log 2>&1; tail -3 gpurun_out/r2z_smoke.log
