python -m pytest tests -m gpu -q --timeout 900 2>&1 | tail -15 > gpurun_out/r2h_tests.log; tail -8 gpurun_out/r2h_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2h_smoke.log 2>&1; tail -3 gpurun_out/r2h_smoke.log
python bench.py > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err; tail -3 gpurun_out/r2h_bench.err
python bench.py --impl reference > gpurun_out/r2h_ref.json 2> gpurun_out/r2h_ref.err; tail -3 gpurun_out/r2h_ref.err
