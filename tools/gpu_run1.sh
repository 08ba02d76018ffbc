set -x
python -m pytest tests -m gpu -q --timeout 600 2>&1 | tail -40 > gpurun_out/r2a_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1
for v in 0 1 2; do
  python bench.py --no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin --render-variant $v > gpurun_out/r2a_c2_v$v.json 2> gpurun_out/r2a_c2_v$v.err
  python bench.py --workload c3 --steps 10 --no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin --render-variant $v > gpurun_out/r2a_c3_v$v.json 2> gpurun_out/r2a_c3_v$v.err
done
python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
python bench.py --impl reference > gpurun_out/r2a_ref.json 2> gpurun_out/r2a_ref.err
tail -5 gpurun_out/r2a_tests.log
