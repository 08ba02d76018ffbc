set -x
python -m pytest tests/test_render_variants_gpu.py -q --timeout 600 2>&1 | tail -15 > gpurun_out/r2c_tests.log
for v in 0x00 0x10 0x20 0x30 0x40 0x50 0x60; do
  python bench.py --no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin --render-variant $v > gpurun_out/r2c_c2_$v.json 2> gpurun_out/r2c_c2_$v.err
  python bench.py --workload c3 --steps 10 --no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin --render-variant $v > gpurun_out/r2c_c3_$v.json 2> gpurun_out/r2c_c3_$v.err
done
cat gpurun_out/r2c_tests.log
