set -x
python -m pytest tests/test_gnn_timed_shape_gpu.py -q --timeout 600 2>&1 | tail -15 > gpurun_out/r2b_tests.log
python tools/gnn_convergence.py --n 7 --graphs 8 --steps 300 > gpurun_out/r2b_conv_7x7.json 2> gpurun_out/r2b_conv_7x7.err
python tools/gnn_convergence.py --n 20 --graphs 2 --steps 300 > gpurun_out/r2b_conv_20x20.json 2> gpurun_out/r2b_conv_20x20.err
B="python bench.py --no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin --steps 3 --warmup 3"
for v in 0 1; do
$B --render-variant $v > gpurun_out/r2b_plain_v$v.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:render_ -s 6 -c 2 -o gpurun_out/r2b_render_v$v -f $B --render-variant $v > gpurun_out/r2b_ncu_v$v.log 2>&1
done
cat gpurun_out/r2b_tests.log
