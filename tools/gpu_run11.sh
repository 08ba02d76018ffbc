# r2k: auto forward variant, C4 variants, compute-sanitizer logs
set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
python -m pytest tests/test_render_variants_gpu.py tests/test_rasterizer_gpu.py -q --timeout 900 2>&1 | tail -8 > gpurun_out/r2k_tests.log; tail -3 gpurun_out/r2k_tests.log
for v in 0x4f 0x43 0x44 0x40; do
  python bench.py $B --workload c4 --steps 3 --warmup 1 --render-variant $v > gpurun_out/r2k_c4_v$v.json 2> gpurun_out/r2k_c4_v$v.err
done
python bench.py $B > gpurun_out/r2k_c2.json 2> gpurun_out/r2k_c2.err
python bench.py $B --workload c3 --steps 10 > gpurun_out/r2k_c3.json 2> gpurun_out/r2k_c3.err
# compute-sanitizer: the rasterizer chain (per-tile binning and global sort paths), kNN, loss, Adam on small scenes
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_target.py > gpurun_out/r2k_sanitizer_$tool.log 2>&1; tail -4 gpurun_out/r2k_sanitizer_$tool.log
done
du -sh gpurun_out
