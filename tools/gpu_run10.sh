# r2j: packed forward variants (parity + timing + ncu table), flaky-test fix, compute-sanitizer logs
set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
python -m pytest tests/test_render_variants_gpu.py tests/test_train_step_gpu.py tests/test_peer_allreduce_gpu.py tests/test_gnn_gpu.py -q --timeout 900 2>&1 | tail -30 > gpurun_out/r2j_tests.log; tail -5 gpurun_out/r2j_tests.log
for v in 0x43 0x44 0x40; do
  python bench.py $B --render-variant $v > gpurun_out/r2j_c2_v$v.json 2> gpurun_out/r2j_c2_v$v.err
  python bench.py $B --workload c3 --steps 10 --render-variant $v > gpurun_out/r2j_c3_v$v.json 2> gpurun_out/r2j_c3_v$v.err
done
for v in 0x43 0x44; do
  ncu --set full --clock-control none --import-source on -k regex:'render_forward' -c 2 -o gpurun_out/r2j_fwd$v -f python bench.py $B --steps 2 --warmup 1 --render-variant $v > gpurun_out/r2j_ncu_fwd$v.log 2>&1
  python tools/ncu_table.py gpurun_out/r2j_fwd$v.ncu-rep gpurun_out/r2j_ncu_fwd${v}_table.json > gpurun_out/r2j_ncu_fwd${v}_table.md
done
python tools/ncu_mix.py gpurun_out/r2j_fwd0x43.ncu-rep render_forward gpurun_out/r2j_fwd0x43_mix_dump.txt > gpurun_out/r2j_fwd0x43_mix.txt 2>&1
rm -f gpurun_out/r2j_fwd0x44.ncu-rep
du -sh gpurun_out
