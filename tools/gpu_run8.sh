# r2i: full GPU suite, both bench arms, forward/backward variant A/B, ncu launch list + full captures (C2, C3, GNN)
set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
( time python -m pytest tests -m gpu -q --timeout 900 ) 2>&1 | tail -25 > gpurun_out/r2i_tests.log; tail -6 gpurun_out/r2i_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2i_smoke.log 2>&1; tail -2 gpurun_out/r2i_smoke.log
( time python bench.py ) > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err; tail -4 gpurun_out/r2i_bench.err
( time python bench.py --impl reference ) > gpurun_out/r2i_ref.json 2> gpurun_out/r2i_ref.err; tail -4 gpurun_out/r2i_ref.err
for v in 0x40 0x41 0x42 0x00 0x10 0x20 0x60; do
  python bench.py $B --render-variant $v > gpurun_out/r2i_c2_v$v.json 2> gpurun_out/r2i_c2_v$v.err
  python bench.py $B --workload c3 --steps 10 --render-variant $v > gpurun_out/r2i_c3_v$v.json 2> gpurun_out/r2i_c3_v$v.err
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2i_launches_c2.csv python bench.py $B --steps 2 --warmup 1 > gpurun_out/r2i_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess|bin_|gaussian_backward' -c 21 -o gpurun_out/r2i_c2 -f python bench.py $B --steps 2 --warmup 1 > gpurun_out/r2i_ncu_c2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'render_forward' -c 3 -o gpurun_out/r2i_c2_fwd2 -f python bench.py $B --steps 2 --warmup 1 --render-variant 0x42 > gpurun_out/r2i_ncu_c2_fwd2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess|sort_|emit|tile_ranges|gaussian_backward' -c 26 -o gpurun_out/r2i_c3 -f python bench.py $B --workload c3 --steps 1 --warmup 1 > gpurun_out/r2i_ncu_c3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'gnn_' -c 40 -o gpurun_out/r2i_gnn_c1 -f python tests/gnn_profile.py 1 20 1 > gpurun_out/r2i_ncu_gnn_c1.log 2>&1
ncu --set full --clock-control none -k regex:'gnn_' --launch-skip 32 -c 6 -o gpurun_out/r2i_gnn_128 -f python tests/gnn_profile.py 128 40 1 > gpurun_out/r2i_ncu_gnn_128.log 2>&1
ls -la gpurun_out | head -60
