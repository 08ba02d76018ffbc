# r2i: full GPU suite, both bench arms, forward/backward variant A/B, ncu launch list + full captures (C2, C3, GNN).
# gpurun_out/ must stay under 64 MiB: reports are reduced to tables on the box and the big ones deleted.
set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
( time python -m pytest tests -m gpu -q --timeout 900 ) 2>&1 | tail -60 > gpurun_out/r2i_tests.log; tail -6 gpurun_out/r2i_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2i_smoke.log 2>&1; tail -2 gpurun_out/r2i_smoke.log
( time python bench.py ) > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err; tail -4 gpurun_out/r2i_bench.err
( time python bench.py --impl reference ) > gpurun_out/r2i_ref.json 2> gpurun_out/r2i_ref.err; tail -4 gpurun_out/r2i_ref.err
for v in 0x40 0x41 0x42 0x00 0x10 0x20 0x60; do
  python bench.py $B --render-variant $v > gpurun_out/r2i_c2_v$v.json 2> gpurun_out/r2i_c2_v$v.err
  python bench.py $B --workload c3 --steps 10 --render-variant $v > gpurun_out/r2i_c3_v$v.json 2> gpurun_out/r2i_c3_v$v.err
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2i_launches_c2.csv python bench.py $B --steps 2 --warmup 1 > gpurun_out/r2i_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess|bin_|gaussian_backward' -c 21 -o gpurun_out/r2i_c2 -f python bench.py $B --steps 2 --warmup 1 > gpurun_out/r2i_ncu_c2.log 2>&1
python tools/ncu_table.py gpurun_out/r2i_c2.ncu-rep gpurun_out/r2i_ncu_c2_table.json > gpurun_out/r2i_ncu_c2_table.md
for v in 0x41 0x42; do
  ncu --set full --clock-control none -k regex:'render_forward' -c 3 -o gpurun_out/tmp_fwd -f python bench.py $B --steps 2 --warmup 1 --render-variant $v > gpurun_out/r2i_ncu_c2_fwd$v.log 2>&1
  python tools/ncu_table.py gpurun_out/tmp_fwd.ncu-rep gpurun_out/r2i_ncu_c2_fwd${v}_table.json > gpurun_out/r2i_ncu_c2_fwd${v}_table.md; rm -f gpurun_out/tmp_fwd.ncu-rep
done
ncu --set full --clock-control none -k regex:'render_|preprocess|sort_|emit|tile_ranges|gaussian_backward' -c 26 -o gpurun_out/tmp_c3 -f python bench.py $B --workload c3 --steps 1 --warmup 1 > gpurun_out/r2i_ncu_c3.log 2>&1
python tools/ncu_table.py gpurun_out/tmp_c3.ncu-rep gpurun_out/r2i_ncu_c3_table.json > gpurun_out/r2i_ncu_c3_table.md; rm -f gpurun_out/tmp_c3.ncu-rep
ncu --set full --clock-control none -k regex:'gnn_' -c 40 -o gpurun_out/tmp_g1 -f python tests/gnn_profile.py 1 20 1 > gpurun_out/r2i_ncu_gnn_c1.log 2>&1
python tools/ncu_table.py gpurun_out/tmp_g1.ncu-rep gpurun_out/r2i_ncu_gnn_c1_table.json > gpurun_out/r2i_ncu_gnn_c1_table.md; rm -f gpurun_out/tmp_g1.ncu-rep
ncu --set full --clock-control none -k regex:'gnn_' --launch-skip 32 -c 6 -o gpurun_out/tmp_g128 -f python tests/gnn_profile.py 128 40 1 > gpurun_out/r2i_ncu_gnn_128.log 2>&1
python tools/ncu_table.py gpurun_out/tmp_g128.ncu-rep gpurun_out/r2i_ncu_gnn_128_table.json > gpurun_out/r2i_ncu_gnn_128_table.md; rm -f gpurun_out/tmp_g128.ncu-rep
python tests/gnn_profile.py 1 20 20 > gpurun_out/r2i_gnn_c1_kernels.txt 2>&1
du -sh gpurun_out; ls -la gpurun_out | sort -k5 -n -r | head -5
