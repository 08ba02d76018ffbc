# r2 final single-GPU evidence: full suite, smoke, both bench arms, launch list, ncu tables (C2, C3, GNN), GNN traces.
# gpurun_out/ stays under 64 MiB: reports are reduced to tables on the box; only the C2 report (with sources) is kept.
set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
( time python -m pytest tests -m gpu -q --timeout 900 ) 2>&1 | tail -40 > gpurun_out/r2z_tests.log; tail -6 gpurun_out/r2z_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2z_smoke.log 2>&1; tail -3 gpurun_out/r2z_smoke.log
( time python bench.py ) > gpurun_out/r2z_bench.json 2> gpurun_out/r2z_bench.err; tail -4 gpurun_out/r2z_bench.err
( time python bench.py --impl reference ) > gpurun_out/r2z_ref.json 2> gpurun_out/r2z_ref.err; tail -4 gpurun_out/r2z_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2z_launches_c2.csv python bench.py $B --steps 2 --warmup 1 > gpurun_out/r2z_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess|bin_|gaussian_backward' -c 14 -o gpurun_out/r2z_c2 -f python bench.py $B --steps 2 --warmup 1 > gpurun_out/r2z_ncu_c2.log 2>&1
python tools/ncu_table.py gpurun_out/r2z_c2.ncu-rep gpurun_out/r2z_ncu_c2_table.json > gpurun_out/r2z_ncu_c2_table.md
python tools/ncu_mix.py gpurun_out/r2z_c2.ncu-rep render_forward gpurun_out/r2z_fwd_mix_dump.txt > gpurun_out/r2z_fwd_mix.txt 2>&1
python tools/ncu_mix.py gpurun_out/r2z_c2.ncu-rep render_backward gpurun_out/r2z_bwd_mix_dump.txt > gpurun_out/r2z_bwd_mix.txt 2>&1
ncu --set full --clock-control none -k regex:'render_|preprocess|sort_|emit|tile_ranges|gaussian_backward' --launch-skip 13 -c 13 -o gpurun_out/tmp_c3 -f python bench.py $B --workload c3 --steps 1 --warmup 1 > gpurun_out/r2z_ncu_c3.log 2>&1
python tools/ncu_table.py gpurun_out/tmp_c3.ncu-rep gpurun_out/r2z_ncu_c3_table.json > gpurun_out/r2z_ncu_c3_table.md; rm -f gpurun_out/tmp_c3.ncu-rep
ncu --set full --clock-control none -k regex:'gnn_' -c 6 -o gpurun_out/tmp_g1 -f python tests/gnn_profile.py 1 20 1 > gpurun_out/r2z_ncu_gnn_c1.log 2>&1
python tools/ncu_table.py gpurun_out/tmp_g1.ncu-rep gpurun_out/r2z_ncu_gnn_c1_table.json > gpurun_out/r2z_ncu_gnn_c1_table.md; rm -f gpurun_out/tmp_g1.ncu-rep
ncu --set full --clock-control none -k regex:'gnn_' --launch-skip 32 -c 6 -o gpurun_out/tmp_g128 -f python tests/gnn_profile.py 128 40 1 > gpurun_out/r2z_ncu_gnn_128.log 2>&1
python tools/ncu_table.py gpurun_out/tmp_g128.ncu-rep gpurun_out/r2z_ncu_gnn_128_table.json > gpurun_out/r2z_ncu_gnn_128_table.md; rm -f gpurun_out/tmp_g128.ncu-rep
ncu --set full --clock-control none -k regex:'gnn_bwd|gnn_wgrad' --launch-skip 40 -c 8 -o gpurun_out/tmp_gt -f python tests/gnn_profile_train.py 32 > gpurun_out/r2z_ncu_gnn_train.log 2>&1
python tools/ncu_table.py gpurun_out/tmp_gt.ncu-rep gpurun_out/r2z_ncu_gnn_train_table.json > gpurun_out/r2z_ncu_gnn_train_table.md; rm -f gpurun_out/tmp_gt.ncu-rep
python tools/gnn_trace.py 20 0 0 > gpurun_out/r2z_gnn_trace_launches_edge.txt 2>&1
python tools/gnn_trace.py 20 0 1 > gpurun_out/r2z_gnn_trace_persistent_edge.txt 2>&1
python tools/gnn_trace.py 20 18 1 > gpurun_out/r2z_gnn_trace_persistent_node.txt 2>&1
python tools/dropin_profile.py > gpurun_out/r2z_dropin_profile.json 2> gpurun_out/r2z_dropin_profile.err
du -sh gpurun_out; ls -la gpurun_out | sort -k5 -n -r | head -4
