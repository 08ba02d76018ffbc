set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
timeout 900 python -m pytest tests/test_rasterizer_gpu.py tests/test_render_variants_gpu.py tests/test_train_step_gpu.py tests/test_gnn_convergence_gpu.py -q --timeout 600 2>&1 | tail -6 > gpurun_out/r3a_tests.log; tail -4 gpurun_out/r3a_tests.log
python bench.py $B > gpurun_out/r3a_c2.json 2> gpurun_out/r3a_c2.err
python bench.py $B --workload c3 --steps 10 > gpurun_out/r3a_c3.json 2> gpurun_out/r3a_c3.err
python bench.py $B --workload c4 --steps 3 --warmup 1 > gpurun_out/r3a_c4.json 2> gpurun_out/r3a_c4.err
python - <<'PY'
import json
for w in ('c2','c3','c4'):
    d=json.loads(open(f'gpurun_out/r3a_{w}.json').read().strip().splitlines()[-1])
    print(w, d['ms_per_step'], d['value'], {k: round(v['ms_per_step']*1e3,1) for k,v in d['kernels'].items()})
PY
ncu --set full --clock-control none -k regex:'render_' -c 4 -o gpurun_out/tmp_r3a -f python bench.py $B --steps 2 --warmup 1 > gpurun_out/r3a_ncu.log 2>&1
python tools/ncu_table.py gpurun_out/tmp_r3a.ncu-rep > gpurun_out/r3a_ncu_table.md; rm -f gpurun_out/tmp_r3a.ncu-rep; cat gpurun_out/r3a_ncu_table.md | cut -c1-200
