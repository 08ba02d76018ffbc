set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
timeout 900 python -m pytest tests/test_rasterizer_gpu.py tests/test_render_variants_gpu.py tests/test_golden.py -q --timeout 600 2>&1 | tail -3
python bench.py $B > gpurun_out/r3d_c2.json 2> gpurun_out/r3d_c2.err
python bench.py $B --workload c3 --steps 10 > gpurun_out/r3d_c3.json 2> gpurun_out/r3d_c3.err
python - <<'PY'
import json
for w in ('c2','c3'):
    d=json.loads(open(f'gpurun_out/r3d_{w}.json').read().strip().splitlines()[-1])
    print(w, d['ms_per_step'], d['value'], {k: round(v['ms_per_step']*1e3,1) for k,v in d['kernels'].items()})
PY
