set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
timeout 900 python -m pytest tests/test_rasterizer_gpu.py tests/test_render_variants_gpu.py -q --timeout 600 2>&1 | tail -3
for v in 0x4f 0x5f 0x3f 0x1f 0x6f; do
  python bench.py $B --render-variant $v > gpurun_out/r3c_c2_v$v.json 2> gpurun_out/r3c_c2_v$v.err
done
python - <<'PY'
import json
for v in ('0x4f','0x5f','0x3f','0x1f','0x6f'):
    d=json.loads(open(f'gpurun_out/r3c_c2_v{v}.json').read().strip().splitlines()[-1])
    print(v, d['ms_per_step'], d['value'], {k: round(x['ms_per_step']*1e3,1) for k,x in d['kernels'].items()})
PY
