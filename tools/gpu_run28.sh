set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
for i in 1 2; do timeout 900 python -m pytest tests/test_rasterizer_gpu.py tests/test_render_variants_gpu.py tests/test_train_step_gpu.py tests/test_golden.py -q --timeout 600 2>&1 | tail -3; done
python bench.py $B > gpurun_out/r3b_c2.json 2> gpurun_out/r3b_c2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3b_c2.json').read().strip().splitlines()[-1])
print('c2', d['ms_per_step'], d['value'], d['e2e']['value'], {k: round(v['ms_per_step']*1e3,1) for k,v in d['kernels'].items()})
PY
python - <<'PY' > gpurun_out/r3b_dropin.json 2> gpurun_out/r3b_dropin.err
import json, sys, torch
sys.argv=['bench.py']
import bench
dev=torch.device('cuda',0); torch.cuda.set_device(dev)
print(json.dumps(bench.dropin_bench(dev, 'b200')))
PY
python -c "import json; d=json.load(open('gpurun_out/r3b_dropin.json')); print(d['ms_per_iteration'], d['resident_ground_truth']['ms_per_iteration'])"; tail -2 gpurun_out/r3b_dropin.err
