set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
timeout 900 python -m pytest tests/test_train_step_gpu.py tests/test_rasterizer_gpu.py tests/test_knn_gpu.py tests/test_gnn_gpu.py -q --timeout 600 2>&1 | tail -8 > gpurun_out/r2u_tests.log; tail -4 gpurun_out/r2u_tests.log
python bench.py $B --workload c3 --steps 10 > gpurun_out/r2u_c3.json 2> gpurun_out/r2u_c3.err
python bench.py $B --workload c4 --steps 3 --warmup 1 > gpurun_out/r2u_c4.json 2> gpurun_out/r2u_c4.err
python - <<'PY'
import json
for w in ('c3','c4'):
    d=json.loads(open(f'gpurun_out/r2u_{w}.json').read().strip().splitlines()[-1])
    print(w, d['ms_per_step'], {k: round(v['ms_per_step']*1e3,1) for k,v in d['kernels'].items()})
PY
python - <<'PY' > gpurun_out/r2u_dropin.json 2> gpurun_out/r2u_dropin.err
import json, sys, torch
sys.argv=['bench.py']
import bench
dev=torch.device('cuda',0); torch.cuda.set_device(dev)
print(json.dumps(bench.dropin_bench(dev, 'b200')))
PY
cat gpurun_out/r2u_dropin.json; tail -3 gpurun_out/r2u_dropin.err
