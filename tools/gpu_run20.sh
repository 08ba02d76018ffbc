set -x
timeout 900 python -m pytest tests/test_train_step_gpu.py tests/test_mesh_transform_gpu.py tests/test_mesh_rotation.py -q --timeout 600 2>&1 | tail -15 > gpurun_out/r2t_tests.log; tail -5 gpurun_out/r2t_tests.log
python - <<'PY' > gpurun_out/r2t_dropin.json 2> gpurun_out/r2t_dropin.err
import json, sys, torch
sys.argv=['bench.py']
import bench
dev=torch.device('cuda',0); torch.cuda.set_device(dev)
print(json.dumps(bench.dropin_bench(dev, 'b200')))
PY
cat gpurun_out/r2t_dropin.json; tail -3 gpurun_out/r2t_dropin.err
python tools/dropin_profile.py > gpurun_out/r2t_dropin_profile.json 2> gpurun_out/r2t_dropin_profile.err; tail -3 gpurun_out/r2t_dropin_profile.err
