# r2l: fused mesh transform tests, drop-in timing with it, smoke
set -x
python -m pytest tests/test_mesh_transform_gpu.py tests/test_mesh_rotation.py tests/test_mesh_xyz_gpu.py tests/test_train_step_gpu.py -q --timeout 900 2>&1 | tail -30 > gpurun_out/r2l_tests.log; tail -5 gpurun_out/r2l_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2l_smoke.log 2>&1; tail -4 gpurun_out/r2l_smoke.log
python - <<'PY' > gpurun_out/r2l_dropin.json 2> gpurun_out/r2l_dropin.err
import json, sys, torch
sys.argv=['bench.py']
import bench
dev=torch.device('cuda',0); torch.cuda.set_device(dev)
print(json.dumps(bench.dropin_bench(dev, 'b200')))
PY
cat gpurun_out/r2l_dropin.json; tail -3 gpurun_out/r2l_dropin.err
python tools/dropin_profile.py > gpurun_out/r2l_dropin_profile.json 2> gpurun_out/r2l_dropin_profile.err; tail -3 gpurun_out/r2l_dropin_profile.err
