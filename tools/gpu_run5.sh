python -m pytest tests/test_train_step_gpu.py tests/test_adam_gpu.py -q --timeout 600 -x 2>&1 | tail -40 > gpurun_out/r2e_tests.log
cat gpurun_out/r2e_tests.log | tail -40
python - <<'PY' > gpurun_out/r2e_dropin.json 2> gpurun_out/r2e_dropin.err
import json, sys, torch
sys.argv=['bench.py']
import bench
dev=torch.device('cuda',0); torch.cuda.set_device(dev)
out={}
for impl in ('b200','reference'):
    try:
        out[impl]=bench.dropin_bench(dev, impl)
    except Exception as e:
        import traceback; out[impl]={'error': traceback.format_exc()}
print(json.dumps(out))
PY
cat gpurun_out/r2e_dropin.json; tail -5 gpurun_out/r2e_dropin.err
