# Per-view rasterizations of the drop-in iteration on side streams: train_step tests, then A/B of the drop-in record.
set -x
( time python -m pytest tests/test_train_step_gpu.py -x -m gpu -q --timeout 600 ) 2>&1 | tail -25 > gpurun_out/r3i_tests.log; tail -5 gpurun_out/r3i_tests.log
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --steps 10 --warmup 3"
for s in 0 1; do
  CSB_VIEW_STREAMS=$s python bench.py $B > gpurun_out/r3i_streams$s.json 2> gpurun_out/r3i_streams$s.err
done
python - <<'PY'
import json
for s in (0, 1):
    try:
        d = json.loads([l for l in open(f"gpurun_out/r3i_streams{s}.json") if l.startswith("{")][-1])
        print("CSB_VIEW_STREAMS", s, json.dumps(d.get("dropin"))[:600])
    except Exception as e:
        print(s, "failed", e)
PY
tail -3 gpurun_out/r3i_streams1.err
