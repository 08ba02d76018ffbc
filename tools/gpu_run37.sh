# Backward warp sums through shared memory (render variant 0x7_): parity, then A/B against the default (0x4F) on C2, C3, C4.
set -x
( time python -m pytest tests/test_render_variants_gpu.py -x -m gpu -q --timeout 900 ) 2>&1 | tail -25 > gpurun_out/r3h_tests.log; tail -5 gpurun_out/r3h_tests.log
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
for v in 0x4F 0x7F; do
  python bench.py $B --render-variant $v > gpurun_out/r3h_c2_$v.json 2> gpurun_out/r3h_c2_$v.err
  python bench.py $B --workload c3 --steps 10 --warmup 3 --render-variant $v > gpurun_out/r3h_c3_$v.json 2> gpurun_out/r3h_c3_$v.err
  python bench.py $B --workload c4 --steps 5 --warmup 3 --render-variant $v > gpurun_out/r3h_c4_$v.json 2> gpurun_out/r3h_c4_$v.err
done
python - <<'PY'
import json
for wl in ("c2", "c3", "c4"):
    for v in ("0x4F", "0x7F"):
        try:
            d = json.loads([l for l in open(f"gpurun_out/r3h_{wl}_{v}.json") if l.startswith("{")][-1])
            k = d["kernels"]
            print(wl, v, "ms/step %.4f" % d["ms_per_step"], "fwd %.1f us" % (k["render_forward"]["ms_per_step"] * 1e3), "bwd %.1f us" % (k["render_backward"]["ms_per_step"] * 1e3))
        except Exception as e:
            print(wl, v, "failed", e)
PY
