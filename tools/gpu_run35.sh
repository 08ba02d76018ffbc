# Exact ellipse-vs-quad visit culling: parity tests, then A/B (csb_debug_cull_exact 0 / 1) on C2, C3, C4.
set -x
( time python -m pytest tests/test_render_variants_gpu.py tests/test_rasterizer_gpu.py -x -m gpu -q --timeout 900 ) 2>&1 | tail -25 > gpurun_out/r3f_tests.log; tail -5 gpurun_out/r3f_tests.log
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
for c in 0 1; do
  python bench.py $B --cull-exact $c > gpurun_out/r3f_c2_cull$c.json 2> gpurun_out/r3f_c2_cull$c.err
  python bench.py $B --workload c3 --steps 10 --warmup 3 --cull-exact $c > gpurun_out/r3f_c3_cull$c.json 2> gpurun_out/r3f_c3_cull$c.err
  python bench.py $B --workload c4 --steps 5 --warmup 3 --cull-exact $c > gpurun_out/r3f_c4_cull$c.json 2> gpurun_out/r3f_c4_cull$c.err
done
python - <<'PY'
import json
for wl in ("c2", "c3", "c4"):
    for c in (0, 1):
        try:
            d = json.loads([l for l in open(f"gpurun_out/r3f_{wl}_cull{c}.json") if l.startswith("{")][-1])
            k = d["kernels"]
            print(wl, "cull", c, "ms/step %.4f" % d["ms_per_step"], "fwd %.1f us" % (k["render_forward"]["ms_per_step"] * 1e3), "bwd %.1f us" % (k["render_backward"]["ms_per_step"] * 1e3))
        except Exception as e:
            print(wl, c, "failed", e)
PY
