set -x
timeout 900 python -m pytest tests/test_rasterizer_gpu.py tests/test_train_step_gpu.py tests/test_render_variants_gpu.py -q --timeout 600 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c2', d['value'], d['ms_per_step'], d['e2e']['value'])"
