set -x
B="--no-gnn --no-loss --no-sub --no-cpu-baseline --no-dropin"
for m in 3 4; do
  CSB_SORT_MINB=$m python bench.py $B --workload c3 --steps 10 > gpurun_out/r2v_c3_m$m.json 2> gpurun_out/r2v_c3_m$m.err
  CSB_SORT_MINB=$m python bench.py $B --workload c4 --steps 3 --warmup 1 > gpurun_out/r2v_c4_m$m.json 2> gpurun_out/r2v_c4_m$m.err
done
python - <<'PY'
import json
for m in (3,4):
  for w in ('c3','c4'):
    d=json.loads(open(f'gpurun_out/r2v_{w}_m{m}.json').read().strip().splitlines()[-1])
    print(m, w, d['ms_per_step'], {k: round(v['ms_per_step']*1e3,1) for k,v in d['kernels'].items()})
PY
