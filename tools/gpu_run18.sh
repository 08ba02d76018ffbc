# 2-GPU check of the default bench line (incl. the C4 strong-scaling and C5 sub-records) and the 2-rank tests
set -x
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/r2r_bench_2gpu.json 2> gpurun_out/r2r_bench_2gpu.err; tail -5 gpurun_out/r2r_bench_2gpu.err
timeout 600 python -m pytest tests/test_peer_allreduce_gpu.py -q --timeout 300 2>&1 | tail -5 > gpurun_out/r2r_tests.log; tail -3 gpurun_out/r2r_tests.log
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2r_bench_2gpu.json').read().strip().splitlines()[-1])
print('C2', d['value'], d['ms_per_step'], d['e2e']['value'], d.get('allreduce'))
print('C4', d['c4']['ms_per_step'], d['c4']['value'], d['c4']['views_per_gpu'], d['c4']['allreduce'])
print('C5', d['gnn']['train_c5']['ms_per_step'], d['gnn']['train_c5']['graphs_per_s'])
PY
