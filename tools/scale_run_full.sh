# the driver's scaling run: default bench.py at N = 8, 4, 2 (torchrun) and 1 on one box.  usage: bash tools/scale_run_full.sh TAG
TAG=${1:-s}
for n in 8 4 2; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 50 --warmup 5 > gpurun_out/${TAG}_scale_n$n.json 2> gpurun_out/${TAG}_scale_n$n.err || tail -5 gpurun_out/${TAG}_scale_n$n.err
done
python bench.py --gpus 1 --steps 50 --warmup 5 > gpurun_out/${TAG}_scale_n1.json 2> gpurun_out/${TAG}_scale_n1.err
python - <<PY
import json
for n in [1,2,4,8]:
    try:
        d=json.loads(open('gpurun_out/${TAG}_scale_n%d.json'%n).read().strip().splitlines()[-1]); t=d['gnn']['train_c5']
        print(n, round(d['value'],1), round(d['ms_per_step'],4), round(d['e2e']['value'],1), d.get('allreduce'), 'gnn train', round(t['ms_per_step'],1), round(t['graphs_per_s'],1))
    except Exception as ex: print(n, 'ERR', ex)
PY
