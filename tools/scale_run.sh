# usage: bash tools/scale_run.sh TAG "8 4 2"   (run on a multi-GPU box: gpurun --gpus 8 -- 'bash tools/scale_run.sh s2 "8 4 2"')
TAG=${1:-s}; NS=${2:-"8 4 2"}
for n in $NS; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 50 --warmup 5 --no-cpu-baseline --no-gnn --no-loss > gpurun_out/${TAG}_scale_n$n.json 2> gpurun_out/${TAG}_scale_n$n.err || tail -5 gpurun_out/${TAG}_scale_n$n.err
done
python bench.py --gpus 1 --steps 50 --warmup 5 --no-cpu-baseline --no-gnn --no-loss > gpurun_out/${TAG}_scale_n1.json 2> gpurun_out/${TAG}_scale_n1.err
python - <<PY
import json
for n in [1,2,4,8]:
    try:
        d=json.loads(open('gpurun_out/${TAG}_scale_n%d.json'%n).read().strip().splitlines()[-1]); print(n, round(d['value'],1), round(d['ms_per_step'],4), round(d['e2e']['value'],1), {k:round(v['ms_per_step']*1000,1) for k,v in d['kernels'].items() if 'peer' in k or 'gaussian_b' in k}, d.get('allreduce'))
    except Exception as ex: print(n, 'ERR', ex)
PY
