# usage: bash tools/run_scale_full.sh <tag> <ngpus> — the driver's own bench command at N ranks (peer-memory exchange, views included)
TAG=${1:-s}; N=${2:-8}
mkdir -p gpurun_out
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py"
timeout 600 $L --gpus $N --steps 200 --warmup 5 > gpurun_out/bench_${TAG}_full_n$N.json 2> gpurun_out/bench_${TAG}_full_n$N.err; echo "full bench n=$N rc=$?"
python -c "
import json;d=json.loads(open('gpurun_out/bench_${TAG}_full_n$N.json').read().strip().splitlines()[-1]);print('n',d['n_gpus'],'fps',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'equal',d.get('sharded_equals_single'),d.get('exchange'),'host',d['host_busy_ms_per_step']);print('views',d['views']['value'],d['views']['e2e']['value'])
for r in d.get('per_rank_pass_ms_last_pipelined_frame') or []: print('  ',r)"
