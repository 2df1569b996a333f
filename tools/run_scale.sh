# usage: bash tools/run_scale.sh <tag> "<list of N>"
TAG=${1:-s}; NS=${2:-"1 2"}
mkdir -p gpurun_out
for n in $NS; do
  if [ $n = 1 ]; then L="python bench.py"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py"; fi
  timeout 600 $L --gpus $n --steps 100 --warmup 5 --no-cpu-baseline --no-views --no-scene-update > gpurun_out/bench_${TAG}_n$n.json 2> gpurun_out/bench_${TAG}_n$n.err; echo "bench n=$n rc=$?"
  python -c "
import json;d=json.loads(open('gpurun_out/bench_${TAG}_n$n.json').read().strip().splitlines()[-1]);print('n',d['n_gpus'],'fps',round(d['value'],1),'e2e',round(d['e2e']['value'],1));print({k:round(v,3) for k,v in d['pass_ms'].items()},'host_issue',round(d.get('host_issue_ms_per_step'),3),round(d.get('host_vkr_draw_call_ms'),3));print('pipelined',{k:round(v,3) for k,v in d['pass_ms_last_pipelined_frame'].items()});print(d.get('cluster_culling'))"
done
