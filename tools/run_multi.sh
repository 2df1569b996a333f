# usage: bash tools/run_multi.sh <tag> <ngpus>
TAG=${1:-m}; N=${2:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_sharding_gpu.py -x -q > gpurun_out/pytest_multi_$TAG.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_multi_$TAG.log
for n in 1 $N; do
  if [ $n = 1 ]; then L="python bench.py"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py"; fi
  timeout 600 $L --gpus $n --steps 50 --warmup 5 --no-cpu-baseline --no-views --no-scene-update > gpurun_out/bench_${TAG}_n$n.json 2> gpurun_out/bench_${TAG}_n$n.err; echo "bench n=$n rc=$?"; tail -2 gpurun_out/bench_${TAG}_n$n.err
  python -c "
import json;d=json.loads(open('gpurun_out/bench_${TAG}_n$n.json').read().strip().splitlines()[-1]);print('n',d['n_gpus'],'fps',d['value'],'e2e',d['e2e']['value']);print(d['pass_ms'],'host_issue',d.get('host_issue_ms_per_step'),d.get('host_vkr_draw_call_ms'));print(d.get('cluster_culling'))"
done
