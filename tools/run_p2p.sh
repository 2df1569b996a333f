# usage: bash tools/run_p2p.sh <tag> <ngpus>  (gpurun --gpus N) — sharding tests, then the bench at N ranks over peer memory and over NCCL
TAG=${1:-p}; N=${2:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$TAG.txt 2>&1
timeout 900 python -m pytest tests/test_sharding_gpu.py -x -q > gpurun_out/pytest_multi_$TAG.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_multi_$TAG.log
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py"
for mode in p2p nccl; do
  if [ $mode = nccl ]; then export VKR_EXCHANGE=nccl; else unset VKR_EXCHANGE; fi
  timeout 420 $L --gpus $N --steps 100 --warmup 5 --no-cpu-baseline --no-views --no-scene-update > gpurun_out/bench_${TAG}_${mode}_n$N.json 2> gpurun_out/bench_${TAG}_${mode}_n$N.err; echo "bench $mode n=$N rc=$?"; tail -3 gpurun_out/bench_${TAG}_${mode}_n$N.err
  python -c "
import json;d=json.loads(open('gpurun_out/bench_${TAG}_${mode}_n$N.json').read().strip().splitlines()[-1]);print('$mode n',d['n_gpus'],'fps',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'equal',d.get('sharded_equals_single'),d.get('exchange'));print({k:round(v,3) for k,v in d['pass_ms_last_pipelined_frame'].items()});print(d.get('per_rank_pass_ms_last_pipelined_frame'))"
done
unset VKR_EXCHANGE
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu-baseline --no-views --no-scene-update > gpurun_out/bench_${TAG}_n1.json 2> gpurun_out/bench_${TAG}_n1.err; echo "bench n=1 rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_${TAG}_n1.json'));print('n1 fps',round(d['value'],1),'e2e',round(d['e2e']['value'],1))"
