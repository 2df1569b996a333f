# usage: bash tools/run_p2p_quick.sh <tag> <ngpus> "<modes: p2p nccl>" [steps] — the bench at N ranks over peer memory and / or NCCL
TAG=${1:-p}; N=${2:-2}; MODES=${3:-"p2p"}; STEPS=${4:-100}
mkdir -p gpurun_out
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py"
for mode in $MODES; do
  if [ $mode = nccl ]; then export VKR_EXCHANGE=nccl; else unset VKR_EXCHANGE; fi
  timeout 420 $L --gpus $N --steps $STEPS --warmup 5 --no-cpu-baseline --no-views --no-scene-update > gpurun_out/bench_${TAG}_${mode}_n$N.json 2> gpurun_out/bench_${TAG}_${mode}_n$N.err; echo "bench $mode n=$N rc=$?"; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/bench_${TAG}_${mode}_n$N.err | grep -i "error\|vkr" | head -8
  python -c "
import json;d=json.loads(open('gpurun_out/bench_${TAG}_${mode}_n$N.json').read().strip().splitlines()[-1]);print('$mode n',d['n_gpus'],'fps',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'equal',d.get('sharded_equals_single'),d.get('exchange'),'host_issue',round(d['host_issue_ms_per_step'],3));print({k:round(v,3) for k,v in d['pass_ms_last_pipelined_frame'].items()})
for r in d.get('per_rank_pass_ms_last_pipelined_frame') or []: print('  ',r)"
done
