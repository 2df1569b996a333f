# usage: bash tools/run_n2_check.sh <tag>   (gpurun --gpus 2): the sharding tests, then the driver's own command line at N = 2
TAG=${1:-n2}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_sharding_gpu.py -x -q > gpurun_out/pytest_multi_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_multi_$TAG.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/bench_${TAG}_n2.json 2> gpurun_out/bench_${TAG}_n2.err; echo "bench n=2 rc=$?"; tail -2 gpurun_out/bench_${TAG}_n2.err
python -c "
import json;d=json.loads(open('gpurun_out/bench_${TAG}_n2.json').read().strip().splitlines()[-1]);print('n',d['n_gpus'],'fps',d['value'],'e2e',d['e2e']['value'],'sharded_equals_single',d.get('sharded_equals_single'),'views',d.get('views',{}).get('value'))"
