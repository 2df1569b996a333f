# usage: bash tools/run_final.sh <tag>   — what the driver runs at the round's end, on one GPU: the GPU tests, smoke(), the bench line, the reference arm
TAG=${1:-final}
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke_$TAG.log
timeout 300 python bench.py --steps 200 --warmup 5 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_$TAG.json'));print('fps',d['value'],'e2e',d['e2e']['value'],'roofline',d['roofline']['frac'],'traffic',d['roofline']['traffic_source']);print('scene_update',{k:(v['value'] if isinstance(v,dict) else None) for k,v in d['scene_update'].items() if k.startswith('update')});print('views',d['views']['value'])"
