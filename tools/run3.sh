TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_$TAG.log
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
python -c "
import json;d=json.load(open('gpurun_out/bench_$TAG.json'));print('fps',d['value'],'e2e',d['e2e']['value']);print(d['pass_ms']);print(d['roofline']);print(d.get('cluster_culling'))"
