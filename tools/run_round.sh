# usage: bash tools/run_round.sh <tag>   — one box, one GPU: gpu tests, smoke, bench (both workloads), launch list, full ncu capture
TAG=${1:-x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_$TAG.log
timeout 600 python bench.py --steps 50 --warmup 5 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
timeout 600 python bench.py --workload views --steps 4 --warmup 3 > gpurun_out/views_$TAG.json 2> gpurun_out/views_$TAG.err; echo "views rc=$?"; tail -3 gpurun_out/views_$TAG.err; cat gpurun_out/views_$TAG.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/ref_$TAG.json 2> gpurun_out/ref_$TAG.err; echo "ref rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-views --no-scene-update"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"geometry_kernel|shade_kernel|large_raster" -s 12 -c 5 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1; echo "full capture rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_$TAG.json'));print('fps',d['value'],'e2e',d['e2e']['value']);print(d['pass_ms']);print(d['roofline']);print(d.get('cluster_culling'));print(d.get('cpu_baseline'))"
