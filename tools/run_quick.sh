# usage: bash tools/run_quick.sh <tag> [pytest]  — one box, one GPU: (optional gpu tests,) bench config 4 and views
TAG=${1:-x}
mkdir -p gpurun_out
if [ "$2" = "pytest" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_$TAG.log
fi
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-views --no-scene-update > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_$TAG.err
timeout 600 python bench.py --workload views --steps 4 --warmup 3 > gpurun_out/views_$TAG.json 2> gpurun_out/views_$TAG.err; echo "views rc=$?"; tail -3 gpurun_out/views_$TAG.err
python -c "
import json;d=json.load(open('gpurun_out/bench_$TAG.json'));print('fps',d['value'],'e2e',d['e2e']);print(d['pass_ms_last_pipelined_frame']);print(d['roofline']['kernel_ms'], d['host_issue_ms_per_step'], d['host_vkr_draw_call_ms'])
d=json.load(open('gpurun_out/views_$TAG.json'));print('views',d['value'],'e2e',d['e2e'])"
